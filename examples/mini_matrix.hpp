// mini_matrix.hpp -- a stand-in for the handful of Eigen types the reference's drivers pass to the element / solver API
// (Eigen is not in this image).  Column-major storage, data() / size() / rows() / cols() / resize() / operator() /
// setZero(): exactly the surface include/tlfea_shim.hpp's templates rely on, so a driver written against these builds
// unchanged against real Eigen::VectorXd / VectorXi / MatrixXd / MatrixXi.
#pragma once
#include <cstddef>
#include <vector>

namespace mini {
template <class T>
struct Vector {
  std::vector<T> v;
  Vector() {}
  explicit Vector(long n) : v((size_t)n) {}
  long size() const { return (long)v.size(); }
  void resize(long n) { v.resize((size_t)n); }
  void setZero() { v.assign(v.size(), T(0)); }
  T *data() { return v.data(); }
  const T *data() const { return v.data(); }
  T &operator()(long i) { return v[(size_t)i]; }
  const T &operator()(long i) const { return v[(size_t)i]; }
};
template <class T>
struct Matrix {
  std::vector<T> v;
  long r = 0, c = 0;
  Matrix() {}
  Matrix(long rows, long cols) : v((size_t)(rows * cols)), r(rows), c(cols) {}
  long rows() const { return r; }
  long cols() const { return c; }
  long size() const { return r * c; }
  void resize(long rows, long cols) { r = rows, c = cols, v.resize((size_t)(rows * cols)); }
  T *data() { return v.data(); }
  const T *data() const { return v.data(); }
  T &operator()(long i, long j) { return v[(size_t)(j * r + i)]; }
  const T &operator()(long i, long j) const { return v[(size_t)(j * r + i)]; }
};
using VectorXd = Vector<double>;
using VectorXi = Vector<int>;
using MatrixXd = Matrix<double>;
using MatrixXi = Matrix<int>;
}  // namespace mini
