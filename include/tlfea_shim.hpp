// tlfea_shim.hpp -- header-only C++ classes with the reference's names and method signatures
// (lib_src/elements/ElementBase.h:21-49, FEAT10Data.cuh:377-779, lib_src/solvers/SolverBase.h:16-22,
// SyncedNewton.cuh:29-356) that forward to the C ABI of libtlfea_b200.so.
// Every array argument is a TEMPLATE on the caller's container: anything with data() / size() / resize(n) works as a
// vector (std::vector, Eigen::VectorXd, Eigen::VectorXi), anything with data() / rows() / cols() / resize(r, c) and
// column-major storage as a matrix (Eigen::MatrixXd, Eigen::MatrixXi) -- so a reference driver that passes Eigen
// objects compiles against this header unchanged, and Eigen itself is not needed to build the header (it is absent from
// this image; tests/test_abi.py compiles it with a 40-line stand-in matrix, examples/mini_matrix.hpp).
// Errors follow the reference: print and exit(EXIT_FAILURE) (lib_utils/cuda_utils.h:12-18).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

#include <initializer_list>

#include "tlfea_b200.h"

#define TLFEA_OR_DIE(call)                                                        \
  do {                                                                            \
    if ((call) != TLFEA_OK) {                                                     \
      std::fprintf(stderr, "%s failed: %s\n", #call, tlfea_last_error());         \
      std::exit(EXIT_FAILURE);                                                    \
    }                                                                             \
  } while (0)

enum ElementType { TYPE_3243 = 0, TYPE_3443 = 1, TYPE_T10 = 2 };

struct ElementBase {
  ElementType type;
  tlfea_elem *d_data = nullptr;  // the reference keeps a device mirror here; we keep the opaque handle
  virtual ~ElementBase() {}
  int get_n_beam() const { return tlfea_elem_n_elem(d_data); }
  int get_n_coef() const { return tlfea_elem_n_coef(d_data); }
  void CalcMassMatrix() { TLFEA_OR_DIE(tlfea_elem_calc_mass(d_data)); }
  void CalcInternalForce() { TLFEA_OR_DIE(tlfea_elem_calc_internal_force(d_data)); }
  void CalcConstraintData() { TLFEA_OR_DIE(tlfea_elem_calc_constraint(d_data)); }
  void CalcP() { TLFEA_OR_DIE(tlfea_elem_calc_p(d_data)); }
  int get_n_constraint() const { return tlfea_elem_n_constraint(d_data); }
  template <class Vec>
  void RetrieveInternalForceToCPU(Vec &f) {
    f.resize(3 * get_n_coef());
    TLFEA_OR_DIE(tlfea_elem_get_internal_force(d_data, f.data(), (int)f.size()));
  }
  template <class Vec>
  void RetrievePositionToCPU(Vec &x, Vec &y, Vec &z) {
    const int n = get_n_coef();
    x.resize(n), y.resize(n), z.resize(n);
    TLFEA_OR_DIE(tlfea_elem_get_positions(d_data, x.data(), y.data(), z.data(), n));
  }
  template <class Vec>
  void RetrieveConstraintDataToCPU(Vec &c) {
    c.resize(tlfea_elem_n_constraint(d_data));
    TLFEA_OR_DIE(tlfea_elem_get_constraint(d_data, c.data(), (int)c.size()));
  }
  // dense n_constraint x 3 n_coef, as the reference returns it (ElementBase.h:21-49)
  template <class Mat>
  void RetrieveConstraintJacobianToCPU(Mat &J) {
    const int nc = tlfea_elem_n_constraint(d_data), nd = 3 * get_n_coef();
    const int64_t nnz = tlfea_elem_constraint_jacobian_nnz(d_data);
    std::vector<int> off((size_t)nc + 1), cols((size_t)(nnz > 0 ? nnz : 0));
    std::vector<double> vals(cols.size());
    if (nc > 0) TLFEA_OR_DIE(tlfea_elem_get_constraint_jacobian_csr(d_data, off.data(), cols.data(), vals.data()));
    J.resize(nc, nd);
    for (int64_t k = 0; k < (int64_t)nc * nd; k++) J.data()[k] = 0.0;
    for (int c = 0; c < nc; c++)
      for (int k = off[c]; k < off[c + 1]; k++) J.data()[(size_t)cols[k] * nc + c] = vals[k];  // column-major
  }
  // The library derives J, J^T and their CSR forms itself when constraints are set; the reference's explicit
  // conversion steps (FEAT10Data.cu:501-601) only check that there is something to convert.
  void ConvertToCSR_ConstraintJac() {}
  void ConvertToCSR_ConstraintJacT() {}
  void BuildConstraintJacobianCSR() {}
  void SetDensity(double rho0) { TLFEA_OR_DIE(tlfea_elem_set_density(d_data, rho0)); }
  void SetDamping(double eta, double lambda) { TLFEA_OR_DIE(tlfea_elem_set_damping(d_data, eta, lambda)); }
  void SetSVK(double E, double nu) { TLFEA_OR_DIE(tlfea_elem_set_svk(d_data, E, nu)); }
  void SetMooneyRivlin(double mu10, double mu01, double kappa) {
    TLFEA_OR_DIE(tlfea_elem_set_mooney_rivlin(d_data, mu10, mu01, kappa));
  }
  template <class Vec>
  void SetExternalForce(const Vec &f) {
    TLFEA_OR_DIE(tlfea_elem_set_external_force(d_data, f.data(), (int)f.size()));
  }
  template <class IVec>
  void SetNodalFixed(const IVec &nodes) {
    TLFEA_OR_DIE(tlfea_elem_set_nodal_fixed(d_data, nodes.data(), (int)nodes.size()));
  }
  void SetNodalFixed(std::initializer_list<int> nodes) { SetNodalFixed(std::vector<int>(nodes)); }
  template <class IVec>
  void UpdateNodalFixed(const IVec &nodes) {
    TLFEA_OR_DIE(tlfea_elem_update_nodal_fixed(d_data, nodes.data(), (int)nodes.size()));
  }
  void SetLinearConstraintsCSR(const std::vector<int> &offsets, const std::vector<int> &columns,
                               const std::vector<double> &values, const std::vector<double> &rhs) {
    TLFEA_OR_DIE(tlfea_elem_set_linear_constraints_csr(d_data, (int)offsets.size() - 1, offsets.data(), columns.data(),
                                                       values.data(), rhs.data()));
  }
  template <class Vec>
  void UpdateConstraintTargets(const Vec &x, const Vec &y, const Vec &z) {
    TLFEA_OR_DIE(tlfea_elem_update_constraint_targets(d_data, x.data(), y.data(), z.data(), (int)x.size()));
  }
  void RetrieveMassCSRToCPU(std::vector<int> &offsets, std::vector<int> &columns, std::vector<double> &values) {
    const int64_t nnz = tlfea_elem_mass_nnz(d_data);
    offsets.assign((size_t)get_n_coef() + 1, 0);
    columns.assign((size_t)(nnz > 0 ? nnz : 0), 0);
    values.assign(columns.size(), 0.0);
    if (nnz < 0) return;
    TLFEA_OR_DIE(tlfea_elem_get_mass_csr(d_data, offsets.data(), columns.data(), values.data()));
    columns.resize((size_t)offsets.back());
    values.resize((size_t)offsets.back());
  }
  void RetrieveDeformationGradientToCPU(std::vector<double> &F) {
    F.resize((size_t)get_n_beam() * tlfea_elem_n_qp(d_data) * 9);
    TLFEA_OR_DIE(tlfea_elem_get_F(d_data, F.data(), (int64_t)F.size()));
  }
  void RetrievePFromFToCPU(std::vector<double> &P) {
    P.resize((size_t)get_n_beam() * tlfea_elem_n_qp(d_data) * 9);
    TLFEA_OR_DIE(tlfea_elem_get_P(d_data, P.data(), (int64_t)P.size()));
  }
  void RetrieveDetJToCPU(std::vector<double> &detJ) {
    detJ.resize((size_t)get_n_beam() * tlfea_elem_n_qp(d_data));
    TLFEA_OR_DIE(tlfea_elem_get_detJ(d_data, detJ.data(), (int64_t)detJ.size()));
  }
  // the reference's nested forms: [element][point] -> 3x3 matrix (column-major, FEAT10Data.cu:665-686) / scalar
  template <class Mat>
  void RetrievePFromFToCPU(std::vector<std::vector<Mat>> &P) {
    std::vector<double> flat;
    RetrievePFromFToCPU(flat);
    unflatten(flat, 3, 3, P);
  }
  template <class Mat>
  void RetrieveDeformationGradientToCPU(std::vector<std::vector<Mat>> &F) {
    std::vector<double> flat;
    RetrieveDeformationGradientToCPU(flat);
    unflatten(flat, 3, 3, F);
  }
  void RetrieveDetJToCPU(std::vector<std::vector<double>> &detJ) {
    std::vector<double> flat;
    RetrieveDetJToCPU(flat);
    const int q = tlfea_elem_n_qp(d_data);
    detJ.assign((size_t)get_n_beam(), std::vector<double>((size_t)q));
    for (size_t e = 0; e < detJ.size(); e++)
      for (int k = 0; k < q; k++) detJ[e][(size_t)k] = flat[e * q + k];
  }
  template <class Vec>
  void UpdatePositions(const Vec &x, const Vec &y, const Vec &z) {
    TLFEA_OR_DIE(tlfea_elem_update_positions(d_data, x.data(), y.data(), z.data(), (int)x.size()));
  }
  // connectivity as the caller gave it: n_elem x nodes-per-element, the reference's storage order
  template <class IMat>
  void RetrieveConnectivityToCPU(IMat &conn) {
    const int ne = get_n_beam(), npe = type == TYPE_T10 ? 10 : (type == TYPE_3243 ? 2 : 4);
    conn.resize(ne, npe);
    std::vector<int> raw((size_t)ne * npe);
    TLFEA_OR_DIE(tlfea_elem_get_connectivity(d_data, raw.data(), (int64_t)raw.size()));
    for (int e = 0; e < ne; e++)
      for (int a = 0; a < npe; a++)  // library layout: column-major for T10 / 3443, row-major for 3243
        conn.data()[(size_t)a * ne + e] = (type == TYPE_3243) ? raw[(size_t)e * npe + a] : raw[(size_t)a * ne + e];
  }
  template <class Mat>
  void unflatten(const std::vector<double> &flat, int r, int c, std::vector<std::vector<Mat>> &out) {
    const int q = tlfea_elem_n_qp(d_data);
    out.resize((size_t)get_n_beam());
    for (size_t e = 0; e < out.size(); e++) {
      out[e].resize((size_t)q);
      for (int k = 0; k < q; k++) {
        out[e][(size_t)k].resize(r, c);
        for (int i = 0; i < r * c; i++) out[e][(size_t)k].data()[i] = flat[(e * q + k) * (size_t)(r * c) + i];
      }
    }
  }
  double *GetExternalForceDevicePtr() { return tlfea_elem_devptr_f_ext(d_data); }
  // contact coupling (CollisionSystemBase.h:17-29, test_item_drop.cc:566-598): [x | y | z] node buffer for the collision
  // system, f_ext = base + alpha * contact -- device to device, on the element's stream
  void PackNodesXYZ(double *d_nodes_xyz) { TLFEA_OR_DIE(tlfea_elem_pack_nodes_xyz(d_data, d_nodes_xyz, get_n_coef())); }
  void SetExternalForceDevice(const double *d_base, const double *d_contact, double alpha = 1.0) {
    TLFEA_OR_DIE(tlfea_elem_set_external_force_device(d_data, d_base, d_contact, alpha, 3 * get_n_coef()));
  }
  const double *GetX12DevicePtr() { return tlfea_elem_devptr_x(d_data); }
  const double *GetY12DevicePtr() { return tlfea_elem_devptr_y(d_data); }
  const double *GetZ12DevicePtr() { return tlfea_elem_devptr_z(d_data); }
  void Destroy() {
    TLFEA_OR_DIE(tlfea_elem_destroy(d_data));
    d_data = nullptr;
  }
};

struct GPU_FEAT10_Data : ElementBase {
  GPU_FEAT10_Data(int num_elements, int num_nodes, int device = 0) {
    type = TYPE_T10;
    TLFEA_OR_DIE(tlfea_elem_create(TLFEA_TYPE_T10, num_elements, num_nodes, device, &d_data));
  }
  void Initialize() {}  // allocation happens at construction (FEAT10Data.cuh:382-433)
  // Setup(tet5pt_x, tet5pt_y, tet5pt_z, tet5pt_weights, h_x12, h_y12, h_z12, element_connectivity)
  // connectivity column-major [10][n_elem] like the reference's Eigen::MatrixXi (FEAT10Data.cuh:435-533)
  // QVec / Vec: vectors of double; Conn: a column-major n_elem x 10 integer matrix (Eigen::MatrixXi) or a flat
  // std::vector<int> in that order
  template <class QVec, class Vec, class Conn>
  void Setup(const QVec &qx, const QVec &qy, const QVec &qz, const QVec &qw, const Vec &x, const Vec &y, const Vec &z,
             const Conn &conn_colmajor) {
    TLFEA_OR_DIE(tlfea_elem_setup_t10(d_data, qx.data(), qy.data(), qz.data(), qw.data(), x.data(), y.data(),
                                      z.data(), conn_colmajor.data()));
  }
  void CalcDnDuPre() { TLFEA_OR_DIE(tlfea_elem_precompute(d_data)); }
  // [element][point] -> 10 x 3 matrix of reference gradients (column-major: (a, j) at j*10 + a, FEAT10Data.cu:614-635)
  template <class Mat>
  void RetrieveDnDuPreToCPU(std::vector<std::vector<Mat>> &g) {
    std::vector<double> flat((size_t)get_n_beam() * 5 * 30);
    TLFEA_OR_DIE(tlfea_elem_get_grad_N(d_data, flat.data(), (int64_t)flat.size()));
    unflatten(flat, 10, 3, g);
  }
  // legacy ASCII VTK, VTK_QUADRATIC_TETRA = 24 (FEAT10Data.cu:841-877)
  void WriteOutputVTK(const std::string &filename) {
    std::vector<double> x, y, z;
    RetrievePositionToCPU(x, y, z);
    const int ne = get_n_beam();
    std::vector<int> conn((size_t)ne * 10);
    TLFEA_OR_DIE(tlfea_elem_get_connectivity(d_data, conn.data(), (int64_t)conn.size()));
    std::ofstream out(filename);
    if (!out) {
      std::fprintf(stderr, "WriteOutputVTK: cannot open %s\n", filename.c_str());
      return;
    }
    out << "# vtk DataFile Version 3.0\nT10 mesh output\nASCII\nDATASET UNSTRUCTURED_GRID\n";
    out << "POINTS " << x.size() << " float\n";
    for (size_t i = 0; i < x.size(); i++) out << x[i] << " " << y[i] << " " << z[i] << "\n";
    out << "CELLS " << ne << " " << ne * 11 << "\n";
    for (int e = 0; e < ne; e++) {
      out << "10 ";
      for (int a = 0; a < 10; a++) out << conn[(size_t)a * ne + e] << " ";
      out << "\n";
    }
    out << "CELL_TYPES " << ne << "\n";
    for (int e = 0; e < ne; e++) out << "24\n";
  }
  void BuildMassCSRPattern() { TLFEA_OR_DIE(tlfea_elem_build_mass_pattern(d_data)); }
};

// ANCF3243Data.cuh:434-440 ctor (n_nodes, n_elem); Setup(L, W, H, x12, y12, z12, connectivity row-major [n_elem][2]).
// The Gauss tables the reference passes to Setup are compile-time constants of the library (quadrature_utils.h).
struct GPU_ANCF3243_Data : ElementBase {
  GPU_ANCF3243_Data(int num_nodes, int num_elements, int device = 0) {
    type = TYPE_3243;
    TLFEA_OR_DIE(tlfea_elem_create(TLFEA_TYPE_3243, num_elements, num_nodes, device, &d_data));
  }
  void Initialize() {}
  void Setup(const std::vector<double> &L, const std::vector<double> &W, const std::vector<double> &H,
             const std::vector<double> &x, const std::vector<double> &y, const std::vector<double> &z,
             const std::vector<int> &conn_rowmajor) {
    TLFEA_OR_DIE(tlfea_elem_setup_3243(d_data, L.data(), W.data(), H.data(), x.data(), y.data(), z.data(),
                                       conn_rowmajor.data()));
  }
  void Setup(double L, double W, double H, const std::vector<double> &x, const std::vector<double> &y,
             const std::vector<double> &z, const std::vector<int> &conn_rowmajor) {  // uniform beam overload
    const size_t n = (size_t)get_n_beam();
    Setup(std::vector<double>(n, L), std::vector<double>(n, W), std::vector<double>(n, H), x, y, z, conn_rowmajor);
  }
  void CalcDsDuPre() { TLFEA_OR_DIE(tlfea_elem_precompute(d_data)); }
  void BuildMassCSRPattern() { TLFEA_OR_DIE(tlfea_elem_build_mass_pattern(d_data)); }
};

// ANCF3443Data.cuh:445-455 ctor (n_nodes, n_elem); connectivity column-major [4][n_elem].
struct GPU_ANCF3443_Data : ElementBase {
  GPU_ANCF3443_Data(int num_nodes, int num_elements, int device = 0) {
    type = TYPE_3443;
    TLFEA_OR_DIE(tlfea_elem_create(TLFEA_TYPE_3443, num_elements, num_nodes, device, &d_data));
  }
  void Initialize() {}
  void Setup(const std::vector<double> &L, const std::vector<double> &W, const std::vector<double> &H,
             const std::vector<double> &x, const std::vector<double> &y, const std::vector<double> &z,
             const std::vector<int> &conn_colmajor) {
    TLFEA_OR_DIE(tlfea_elem_setup_3443(d_data, L.data(), W.data(), H.data(), x.data(), y.data(), z.data(),
                                       conn_colmajor.data()));
  }
  void Setup(double L, double W, double H, const std::vector<double> &x, const std::vector<double> &y,
             const std::vector<double> &z, const std::vector<int> &conn_colmajor) {
    const size_t n = (size_t)get_n_beam();
    Setup(std::vector<double>(n, L), std::vector<double>(n, W), std::vector<double>(n, H), x, y, z, conn_colmajor);
  }
  void CalcDsDuPre() { TLFEA_OR_DIE(tlfea_elem_precompute(d_data)); }
  void BuildMassCSRPattern() { TLFEA_OR_DIE(tlfea_elem_build_mass_pattern(d_data)); }
};

struct SyncedNewtonParams {  // SyncedNewton.cuh:29-33, bit-compatible with tlfea_newton_params
  double inner_atol, inner_rtol, outer_tol, rho;
  int max_outer, max_inner;
  double time_step;
};

struct SolverBase {
  virtual ~SolverBase() {}
  virtual void Solve() = 0;
  virtual void SetParameters(void *params) = 0;
};

struct SyncedNewtonSolver : SolverBase {
  tlfea_newton *s = nullptr;
  SyncedNewtonSolver(ElementBase *data, int n_constraints) {
    TLFEA_OR_DIE(tlfea_newton_create(data->d_data, n_constraints, &s));
  }
  ~SyncedNewtonSolver() override { tlfea_newton_destroy(s); }
  void Setup() { TLFEA_OR_DIE(tlfea_newton_setup(s)); }
  void SetParameters(void *params) override {
    TLFEA_OR_DIE(tlfea_newton_set_params(s, static_cast<const tlfea_newton_params *>(params)));
  }
  void AnalyzeHessianSparsity() { TLFEA_OR_DIE(tlfea_newton_analyze_sparsity(s)); }
  void SetFixedSparsityPattern(bool fixed) { TLFEA_OR_DIE(tlfea_newton_set_fixed_pattern(s, fixed ? 1 : 0)); }
  void Solve() override { TLFEA_OR_DIE(tlfea_newton_solve_step(s)); }
  double *GetVelocityGuessDevicePtr() { return tlfea_newton_devptr_v(s); }
  // Not in the reference: settings of the block-Jacobi PCG that stands in for the cuDSS solve
  void SetPCG(int max_iters, double rel_tol) { TLFEA_OR_DIE(tlfea_newton_set_pcg(s, max_iters, rel_tol)); }
  // Not in the reference: the per-iteration kernels driven from host buffers (UpdatePositions + force / tangent
  // assembly + residual norm + RetrieveInternalForceToCPU) with the PCIe copies hidden under the kernels.
  // x, y, z (n_coef each) must stay valid until AssembleFromHostEnd returns; f_int holds 3 * n_coef doubles.
  void AssembleFromHostBegin(const double *x, const double *y, const double *z, int n_coef) {
    TLFEA_OR_DIE(tlfea_newton_assemble_host_begin(s, x, y, z, n_coef));
  }
  double AssembleFromHostEnd(double *f_int, int n_dofs) {
    double norm_g = 0.0;
    TLFEA_OR_DIE(tlfea_newton_assemble_host_end(s, f_int, n_dofs, &norm_g));
    return norm_g;
  }
};
