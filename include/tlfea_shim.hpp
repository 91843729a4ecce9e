// tlfea_shim.hpp -- header-only C++ classes with the reference's names and method signatures
// (lib_src/elements/ElementBase.h:21-49, FEAT10Data.cuh:377-779, lib_src/solvers/SolverBase.h:16-22,
// SyncedNewton.cuh:29-356) that forward to the C ABI of libtlfea_b200.so.  Plain std::vector in place of Eigen
// types; define TLFEA_HAVE_EIGEN before including to get Eigen::VectorXd / MatrixXi overloads.
// Errors follow the reference: print and exit(EXIT_FAILURE) (lib_utils/cuda_utils.h:12-18).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <vector>

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
  void RetrieveInternalForceToCPU(std::vector<double> &f) {
    f.resize(3 * (size_t)get_n_coef());
    TLFEA_OR_DIE(tlfea_elem_get_internal_force(d_data, f.data(), (int)f.size()));
  }
  void RetrievePositionToCPU(std::vector<double> &x, std::vector<double> &y, std::vector<double> &z) {
    const int n = get_n_coef();
    x.resize(n), y.resize(n), z.resize(n);
    TLFEA_OR_DIE(tlfea_elem_get_positions(d_data, x.data(), y.data(), z.data(), n));
  }
  void RetrieveConstraintDataToCPU(std::vector<double> &c) {
    c.resize((size_t)tlfea_elem_n_constraint(d_data));
    TLFEA_OR_DIE(tlfea_elem_get_constraint(d_data, c.data(), (int)c.size()));
  }
  void SetDensity(double rho0) { TLFEA_OR_DIE(tlfea_elem_set_density(d_data, rho0)); }
  void SetDamping(double eta, double lambda) { TLFEA_OR_DIE(tlfea_elem_set_damping(d_data, eta, lambda)); }
  void SetSVK(double E, double nu) { TLFEA_OR_DIE(tlfea_elem_set_svk(d_data, E, nu)); }
  void SetMooneyRivlin(double mu10, double mu01, double kappa) {
    TLFEA_OR_DIE(tlfea_elem_set_mooney_rivlin(d_data, mu10, mu01, kappa));
  }
  void SetExternalForce(const std::vector<double> &f) {
    TLFEA_OR_DIE(tlfea_elem_set_external_force(d_data, f.data(), (int)f.size()));
  }
  void SetNodalFixed(const std::vector<int> &nodes) {
    TLFEA_OR_DIE(tlfea_elem_set_nodal_fixed(d_data, nodes.data(), (int)nodes.size()));
  }
  void UpdateNodalFixed(const std::vector<int> &nodes) {
    TLFEA_OR_DIE(tlfea_elem_update_nodal_fixed(d_data, nodes.data(), (int)nodes.size()));
  }
  void SetLinearConstraintsCSR(const std::vector<int> &offsets, const std::vector<int> &columns,
                               const std::vector<double> &values, const std::vector<double> &rhs) {
    TLFEA_OR_DIE(tlfea_elem_set_linear_constraints_csr(d_data, (int)offsets.size() - 1, offsets.data(), columns.data(),
                                                       values.data(), rhs.data()));
  }
  void UpdateConstraintTargets(const std::vector<double> &x, const std::vector<double> &y,
                               const std::vector<double> &z) {
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
  void UpdatePositions(const std::vector<double> &x, const std::vector<double> &y, const std::vector<double> &z) {
    TLFEA_OR_DIE(tlfea_elem_update_positions(d_data, x.data(), y.data(), z.data(), (int)x.size()));
  }
  double *GetExternalForceDevicePtr() { return tlfea_elem_devptr_f_ext(d_data); }
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
  void Setup(const std::vector<double> &qx, const std::vector<double> &qy, const std::vector<double> &qz,
             const std::vector<double> &qw, const std::vector<double> &x, const std::vector<double> &y,
             const std::vector<double> &z, const std::vector<int> &conn_colmajor) {
    TLFEA_OR_DIE(tlfea_elem_setup_t10(d_data, qx.data(), qy.data(), qz.data(), qw.data(), x.data(), y.data(),
                                      z.data(), conn_colmajor.data()));
  }
  void CalcDnDuPre() { TLFEA_OR_DIE(tlfea_elem_precompute(d_data)); }
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
