/*
 * tlfea_b200.h -- C ABI of libtlfea_b200.so
 *
 * B200-native (sm_100a) replacement for ONE hot path of
 * uwsbel/Total-Lagrangian-FEA: per-element Gauss-point quadrature of internal
 * force and consistent tangent for FEAT10 / ANCF3243 / ANCF3443 under SVK,
 * Mooney-Rivlin and Kelvin-Voigt damping, plus deterministic (atomic-free)
 * assembly into the fixed CSR pattern of  H = M/h + h K_t (+C_visc) + h^2 rho J^T J.
 *
 * The reference has no FFI: its boundary is two abstract C++ classes that
 * drivers include directly.  Every entry point below names the reference
 * method (file:line, relative to the reference root) whose behaviour it
 * replaces.  Plain pointers and sizes only -- no Eigen, no torch types.
 *
 * Conventions
 *   - every call returns TLFEA_OK (0) or a negative tlfea_status;
 *     tlfea_last_error() gives a message for the calling thread.  (The
 *     reference prints and exit()s: lib_utils/cuda_utils.h:12-18.)
 *   - host input arrays are copied; `_devptr_` getters return borrowed device
 *     pointers that stay valid until destroy.
 *   - connectivity layout is the reference's: FEAT10 and ANCF3443 column-major
 *     conn[n*n_elem + e] (Eigen::MatrixXi, FEAT10Data.cuh:36-39,
 *     ANCF3443Data.cuh:212-213); ANCF3243 row-major conn[e*2 + n]
 *     (ANCF3243Data.cuh:310-312).
 *   - per-Gauss-point 3x3 outputs are column-major like the reference's
 *     Eigen::Map (F(i,j) -> p[j*3+i]); grad_N is [(e*Q+q)*3S + j*S + a].
 *   - all work is enqueued on the handle's stream; getters synchronise it.
 */
#ifndef TLFEA_B200_H
#define TLFEA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  TLFEA_OK = 0,
  TLFEA_ERR_INVALID = -1,   /* bad argument / call order (reference: std::cerr + return) */
  TLFEA_ERR_CUDA = -2,      /* CUDA runtime failure (reference: exit(EXIT_FAILURE)) */
  TLFEA_ERR_UNSUPPORTED = -3,
  TLFEA_ERR_NO_DEVICE = -4  /* no CUDA device: there is NO CPU fallback */
} tlfea_status;

/* Values match the reference enum ElementType (lib_src/elements/ElementBase.h:19). */
typedef enum { TLFEA_TYPE_3243 = 0, TLFEA_TYPE_3443 = 1, TLFEA_TYPE_T10 = 2 } tlfea_elem_type;

/* Bit-compatible with SyncedNewtonParams (lib_src/solvers/SyncedNewton.cuh:29-33). */
typedef struct {
  double inner_atol, inner_rtol, outer_tol, rho;
  int max_outer, max_inner;
  double time_step;
} tlfea_newton_params;

typedef struct tlfea_elem tlfea_elem;
typedef struct tlfea_newton tlfea_newton;

const char *tlfea_last_error(void);
int tlfea_version(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
int64_t tlfea_launch_count(void);
int tlfea_device_count(void);
/* Host threads (OpenMP) of the one-time host-side analyses (the patch plan of the single-pass FEAT10 assembly):
 * 0 = all processors.  Launchers that export OMP_NUM_THREADS=1 for every rank (torchrun) otherwise make every rank build
 * its plan on one core.  New design, no reference line. */
int tlfea_set_host_threads(int n);

/* ---------------------------------------------------------------- elements */

/* GPU_FEAT10_Data(n_elem,n_nodes)+Initialize()  FEAT10Data.cuh:377-433
 * GPU_ANCF3243_Data(n_nodes,n_elem)             ANCF3243Data.cuh:434-440
 * GPU_ANCF3443_Data(n_nodes,n_elem)             ANCF3443Data.cuh:445-455
 * `device` is the CUDA ordinal the handle lives on. */
int tlfea_elem_create(int type, int n_elem, int n_nodes, int device, tlfea_elem **out);
/* Destroy()  FEAT10Data.cuh:714-779 */
int tlfea_elem_destroy(tlfea_elem *h);
int tlfea_elem_sync(tlfea_elem *h);
/* Run all of the handle's work on a caller-owned cudaStream_t (the reference uses the default
 * stream throughout; a host framework passes its current stream).  NULL = legacy default stream. */
int tlfea_elem_set_stream(tlfea_elem *h, void *cuda_stream);

/* Setup(tet5pt_x,y,z,w, x,y,z, connectivity)  FEAT10Data.cuh:435-533.
 * Quadrature tables are taken from the caller exactly as the reference does. */
int tlfea_elem_setup_t10(tlfea_elem *h, const double *qx, const double *qy, const double *qz, const double *qw,
                         const double *x, const double *y, const double *z, const int *conn_colmajor);

/* Setup(L,W,H, gauss_xi_m,..., weights..., x12,y12,z12, connectivity)  ANCF3243Data.cuh:511-522,653-664.
 * x,y,z hold 4 coefficients per node (position, r_u, r_v, r_w): length 4*n_nodes.
 * B_inv is built inside (cpu_utils.cc:125-209 restated with our own LU). */
int tlfea_elem_setup_3243(tlfea_elem *h, const double *L, const double *W, const double *H, const double *x,
                          const double *y, const double *z, const int *conn_rowmajor);
/* ANCF3443Data.cuh Setup; conn column-major [4][n_elem]. */
int tlfea_elem_setup_3443(tlfea_elem *h, const double *L, const double *W, const double *H, const double *x,
                          const double *y, const double *z, const int *conn_colmajor);

/* SetDensity / SetDamping / SetSVK(E,nu) / SetMooneyRivlin  FEAT10Data.cuh:538-634 */
int tlfea_elem_set_density(tlfea_elem *h, double rho0);
int tlfea_elem_set_damping(tlfea_elem *h, double eta_damp, double lambda_damp);
int tlfea_elem_set_svk(tlfea_elem *h, double E, double nu);
int tlfea_elem_set_mooney_rivlin(tlfea_elem *h, double mu10, double mu01, double kappa);
/* SetExternalForce(VectorXd 3*n_coef)  FEAT10Data.cuh:636-646 */
int tlfea_elem_set_external_force(tlfea_elem *h, const double *f_ext, int n);
/* SetNodalFixed(VectorXi)  FEAT10Data.cu:728-749 (a second call is an error, as in the reference) */
int tlfea_elem_set_nodal_fixed(tlfea_elem *h, const int *fixed_nodes, int n_fixed);
/* UpdateNodalFixed (FEAT10Data.cu:751-839): replace the pinned set.  A solver already bound to the element keeps
 * working if the number of pinned nodes is unchanged (its constraint slots are re-derived on the next call). */
int tlfea_elem_update_nodal_fixed(tlfea_elem *h, const int *fixed_nodes, int n_fixed);
/* SetLinearConstraintsCSR(offsets,columns,values,rhs)  ANCF3243Data.cuh:810-940: rows = constraints, cols = DOFs */
int tlfea_elem_set_linear_constraints_csr(tlfea_elem *h, int n_constraints, const int *offsets, const int *columns,
                                          const double *values, const double *rhs);
/* UpdatePositions / UpdateConstraintTargets  FEAT10Data.cuh:672-712 (n = n_coef) */
int tlfea_elem_update_positions(tlfea_elem *h, const double *x, const double *y, const double *z, int n);
int tlfea_elem_update_constraint_targets(tlfea_elem *h, const double *x, const double *y, const double *z, int n);

/* CalcDnDuPre / CalcDsDuPre  FEAT10Data.cu:284-292, ANCF3243Data.cu:102-198, ANCF3443Data.cu:96-182 */
int tlfea_elem_precompute(tlfea_elem *h);
/* BuildMassCSRPattern  FEAT10Data.cu:372-440 (sorted unique coefficient pairs) */
int tlfea_elem_build_mass_pattern(tlfea_elem *h);
/* CalcMassMatrix  FEAT10Data.cu:351-370 */
int tlfea_elem_calc_mass(tlfea_elem *h);
/* CalcP  FEAT10Data.cu:294-312 (no viscous part: v_guess = nullptr) */
int tlfea_elem_calc_p(tlfea_elem *h);
/* CalcInternalForce  FEAT10Data.cu:314-333 -- ACCUMULATES into f_int from the stored P, like the reference */
int tlfea_elem_calc_internal_force(tlfea_elem *h);
int tlfea_elem_clear_internal_force(tlfea_elem *h);
/* CalcConstraintData  FEAT10Data.cu:335-349 */
int tlfea_elem_calc_constraint(tlfea_elem *h);

int tlfea_elem_n_elem(const tlfea_elem *h);
int tlfea_elem_n_coef(const tlfea_elem *h);        /* get_n_coef() */
int tlfea_elem_n_shape(const tlfea_elem *h);
int tlfea_elem_n_qp(const tlfea_elem *h);
int tlfea_elem_n_constraint(const tlfea_elem *h);
int64_t tlfea_elem_mass_nnz(const tlfea_elem *h); /* with linear CSR constraints: an upper bound (the stored pattern
                                                    * also holds the J^T J couplings; get_mass_csr filters them, and
                                                    * offsets[n_coef] is the exact count) */

/* Retrieve*ToCPU  FEAT10Data.cu:603-726; out buffers are caller-allocated */
int tlfea_elem_get_internal_force(tlfea_elem *h, double *out, int n);                    /* 3*n_coef */
int tlfea_elem_get_positions(tlfea_elem *h, double *x, double *y, double *z, int n);     /* n_coef each */
int tlfea_elem_get_F(tlfea_elem *h, double *out, int64_t n);                              /* n_elem*Q*9 */
int tlfea_elem_get_P(tlfea_elem *h, double *out, int64_t n);                              /* n_elem*Q*9 */
int tlfea_elem_get_grad_N(tlfea_elem *h, double *out, int64_t n);                         /* n_elem*Q*3S */
int tlfea_elem_get_detJ(tlfea_elem *h, double *out, int64_t n);                           /* n_elem*Q */
int tlfea_elem_get_mass_csr(tlfea_elem *h, int *offsets, int *columns, double *values);  /* n_coef+1, nnz, nnz */
int tlfea_elem_get_constraint(tlfea_elem *h, double *out, int n);
int tlfea_elem_get_connectivity(tlfea_elem *h, int *out, int64_t n);
/* RetrieveConstraintJacobianToCPU (ElementBase.h:21-49) in CSR form, rows = constraints, columns = DOFs: selection rows
 * for SetNodalFixed (FEAT10Data.cu:501-536), the caller's matrix for SetLinearConstraintsCSR.  offsets holds
 * n_constraint + 1 ints; columns / values tlfea_elem_constraint_jacobian_nnz entries. */
int64_t tlfea_elem_constraint_jacobian_nnz(tlfea_elem *h);
int tlfea_elem_get_constraint_jacobian_csr(tlfea_elem *h, int *offsets, int *columns, double *values);

/* GetX12DevicePtr.. / GetExternalForceDevicePtr  FEAT10Data.cuh:648-670 */
const double *tlfea_elem_devptr_x(tlfea_elem *h);
const double *tlfea_elem_devptr_y(tlfea_elem *h);
const double *tlfea_elem_devptr_z(tlfea_elem *h);
double *tlfea_elem_devptr_f_ext(tlfea_elem *h);
double *tlfea_elem_devptr_f_int(tlfea_elem *h);

/* Contact coupling adapter (the collision system itself stays the reference's: it only produces f_ext).
 * tlfea_elem_pack_nodes_xyz fills the caller's CollisionSystemInput::d_nodes_xyz = [x(0..n-1), y(0..n-1), z(0..n-1)]
 * (lib_src/collision/CollisionSystemBase.h:17-29) from the current positions, device to device, replacing the three
 * cudaMemcpy of lib_bin/collision_system/test_item_drop.cc:566-575.  tlfea_elem_set_external_force_device writes
 * f_ext = base + alpha * contact on the device (d_contact may be NULL), replacing the gravity copy + cublasDaxpy of
 * test_item_drop.cc:589-598.  Both run on the handle's stream. */
int tlfea_elem_pack_nodes_xyz(tlfea_elem *h, double *d_nodes_xyz, int n_nodes);
int tlfea_elem_set_external_force_device(tlfea_elem *h, const double *d_base, const double *d_contact, double alpha,
                                         int n);

/* ------------------------------------------------------------------ solver */

/* SyncedNewtonSolver(ElementBase*, n_constraints)  SyncedNewton.cuh:37-149 */
int tlfea_newton_create(tlfea_elem *e, int n_constraints, tlfea_newton **out);
int tlfea_newton_destroy(tlfea_newton *s);
/* SetParameters(void*)  SyncedNewton.cuh:206-229 */
int tlfea_newton_set_params(tlfea_newton *s, const tlfea_newton_params *p);
/* Setup()  SyncedNewton.cuh:231-245 (zero v, v_prev, lambda) */
int tlfea_newton_setup(tlfea_newton *s);
/* AnalyzeHessianSparsity()  SyncedNewton.cu:546-907: DOF-level pattern + our slot maps */
int tlfea_newton_analyze_sparsity(tlfea_newton *s);
/* SetFixedSparsityPattern(bool)  SyncedNewton.cuh:351-356 */
int tlfea_newton_set_fixed_pattern(tlfea_newton *s, int fixed);

/* THE BENCHMARKED UNIT.  From current x (and v_guess): f_int, and
 * H.values = M/h + h*K_t (+C_visc) + h^2*rho*J^T J, from zero, pattern fixed.
 * Replaces, per Newton iteration, SyncedNewton.cu:1046-1052 (compute_p, clear,
 * compute_internal_force) and :1080-1097 (memset, hessian mass / tangent /
 * constraints). */
int tlfea_newton_assemble(tlfea_newton *s);
/* force half only (what AdamW/Nesterov consume): SyncedNewton.cu:1046-1052 */
int tlfea_newton_assemble_force(tlfea_newton *s);
/* The same unit from HOST buffers, with the copies overlapped with the kernels:
 *   begin = UpdatePositions(x,y,z) (FEAT10Data.cuh:672-712) + tlfea_newton_assemble.  The coordinates go up in chunks
 *           on a copy stream and the element kernel starts on each element chunk as soon as the coordinates it reads
 *           have landed; then f_int is summed.  Returns without waiting (the H rows are reduced by end).  Between
 *           begin and end the caller may work on f_int on the element's stream (the interface halo sum of a
 *           partitioned mesh) but must not touch H.  x, y, z (n = n_coef each; pinned memory
 *           for a truly asynchronous copy) must stay valid until end returns.
 *   end   = tlfea_newton_residual (if norm_g) + RetrieveInternalForceToCPU into f_int[n = 3 n_coef] (if f_int) on the
 *           main stream, then the H rows are reduced on a second stream WHILE f_int and ||g|| travel back to the host;
 *           returns when f_int, ||g|| AND H are complete. */
int tlfea_newton_assemble_host_begin(tlfea_newton *s, const double *x, const double *y, const double *z, int n);
int tlfea_newton_assemble_host_end(tlfea_newton *s, double *f_int, int n, double *norm_g);
/* Optional: the host buffer (3*n_coef doubles, pinned for overlap) that the NEXT _host_begin/_host_end pair will return
 * f_int in.  Known before _begin, the single-pass FEAT10 path copies each coefficient-id range back as soon as its last
 * chunk of patches is done, under the remaining kernels; the caller must not touch f_int on the device between _begin
 * and _end, except through tlfea_dist_halo_sum: _end sends the id ranges shared with the neighbours again.  NULL clears. */
int tlfea_newton_set_host_result_buffer(tlfea_newton *s, double *f_int_host, int n);
/* constraints_eval + compute_grad_l + ||g||  SyncedNewton.cu:1056-1064; writes ||g||_2 to *norm_g */
int tlfea_newton_residual(tlfea_newton *s, double *norm_g);
/* One implicit step: OneStepNewtonCuDSS (SyncedNewton.cu:909-1394) with the
 * cuDSS SPD solve replaced by block-Jacobi PCG (cuDSS is not in this image). */
int tlfea_newton_solve_step(tlfea_newton *s);
/* last step's statistics: Newton iterations, total PCG iterations, final ||g||, final ||c|| */
int tlfea_newton_last_stats(tlfea_newton *s, int *newton_iters, int *pcg_iters, double *norm_g, double *norm_c);
int tlfea_newton_set_pcg(tlfea_newton *s, int max_iters, double rel_tol);
/* The reference's linear solve is exact (cuDSS Cholesky, SyncedNewton.cu:1103-1114); the PCG stand-in reports how far it
 * got: ||r|| / ||r0|| of the last linear solve of the last step, and converged = 0 if any linear solve of that step
 * stopped at max_iters above rel_tol (with rel_tol = 0 the count is fixed and converged stays 1). */
int tlfea_newton_last_linear_solve(tlfea_newton *s, double *rel_residual, int *converged);

int64_t tlfea_newton_nnz(const tlfea_newton *s);
int tlfea_newton_n_dofs(const tlfea_newton *s);
/* DOF-level CSR as the reference exposes it to cuDSS (SyncedNewton.cu:1010-1013).
 * row_offsets is 64-bit: nnz(H) of the 10M-tet block exceeds 2^31 (the
 * reference's int would overflow).  col_indices/values may be NULL. */
int tlfea_newton_get_hessian_csr(tlfea_newton *s, int64_t *row_offsets, int *col_indices, double *values);
int tlfea_newton_get_g(tlfea_newton *s, double *out, int n);
int tlfea_newton_get_v(tlfea_newton *s, double *out, int n);
int tlfea_newton_set_v(tlfea_newton *s, const double *v, int n);
int tlfea_newton_set_v_prev(tlfea_newton *s, const double *v, int n);
/* GetVelocityGuessDevicePtr()  SyncedNewton.cuh:341-343 */
double *tlfea_newton_devptr_v(tlfea_newton *s);
double *tlfea_newton_devptr_hessian_values(tlfea_newton *s);
double *tlfea_newton_devptr_g(tlfea_newton *s);
/* y = H x on the device with the assembled values (tests, PCG) */
int tlfea_newton_spmv(tlfea_newton *s, const double *d_x, double *d_y);
/* sum of the CUDA-event time (ms) of assemble calls since the last reset, and their count */
int tlfea_newton_timing(tlfea_newton *s, double *ms_total, int *calls, int reset);
/* Assembly kernel selection: 2 (default) = single-pass patch assembly for FEAT10 (csrc/pk_kernel.cuh: on-chip
 * accumulation per patch, in-place update of H colour by colour) for FEAT10 + SVK (+ Kelvin-Voigt), the two-phase
 * path for FEAT10 + Mooney-Rivlin and the ANCF elements; 3 = the patch assembly for every FEAT10 material;
 * 1 = element kernel + row reduce for every element type (csrc/ek_kernel.cuh, csrc/ancf_ek.cuh); 0 = generic
 * warp-per-row gather.  All are deterministic (no atomics); results agree to rounding. */
int tlfea_newton_set_assembly(tlfea_newton *s, int mode);
/* Warps (= sub-patches) per patch of the single-pass assembly: 4 (two CTAs per SM, default) or 8.  Takes effect
 * when the plan is (re)built, i.e. before the first assemble.  New design, no reference line. */
int tlfea_newton_set_patch_warps(tlfea_newton *s, int warps);
/* Id-range chunks of the patch launch order (the granularity at which a host-buffer step overlaps its copies with the
 * kernels): 0 = automatic (at most 4, fewer for meshes too small to fill the launches), 1..16 = exactly that many.
 * Takes effect when the plan is (re)built.  New design, no reference line. */
int tlfea_newton_set_patch_chunks(tlfea_newton *s, int chunks);
/* Affine-element compression of the single-pass FEAT10 assembly: straight-sided tets are integrated from
 * J^-T | detJ (80 B per element, found at CalcDnDuPre) instead of the 150 stored gradients + 5 detJ (1 240 B); curved
 * elements keep the stored gradients.  Results agree to rounding.  Off by default (it trades HBM bytes for issue slots
 * and the kernel is issue-bound); TLFEA_PK_AFFINE=1 turns it on for solvers that never call this. */
int tlfea_newton_set_affine_compression(tlfea_newton *s, int on);
/* out[8] = patches, colours, warps per patch, sum over patches of their distinct node pairs, bytes of write-out
 * lists, pattern blocks no element touches, 0, 0 (valid after the first FEAT10 assemble in mode 2) */
int tlfea_newton_patch_stats(tlfea_newton *s, int64_t *out);
/* per-kernel CUDA-event timing of the next assemble calls (0/1) */
int tlfea_newton_set_profile(tlfea_newton *s, int on);
/* element-phase, row-phase and remaining (constraint term) milliseconds accumulated while profiling */
int tlfea_newton_kernel_ms(tlfea_newton *s, double *gp_ms, double *row_ms, double *other_ms);

/* --------------------------------------------------------------- multi-GPU */
/* Element-partitioned assembly and Newton iteration (new design; the reference is single-GPU).  One rank
 * per GPU, each with its own element / solver handles over its part of the mesh; every rank assembles its own
 * elements (sub-assembled H, no matrix exchange).  The DATA PLANE is in the library, over NCCL loaded at run time
 * (libnccl.so.2; inside a torch process the copy torch uses), so a C++ host needs nothing but these calls:
 *   tlfea_dist_unique_id   rank 0 creates the 128-byte ncclUniqueId; the host hands it to every rank (file, MPI, torch)
 *   tlfea_dist_init        ncclCommInitRank on the solver's device (collective over the ranks)
 *   tlfea_dist_set_halo    the coefficient-id ranges [coef_lo, coef_hi) this rank shares with each neighbour rank (the
 *                          partitioner numbers interface planes contiguously); each shared coefficient has exactly
 *                          two holders
 *   tlfea_dist_halo_sum    d_vec (width doubles per coefficient: 3 = nodal vector, 9 = diagonal blocks) gets, on the
 *                          shared ranges, the sum over both holders: pairwise ncclSend / ncclRecv in one group call on
 *                          the handle's stream, added in neighbour order (bitwise the same on both sides)
 *   tlfea_dist_allreduce_sum   n device-resident doubles (||g||^2, r.z, p.Ap), in place
 * All asynchronous on the handle's stream. */
int tlfea_dist_unique_id(void *out128);
int tlfea_dist_init(tlfea_newton *s, const void *id128, int rank, int world);
int tlfea_dist_set_halo(tlfea_newton *s, int n_nbr, const int *nbr_rank, const int64_t *coef_lo, const int64_t *coef_hi);
int tlfea_dist_halo_sum(tlfea_newton *s, double *d_vec, int width);
int tlfea_dist_allreduce_sum(tlfea_newton *s, double *d_buf, int n);
int64_t tlfea_dist_halo_bytes(tlfea_newton *s); /* bytes this rank sends per halo sum of a nodal vector */
/* Optional peer-memory path of the halo sum for the GPUs of one node (NVLink / NVSwitch): every rank exports one
 * mailbox per neighbour (tlfea_dist_p2p_export: the 64-byte cudaIpcMemHandle_t of the buffer neighbour k writes into),
 * the host hands the handle to that neighbour, which opens it (tlfea_dist_p2p_connect with ITS index of the exporting
 * rank).  Once every neighbour is connected tlfea_dist_halo_sum copies the shared range straight into the neighbour's
 * memory and waits for the neighbour's message with a system-scope flag: two small kernels per neighbour on the
 * handle's stream, no NCCL kernel in the step (which cost 0.85 ms per step between the assembly kernels, measured).
 * Same sums, same order as the NCCL path.  New design; the reference is single-GPU. */
int tlfea_dist_p2p_export(tlfea_newton *s, int k, void *out64);
int tlfea_dist_p2p_connect(tlfea_newton *s, int k, const void *handle64);
int tlfea_dist_finalize(tlfea_newton *s);

/* Distributed Newton/Krylov building blocks (north_star: "distributed Newton/Krylov residual norms use allreduce").
 * Every rank owns a sub-assembled H; between these calls the host side halo-sums interface entries of the vector
 * the op names and all-reduces the 4 partial scalars at TLFEA_KVEC_SCAL (r.z, p.Ap, new r.z, r.r | g.g), which
 * stay in device memory: alpha and beta are formed on the device.  Replaces the single-GPU cuDSS solve of
 * OneStepNewtonCuDSS (SyncedNewton.cu:995-1029,1103-1119) for the partitioned mesh. */
enum {
  TLFEA_KOP_SAVE_POSITIONS = 0, /* x_prev = x (start of a time step) */
  TLFEA_KOP_RESIDUAL = 1,       /* local g; halo-sum the vector at tlfea_newton_devptr_g afterwards */
  TLFEA_KOP_NORM_G = 2,         /* scal[3] = g.g over owned DOFs */
  TLFEA_KOP_BEGIN = 3,          /* r = -g, dv = 0, raw 3x3 diagonal blocks -> KVEC_DIAG [n_coef][9]; halo-sum DIAG */
  TLFEA_KOP_PRECOND = 4,        /* invert the blocks, z = Dinv r, p = z, scal[0] = r.z, scal[3] = r.r */
  TLFEA_KOP_SPMV = 5,           /* Ap = H_local p; halo-sum KVEC_AP */
  TLFEA_KOP_PAP = 6,            /* scal[1] = p.Ap */
  TLFEA_KOP_UPDATE_XR = 7,      /* dv += alpha p, r -= alpha Ap, z = Dinv r, scal[2] = r.z, scal[3] = r.r */
  TLFEA_KOP_UPDATE_P = 8,       /* p = z + beta p, scal[0] = scal[2] */
  TLFEA_KOP_APPLY = 9           /* v += dv, x = x_prev + h v */
};
enum { TLFEA_KVEC_Z = 0, TLFEA_KVEC_P = 1, TLFEA_KVEC_AP = 2, TLFEA_KVEC_DIAG = 3, TLFEA_KVEC_R = 4, TLFEA_KVEC_DV = 5,
       TLFEA_KVEC_SCAL = 6 };
/* 3x3 diagonal blocks of the assembled H, out[n_coef][9] row-major (n = 9*n_coef): the per-node Hessian block of the
 * reference's VBD sweep (vbd_accumulate_residual_and_hessian_diag, FEAT10DataFunc.cuh:295-395); together with
 * tlfea_newton_get_g it feeds VBD / block-Jacobi consumers without re-integrating the elements. */
int tlfea_newton_get_diag_blocks(tlfea_newton *s, double *out, int64_t n);
int tlfea_newton_set_owned_range(tlfea_newton *s, int64_t dof_lo, int64_t dof_hi); /* dot products count [lo, hi) */
int tlfea_newton_krylov_op(tlfea_newton *s, int op);
double *tlfea_newton_devptr_krylov(tlfea_newton *s, int which);

/* Allocation-free deterministic dot product: *d_out = sum_i d_a[i] * d_b[i] (d_a == d_b gives ||a||^2), all pointers
 * DEVICE pointers on the current device, asynchronous on `stream` (a cudaStream_t, NULL = default stream).
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It runs the same two
 * kernels with the same fixed 592 x 256 grid and reduction tree as the handle-based path (tlfea_newton_residual's
 * ||g||), so the results agree bitwise; the 592 partial sums live in a static __device__ array (4 736 B, held by every
 * GPU that loads the library until exit) instead of a cudaMalloc'ed buffer.  The call makes no cudaMalloc / cudaFree
 * and needs no handle.  Limits: n <= TLFEA_NOALLOC_DOT_MAX_N, otherwise TLFEA_ERR_INVALID with a message.  All calls on
 * one device share the array: calls that overlap on the GPU (different streams, no ordering) RACE -- serialise them.
 * Every call overwrites all partials, so nothing carries over from the previous call. */
#define TLFEA_NOALLOC_DOT_MAX_N (((int64_t)1) << 32) /* 4 294 967 296 doubles = 32 GiB per vector */
int tlfea_dot_noalloc(const double *d_a, const double *d_b, int64_t n, double *d_out, void *stream);

/* FP64 FMA peak micro-benchmark (roofline denominator for the tangent kernel): TFLOP/s */
int tlfea_fp64_peak(int device, double *tflops);

/* Test hooks for the single-pass FEAT10 assembly plan (csrc/patch_plan.cuh; new design, no reference line):
 * the plan is pure host data, built without a GPU, so the CPU test-suite can replay the kernel's summation
 * order.  ecoef is [n_elem][10] row-major coefficient ids; off/cols the coefficient-level CSR pattern; n_chunks the
 * id-range chunks of the launch order (color_ptr then has n_chunks * colours + 1 entries). */
int tlfea_debug_patch_plan_build(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y,
                                 const double *z, const int *off, const int *cols, int W, int n_chunks, void **plan_out);
/* sizes[8] = patches, launches (chunks * colours), elements, rows, blob bytes, largest patch blob, uncovered pattern blocks, sum of local blocks */
int tlfea_debug_patch_plan_sizes(void *plan, int64_t *sizes);
int tlfea_debug_patch_plan_copy(void *plan, int *color_ptr, void *desc, void *elems, int *row_node, int64_t *row_ptr,
                                void *blob, int64_t *uncovered);
int tlfea_debug_patch_plan_free(void *plan);

#ifdef __cplusplus
}
#endif
#endif /* TLFEA_B200_H */
