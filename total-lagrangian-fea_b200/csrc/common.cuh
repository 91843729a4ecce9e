// common.cuh -- shared state, error handling and launch accounting for libtlfea_b200.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <utility>
#include <vector>

#include "../../include/tlfea_b200.h"

namespace tlfea {

extern thread_local std::string g_last_error;
extern std::atomic<int64_t> g_launches;
// element handles that are still alive: a solver may be destroyed after its element (garbage collectors
// finalise in arbitrary order) and must not touch the freed handle then
bool elem_is_live(const void *h);

inline int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

#define TL_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (call);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return tlfea::fail(TLFEA_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call,              \
                         cudaGetErrorString(_e));                                                  \
  } while (0)

#define TL_CHECK(cond, ...)                                          \
  do {                                                               \
    if (!(cond)) return tlfea::fail(TLFEA_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define TL_TRY(expr)          \
  do {                        \
    int _s = (expr);          \
    if (_s != TLFEA_OK) return _s; \
  } while (0)

// every kernel launch goes through this so bench.py can report gpu_launches
#define TL_LAUNCH(kernel, grid, block, smem, stream, ...)      \
  do {                                                         \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__); \
    tlfea::g_launches.fetch_add(1, std::memory_order_relaxed);  \
  } while (0)

template <typename T>
inline int dmalloc(T **p, size_t n) {
  *p = nullptr;
  if (n == 0) n = 1;
  TL_CUDA(cudaMalloc((void **)p, n * sizeof(T)));
  return TLFEA_OK;
}

template <typename T>
inline void dfree(T *&p) {
  if (p) cudaFree(p);
  p = nullptr;
}

inline unsigned blocks_for(int64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

// SM count of a device (grids of the persistent kernels are sized in multiples of it), looked up once per thread
inline int sm_count(int device) {
  static thread_local int dev = -1, n = 148;
  if (dev != device) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && v > 0) n = v;
    dev = device;
  }
  return n;
}

constexpr int MAX_Q = 48;
struct QuadW {
  double w[MAX_Q];  // combined quadrature weight per stiffness Gauss point
};

// Coefficient-level sparsity + the maps that replace the reference's per-entry
// binary search (FEAT10DataFunc.cuh:487-502) and its FP64 atomics.
struct Pattern {
  int n_coef = 0;
  int S = 0;
  int64_t n_visits = 0;           // n_elem * S
  int64_t nnz = 0;                // coefficient-level nnz (mass pattern)
  int max_deg = 0;                // longest coefficient row
  int max_inc = 0;                // most elements incident to one coefficient
  int *d_off = nullptr;           // [n_coef+1]  sorted-unique adjacency, == reference d_csr_offsets
  int *d_cols = nullptr;          // [nnz]       == reference d_csr_columns
  int *d_inc_ptr = nullptr;       // [n_coef+1]  coefficient -> its (element, local) visits
  int *d_inc_ea = nullptr;        // [n_visits]  e*S + a, ascending within a coefficient
  uint16_t *d_slot = nullptr;     // [n_visits*S] position of column ecoef[e][b] in the visit's row
  void release() {
    dfree(d_off);
    dfree(d_cols);
    dfree(d_inc_ptr);
    dfree(d_inc_ea);
    dfree(d_slot);
  }
};

int build_pattern(const int *d_ecoef, int n_elem, int S, int n_coef, cudaStream_t st, Pattern *out,
                  const unsigned long long *h_extra = nullptr, int64_t n_extra = 0);

}  // namespace tlfea

enum { TL_MODEL_SVK = 0, TL_MODEL_MR = 1 };
enum { TL_CONS_NONE = 0, TL_CONS_FIXED_NODES = 1, TL_CONS_LINEAR_CSR = 2 };

struct tlfea_elem {
  int type = 0, n_elem = 0, n_nodes = 0, n_coef = 0, S = 0, Q = 0, device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  bool is_setup = false, is_precomputed = false, is_pattern = false, is_mass = false;
  int pattern_version = 0;            // bumped whenever the coefficient pattern is (re)built: bound solvers must re-analyse
  int mass_version = 0;               // bumped by CalcMassMatrix: cached copies of the mass values refresh
  int cons_version = 0;               // bumped by UpdateNodalFixed: solvers refresh their constraint slots
  bool pat_extra = false;             // the pattern also holds J^T J couplings of linear constraints
  std::vector<int> h_j_off, h_j_cols;  // host copy of the constraint Jacobian structure

  int *d_conn = nullptr;    // caller layout (reference layout), n_elem * nodes_per_elem
  int nodes_per_elem = 0;
  int *d_ecoef = nullptr;   // [n_elem][S] coefficient ids, row-major
  double *d_x = nullptr, *d_y = nullptr, *d_z = nullptr;           // current coefficients   (x12,y12,z12)
  double *d_xr = nullptr, *d_yr = nullptr, *d_zr = nullptr;        // reference coefficients (x12_jac,..)
  double *d_gradN = nullptr;  // [n_elem][Q][3S]
  double *d_detJ = nullptr;   // [n_elem][Q]
  // FEAT10 affine-element compression: straight-sided tets have a constant Jacobian, so the 150 reference
  // gradients of an element are J^-T times 150 mesh-independent numbers.  d_aff[e] = J^-T (9, row-major) | detJ;
  // d_aff_flag[e] = 1 where the element is affine (else the kernels read d_gradN / d_detJ as before)
  double *d_aff = nullptr;
  unsigned char *d_aff_flag = nullptr;
  int aff_version = 0;        // bumped by CalcDnDuPre
  double t10_dn[150];         // dN_a/dxi_k at the caller's quadrature points: [q][a][k]
  double *d_F = nullptr, *d_P = nullptr;  // [n_elem][Q][9], allocated on first CalcP
  double *d_f_int = nullptr, *d_f_ext = nullptr;
  double *d_fe = nullptr;     // [n_elem][S][3] element force staging
  double *d_mass = nullptr;   // [pattern.nnz]
  double *d_LWH = nullptr;    // ANCF: [n_elem][3]
  double *d_Binv = nullptr;   // ANCF: [n_elem][S*S]
  tlfea::QuadW qw;
  double t10_q[4][5];         // FEAT10 quadrature points/weights as given by the caller
  tlfea::Pattern pat;

  // material (FEAT10Data.cuh:538-634)
  int model = TL_MODEL_SVK;
  double rho0 = 0, E = 0, nu = 0, lambda = 0, mu = 0, mu10 = 0, mu01 = 0, kappa = 0, eta = 0, lambda_d = 0;

  // constraints
  int cons_mode = TL_CONS_NONE;
  int n_constraint = 0;
  int n_fixed = 0;
  int *d_fixed = nullptr;
  double *d_constraint = nullptr;
  // general linear constraints (rows = constraints) and transpose (rows = DOFs)
  int *d_j_off = nullptr, *d_j_cols = nullptr;
  double *d_j_vals = nullptr, *d_j_rhs = nullptr;
  int *d_jt_off = nullptr, *d_jt_cols = nullptr;
  double *d_jt_vals = nullptr;
  int64_t j_nnz = 0;
};

struct tlfea_newton;
namespace tlfea {
int build_row_order(const tlfea_elem *h, int **d_order);
int build_completion_order(const tlfea_elem *h, int chunk, int **d_order, std::vector<int> *chunk_ptr);
// callbacks of the chunked launch order (host-buffer step): before the first and after the last kernel of chunk c
struct PatchHooks {
  void *ctx;
  int (*before_chunk)(void *ctx, int chunk);
  int (*after_chunk)(void *ctx, int chunk);
};
int patch_prepare(tlfea_newton *s);
int patch_assemble(tlfea_newton *s, bool tangent, bool damp, const PatchHooks *hooks = nullptr);
int patch_chunks(tlfea_newton *s, int *n_chunks, const int **range_ready);
// coefficient-id ranges [lo, hi) this rank shares with its neighbours (csrc/dist.cu); empty without a communicator
void dist_halo_ranges(const tlfea_newton *s, std::vector<std::pair<int64_t, int64_t>> &out);
void patch_release(tlfea_newton *s);
void dist_release(tlfea_newton *s);
int patch_stats(tlfea_newton *s, int64_t *out);
}

struct tlfea_newton {
  tlfea_elem *e = nullptr;
  int device = 0;
  // 0 = generic row gather, 1 = element kernel + row reduce (two-phase), 2 = single-pass patch assembly for FEAT10
  // (pk_kernel.cuh; the ANCF elements use the two-phase path)
  int asm_mode = 2;
  void *patch = nullptr;  // PatchPlanDev (patch_asm.cu)
  void *dist = nullptr;   // DistCtx (dist.cu): NCCL communicator + halo neighbours of a partitioned mesh
  int pk_w = 0;           // warps per patch (0 = TLFEA_PK_W or 4)
  int pk_chunks = 0;      // id-range chunks of the launch order (0 = TLFEA_PK_CHUNKS or at most 4, by mesh size)
  int pk_affine = -1;     // affine-element compression: -1 = TLFEA_PK_AFFINE (default off), 0 off, 1 on
  int n_constraints = 0;
  tlfea_newton_params prm{};
  bool analyzed = false, fixed_pattern = false;
  int pattern_version_seen = -1;  // the element pattern the solver's H, slots and plans were built for
  int64_t nnz = 0;  // DOF-level nnz = 9 * pattern.nnz
  double *d_H = nullptr;
  double *d_gp = nullptr;  // per-Gauss-point records
  double *d_stage = nullptr;  // symmetric element block records of the two-phase assembly (ek_kernel.cuh)
  void *d_rowdesc = nullptr;  // RowDesc[n_coef] in processing order (completion order, Morton ties; ek_kernel.cuh)
  std::vector<int> chunk_ptr;  // rows completed by element chunk k: [chunk_ptr[k], chunk_ptr[k+1])
  int ek_chunk = 0;            // elements per chunk of the interleaved E/R schedule
  int *d_tickets = nullptr;    // one ticket counter per chunk
  int rec_doubles = 0;
  double *d_v = nullptr, *d_v_prev = nullptr, *d_lambda = nullptr, *d_g = nullptr, *d_dv = nullptr, *d_r = nullptr;
  double *d_xp = nullptr, *d_yp = nullptr, *d_zp = nullptr;
  // constraint term h^2 rho J^T J as a (slot, coefficient) list with unique slots
  int cons_version_seen = 0;
  int64_t cons_own_lo = 0, cons_own_hi = -2;  // owned range the constraint slots were built for
  int64_t own_lo = 0, own_hi = -1;  // DOF range owned by this rank (dot products of the distributed solve)
  int64_t n_cons_slots = 0;
  int64_t *d_cons_slot = nullptr;
  double *d_cons_val = nullptr;
  // PCG workspace
  double *d_pcg[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  double *d_scal = nullptr;  // device scalars
  double *h_scal = nullptr;  // pinned
  int pcg_max_iters = 200;
  double pcg_rel_tol = 1e-8;
  int last_newton = 0, last_pcg = 0;
  double last_pcg_rel = 0;     // ||r|| / ||r0|| reached by the last linear solve
  bool last_pcg_converged = true;  // false: the last linear solve stopped at max_iters above rel_tol
  double last_norm_g = 0, last_norm_c = 0;
  // pipelined host-buffer step (tlfea_newton_assemble_host_begin / _end): copy stream, row-reduce stream,
  // events 0..15 = coordinate chunk k is on the device, 16 = tail of the main stream, 17 = residual kernels queued,
  // 18 = H rows done
  cudaStream_t st_copy = nullptr, st_rows = nullptr;
  cudaEvent_t ev_pipe[19] = {};
  std::vector<int> pipe_need;  // last coordinate chunk the element chunk j reads
  bool pipe_open = false, pipe_overlapped = false;
  // single-pass (patch) host-buffer step: per id range, the event after its force gather; the caller's result buffer
  // (registered before begin) receives finished ranges while later chunks still compute
  cudaEvent_t ev_g[16] = {};
  double *h_f_out = nullptr;
  bool pipe_pk = false, pipe_f_copied = false;
  // timing
  // A ring of event sets (start, end, two phase marks): tlfea_newton_assemble records and returns; elapsed times are
  // read when the ring is full or when a timing getter is called -- no host synchronisation per assemble
  static constexpr int EV_RING = 16;
  cudaEvent_t ev_ring[EV_RING][4] = {};
  cudaEvent_t *ev = ev_ring[0];
  int ev_pending = 0;
  bool ev_tangent[EV_RING] = {};
  double ms_total = 0;
  int calls = 0;
  bool profile = false;
  double k_ms[3] = {0, 0, 0};
};
