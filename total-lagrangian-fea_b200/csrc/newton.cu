// newton.cu -- host-side mirror of SyncedNewtonSolver (lib_src/solvers/SyncedNewton.{cu,cuh})
// for the hot path: sparsity analysis, the fused force + tangent + H assembly, the residual,
// and one implicit ALM/Newton step whose SPD solve is a block-Jacobi PCG (cuDSS is absent here).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "patch_plan.cuh"
#include "kernels.cuh"
#include "ek_kernel.cuh"
#include "ancf_ek.cuh"

namespace tlfea {
int launch_gp(tlfea_elem *h, const double *v, double *rec, double *fe, double *F_out, double *P_out);
int ensure_pattern(tlfea_elem *h);

// ---------------------------------------------------------------------------
// small vector kernels (SyncedNewton.cu:412-534)
// ---------------------------------------------------------------------------
__global__ void k_grad_L(int n_coef, const int *__restrict__ off, const int *__restrict__ cols,
                         const double *__restrict__ mass, const double *__restrict__ v,
                         const double *__restrict__ v_prev, const double *__restrict__ f_int,
                         const double *__restrict__ f_ext, double inv_h, double h, const int *jt_off,
                         const int *jt_cols, const double *jt_vals, const double *lam, const double *con, double rho,
                         int64_t own_lo, int64_t own_hi, double *g) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_coef * 3) return;
  const int ni = (int)(tid / 3), di = (int)(tid % 3);
  double res = 0.0;
  // (M (v - v_prev)) / h, summed in the reference order; the loads of four entries are issued together (the
  // gathers depend on the column loads: one entry at a time leaves the memory system idle), the adds stay in order
  const int r0 = off[ni], r1 = off[ni + 1];
  int idx = r0;
  for (; idx + 4 <= r1; idx += 4) {
    int64_t tj[4];
    double m[4], a[4], b[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      tj[k] = (int64_t)cols[idx + k] * 3 + di;
      m[k] = mass[idx + k];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      a[k] = v[tj[k]];
      b[k] = v_prev[tj[k]];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) res += m[k] * (a[k] - b[k]) * inv_h;
  }
  for (; idx < r1; idx++) {
    const int64_t tj = (int64_t)cols[idx] * 3 + di;
    res += mass[idx] * (v[tj] - v_prev[tj]) * inv_h;
  }
  res -= (-f_int[tid]);
  // distributed solve: M and f_int are sub-assembled (summed over ranks by the halo sum), f_ext and the constraint
  // terms are NOT -- only the rank that owns the DOF adds them
  const bool own = tid >= own_lo && tid < own_hi;
  if (own) res -= f_ext[tid];
  if (jt_off && own) {
    for (int idx = jt_off[tid]; idx < jt_off[tid + 1]; idx++) {
      const int c = jt_cols[idx];
      res += h * jt_vals[idx] * (lam[c] + rho * con[c]);
    }
  }
  g[tid] = res;
}

// fixed-node constraints: J^T has exactly one unit entry per constrained DOF
__global__ void k_grad_L_fixed(int n_fixed, const int *__restrict__ fixed, const double *__restrict__ lam,
                               const double *__restrict__ con, double h, double rho, int64_t own_lo, int64_t own_hi,
                               double *g) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 3 * n_fixed) return;
  const int64_t dof = (int64_t)fixed[c / 3] * 3 + c % 3;
  if (dof < own_lo || dof >= own_hi) return;  // the owner rank adds the term (see k_grad_L)
  g[dof] += h * 1.0 * (lam[c] + rho * con[c]);
}

__global__ void k_fixed_cons_slots(int n_fixed, const int *__restrict__ fixed, const int *__restrict__ off,
                                   const int *__restrict__ cols, int64_t own_lo, int64_t own_hi, int64_t *slot,
                                   double *val) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 3 * n_fixed) return;
  const int n = fixed[c / 3], d = c % 3;
  const int rs = off[n], deg = off[n + 1] - rs;
  int lo = 0, hi = deg - 1, pos = 0;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1, cc = cols[rs + mid];
    if (cc == n) {
      pos = mid;
      break;
    }
    if (cc < n)
      lo = mid + 1;
    else
      hi = mid - 1;
  }
  slot[c] = 9 * (int64_t)rs + (int64_t)d * 3 * deg + 3 * pos + d;
  const int64_t dof = 3 * (int64_t)n + d;
  val[c] = (dof >= own_lo && dof < own_hi) ? 1.0 : 0.0;  // h^2 rho J^T J goes into the owner rank's sub-assembled H
}

// H += h^2 rho * (J^T J) on its precomputed unique slots   (SyncedNewton.cu:291-341)
__global__ void k_add_constraint_term(int64_t n, const int64_t *__restrict__ slot, const double *__restrict__ val,
                                      double factor, double *H) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) H[slot[i]] += factor * val[i];
}

__global__ void k_axpy(int64_t n, double a, const double *x, double *y) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] += a * x[i];
}
__global__ void k_update_pos(int n, const double *xp, const double *yp, const double *zp, const double *v, double h,
                             double *x, double *y, double *z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  x[i] = xp[i] + v[3 * (int64_t)i + 0] * h;
  y[i] = yp[i] + v[3 * (int64_t)i + 1] * h;
  z[i] = zp[i] + v[3 * (int64_t)i + 2] * h;
}
__global__ void k_dual_update(int n, const double *con, double rho, double *lam) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lam[i] += rho * con[i];
}

// deterministic dot product: fixed grid, fixed tree
constexpr int DOT_BLOCKS = 592;  // 4 x 148 SMs
__global__ void __launch_bounds__(256) k_dot_partial(int64_t n, const double *__restrict__ a,
                                                     const double *__restrict__ b, double *partial) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) s += a[i] * b[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(1024) k_dot_final(int nb, const double *__restrict__ partial, double *out) {
  __shared__ double sh[1024];
  double s = 0.0;
  for (int i = threadIdx.x; i < nb; i += 1024) s += partial[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sh[0];
}

// y = H x; one warp per coefficient (3 scalar rows), fixed shuffle tree
__global__ void __launch_bounds__(256) k_spmv(int n_coef, const int *__restrict__ off, const int *__restrict__ cols,
                                              const double *__restrict__ H, const double *__restrict__ x, double *y) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= n_coef) return;
  const int rs = off[row], deg = off[row + 1] - rs;
  const int64_t base = 9 * (int64_t)rs;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0;
  // the three scalar rows of the block row are contiguous runs of 3*deg doubles: lane-linear (coalesced) reads,
  // x gathered per (column, component)
  const double *h0 = H + base, *h1 = h0 + 3 * deg, *h2 = h1 + 3 * deg;
  for (int t = lane; t < 3 * deg; t += 32) {
    const int sl = t / 3;
    const double xv = x[3 * (int64_t)cols[rs + sl] + (t - 3 * sl)];
    s0 += h0[t] * xv;
    s1 += h1[t] * xv;
    s2 += h2[t] * xv;
  }
  for (int o = 16; o > 0; o >>= 1) {
    s0 += __shfl_xor_sync(0xffffffffu, s0, o);
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (lane == 0) {
    y[3 * (int64_t)row] = s0;
    y[3 * (int64_t)row + 1] = s1;
    y[3 * (int64_t)row + 2] = s2;
  }
}

// block-Jacobi preconditioner: Dinv = inverse of the 3x3 diagonal blocks of H, built once per linear solve
// (layout [9][n_coef] so the apply kernel reads it coalesced); identity where the block is singular
__global__ void k_block_jacobi_setup(int n_coef, const int *__restrict__ off, const int *__restrict__ cols,
                                     const double *__restrict__ H, double *__restrict__ Dinv) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_coef) return;
  const int rs = off[i], deg = off[i + 1] - rs;
  int lo = 0, hi = deg - 1, pos = 0;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1, c = cols[rs + mid];
    if (c == i) {
      pos = mid;
      break;
    }
    if (c < i)
      lo = mid + 1;
    else
      hi = mid - 1;
  }
  double m[9] = {1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0};
  if (deg == 0) {  // an isolated coefficient has no diagonal block to read
#pragma unroll
    for (int q = 0; q < 9; q++) Dinv[(int64_t)q * n_coef + i] = m[q];
    return;
  }
  const double *h0 = H + 9 * (int64_t)rs + 3 * pos, *h1 = h0 + 3 * deg, *h2 = h1 + 3 * deg;
  const double a = h0[0], b = h0[1], c = h0[2], d = h1[0], e = h1[1], f = h1[2], g = h2[0], hh = h2[1], k = h2[2];
  const double A = e * k - f * hh, B = -(d * k - f * g), C = d * hh - e * g;
  const double det = a * A + b * B + c * C;
  if (fabs(det) >= 1e-300) {
    const double id = 1.0 / det;
    m[0] = A * id;
    m[1] = -(b * k - c * hh) * id;
    m[2] = (b * f - c * e) * id;
    m[3] = B * id;
    m[4] = (a * k - c * g) * id;
    m[5] = -(a * f - c * d) * id;
    m[6] = C * id;
    m[7] = -(a * hh - b * g) * id;
    m[8] = (a * e - b * d) * id;
  }
#pragma unroll
  for (int q = 0; q < 9; q++) Dinv[(int64_t)q * n_coef + i] = m[q];
}
// distributed solve: raw diagonal blocks [n_coef][9] (interface nodes are completed by a halo sum before inversion)
__global__ void k_diag_extract(int n_coef, const int *__restrict__ off, const int *__restrict__ cols,
                               const double *__restrict__ H, double *__restrict__ diag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_coef) return;
  const int rs = off[i], deg = off[i + 1] - rs;
  int lo = 0, hi = deg - 1, pos = -1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1, c = cols[rs + mid];
    if (c == i) {
      pos = mid;
      break;
    }
    if (c < i)
      lo = mid + 1;
    else
      hi = mid - 1;
  }
  double *o = diag + 9 * (int64_t)i;
  if (pos < 0) {
#pragma unroll
    for (int q = 0; q < 9; q++) o[q] = 0.0;
    return;
  }
  const double *h0 = H + 9 * (int64_t)rs + 3 * pos, *h1 = h0 + 3 * deg, *h2 = h1 + 3 * deg;
  o[0] = h0[0], o[1] = h0[1], o[2] = h0[2];
  o[3] = h1[0], o[4] = h1[1], o[5] = h1[2];
  o[6] = h2[0], o[7] = h2[1], o[8] = h2[2];
}
__global__ void k_diag_invert(int n_coef, const double *__restrict__ diag, double *__restrict__ Dinv) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_coef) return;
  const double *o = diag + 9 * (int64_t)i;
  const double a = o[0], b = o[1], c = o[2], d = o[3], e = o[4], f = o[5], g = o[6], hh = o[7], k = o[8];
  const double A = e * k - f * hh, B = -(d * k - f * g), C = d * hh - e * g;
  const double det = a * A + b * B + c * C;
  double m[9] = {1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0};
  if (fabs(det) >= 1e-300) {
    const double id = 1.0 / det;
    m[0] = A * id;
    m[1] = -(b * k - c * hh) * id;
    m[2] = (b * f - c * e) * id;
    m[3] = B * id;
    m[4] = (a * k - c * g) * id;
    m[5] = -(a * f - c * d) * id;
    m[6] = C * id;
    m[7] = -(a * hh - b * g) * id;
    m[8] = (a * e - b * d) * id;
  }
#pragma unroll
  for (int q = 0; q < 9; q++) Dinv[(int64_t)q * n_coef + i] = m[q];
}
// z = Dinv r
__global__ void k_block_jacobi(int n_coef, const double *__restrict__ Dinv, const double *__restrict__ r, double *z) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_coef) return;
  const double r0 = r[3 * (int64_t)i], r1 = r[3 * (int64_t)i + 1], r2 = r[3 * (int64_t)i + 2];
  const double *m = Dinv + i;
  const int64_t n = n_coef;
  z[3 * (int64_t)i] = m[0] * r0 + m[n] * r1 + m[2 * n] * r2;
  z[3 * (int64_t)i + 1] = m[3 * n] * r0 + m[4 * n] * r1 + m[5 * n] * r2;
  z[3 * (int64_t)i + 2] = m[6 * n] * r0 + m[7 * n] * r1 + m[8 * n] * r2;
}

__global__ void k_pcg_update_xr(int64_t n, const double *scal, const double *p, const double *Ap, double *x,
                                double *r) {
  // alpha = rz / pAp
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double alpha = scal[0] / scal[1];
  x[i] += alpha * p[i];
  r[i] -= alpha * Ap[i];
}
__global__ void k_pcg_update_p(int64_t n, const double *scal, const double *z, double *p) {
  // beta = rz_new / rz_old
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double beta = scal[2] / scal[0];
  p[i] = z[i] + beta * p[i];
}
__global__ void k_neg_copy(int64_t n, const double *g, double *r) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) r[i] = -g[i];
}
__global__ void k_copy_scalar(double *dst, const double *src) { *dst = *src; }

template <int S, int Q>
static int launch_row_t(tlfea_newton *s, const RowArgs &a, int grid, size_t smem, bool damp) {
  tlfea_elem *e = s->e;
#define TL_ROW(MODEL, DAMP)                                                                                     \
  do {                                                                                                          \
    if (smem > 48 * 1024)                                                                                       \
      TL_CUDA(cudaFuncSetAttribute(row_kernel<S, Q, MODEL, DAMP>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                   (int)smem));                                                                 \
    TL_LAUNCH((row_kernel<S, Q, MODEL, DAMP>), grid, ROW_WARPS * 32, smem, e->stream, a);                        \
  } while (0)
  if (e->model == TL_MODEL_SVK) {
    if (damp)
      TL_ROW(TL_MODEL_SVK, true);
    else
      TL_ROW(TL_MODEL_SVK, false);
  } else {
    if (damp)
      TL_ROW(TL_MODEL_MR, true);
    else
      TL_ROW(TL_MODEL_MR, false);
  }
#undef TL_ROW
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

static int dot(tlfea_newton *s, const double *a, const double *b, double *d_out) {
  const int64_t n = 3 * (int64_t)s->e->n_coef;
  double *partial = s->d_scal + 8;
  TL_LAUNCH(k_dot_partial, DOT_BLOCKS, 256, 0, s->e->stream, n, a, b, partial);
  TL_LAUNCH(k_dot_final, 1, 1024, 0, s->e->stream, DOT_BLOCKS, partial, d_out);
  return TLFEA_OK;
}

// partial dot over the DOF range this rank owns (distributed solve: shared DOFs are counted once)
static int dot_owned(tlfea_newton *s, const double *a, const double *b, double *d_out) {
  const int64_t n = 3 * (int64_t)s->e->n_coef;
  const int64_t lo = std::max<int64_t>(0, s->own_lo), hi = (s->own_hi < 0) ? n : std::min<int64_t>(n, s->own_hi);
  double *partial = s->d_scal + 8;
  TL_LAUNCH(k_dot_partial, DOT_BLOCKS, 256, 0, s->e->stream, std::max<int64_t>(hi - lo, 0), a + lo, b + lo, partial);
  TL_LAUNCH(k_dot_final, 1, 1024, 0, s->e->stream, DOT_BLOCKS, partial, d_out);
  return TLFEA_OK;
}

static int norm2(tlfea_newton *s, const double *a, int64_t n, double *out) {
  double *partial = s->d_scal + 8;
  TL_LAUNCH(k_dot_partial, DOT_BLOCKS, 256, 0, s->e->stream, n, a, a, partial);
  TL_LAUNCH(k_dot_final, 1, 1024, 0, s->e->stream, DOT_BLOCKS, partial, s->d_scal + 4);
  TL_CUDA(cudaMemcpyAsync(s->h_scal, s->d_scal + 4, sizeof(double), cudaMemcpyDeviceToHost, s->e->stream));
  TL_CUDA(cudaStreamSynchronize(s->e->stream));
  *out = std::sqrt(s->h_scal[0]);
  return TLFEA_OK;
}

static int ensure_records(tlfea_newton *s) {
  const int need = (s->e->model == TL_MODEL_MR) ? REC_MR : REC_SVK;
  if (need != s->rec_doubles) {
    dfree(s->d_gp);
    s->rec_doubles = need;
    TL_TRY(dmalloc(&s->d_gp, (size_t)s->e->n_elem * s->e->Q * need));
  }
  return TLFEA_OK;
}

// ---------------------------------------------------------------------------
// two-phase assembly (ek_kernel.cuh): element kernel -> symmetric block records -> row reduce
// ---------------------------------------------------------------------------
struct EKCfg {
  int g, w;
};
static EKCfg ek_cfg() {
  static EKCfg c = [] {
    EKCfg t{2, 8};
    const char *e = getenv("TLFEA_EK_CFG");  // "65" = groups of 6, 5 warps ... (tuning aid)
    if (e && strlen(e) == 2) t = EKCfg{e[0] - '0', e[1] - '0'};
    return t;
  }();
  return c;
}

static bool ek_supported(const tlfea_newton *s) {
  return s->asm_mode >= 1;  // every element type, SVK and Mooney-Rivlin, with or without Kelvin-Voigt damping
}
// mode 2: FEAT10 + SVK (with or without Kelvin-Voigt damping) takes the single-pass patch kernel; Mooney-Rivlin keeps
// the two-phase path, which is faster for it (its 81-entry tangent tensor leaves room for one 4-warp CTA per SM only:
// 70.8 ms against 65.7 ms per 10.1 M-tet assemble).  mode 3 forces the patch kernel for every FEAT10 material.
static bool pk_selected(const tlfea_newton *s) {
  if (s->e->type != TLFEA_TYPE_T10) return false;
  return s->asm_mode == 3 || (s->asm_mode == 2 && s->e->model == TL_MODEL_SVK);
}

// single-pass patch assembly (pk_kernel.cuh) + the ordered nodal force gather
static int assemble_pk(tlfea_newton *s, bool tangent, bool damp) {
  tlfea_elem *e = s->e;
  TL_TRY(patch_assemble(s, tangent, damp));
  if (s->profile) TL_CUDA(cudaEventRecord(s->ev[2], e->stream));
  TL_LAUNCH(force_gather_kernel, blocks_for((int64_t)e->n_coef * 3, 128), 128, 0, e->stream, e->n_coef,
            e->pat.d_inc_ptr, e->pat.d_inc_ea, e->d_fe, e->d_f_int, 0);
  if (s->profile) TL_CUDA(cudaEventRecord(s->ev[3], e->stream));
  return TLFEA_OK;
}

template <int G, int W, bool DAMP, bool TANGENT, int MODEL>
static int launch_ek_t(tlfea_elem *e, const EKArgs &a) {
  const size_t smem = EKShape<G, DAMP, MODEL>::smem_bytes(W);
  TL_CHECK(smem <= 227 * 1024, "element kernel configuration needs %zu bytes of shared memory", smem);
  static thread_local int attr_dev = -1;  // the opt-in is per function and device: set it once
  if (attr_dev != e->device) {
    TL_CUDA(cudaFuncSetAttribute(ek_kernel<G, W, DAMP, TANGENT, MODEL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    attr_dev = e->device;
  }
  const int n_groups = (a.n_elem + G - 1) / G;
  const int grid = std::max(1, std::min(sm_count(e->device), (n_groups + W - 1) / W));
  TL_LAUNCH((ek_kernel<G, W, DAMP, TANGENT, MODEL>), grid, W * 32, smem, e->stream, a);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

template <bool DAMP, bool TANGENT>
static int launch_ek_cfg(tlfea_elem *e, const EKArgs &a) {
  // Mooney-Rivlin keeps an 81-entry tangent tensor per Gauss point in shared memory: 7 warps fit
  if (e->model == TL_MODEL_MR) return launch_ek_t<2, 7, DAMP, TANGENT, TL_MODEL_MR>(e, a);
  const EKCfg c = ek_cfg();
  if (!DAMP && c.g == 2 && c.w == 9) return launch_ek_t<2, 9, DAMP, TANGENT, TL_MODEL_SVK>(e, a);
  return launch_ek_t<2, 8, DAMP, TANGENT, TL_MODEL_SVK>(e, a);  // most warps per SM: the kernel is latency / LSU bound
}

template <int S, int Q, int EPB, bool DAMP, bool TANGENT, int MODEL>
static int launch_aek_t(tlfea_elem *e, const AEKArgs &a) {
  using Sh = AEKShape<S, Q, EPB, DAMP, MODEL>;
  static_assert(Sh::SMEM <= 227 * 1024, "fits the shared memory of one SM");
  if (Sh::SMEM > 48 * 1024)
    TL_CUDA(cudaFuncSetAttribute(aek_kernel<S, Q, EPB, DAMP, TANGENT, MODEL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)Sh::SMEM));
  const int n_groups = (a.n_elem + EPB - 1) / EPB;
  const int per_sm = std::max(1, std::min((int)((227 * 1024) / (Sh::SMEM + 1024)), 2048 / Sh::THREADS));
  const int grid = std::max(1, std::min(n_groups, sm_count(e->device) * per_sm));
  TL_LAUNCH((aek_kernel<S, Q, EPB, DAMP, TANGENT, MODEL>), grid, Sh::THREADS, Sh::SMEM, e->stream, a);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}
template <bool DAMP, bool TANGENT>
static int launch_aek(tlfea_elem *e, const AEKArgs &a) {
  if (e->model == TL_MODEL_MR) {  // 3 elements per CTA for the beam: the tangent tensors take 90 doubles per point
    if (e->type == TLFEA_TYPE_3243) return launch_aek_t<8, 12, 3, DAMP, TANGENT, TL_MODEL_MR>(e, a);
    return launch_aek_t<16, 48, 1, DAMP, TANGENT, TL_MODEL_MR>(e, a);
  }
  if (e->type == TLFEA_TYPE_3243) return launch_aek_t<8, 12, 7, DAMP, TANGENT, TL_MODEL_SVK>(e, a);
  return launch_aek_t<16, 48, 1, DAMP, TANGENT, TL_MODEL_SVK>(e, a);
}

static int ek_chunk_elems() {
  static int c = [] {
    const char *e = getenv("TLFEA_EK_CHUNK");  // elements per interleaved E/R chunk (0 = two full passes)
    return e ? atoi(e) : 0;
  }();
  return c;
}

// Arguments and launch helpers of the two-phase assembly, shared by the plain schedule (assemble_ek) and the
// pipelined host-buffer schedule (assemble_host_begin).
struct EkPlan {
  tlfea_newton *s = nullptr;
  bool tangent = true, damp = false, t10 = true;
  int64_t rec = 0;
  EKArgs a;
  AEKArgs aa;
  RKArgs r;
  size_t smem = 0;
  int per_sm = 1, rk_cfg = 0;

  int init(tlfea_newton *s_, bool tangent_, bool damp_) {
    s = s_, tangent = tangent_, damp = damp_;
    tlfea_elem *e = s->e;
    const double h = s->prm.time_step;
    t10 = (e->type == TLFEA_TYPE_T10);
    rec = (int64_t)e->S * e->S * 9;  // doubles of one element's block records [a][b][9]
    if (tangent && !s->d_stage) TL_TRY(dmalloc(&s->d_stage, (size_t)e->n_elem * rec));
    if (tangent && !s->d_rowdesc) {
      int chunk = ek_chunk_elems() & ~1;  // element pairs: an odd chunk would misalign the 16-byte bulk copies
      if (chunk <= 0 || chunk >= e->n_elem) chunk = e->n_elem;
      s->ek_chunk = chunk;
      int *d_order = nullptr;
      TL_TRY(build_completion_order(e, chunk, &d_order, &s->chunk_ptr));
      RowDesc *desc = nullptr;
      TL_TRY(dmalloc(&desc, (size_t)e->n_coef));
      TL_LAUNCH(k_row_desc, blocks_for(e->n_coef, 256), 256, 0, e->stream, e->n_coef, d_order, e->pat.d_off,
                e->pat.d_inc_ptr, desc);
      TL_CUDA(cudaStreamSynchronize(e->stream));
      cudaFree(d_order);
      s->d_rowdesc = desc;
      TL_TRY(dmalloc(&s->d_tickets, s->chunk_ptr.size()));
    }
    a.x = e->d_x;
    a.y = e->d_y;
    a.z = e->d_z;
    a.v = damp ? s->d_v : nullptr;
    a.qw = e->qw;
    a.lambda = e->lambda;
    a.mu = e->mu;
    a.eta = e->eta;
    a.lambda_d = e->lambda_d;
    a.c1 = h * e->lambda + e->lambda_d;
    a.c2 = h * e->mu + e->eta;
    a.c3 = h * e->mu + e->eta;
    a.c4 = h;
    a.mu10 = e->mu10;
    a.mu01 = e->mu01;
    a.kappa = e->kappa;
    aa.x = e->d_x;
    aa.y = e->d_y;
    aa.z = e->d_z;
    aa.v = damp ? s->d_v : nullptr;
    aa.qw = e->qw;
    aa.lambda = e->lambda;
    aa.mu = e->mu;
    aa.eta = e->eta;
    aa.lambda_d = e->lambda_d;
    aa.c1 = a.c1;
    aa.c2 = a.c2;
    aa.c4 = h;
    aa.mu10 = e->mu10;
    aa.mu01 = e->mu01;
    aa.kappa = e->kappa;
    if (!tangent) return TLFEA_OK;
    r.inc_ea = e->pat.d_inc_ea;
    r.slot = e->pat.d_slot;
    r.mass = e->d_mass;
    r.stage = s->d_stage;
    r.fe = e->d_fe;
    r.inv_h = 1.0 / h;
    r.H = s->d_H;
    r.f_int = e->d_f_int;
    r.row_stride = ((e->pat.max_deg * 9 + 32 + 1) | 1);  // block row + 32 doubles for the force gather
    smem = (size_t)RK_WARPS * r.row_stride * sizeof(double);
    TL_CHECK(smem <= 220 * 1024, "coefficient row too long for the row-reduce kernel (max_deg=%d)", e->pat.max_deg);
    static const int cfg = [] {
      const char *v = getenv("TLFEA_RK_CFG");  // tuning aid: 0 = (U 5, 3 blocks/SM), 1 = (4, 4), 2 = (8, 2)
      return v ? atoi(v) : 0;
    }();
    rk_cfg = cfg;
    const int want_sm = rk_cfg == 1 ? 4 : (rk_cfg == 2 ? 2 : 3);
    if (smem > 48 * 1024) {
      TL_CUDA(cudaFuncSetAttribute(rk_kernel<10, 5, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      TL_CUDA(cudaFuncSetAttribute(rk_kernel<10, 4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      TL_CUDA(cudaFuncSetAttribute(rk_kernel<10, 8, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      TL_CUDA(cudaFuncSetAttribute(rk_kernel<8, 5, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      TL_CUDA(cudaFuncSetAttribute(rk_kernel<16, 3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    per_sm = std::max(1, std::min(want_sm, (int)((227 * 1024) / (smem + 2048))));
    return TLFEA_OK;
  }

  // element kernel over elements [e0, e0 + ne) on the element's stream
  int elems(int e0, int ne) {
    tlfea_elem *e = s->e;
    if (!t10) {
      aa.n_elem = ne;
      aa.ecoef = e->d_ecoef + (int64_t)e0 * e->S;
      aa.gradN = e->d_gradN + (int64_t)e0 * e->Q * 3 * e->S;
      aa.detJ = e->d_detJ + (int64_t)e0 * e->Q;
      aa.fe = e->d_fe + (int64_t)e0 * e->S * 3;
      aa.stage = tangent ? s->d_stage + (int64_t)e0 * rec : nullptr;
      if (tangent) return damp ? launch_aek<true, true>(e, aa) : launch_aek<false, true>(e, aa);
      return damp ? launch_aek<true, false>(e, aa) : launch_aek<false, false>(e, aa);
    }
    a.n_elem = ne;
    a.ecoef = e->d_ecoef + (int64_t)e0 * 10;
    a.gradN = e->d_gradN + (int64_t)e0 * 150;
    a.detJ = e->d_detJ + (int64_t)e0 * 5;
    a.fe = e->d_fe + (int64_t)e0 * 30;
    a.stage = tangent ? s->d_stage + (int64_t)e0 * EK_ELEM : nullptr;
    if (tangent) return damp ? launch_ek_cfg<true, true>(e, a) : launch_ek_cfg<false, true>(e, a);
    return damp ? launch_ek_cfg<true, false>(e, a) : launch_ek_cfg<false, false>(e, a);
  }

  // row-reduce kernel over the rows completed by element chunk k, on stream st; with_force = false leaves f_int
  // to a separate force gather (pipelined schedule)
  int rows(int k, cudaStream_t st, bool with_force) {
    tlfea_elem *e = s->e;
    const int r0 = s->chunk_ptr[k], nr = s->chunk_ptr[k + 1] - r0;
    if (nr <= 0) return TLFEA_OK;
    r.f_int = with_force ? e->d_f_int : nullptr;
    r.n_coef = nr;
    r.desc = static_cast<const RowDesc *>(s->d_rowdesc) + r0;
    r.counter = s->d_tickets + k;
    const int blocks = (nr + RK_WARPS * RK_T - 1) / (RK_WARPS * RK_T);
    const int grid = std::min(blocks, sm_count(e->device) * per_sm);
    if (e->type == TLFEA_TYPE_3243)
      TL_LAUNCH((rk_kernel<8, 5, 3>), grid, RK_WARPS * 32, smem, st, r);
    else if (e->type == TLFEA_TYPE_3443)
      TL_LAUNCH((rk_kernel<16, 3, 3>), grid, RK_WARPS * 32, smem, st, r);
    else if (rk_cfg == 1)
      TL_LAUNCH((rk_kernel<10, 4, 4>), grid, RK_WARPS * 32, smem, st, r);
    else if (rk_cfg == 2)
      TL_LAUNCH((rk_kernel<10, 8, 2>), grid, RK_WARPS * 32, smem, st, r);
    else
      TL_LAUNCH((rk_kernel<10, 5, 3>), grid, RK_WARPS * 32, smem, st, r);
    return TLFEA_OK;
  }
};

// Schedule: the element range is cut into chunks; for chunk k the element kernel writes the chunk's block
// records and the row-reduce kernel consumes them for the rows that chunk completes (row order = completion
// order, see build_completion_order).  Default: one chunk = two full passes.
static int assemble_ek(tlfea_newton *s, bool tangent, bool damp) {
  tlfea_elem *e = s->e;
  EkPlan pl;
  TL_TRY(pl.init(s, tangent, damp));
  if (!tangent) {
    TL_TRY(pl.elems(0, e->n_elem));
    if (s->profile) TL_CUDA(cudaEventRecord(s->ev[2], e->stream));
    TL_LAUNCH(force_gather_kernel, blocks_for((int64_t)e->n_coef * 3, 128), 128, 0, e->stream, e->n_coef,
              e->pat.d_inc_ptr, e->pat.d_inc_ea, e->d_fe, e->d_f_int, 0);
    return TLFEA_OK;
  }
  const int n_chunks = (int)s->chunk_ptr.size() - 1;
  TL_CUDA(cudaMemsetAsync(s->d_tickets, 0, s->chunk_ptr.size() * sizeof(int), e->stream));
  for (int k = 0; k < n_chunks; k++) {
    const int e0 = k * s->ek_chunk, ne = std::min(s->ek_chunk, e->n_elem - e0);
    TL_TRY(pl.elems(e0, ne));
    if (s->profile && k == n_chunks - 1) TL_CUDA(cudaEventRecord(s->ev[2], e->stream));
    TL_TRY(pl.rows(k, e->stream, true));
  }
  TL_CUDA(cudaGetLastError());
  if (s->profile) TL_CUDA(cudaEventRecord(s->ev[3], e->stream));
  return TLFEA_OK;
}

static int64_t own_lo_of(const tlfea_newton *s) { return std::max<int64_t>(0, s->own_lo); }
static int64_t own_hi_of(const tlfea_newton *s) {
  const int64_t n = 3 * (int64_t)s->e->n_coef;
  return (s->own_hi < 0) ? n : std::min<int64_t>(n, s->own_hi);
}
// slots of the h^2 rho J^T J diagonal entries of pinned nodes (J = selection rows)
static int fixed_cons_slots(tlfea_newton *s) {
  tlfea_elem *e = s->e;
  TL_CHECK(s->n_constraints == e->n_constraint, "n_constraints (%d) != element constraints (%d)", s->n_constraints,
           e->n_constraint);
  dfree(s->d_cons_slot);
  dfree(s->d_cons_val);
  s->d_cons_slot = nullptr;
  s->d_cons_val = nullptr;
  s->n_cons_slots = 3 * (int64_t)e->n_fixed;
  TL_TRY(dmalloc(&s->d_cons_slot, (size_t)std::max<int64_t>(s->n_cons_slots, 1)));
  TL_TRY(dmalloc(&s->d_cons_val, (size_t)std::max<int64_t>(s->n_cons_slots, 1)));
  if (e->n_fixed > 0)
    TL_LAUNCH(k_fixed_cons_slots, blocks_for(3 * e->n_fixed, 256), 256, 0, e->stream, e->n_fixed, e->d_fixed,
              e->pat.d_off, e->pat.d_cols, own_lo_of(s), own_hi_of(s), s->d_cons_slot, s->d_cons_val);
  s->cons_version_seen = e->cons_version;
  s->cons_own_lo = own_lo_of(s), s->cons_own_hi = own_hi_of(s);
  return TLFEA_OK;
}
static int refresh_constraints(tlfea_newton *s) {  // the element's pinned set changed (UpdateNodalFixed)
  if (!s->analyzed || s->n_constraints == 0) return TLFEA_OK;
  const bool own_same = s->e->cons_mode != TL_CONS_FIXED_NODES ||
                        (s->cons_own_lo == own_lo_of(s) && s->cons_own_hi == own_hi_of(s));
  if (s->cons_version_seen == s->e->cons_version && own_same) return TLFEA_OK;
  TL_CHECK(s->e->cons_mode == TL_CONS_FIXED_NODES, "constraint kind changed under a live solver");
  return fixed_cons_slots(s);
}

static int do_assemble(tlfea_newton *s, bool tangent) {
  tlfea_elem *e = s->e;
  TL_CHECK(e->is_precomputed, "CalcDnDuPre/CalcDsDuPre must run before assembling");
  if (tangent) {
    TL_CHECK(s->analyzed, "AnalyzeHessianSparsity must run before assembling");
    TL_CHECK(e->is_mass, "CalcMassMatrix must run before assembling H = M/h + ...");
    TL_CHECK(s->prm.time_step > 0.0, "time_step must be positive (SetParameters)");
  }
  TL_CUDA(cudaSetDevice(e->device));
  TL_TRY(ensure_pattern(e));
  if (tangent)
    TL_CHECK(s->pattern_version_seen == e->pattern_version,
             "the element's sparsity pattern was rebuilt after AnalyzeHessianSparsity (SetLinearConstraintsCSR on a "
             "live solver): create a new solver");
  TL_TRY(refresh_constraints(s));
  const bool damp = (e->eta != 0.0 || e->lambda_d != 0.0);
  TL_CUDA(cudaEventRecord(s->ev[0], e->stream));
  if (ek_supported(s)) {
    if (pk_selected(s))
      TL_TRY(assemble_pk(s, tangent, damp));
    else
      TL_TRY(assemble_ek(s, tangent, damp));
    if (tangent && s->n_cons_slots > 0) {
      TL_LAUNCH(k_add_constraint_term, blocks_for(s->n_cons_slots, 256), 256, 0, e->stream, s->n_cons_slots,
                s->d_cons_slot, s->d_cons_val, s->prm.time_step * s->prm.time_step * s->prm.rho, s->d_H);
    }
    TL_CUDA(cudaGetLastError());
    TL_CUDA(cudaEventRecord(s->ev[1], e->stream));
    return TLFEA_OK;
  }
  if (tangent) TL_TRY(ensure_records(s));
  TL_TRY(launch_gp(e, damp ? s->d_v : nullptr, tangent ? s->d_gp : nullptr, e->d_fe, nullptr, nullptr));
  if (s->profile) TL_CUDA(cudaEventRecord(s->ev[2], e->stream));
  if (!tangent) {
    TL_LAUNCH(force_gather_kernel, blocks_for((int64_t)e->n_coef * 3, 128), 128, 0, e->stream, e->n_coef,
              e->pat.d_inc_ptr, e->pat.d_inc_ea, e->d_fe, e->d_f_int, 0);
  } else {
    const double h = s->prm.time_step;
    RowArgs a;
    a.n_coef = e->n_coef;
    a.off = e->pat.d_off;
    a.inc_ptr = e->pat.d_inc_ptr;
    a.inc_ea = e->pat.d_inc_ea;
    a.slot = e->pat.d_slot;
    a.mass = e->d_mass;
    a.rec = s->d_gp;
    a.gradN = e->d_gradN;
    a.detJ = e->d_detJ;
    a.fe = e->d_fe;
    a.qw = e->qw;
    a.h = h;
    a.inv_h = 1.0 / h;
    a.c1 = h * e->lambda + e->lambda_d;
    a.c2 = h * e->mu + e->eta;
    a.c3 = h * e->mu + e->eta;
    a.c4 = h;
    a.mu = e->mu;
    a.lambda_d = e->lambda_d;
    a.eta = e->eta;
    a.H = s->d_H;
    a.f_int = e->d_f_int;
    a.row_stride = ((e->pat.max_deg * 9 + 1) | 1);  // odd stride: warps start on different banks
    const size_t smem = (size_t)ROW_WARPS * a.row_stride * sizeof(double);
    TL_CHECK(smem <= 200 * 1024, "coefficient row too long for the shared-memory row buffer (max_deg=%d)",
             e->pat.max_deg);
    const int rows_blocks = (e->n_coef + ROW_WARPS - 1) / ROW_WARPS;
    const int grid = std::min(rows_blocks, sm_count(e->device) * 32);
    a.perm = nullptr;
    a.chunk = ROW_WARPS;
    {
      switch (e->type) {
        case TLFEA_TYPE_T10: TL_TRY((launch_row_t<10, 5>(s, a, grid, smem, damp))); break;
        case TLFEA_TYPE_3243: TL_TRY((launch_row_t<8, 12>(s, a, grid, smem, damp))); break;
        default: TL_TRY((launch_row_t<16, 48>(s, a, grid, smem, damp)));
      }
    }
    if (s->profile) TL_CUDA(cudaEventRecord(s->ev[3], e->stream));
    if (s->n_cons_slots > 0) {
      TL_LAUNCH(k_add_constraint_term, blocks_for(s->n_cons_slots, 256), 256, 0, e->stream, s->n_cons_slots,
                s->d_cons_slot, s->d_cons_val, h * h * s->prm.rho, s->d_H);
    }
  }
  TL_CUDA(cudaGetLastError());
  TL_CUDA(cudaEventRecord(s->ev[1], e->stream));
  return TLFEA_OK;
}

// reads the elapsed times of every pending event set (one host wait, on the newest set)
static int drain_timing(tlfea_newton *s) {
  if (s->ev_pending == 0) return TLFEA_OK;
  TL_CUDA(cudaEventSynchronize(s->ev_ring[s->ev_pending - 1][1]));
  for (int k = 0; k < s->ev_pending; k++) {
    cudaEvent_t *ev = s->ev_ring[k];
    float ms = 0;
    TL_CUDA(cudaEventElapsedTime(&ms, ev[0], ev[1]));
    s->ms_total += ms;
    s->calls++;
    if (s->profile && s->ev_tangent[k]) {
      float a = 0, b = 0, c = 0;
      cudaEventElapsedTime(&a, ev[0], ev[2]);
      cudaEventElapsedTime(&b, ev[2], ev[3]);
      cudaEventElapsedTime(&c, ev[3], ev[1]);
      s->k_ms[0] += a;
      s->k_ms[1] += b;
      s->k_ms[2] += c;
    }
  }
  s->ev_pending = 0;
  s->ev = s->ev_ring[0];
  return TLFEA_OK;
}

// the assemble just queued used the current event set: keep it pending, move on to the next set
static int collect_timing(tlfea_newton *s, bool tangent) {
  s->ev_tangent[s->ev_pending] = tangent;
  s->ev_pending++;
  if (s->ev_pending == tlfea_newton::EV_RING) return drain_timing(s);
  s->ev = s->ev_ring[s->ev_pending];
  return TLFEA_OK;
}

static int eval_constraints(tlfea_newton *s) {
  if (s->n_constraints > 0 && s->e->cons_mode != TL_CONS_NONE) return tlfea_elem_calc_constraint(s->e);
  return TLFEA_OK;
}

static int compute_residual(tlfea_newton *s) {
  tlfea_elem *e = s->e;
  const double h = s->prm.time_step;
  TL_CHECK(h > 0.0, "time_step must be positive (SetParameters)");
  TL_CHECK(e->is_mass, "CalcMassMatrix must run before the residual");
  TL_TRY(refresh_constraints(s));
  TL_TRY(eval_constraints(s));
  const bool lin = (s->n_constraints > 0 && e->cons_mode == TL_CONS_LINEAR_CSR);
  TL_LAUNCH(k_grad_L, blocks_for((int64_t)e->n_coef * 3, 256), 256, 0, e->stream, e->n_coef, e->pat.d_off,
            e->pat.d_cols, e->d_mass, s->d_v, s->d_v_prev, e->d_f_int, e->d_f_ext, 1.0 / h, h,
            lin ? e->d_jt_off : nullptr, e->d_jt_cols, e->d_jt_vals, s->d_lambda, e->d_constraint, s->prm.rho,
            own_lo_of(s), own_hi_of(s), s->d_g);
  if (s->n_constraints > 0 && e->cons_mode == TL_CONS_FIXED_NODES && e->n_fixed > 0) {
    TL_LAUNCH(k_grad_L_fixed, blocks_for(3 * e->n_fixed, 256), 256, 0, e->stream, e->n_fixed, e->d_fixed,
              s->d_lambda, e->d_constraint, h, s->prm.rho, own_lo_of(s), own_hi_of(s), s->d_g);
  }
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

// block-Jacobi PCG on H dv = r, dv starts at 0 (SyncedNewton.cu:1112-1114 zeroes delta_v before the solve)
static int pcg_solve(tlfea_newton *s, int *iters_out) {
  tlfea_elem *e = s->e;
  const int64_t n = 3 * (int64_t)e->n_coef;
  double *x = s->d_dv, *r = s->d_r, *z = s->d_pcg[0], *p = s->d_pcg[1], *Ap = s->d_pcg[2];
  const unsigned gb = blocks_for(n, 256), nb = blocks_for(e->n_coef, 256), sb = blocks_for((int64_t)e->n_coef * 32, 256);
  TL_CUDA(cudaMemsetAsync(x, 0, n * sizeof(double), e->stream));
  double *Dinv = s->d_pcg[3];
  TL_LAUNCH(k_block_jacobi_setup, nb, 256, 0, e->stream, e->n_coef, e->pat.d_off, e->pat.d_cols, s->d_H, Dinv);
  TL_LAUNCH(k_block_jacobi, nb, 256, 0, e->stream, e->n_coef, Dinv, r, z);
  TL_CUDA(cudaMemcpyAsync(p, z, n * sizeof(double), cudaMemcpyDeviceToDevice, e->stream));
  TL_TRY(dot(s, r, z, s->d_scal + 0));  // rz
  double r0 = 0;
  TL_TRY(norm2(s, r, n, &r0));
  int it = 0;
  double rn = r0;
  bool hit = (r0 == 0.0);
  if (r0 > 0.0) {
    for (it = 1; it <= s->pcg_max_iters; it++) {
      TL_LAUNCH(k_spmv, sb, 256, 0, e->stream, e->n_coef, e->pat.d_off, e->pat.d_cols, s->d_H, p, Ap);
      TL_TRY(dot(s, p, Ap, s->d_scal + 1));
      TL_LAUNCH(k_pcg_update_xr, gb, 256, 0, e->stream, n, s->d_scal, p, Ap, x, r);
      if (s->pcg_rel_tol > 0.0) {  // rel_tol = 0: a fixed iteration count, no host synchronisation inside the loop
        TL_TRY(norm2(s, r, n, &rn));
        if (rn <= s->pcg_rel_tol * r0) {
          hit = true;
          break;
        }
      }
      TL_LAUNCH(k_block_jacobi, nb, 256, 0, e->stream, e->n_coef, Dinv, r, z);
      TL_TRY(dot(s, r, z, s->d_scal + 2));
      TL_LAUNCH(k_pcg_update_p, gb, 256, 0, e->stream, n, s->d_scal, z, p);
      TL_LAUNCH(k_copy_scalar, 1, 1, 0, e->stream, s->d_scal + 0, s->d_scal + 2);
    }
    if (s->pcg_rel_tol <= 0.0) TL_TRY(norm2(s, r, n, &rn));  // what the fixed count reached
  }
  // the reference's cuDSS solve is exact; an iterative stand-in must say how far it got
  s->last_pcg_rel = (r0 > 0.0) ? rn / r0 : 0.0;
  if (s->pcg_rel_tol > 0.0 && !hit) s->last_pcg_converged = false;
  TL_CUDA(cudaGetLastError());
  *iters_out = std::min(it, s->pcg_max_iters);
  return TLFEA_OK;
}

// ---------------------------------------------------------------------------
// Pipelined host-buffer step: H2D of the coordinates in PIPE_K chunks on a copy stream, the element kernel in
// PIPE_K element chunks on the main stream (chunk j starts as soon as the last coordinate chunk it reads has
// landed), then the row reduce on its own stream while the main stream gathers f_int, evaluates the residual
// and copies f_int back to the host.
// ---------------------------------------------------------------------------
constexpr int PIPE_MAX = 16, PIPE_EV_TAIL = 16, PIPE_EV_RES = 17, PIPE_EV_ROWS = 18;
static const int PIPE_K = [] {
  const char *v = getenv("TLFEA_PIPE_K");  // coordinate / element chunks of the pipelined host-buffer step
  const int k = v ? atoi(v) : 8;
  return std::min(std::max(k, 1), PIPE_MAX);
}();

struct PipeBounds {
  int b[PIPE_MAX + 1];
};
__global__ void k_chunk_max_coef(int64_t n_entries, int S, const int *__restrict__ ecoef, PipeBounds eb, int nk,
                                 int *out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_entries) return;
  const int el = (int)(i / S);
  int j = 0;
  for (int k = 1; k < nk; k++) j += (el >= eb.b[k]);
  // one atomic per warp where the warp sits inside one chunk (nearly always)
  const unsigned act = __activemask();
  const int j0 = __shfl_sync(act, j, __ffs(act) - 1);
  if (__all_sync(act, j == j0)) {
    const int mx = __reduce_max_sync(act, ecoef[i]);
    if ((threadIdx.x & 31) == __ffs(act) - 1) atomicMax(out + j, mx);
  } else {
    atomicMax(out + j, ecoef[i]);
  }
}

static void pipe_bounds(int n, bool even, int *b) {
  for (int k = 0; k <= PIPE_K; k++) {
    int64_t v = (int64_t)n * k / PIPE_K;
    if (even && k < PIPE_K) v &= ~(int64_t)1;
    b[k] = (int)v;
  }
}

static int pipe_prepare(tlfea_newton *s) {
  tlfea_elem *e = s->e;
  if (!s->st_copy) TL_CUDA(cudaStreamCreateWithFlags(&s->st_copy, cudaStreamNonBlocking));
  if (!s->st_rows) TL_CUDA(cudaStreamCreateWithFlags(&s->st_rows, cudaStreamNonBlocking));
  for (auto &ev : s->ev_pipe)
    if (!ev) TL_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  if (s->pipe_need.empty()) {
    int eb[PIPE_MAX + 1], cb[PIPE_MAX + 1];
    pipe_bounds(e->n_elem, true, eb);
    pipe_bounds(e->n_coef, false, cb);
    int *d_max = nullptr;
    TL_TRY(dmalloc(&d_max, (size_t)PIPE_MAX));
    TL_CUDA(cudaMemsetAsync(d_max, 0, PIPE_K * sizeof(int), e->stream));
    const int64_t n_entries = (int64_t)e->n_elem * e->S;
    PipeBounds pb;
    for (int k = 0; k <= PIPE_K; k++) pb.b[k] = eb[k];
    TL_LAUNCH(k_chunk_max_coef, blocks_for(n_entries, 256), 256, 0, e->stream, n_entries, e->S, e->d_ecoef, pb, PIPE_K,
              d_max);
    int h_max[PIPE_MAX];
    TL_CUDA(cudaMemcpyAsync(h_max, d_max, PIPE_K * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    TL_CUDA(cudaStreamSynchronize(e->stream));
    dfree(d_max);
    s->pipe_need.assign(PIPE_K, PIPE_K - 1);
    for (int j = 0; j < PIPE_K; j++) {
      int k = 0;
      while (k < PIPE_K - 1 && h_max[j] >= cb[k + 1]) k++;
      s->pipe_need[j] = k;
    }
  }
  return TLFEA_OK;
}

// Host-buffer step of the single-pass patch assembly.  The launch order is chunked by coefficient-id range
// (patch_plan.cu): chunk c starts as soon as the c-th part of x, y, z is on the device, and after the last chunk that
// touches an id range its nodal forces are gathered and (if the caller registered its result buffer) copied back while
// the remaining chunks compute.  Same kernels, same order as the plain sequence: bit-identical H, f_int, ||g||.
struct PkPipe {
  tlfea_newton *s;
  int K;
  const int *ready;
};
static void pk_range(const tlfea_elem *e, int K, int r, int *i0, int *i1) {
  *i0 = (int)pk_chunk_bound(e->n_coef, K, r);
  *i1 = (int)pk_chunk_bound(e->n_coef, K, r + 1);
}
static int pk_before_chunk(void *ctx, int c) {
  PkPipe *pp = static_cast<PkPipe *>(ctx);
  TL_CUDA(cudaStreamWaitEvent(pp->s->e->stream, pp->s->ev_pipe[c], 0));
  return TLFEA_OK;
}
// nodal forces of the id sub-ranges that chunk c completes: ordered gather, then (if the caller registered its result
// buffer) one event and the copies back, contiguous sub-ranges merged
static int pk_after_chunk(void *ctx, int c) {
  PkPipe *pp = static_cast<PkPipe *>(ctx);
  tlfea_newton *s = pp->s;
  tlfea_elem *e = s->e;
  const int NQ = pp->K * PK_GSUB;
  std::vector<std::pair<int, int>> runs;  // [i0, i1) of coefficient ids
  for (int q = 0; q < NQ; q++) {
    if (!(pp->ready[q] == c || (c == 0 && pp->ready[q] < 0))) continue;
    const int i0 = (int)pk_sub_bound(e->n_coef, pp->K, q), i1 = (int)pk_sub_bound(e->n_coef, pp->K, q + 1);
    if (i1 <= i0) continue;
    if (!runs.empty() && runs.back().second == i0) runs.back().second = i1;
    else runs.push_back({i0, i1});
  }
  for (auto &r : runs)
    TL_LAUNCH(force_gather_kernel, blocks_for((int64_t)(r.second - r.first) * 3, 128), 128, 0, e->stream,
              r.second - r.first, e->pat.d_inc_ptr + r.first, e->pat.d_inc_ea, e->d_fe, e->d_f_int + 3 * (int64_t)r.first, 0);
  if (s->h_f_out && !runs.empty()) {
    TL_CUDA(cudaEventRecord(s->ev_g[c], e->stream));
    TL_CUDA(cudaStreamWaitEvent(s->st_copy, s->ev_g[c], 0));
    for (auto &r : runs)
      TL_CUDA(cudaMemcpyAsync(s->h_f_out + 3 * (int64_t)r.first, e->d_f_int + 3 * (int64_t)r.first,
                              (size_t)(r.second - r.first) * 3 * sizeof(double), cudaMemcpyDeviceToHost, s->st_copy));
  }
  return TLFEA_OK;
}

static int assemble_host_begin_pk(tlfea_newton *s, const double *hx, const double *hy, const double *hz, bool damp) {
  tlfea_elem *e = s->e;
  int K = 1;
  const int *ready = nullptr;
  TL_TRY(patch_chunks(s, &K, &ready));
  if (!s->st_copy) TL_CUDA(cudaStreamCreateWithFlags(&s->st_copy, cudaStreamNonBlocking));
  for (int k = 0; k < K; k++) {
    if (!s->ev_pipe[k]) TL_CUDA(cudaEventCreateWithFlags(&s->ev_pipe[k], cudaEventDisableTiming));
    if (!s->ev_g[k]) TL_CUDA(cudaEventCreateWithFlags(&s->ev_g[k], cudaEventDisableTiming));
  }
  for (int k : {PIPE_EV_TAIL, PIPE_EV_RES})
    if (!s->ev_pipe[k]) TL_CUDA(cudaEventCreateWithFlags(&s->ev_pipe[k], cudaEventDisableTiming));
  cudaStream_t m = e->stream, c = s->st_copy;
  TL_CUDA(cudaEventRecord(s->ev_pipe[PIPE_EV_TAIL], m));  // the copies may not overtake queued readers of x, y, z
  TL_CUDA(cudaStreamWaitEvent(c, s->ev_pipe[PIPE_EV_TAIL], 0));
  for (int r = 0; r < K; r++) {
    int i0, i1;
    pk_range(e, K, r, &i0, &i1);
    const size_t bytes = (size_t)(i1 - i0) * sizeof(double);
    if (bytes) {
      TL_CUDA(cudaMemcpyAsync(e->d_x + i0, hx + i0, bytes, cudaMemcpyHostToDevice, c));
      TL_CUDA(cudaMemcpyAsync(e->d_y + i0, hy + i0, bytes, cudaMemcpyHostToDevice, c));
      TL_CUDA(cudaMemcpyAsync(e->d_z + i0, hz + i0, bytes, cudaMemcpyHostToDevice, c));
    }
    TL_CUDA(cudaEventRecord(s->ev_pipe[r], c));
  }
  PkPipe ctx{s, K, ready};
  PatchHooks hooks{&ctx, pk_before_chunk, pk_after_chunk};
  TL_CUDA(cudaEventRecord(s->ev[0], m));
  TL_TRY(patch_assemble(s, true, damp, &hooks));
  if (s->n_cons_slots > 0)
    TL_LAUNCH(k_add_constraint_term, blocks_for(s->n_cons_slots, 256), 256, 0, m, s->n_cons_slots, s->d_cons_slot,
              s->d_cons_val, s->prm.time_step * s->prm.time_step * s->prm.rho, s->d_H);
  TL_CUDA(cudaEventRecord(s->ev[1], m));
  TL_CUDA(cudaGetLastError());
  s->pipe_open = true;
  s->pipe_overlapped = false;
  s->pipe_pk = true;
  s->pipe_f_copied = (s->h_f_out != nullptr);
  return TLFEA_OK;
}

static int assemble_host_begin(tlfea_newton *s, const double *hx, const double *hy, const double *hz) {
  tlfea_elem *e = s->e;
  TL_CHECK(!s->pipe_open, "assemble_host_begin called twice without assemble_host_end");
  TL_CHECK(e->is_precomputed, "CalcDnDuPre/CalcDsDuPre must run before assembling");
  TL_CHECK(s->analyzed, "AnalyzeHessianSparsity must run before assembling");
  TL_CHECK(e->is_mass, "CalcMassMatrix must run before assembling H = M/h + ...");
  TL_CHECK(s->prm.time_step > 0.0, "time_step must be positive (SetParameters)");
  TL_CUDA(cudaSetDevice(e->device));
  TL_TRY(ensure_pattern(e));
  TL_CHECK(s->pattern_version_seen == e->pattern_version,
           "the element's sparsity pattern was rebuilt after AnalyzeHessianSparsity: create a new solver");
  TL_TRY(refresh_constraints(s));
  const bool damp = (e->eta != 0.0 || e->lambda_d != 0.0);
  s->pipe_pk = false;
  if (pk_selected(s)) return assemble_host_begin_pk(s, hx, hy, hz, damp);
  bool overlap = ek_supported(s);
  EkPlan pl;
  if (overlap) {
    TL_TRY(pl.init(s, true, damp));
    overlap = (s->chunk_ptr.size() == 2);  // the interleaved E/R chunk schedule (tuning aid) keeps the plain order
  }
  if (!overlap) {  // generic row-gather path: same result, no overlap
    TL_TRY(tlfea_elem_update_positions(e, hx, hy, hz, e->n_coef));
    TL_TRY(do_assemble(s, true));
    s->pipe_open = true;
    s->pipe_overlapped = false;
    return TLFEA_OK;
  }
  TL_TRY(pipe_prepare(s));
  cudaStream_t m = e->stream, c = s->st_copy;
  int eb[PIPE_MAX + 1], cb[PIPE_MAX + 1];
  pipe_bounds(e->n_elem, true, eb);
  pipe_bounds(e->n_coef, false, cb);
  // the copies may not overtake work already queued on the main stream that still reads the old coordinates
  TL_CUDA(cudaEventRecord(s->ev_pipe[PIPE_EV_TAIL], m));
  TL_CUDA(cudaStreamWaitEvent(c, s->ev_pipe[PIPE_EV_TAIL], 0));
  for (int k = 0; k < PIPE_K; k++) {
    const size_t off = (size_t)cb[k], bytes = (size_t)(cb[k + 1] - cb[k]) * sizeof(double);
    if (bytes) {
      TL_CUDA(cudaMemcpyAsync(e->d_x + off, hx + off, bytes, cudaMemcpyHostToDevice, c));
      TL_CUDA(cudaMemcpyAsync(e->d_y + off, hy + off, bytes, cudaMemcpyHostToDevice, c));
      TL_CUDA(cudaMemcpyAsync(e->d_z + off, hz + off, bytes, cudaMemcpyHostToDevice, c));
    }
    TL_CUDA(cudaEventRecord(s->ev_pipe[k], c));
  }
  int waited = -1;
  for (int j = 0; j < PIPE_K; j++) {
    while (waited < s->pipe_need[j]) TL_CUDA(cudaStreamWaitEvent(m, s->ev_pipe[++waited], 0));
    if (eb[j + 1] > eb[j]) TL_TRY(pl.elems(eb[j], eb[j + 1] - eb[j]));
  }
  while (waited < PIPE_K - 1) TL_CUDA(cudaStreamWaitEvent(m, s->ev_pipe[++waited], 0));
  // the nodal forces (same ascending-element order as the row reduce); the H rows are reduced by _end, AFTER the
  // small residual kernels have gone through: the row-reduce kernel is persistent and leaves no room on the SMs
  TL_LAUNCH(force_gather_kernel, blocks_for((int64_t)e->n_coef * 3, 128), 128, 0, m, e->n_coef, e->pat.d_inc_ptr,
            e->pat.d_inc_ea, e->d_fe, e->d_f_int, 0);
  TL_CUDA(cudaGetLastError());
  s->pipe_open = true;
  s->pipe_overlapped = true;
  return TLFEA_OK;
}

static int assemble_host_end(tlfea_newton *s, double *h_f_int, double *norm_g) {
  tlfea_elem *e = s->e;
  TL_CHECK(s->pipe_open, "assemble_host_end without assemble_host_begin");
  TL_CUDA(cudaSetDevice(e->device));
  s->pipe_open = false;
  cudaStream_t m = e->stream, rw = s->st_rows;
  const int64_t n = 3 * (int64_t)e->n_coef;
  // residual and its norm: kernels only, nothing waits yet
  if (norm_g) {
    TL_TRY(compute_residual(s));
    double *partial = s->d_scal + 8;
    TL_LAUNCH(k_dot_partial, DOT_BLOCKS, 256, 0, m, n, s->d_g, s->d_g, partial);
    TL_LAUNCH(k_dot_final, 1, 1024, 0, m, DOT_BLOCKS, partial, s->d_scal + 4);
  }
  if (s->pipe_overlapped) {  // H rows on their own stream, behind everything queued on the main stream so far
    EkPlan pl;
    TL_TRY(pl.init(s, true, (e->eta != 0.0 || e->lambda_d != 0.0)));
    TL_CUDA(cudaEventRecord(s->ev_pipe[PIPE_EV_RES], m));
    TL_CUDA(cudaStreamWaitEvent(rw, s->ev_pipe[PIPE_EV_RES], 0));
    TL_CUDA(cudaMemsetAsync(s->d_tickets, 0, s->chunk_ptr.size() * sizeof(int), rw));
    TL_TRY(pl.rows(0, rw, false));
    if (s->n_cons_slots > 0) {
      TL_LAUNCH(k_add_constraint_term, blocks_for(s->n_cons_slots, 256), 256, 0, rw, s->n_cons_slots, s->d_cons_slot,
                s->d_cons_val, s->prm.time_step * s->prm.time_step * s->prm.rho, s->d_H);
    }
    TL_CUDA(cudaEventRecord(s->ev_pipe[PIPE_EV_ROWS], rw));
  }
  // the copies back to the host run on the copy engines while the rows are reduced
  if (norm_g) TL_CUDA(cudaMemcpyAsync(s->h_scal, s->d_scal + 4, sizeof(double), cudaMemcpyDeviceToHost, m));
  const bool f_done = s->pipe_pk && s->pipe_f_copied && h_f_int == s->h_f_out;  // ranges went back as they finished
  if (h_f_int && !f_done)
    TL_CUDA(cudaMemcpyAsync(h_f_int, e->d_f_int, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, m));
  if (h_f_int && f_done) {
    // partitioned mesh: the halo sum (tlfea_dist_halo_sum, queued on the main stream between begin and end) completed the
    // entries of the shared id ranges after their first copy went out -- send those ranges again, behind the early copies
    std::vector<std::pair<int64_t, int64_t>> shared;
    dist_halo_ranges(s, shared);
    if (!shared.empty()) {
      TL_CUDA(cudaStreamSynchronize(s->st_copy));
      for (auto &r : shared)
        TL_CUDA(cudaMemcpyAsync(h_f_int + 3 * r.first, e->d_f_int + 3 * r.first,
                                (size_t)(r.second - r.first) * 3 * sizeof(double), cudaMemcpyDeviceToHost, m));
    }
  }
  TL_CUDA(cudaStreamSynchronize(m));
  if (s->pipe_pk && s->pipe_f_copied) TL_CUDA(cudaStreamSynchronize(s->st_copy));
  s->pipe_pk = false;
  if (norm_g) *norm_g = std::sqrt(s->h_scal[0]);
  if (s->pipe_overlapped) {  // H is complete, and later main-stream work is ordered after it
    TL_CUDA(cudaEventSynchronize(s->ev_pipe[PIPE_EV_ROWS]));
    TL_CUDA(cudaStreamWaitEvent(m, s->ev_pipe[PIPE_EV_ROWS], 0));
  }
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

}  // namespace tlfea

using namespace tlfea;

extern "C" {

int tlfea_newton_create(tlfea_elem *e, int n_constraints, tlfea_newton **out) {
  TL_CHECK(e && out, "null input");
  TL_CHECK(n_constraints >= 0, "negative n_constraints");
  TL_CUDA(cudaSetDevice(e->device));
  tlfea_newton *s = new tlfea_newton();
  s->e = e;
  s->device = e->device;
  s->n_constraints = n_constraints;
  const size_t nd = 3 * (size_t)e->n_coef;
  TL_TRY(dmalloc(&s->d_v, nd));
  TL_TRY(dmalloc(&s->d_v_prev, nd));
  TL_TRY(dmalloc(&s->d_lambda, (size_t)std::max(n_constraints, 1)));
  TL_TRY(dmalloc(&s->d_g, nd));
  TL_TRY(dmalloc(&s->d_dv, nd));
  TL_TRY(dmalloc(&s->d_r, nd));
  TL_TRY(dmalloc(&s->d_xp, (size_t)e->n_coef));
  TL_TRY(dmalloc(&s->d_yp, (size_t)e->n_coef));
  TL_TRY(dmalloc(&s->d_zp, (size_t)e->n_coef));
  TL_TRY(dmalloc(&s->d_scal, 8 + DOT_BLOCKS));
  TL_CUDA(cudaMallocHost((void **)&s->h_scal, 8 * sizeof(double)));
  for (auto &set : s->ev_ring)
    for (auto &ev : set) TL_CUDA(cudaEventCreate(&ev));
  TL_CUDA(cudaMemsetAsync(s->d_v, 0, nd * sizeof(double), e->stream));
  TL_CUDA(cudaMemsetAsync(s->d_v_prev, 0, nd * sizeof(double), e->stream));
  TL_CUDA(cudaMemsetAsync(s->d_lambda, 0, std::max(n_constraints, 1) * sizeof(double), e->stream));
  TL_CUDA(cudaMemsetAsync(s->d_g, 0, nd * sizeof(double), e->stream));
  TL_CUDA(cudaStreamSynchronize(e->stream));
  *out = s;
  return TLFEA_OK;
}

int tlfea_newton_destroy(tlfea_newton *s) {
  if (!s) return TLFEA_OK;
  cudaSetDevice(s->device);
  if (elem_is_live(s->e))
    cudaStreamSynchronize(s->e->stream);
  else
    cudaDeviceSynchronize();  // the element handle went first: its stream is gone
  patch_release(s);
  dist_release(s);
  dfree(s->d_H);
  dfree(s->d_gp);
  dfree(s->d_stage);
  if (s->d_rowdesc) cudaFree(s->d_rowdesc);
  s->d_rowdesc = nullptr;
  dfree(s->d_tickets);
  dfree(s->d_v);
  dfree(s->d_v_prev);
  dfree(s->d_lambda);
  dfree(s->d_g);
  dfree(s->d_dv);
  dfree(s->d_r);
  dfree(s->d_xp);
  dfree(s->d_yp);
  dfree(s->d_zp);
  dfree(s->d_cons_slot);
  dfree(s->d_cons_val);
  for (auto &p : s->d_pcg) dfree(p);
  if (s->st_copy) cudaStreamDestroy(s->st_copy);
  if (s->st_rows) cudaStreamDestroy(s->st_rows);
  for (auto &ev : s->ev_pipe)
    if (ev) cudaEventDestroy(ev);
  for (auto &ev : s->ev_g)
    if (ev) cudaEventDestroy(ev);
  dfree(s->d_scal);
  if (s->h_scal) cudaFreeHost(s->h_scal);
  for (auto &set : s->ev_ring)
    for (auto &ev : set)
      if (ev) cudaEventDestroy(ev);
  delete s;
  return TLFEA_OK;
}

int tlfea_newton_set_params(tlfea_newton *s, const tlfea_newton_params *p) {
  TL_CHECK(s && p, "null input");
  s->prm = *p;
  return TLFEA_OK;
}

int tlfea_newton_setup(tlfea_newton *s) {
  TL_CHECK(s, "null handle");
  tlfea_elem *e = s->e;
  TL_CUDA(cudaSetDevice(e->device));
  const size_t nd = 3 * (size_t)e->n_coef;
  TL_CUDA(cudaMemsetAsync(s->d_v, 0, nd * sizeof(double), e->stream));
  TL_CUDA(cudaMemsetAsync(s->d_v_prev, 0, nd * sizeof(double), e->stream));
  TL_CUDA(cudaMemsetAsync(s->d_lambda, 0, std::max(s->n_constraints, 1) * sizeof(double), e->stream));
  TL_CUDA(cudaMemsetAsync(s->d_g, 0, nd * sizeof(double), e->stream));
  return TLFEA_OK;
}

int tlfea_newton_analyze_sparsity(tlfea_newton *s) {
  TL_CHECK(s, "null handle");
  if (s->analyzed) return TLFEA_OK;
  tlfea_elem *e = s->e;
  TL_CUDA(cudaSetDevice(e->device));
  TL_TRY(ensure_pattern(e));
  s->pattern_version_seen = e->pattern_version;
  s->nnz = 9 * e->pat.nnz;
  TL_TRY(dmalloc(&s->d_H, (size_t)s->nnz));
  s->rec_doubles = 0;  // Gauss-point records (row-gather path only) are allocated on first use
  if (s->n_constraints > 0) {
    TL_CHECK(e->cons_mode != TL_CONS_NONE, "solver has constraints but the element has none set up");
    TL_CHECK(s->n_constraints == e->n_constraint, "n_constraints (%d) != element constraints (%d)", s->n_constraints,
             e->n_constraint);
    if (e->cons_mode == TL_CONS_FIXED_NODES) {
      TL_TRY(fixed_cons_slots(s));
    } else {
      // general linear constraints: J^T J on the host (small systems), merged per slot in row order
      std::vector<int> off(e->n_coef + 1), cols(e->pat.nnz);
      TL_CUDA(cudaMemcpy(off.data(), e->pat.d_off, off.size() * sizeof(int), cudaMemcpyDeviceToHost));
      TL_CUDA(cudaMemcpy(cols.data(), e->pat.d_cols, cols.size() * sizeof(int), cudaMemcpyDeviceToHost));
      std::vector<int> joff(e->n_constraint + 1), jcols(e->j_nnz);
      std::vector<double> jvals(e->j_nnz);
      TL_CUDA(cudaMemcpy(joff.data(), e->d_j_off, joff.size() * sizeof(int), cudaMemcpyDeviceToHost));
      TL_CUDA(cudaMemcpy(jcols.data(), e->d_j_cols, jcols.size() * sizeof(int), cudaMemcpyDeviceToHost));
      TL_CUDA(cudaMemcpy(jvals.data(), e->d_j_vals, jvals.size() * sizeof(double), cudaMemcpyDeviceToHost));
      std::vector<std::pair<int64_t, double>> entries;
      for (int c = 0; c < e->n_constraint; c++)
        for (int a = joff[c]; a < joff[c + 1]; a++)
          for (int b = joff[c]; b < joff[c + 1]; b++) {
            const int di = jcols[a], dj = jcols[b], ci = di / 3, cj = dj / 3;
            const int rs = off[ci], deg = off[ci + 1] - rs;
            auto it = std::lower_bound(cols.begin() + rs, cols.begin() + rs + deg, cj);
            // every (ci, cj) of a constraint row is in the pattern (ensure_pattern adds the J^T J couplings);
            // an entry outside it would be dropped as the reference does (SyncedNewton.cu:336)
            if (it == cols.begin() + rs + deg || *it != cj) continue;
            const int pos = (int)(it - (cols.begin() + rs));
            entries.emplace_back(9 * (int64_t)rs + (int64_t)(di % 3) * 3 * deg + 3 * pos + dj % 3,
                                 jvals[a] * jvals[b]);
          }
      std::stable_sort(entries.begin(), entries.end(),
                       [](const std::pair<int64_t, double> &x, const std::pair<int64_t, double> &y) {
                         return x.first < y.first;
                       });
      std::vector<int64_t> us;
      std::vector<double> uv;
      for (auto &en : entries) {
        if (!us.empty() && us.back() == en.first)
          uv.back() += en.second;
        else {
          us.push_back(en.first);
          uv.push_back(en.second);
        }
      }
      s->n_cons_slots = (int64_t)us.size();
      TL_TRY(dmalloc(&s->d_cons_slot, us.size()));
      TL_TRY(dmalloc(&s->d_cons_val, us.size()));
      TL_CUDA(cudaMemcpy(s->d_cons_slot, us.data(), us.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
      TL_CUDA(cudaMemcpy(s->d_cons_val, uv.data(), us.size() * sizeof(double), cudaMemcpyHostToDevice));
      TL_CUDA(cudaDeviceSynchronize());  // pageable copies may return before their DMA lands; the handle's stream is non-blocking
    }
  }
  TL_CUDA(cudaGetLastError());
  TL_CUDA(cudaStreamSynchronize(e->stream));
  s->analyzed = true;
  return TLFEA_OK;
}

int tlfea_newton_set_fixed_pattern(tlfea_newton *s, int fixed) {
  TL_CHECK(s, "null handle");
  s->fixed_pattern = fixed != 0;
  return TLFEA_OK;
}

int tlfea_newton_assemble(tlfea_newton *s) {
  TL_CHECK(s, "null handle");
  TL_CHECK(s->analyzed, "AnalyzeHessianSparsity must run before assembling");
  TL_TRY(do_assemble(s, true));
  return collect_timing(s, true);
}

int tlfea_newton_assemble_force(tlfea_newton *s) {
  TL_CHECK(s, "null handle");
  TL_TRY(do_assemble(s, false));
  return collect_timing(s, false);
}

int tlfea_newton_assemble_host_begin(tlfea_newton *s, const double *x, const double *y, const double *z, int n) {
  TL_CHECK(s && x && y && z, "null input");
  TL_CHECK(n == s->e->n_coef, "Position vector size mismatch.");
  return assemble_host_begin(s, x, y, z);
}

int tlfea_newton_set_host_result_buffer(tlfea_newton *s, double *f_int_host, int n) {
  TL_CHECK(s, "null handle");
  TL_CHECK(!f_int_host || n == 3 * s->e->n_coef, "size mismatch");
  TL_CHECK(!s->pipe_open, "a host-buffer step is in flight");
  s->h_f_out = f_int_host;
  return TLFEA_OK;
}

int tlfea_newton_assemble_host_end(tlfea_newton *s, double *f_int, int n, double *norm_g) {
  TL_CHECK(s, "null handle");
  TL_CHECK(!f_int || n == 3 * s->e->n_coef, "size mismatch");
  return assemble_host_end(s, f_int, norm_g);
}

int tlfea_newton_residual(tlfea_newton *s, double *norm_g) {
  TL_CHECK(s, "null handle");
  TL_CUDA(cudaSetDevice(s->e->device));
  TL_TRY(compute_residual(s));
  if (norm_g) TL_TRY(norm2(s, s->d_g, 3 * (int64_t)s->e->n_coef, norm_g));
  return TLFEA_OK;
}

int tlfea_newton_set_pcg(tlfea_newton *s, int max_iters, double rel_tol) {
  TL_CHECK(s && max_iters > 0 && rel_tol >= 0.0, "bad PCG settings");
  s->pcg_max_iters = max_iters;
  s->pcg_rel_tol = rel_tol;
  return TLFEA_OK;
}

int tlfea_newton_solve_step(tlfea_newton *s) {
  TL_CHECK(s, "null handle");
  tlfea_elem *e = s->e;
  TL_CUDA(cudaSetDevice(e->device));
  if (!s->analyzed) TL_TRY(tlfea_newton_analyze_sparsity(s));
  const int64_t n = 3 * (int64_t)e->n_coef;
  const size_t cb = (size_t)e->n_coef * sizeof(double);
  for (int i = 0; i < 4; i++)  // z, p, Ap, and the inverse diagonal blocks (9 per coefficient = 3n)
    if (!s->d_pcg[i]) TL_TRY(dmalloc(&s->d_pcg[i], (size_t)n * (i == 3 ? 3 : 1)));
  const double h = s->prm.time_step;
  // update_pos_prev  (SyncedNewton.cu:1035)
  TL_CUDA(cudaMemcpyAsync(s->d_xp, e->d_x, cb, cudaMemcpyDeviceToDevice, e->stream));
  TL_CUDA(cudaMemcpyAsync(s->d_yp, e->d_y, cb, cudaMemcpyDeviceToDevice, e->stream));
  TL_CUDA(cudaMemcpyAsync(s->d_zp, e->d_z, cb, cudaMemcpyDeviceToDevice, e->stream));
  s->last_newton = 0;
  s->last_pcg = 0;
  s->last_norm_c = 0;
  s->last_pcg_rel = 0;
  s->last_pcg_converged = true;
  for (int outer = 0; outer < s->prm.max_outer; outer++) {
    double norm_g0 = -1.0;
    for (int it = 0; it < s->prm.max_inner; it++) {
      // force + H in one fused pass; the reference evaluates the force first and skips H on
      // convergence, we assemble both and discard H in that case (one extra pass at most).
      TL_TRY(do_assemble(s, true));
      TL_TRY(collect_timing(s, true));
      TL_TRY(compute_residual(s));
      double norm_g = 0;
      TL_TRY(norm2(s, s->d_g, n, &norm_g));
      s->last_norm_g = norm_g;
      if (norm_g0 < 0.0) norm_g0 = norm_g;
      if (norm_g < s->prm.inner_atol ||
          (s->prm.inner_rtol > 0.0 && norm_g0 > 0.0 && norm_g <= s->prm.inner_rtol * norm_g0))
        break;
      TL_LAUNCH(k_neg_copy, blocks_for(n, 256), 256, 0, e->stream, n, s->d_g, s->d_r);
      int pit = 0;
      TL_TRY(pcg_solve(s, &pit));
      s->last_pcg += pit;
      s->last_newton++;
      TL_LAUNCH(k_axpy, blocks_for(n, 256), 256, 0, e->stream, n, 1.0, s->d_dv, s->d_v);
      TL_LAUNCH(k_update_pos, blocks_for(e->n_coef, 256), 256, 0, e->stream, e->n_coef, s->d_xp, s->d_yp, s->d_zp,
                s->d_v, h, e->d_x, e->d_y, e->d_z);
    }
    TL_CUDA(cudaMemcpyAsync(s->d_v_prev, s->d_v, n * sizeof(double), cudaMemcpyDeviceToDevice, e->stream));
    if (s->n_constraints > 0) {
      TL_TRY(eval_constraints(s));
      TL_LAUNCH(k_dual_update, blocks_for(s->n_constraints, 256), 256, 0, e->stream, s->n_constraints,
                e->d_constraint, s->prm.rho, s->d_lambda);
      double nc = 0;
      TL_TRY(norm2(s, e->d_constraint, s->n_constraints, &nc));
      s->last_norm_c = nc;
      if (nc < s->prm.outer_tol) break;
    }
  }
  TL_CUDA(cudaGetLastError());
  TL_CUDA(cudaStreamSynchronize(e->stream));
  return TLFEA_OK;
}

int tlfea_newton_last_stats(tlfea_newton *s, int *newton_iters, int *pcg_iters, double *norm_g, double *norm_c) {
  TL_CHECK(s, "null handle");
  if (newton_iters) *newton_iters = s->last_newton;
  if (pcg_iters) *pcg_iters = s->last_pcg;
  if (norm_g) *norm_g = s->last_norm_g;
  if (norm_c) *norm_c = s->last_norm_c;
  return TLFEA_OK;
}

int tlfea_newton_last_linear_solve(tlfea_newton *s, double *rel_residual, int *converged) {
  TL_CHECK(s, "null handle");
  if (rel_residual) *rel_residual = s->last_pcg_rel;
  if (converged) *converged = s->last_pcg_converged ? 1 : 0;
  return TLFEA_OK;
}

int64_t tlfea_newton_nnz(const tlfea_newton *s) { return s ? s->nnz : -1; }
int tlfea_newton_n_dofs(const tlfea_newton *s) { return s ? 3 * s->e->n_coef : -1; }

int tlfea_newton_get_hessian_csr(tlfea_newton *s, int64_t *row_offsets, int *col_indices, double *values) {
  TL_CHECK(s && s->analyzed, "AnalyzeHessianSparsity must run first");
  tlfea_elem *e = s->e;
  TL_CUDA(cudaSetDevice(e->device));
  TL_CUDA(cudaStreamSynchronize(e->stream));
  if (row_offsets || col_indices) {
    // expansion of the coefficient pattern: row 3i+d gets 3*deg(i) columns {3j,3j+1,3j+2}
    // (build_dof_row_nnz_from_coef_csr / fill_dof_cols_from_coef_csr, SyncedNewton.cu:162-205)
    std::vector<int> off(e->n_coef + 1), cols(e->pat.nnz);
    TL_CUDA(cudaMemcpy(off.data(), e->pat.d_off, off.size() * sizeof(int), cudaMemcpyDeviceToHost));
    TL_CUDA(cudaMemcpy(cols.data(), e->pat.d_cols, cols.size() * sizeof(int), cudaMemcpyDeviceToHost));
    if (row_offsets) {
      row_offsets[0] = 0;
      for (int i = 0; i < e->n_coef; i++) {
        const int64_t deg = off[i + 1] - off[i];
        for (int d = 0; d < 3; d++) row_offsets[3 * (size_t)i + d + 1] = row_offsets[3 * (size_t)i + d] + 3 * deg;
      }
    }
    if (col_indices) {
      for (int i = 0; i < e->n_coef; i++) {
        const int rs = off[i], deg = off[i + 1] - rs;
        for (int d = 0; d < 3; d++) {
          int64_t o = 9 * (int64_t)rs + (int64_t)d * 3 * deg;
          for (int k = 0; k < deg; k++) {
            const int b = 3 * cols[rs + k];
            col_indices[o++] = b;
            col_indices[o++] = b + 1;
            col_indices[o++] = b + 2;
          }
        }
      }
    }
  }
  if (values) TL_CUDA(cudaMemcpy(values, s->d_H, (size_t)s->nnz * sizeof(double), cudaMemcpyDeviceToHost));
  return TLFEA_OK;
}

static int vec_d2h(tlfea_newton *s, double *out, const double *src, int n) {
  TL_CHECK(s && out && n == 3 * s->e->n_coef, "size mismatch");
  TL_CUDA(cudaSetDevice(s->e->device));
  TL_CUDA(cudaMemcpyAsync(out, src, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s->e->stream));
  TL_CUDA(cudaStreamSynchronize(s->e->stream));
  return TLFEA_OK;
}
static int vec_h2d(tlfea_newton *s, double *dst, const double *in, int n) {
  TL_CHECK(s && in && n == 3 * s->e->n_coef, "size mismatch");
  TL_CUDA(cudaSetDevice(s->e->device));
  TL_CUDA(cudaMemcpyAsync(dst, in, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s->e->stream));
  TL_CUDA(cudaStreamSynchronize(s->e->stream));
  return TLFEA_OK;
}
int tlfea_newton_get_g(tlfea_newton *s, double *out, int n) { return vec_d2h(s, out, s ? s->d_g : nullptr, n); }
int tlfea_newton_get_v(tlfea_newton *s, double *out, int n) { return vec_d2h(s, out, s ? s->d_v : nullptr, n); }
int tlfea_newton_set_v(tlfea_newton *s, const double *v, int n) { return vec_h2d(s, s ? s->d_v : nullptr, v, n); }
int tlfea_newton_set_v_prev(tlfea_newton *s, const double *v, int n) {
  return vec_h2d(s, s ? s->d_v_prev : nullptr, v, n);
}
double *tlfea_newton_devptr_v(tlfea_newton *s) { return s ? s->d_v : nullptr; }
double *tlfea_newton_devptr_hessian_values(tlfea_newton *s) { return s ? s->d_H : nullptr; }
double *tlfea_newton_devptr_g(tlfea_newton *s) { return s ? s->d_g : nullptr; }

int tlfea_newton_spmv(tlfea_newton *s, const double *d_x, double *d_y) {
  TL_CHECK(s && s->analyzed && d_x && d_y, "bad input");
  tlfea_elem *e = s->e;
  TL_CUDA(cudaSetDevice(e->device));
  TL_LAUNCH(k_spmv, blocks_for((int64_t)e->n_coef * 32, 256), 256, 0, e->stream, e->n_coef, e->pat.d_off,
            e->pat.d_cols, s->d_H, d_x, d_y);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

// ---------------------------------------------------------------------------------------------
// Distributed Newton/Krylov building blocks.  Every rank holds a sub-assembled H and consistent copies of the
// vectors; the host side (tlfea/dist_newton.py) completes interface entries with a halo sum and the partial
// scalars with an all-reduce between these calls.  Scalars live in device memory (scal[0] = r.z, [1] = p.Ap,
// [2] = new r.z, [3] = r.r or g.g): alpha and beta are formed on the device, so an iteration needs no host sync.
// ---------------------------------------------------------------------------------------------
int tlfea_newton_set_owned_range(tlfea_newton *s, int64_t dof_lo, int64_t dof_hi) {
  TL_CHECK(s, "null handle");
  s->own_lo = dof_lo;
  s->own_hi = dof_hi;
  return TLFEA_OK;
}

int tlfea_newton_krylov_op(tlfea_newton *s, int op) {
  TL_CHECK(s && s->analyzed, "AnalyzeHessianSparsity must run first");
  tlfea_elem *e = s->e;
  TL_CUDA(cudaSetDevice(e->device));
  const int64_t n = 3 * (int64_t)e->n_coef;
  for (int i = 0; i < 5; i++)  // z, p, Ap, inverse diagonal blocks, raw diagonal blocks
    if (!s->d_pcg[i]) TL_TRY(dmalloc(&s->d_pcg[i], (size_t)n * (i >= 3 ? 3 : 1)));
  double *x = s->d_dv, *r = s->d_r, *z = s->d_pcg[0], *p = s->d_pcg[1], *Ap = s->d_pcg[2], *Dinv = s->d_pcg[3],
         *diag = s->d_pcg[4];
  const unsigned gb = blocks_for(n, 256), nb = blocks_for(e->n_coef, 256);
  const double h = s->prm.time_step;
  switch (op) {
    case TLFEA_KOP_SAVE_POSITIONS: {  // update_pos_prev (SyncedNewton.cu:1035)
      const size_t cb = (size_t)e->n_coef * sizeof(double);
      TL_CUDA(cudaMemcpyAsync(s->d_xp, e->d_x, cb, cudaMemcpyDeviceToDevice, e->stream));
      TL_CUDA(cudaMemcpyAsync(s->d_yp, e->d_y, cb, cudaMemcpyDeviceToDevice, e->stream));
      TL_CUDA(cudaMemcpyAsync(s->d_zp, e->d_z, cb, cudaMemcpyDeviceToDevice, e->stream));
      break;
    }
    case TLFEA_KOP_RESIDUAL:  // local g from the rank's own f_int, mass and f_ext: halo-sum g afterwards
      TL_TRY(compute_residual(s));
      break;
    case TLFEA_KOP_NORM_G:
      TL_TRY(dot_owned(s, s->d_g, s->d_g, s->d_scal + 3));
      break;
    case TLFEA_KOP_BEGIN:  // r = -g, dv = 0, raw diagonal blocks of the local H
      TL_LAUNCH(k_neg_copy, gb, 256, 0, e->stream, n, s->d_g, r);
      TL_CUDA(cudaMemsetAsync(x, 0, n * sizeof(double), e->stream));
      TL_LAUNCH(k_diag_extract, nb, 256, 0, e->stream, e->n_coef, e->pat.d_off, e->pat.d_cols, s->d_H, diag);
      break;
    case TLFEA_KOP_PRECOND:  // after the halo sum of the diagonal blocks
      TL_LAUNCH(k_diag_invert, nb, 256, 0, e->stream, e->n_coef, diag, Dinv);
      TL_LAUNCH(k_block_jacobi, nb, 256, 0, e->stream, e->n_coef, Dinv, r, z);
      TL_CUDA(cudaMemcpyAsync(p, z, n * sizeof(double), cudaMemcpyDeviceToDevice, e->stream));
      TL_TRY(dot_owned(s, r, z, s->d_scal + 0));
      TL_TRY(dot_owned(s, r, r, s->d_scal + 3));
      break;
    case TLFEA_KOP_SPMV:  // Ap = H_local p: halo-sum Ap afterwards
      TL_LAUNCH(k_spmv, blocks_for((int64_t)e->n_coef * 32, 256), 256, 0, e->stream, e->n_coef, e->pat.d_off,
                e->pat.d_cols, s->d_H, p, Ap);
      break;
    case TLFEA_KOP_PAP:
      TL_TRY(dot_owned(s, p, Ap, s->d_scal + 1));
      break;
    case TLFEA_KOP_UPDATE_XR:  // alpha = r.z / p.Ap; dv += alpha p; r -= alpha Ap; z = Dinv r
      TL_LAUNCH(k_pcg_update_xr, gb, 256, 0, e->stream, n, s->d_scal, p, Ap, x, r);
      TL_LAUNCH(k_block_jacobi, nb, 256, 0, e->stream, e->n_coef, Dinv, r, z);
      TL_TRY(dot_owned(s, r, z, s->d_scal + 2));
      TL_TRY(dot_owned(s, r, r, s->d_scal + 3));
      break;
    case TLFEA_KOP_UPDATE_P:  // beta = new r.z / old r.z; p = z + beta p
      TL_LAUNCH(k_pcg_update_p, gb, 256, 0, e->stream, n, s->d_scal, z, p);
      TL_LAUNCH(k_copy_scalar, 1, 1, 0, e->stream, s->d_scal + 0, s->d_scal + 2);
      break;
    case TLFEA_KOP_APPLY:  // v += dv; x = x_prev + h v   (SyncedNewton.cu:1116-1119)
      TL_LAUNCH(k_axpy, gb, 256, 0, e->stream, n, 1.0, x, s->d_v);
      TL_LAUNCH(k_update_pos, nb, 256, 0, e->stream, e->n_coef, s->d_xp, s->d_yp, s->d_zp, s->d_v, h, e->d_x, e->d_y,
                e->d_z);
      break;
    default:
      return fail(TLFEA_ERR_INVALID, "unknown Krylov op %d", op);
  }
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

// 3x3 diagonal blocks of the assembled H, [n_coef][9] row-major: what the reference's VBD sweep builds per node
// (vbd_accumulate_residual_and_hessian_diag, FEAT10DataFunc.cuh:295-395) -- here they are a by-product of the
// deterministic assembly, so the VBD / block-Jacobi consumers read them instead of re-integrating.
int tlfea_newton_get_diag_blocks(tlfea_newton *s, double *out, int64_t n) {
  TL_CHECK(s && s->analyzed && out, "AnalyzeHessianSparsity / Assemble must run first");
  tlfea_elem *e = s->e;
  TL_CHECK(n == 9 * (int64_t)e->n_coef, "size mismatch: need 9 * n_coef doubles");
  TL_CUDA(cudaSetDevice(e->device));
  double *d_tmp = nullptr;
  TL_TRY(dmalloc(&d_tmp, (size_t)n));
  TL_LAUNCH(k_diag_extract, blocks_for(e->n_coef, 256), 256, 0, e->stream, e->n_coef, e->pat.d_off, e->pat.d_cols, s->d_H,
            d_tmp);
  cudaError_t err = cudaMemcpyAsync(out, d_tmp, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, e->stream);
  if (err == cudaSuccess) err = cudaStreamSynchronize(e->stream);
  cudaFree(d_tmp);
  TL_CUDA(err);
  return TLFEA_OK;
}

double *tlfea_newton_devptr_krylov(tlfea_newton *s, int which) {
  if (!s) return nullptr;
  switch (which) {
    case TLFEA_KVEC_Z: return s->d_pcg[0];
    case TLFEA_KVEC_P: return s->d_pcg[1];
    case TLFEA_KVEC_AP: return s->d_pcg[2];
    case TLFEA_KVEC_DIAG: return s->d_pcg[4];
    case TLFEA_KVEC_R: return s->d_r;
    case TLFEA_KVEC_DV: return s->d_dv;
    case TLFEA_KVEC_SCAL: return s->d_scal;
    default: return nullptr;
  }
}

int tlfea_newton_timing(tlfea_newton *s, double *ms_total, int *calls, int reset) {
  TL_CHECK(s, "null handle");
  TL_TRY(drain_timing(s));
  if (ms_total) *ms_total = s->ms_total;
  if (calls) *calls = s->calls;
  if (reset) {
    s->ms_total = 0;
    s->calls = 0;
    s->k_ms[0] = s->k_ms[1] = s->k_ms[2] = 0;
  }
  return TLFEA_OK;
}
int tlfea_newton_set_assembly(tlfea_newton *s, int mode) {
  TL_CHECK(s && mode >= 0 && mode <= 3,
           "mode must be 0 (generic row gather), 1 (element kernel + row reduce), 2 (single-pass patches for FEAT10 + "
           "SVK, two-phase otherwise) or 3 (single-pass patches for every FEAT10 material)");
  s->asm_mode = mode;
  return TLFEA_OK;
}
int tlfea_newton_set_patch_warps(tlfea_newton *s, int warps) {
  TL_CHECK(s && (warps == 4 || warps == 8), "warps per patch must be 4 or 8");
  if (s->pk_w != warps) patch_release(s);
  s->pk_w = warps;
  return TLFEA_OK;
}
int tlfea_newton_set_patch_chunks(tlfea_newton *s, int chunks) {
  TL_CHECK(s && chunks >= 0 && chunks <= 16, "id-range chunks must be 0 (automatic) or 1..16");
  if (s->pk_chunks != chunks) patch_release(s);
  s->pk_chunks = chunks;
  return TLFEA_OK;
}
int tlfea_newton_set_affine_compression(tlfea_newton *s, int on) {
  TL_CHECK(s, "null handle");
  s->pk_affine = on ? 1 : 0;
  return TLFEA_OK;
}
int tlfea_newton_patch_stats(tlfea_newton *s, int64_t *out) {
  TL_CHECK(s && out, "null input");
  for (int i = 0; i < 8; i++) out[i] = 0;
  return patch_stats(s, out);
}
int tlfea_newton_set_profile(tlfea_newton *s, int on) {
  TL_CHECK(s, "null handle");
  TL_TRY(drain_timing(s));
  s->profile = on != 0;
  return TLFEA_OK;
}
int tlfea_newton_kernel_ms(tlfea_newton *s, double *gp_ms, double *row_ms, double *other_ms) {
  TL_CHECK(s, "null handle");
  TL_TRY(drain_timing(s));
  if (gp_ms) *gp_ms = s->k_ms[0];
  if (row_ms) *row_ms = s->k_ms[1];
  if (other_ms) *other_ms = s->k_ms[2];
  return TLFEA_OK;
}

// FP64 FMA peak: 8 independent chains per thread, 4096 dependent steps each
__global__ void __launch_bounds__(256) k_fp64_peak(double *out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
         a7 = a0 + 7;
  const double m = 1.0000001, c = 1e-9;
  for (int i = 0; i < iters; i++) {
    a0 = a0 * m + c; a1 = a1 * m + c; a2 = a2 * m + c; a3 = a3 * m + c;
    a4 = a4 * m + c; a5 = a5 * m + c; a6 = a6 * m + c; a7 = a7 * m + c;
  }
  out[(int64_t)blockIdx.x * 256 + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

int tlfea_fp64_peak(int device, double *tflops) {
  TL_CHECK(tflops, "null out");
  TL_CUDA(cudaSetDevice(device));
  const int blocks = sm_count(device) * 16, iters = 8192;
  double *d = nullptr;
  TL_TRY(dmalloc(&d, (size_t)blocks * 256));
  cudaEvent_t a, b;
  TL_CUDA(cudaEventCreate(&a));
  TL_CUDA(cudaEventCreate(&b));
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    TL_CUDA(cudaEventRecord(a));
    k_fp64_peak<<<blocks, 256>>>(d, iters);
    g_launches.fetch_add(1);
    TL_CUDA(cudaEventRecord(b));
    TL_CUDA(cudaEventSynchronize(b));
    float ms = 0;
    TL_CUDA(cudaEventElapsedTime(&ms, a, b));
    const double fl = 2.0 * 8 * (double)iters * blocks * 256;
    best = std::max(best, fl / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(d);
  *tflops = best;
  return TLFEA_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the deterministic dot product (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).  Same two
// kernels, same fixed grid and tree as dot()/norm2() above, so the result is bitwise the one tlfea_newton_residual
// forms -- only the partial sums live in a file-scope device array instead of the handle's cudaMalloc'ed d_scal.
// Every GPU that loads the library holds the array until exit.  All calls on a device share it: calls that overlap on
// the GPU (different streams) race.  k_dot_partial overwrites every one of the DOT_BLOCKS partials on each call, so
// nothing has to be cleared between calls.
namespace tlfea {
constexpr int64_t NOALLOC_DOT_MAX_N = TLFEA_NOALLOC_DOT_MAX_N;
constexpr int NOALLOC_MAX_DEVICES = 64;
static __device__ __align__(256) double g_noalloc_partial[DOT_BLOCKS];  // 592 doubles = 4 736 B per GPU
// device address of g_noalloc_partial per device (the host-side name is NOT a device address)
static std::atomic<double *> g_noalloc_ptr[NOALLOC_MAX_DEVICES];
}  // namespace tlfea

extern "C" int tlfea_dot_noalloc(const double *d_a, const double *d_b, int64_t n, double *d_out, void *stream) {
  TL_CHECK(d_a && d_b && d_out, "tlfea_dot_noalloc: null device pointer");
  TL_CHECK(n >= 0, "tlfea_dot_noalloc: negative length");
  TL_CHECK(n <= NOALLOC_DOT_MAX_N, "tlfea_dot_noalloc: n = %lld exceeds TLFEA_NOALLOC_DOT_MAX_N = %lld", (long long)n,
           (long long)NOALLOC_DOT_MAX_N);
  int dev = -1;
  TL_CUDA(cudaGetDevice(&dev));
  TL_CHECK(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "tlfea_dot_noalloc: device ordinal %d out of range", dev);
  double *partial = g_noalloc_ptr[dev].load(std::memory_order_acquire);
  if (!partial) {  // once per device, with that device current
    TL_CUDA(cudaGetSymbolAddress((void **)&partial, g_noalloc_partial));
    TL_CHECK(partial != nullptr, "tlfea_dot_noalloc: cudaGetSymbolAddress gave a null address");
    g_noalloc_ptr[dev].store(partial, std::memory_order_release);
  }
  cudaStream_t st = (cudaStream_t)stream;
  TL_LAUNCH(k_dot_partial, DOT_BLOCKS, 256, 0, st, n, d_a, d_b, partial);
  TL_LAUNCH(k_dot_final, 1, 1024, 0, st, DOT_BLOCKS, partial, d_out);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}
