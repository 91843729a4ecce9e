// elem.cu -- element handles: set-up, reference-gradient precompute, mass, CalcP /
// CalcInternalForce, constraints, retrieval.  Host-side mirror of the reference's
// GPU_FEAT10_Data / GPU_ANCF3243_Data / GPU_ANCF3443_Data objects behind the C ABI.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <set>
#include <vector>

#include "ancf.cuh"
#include "common.cuh"
#include "kernels.cuh"

namespace tlfea {
thread_local std::string g_last_error;
std::atomic<int64_t> g_launches{0};
static std::mutex g_live_mu;
static std::set<const void *> g_live;
bool elem_is_live(const void *h) {
  std::lock_guard<std::mutex> lk(g_live_mu);
  return g_live.count(h) != 0;
}
static void elem_set_live(const void *h, bool live) {
  std::lock_guard<std::mutex> lk(g_live_mu);
  if (live)
    g_live.insert(h);
  else
    g_live.erase(h);
}

// ---------------------------------------------------------------------------
// FEAT10 reference gradients: dn_du_pre_kernel (FEAT10Data.cu:97-204)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void solve3_pp(const double A[3][3], const double b[3], double x[3]) {
  // Gaussian elimination with partial pivoting; |pivot| < 1e-14 => zero solution
  // (same singular-element behaviour as FEAT10DataFunc.cuh:30-83).
  double m[3][4];
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) m[i][j] = A[i][j];
    m[i][3] = b[i];
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
    int piv = k;
    double mx = fabs(m[k][k]);
#pragma unroll
    for (int i = k + 1; i < 3; i++)
      if (fabs(m[i][k]) > mx) {
        mx = fabs(m[i][k]);
        piv = i;
      }
    if (piv != k) {
#pragma unroll
      for (int j = 0; j < 4; j++) {
        double t = m[k][j];
        m[k][j] = m[piv][j];
        m[piv][j] = t;
      }
    }
    if (fabs(m[k][k]) < 1e-14) {
      x[0] = x[1] = x[2] = 0.0;
      return;
    }
#pragma unroll
    for (int i = k + 1; i < 3; i++) {
      double f = m[i][k] / m[k][k];
#pragma unroll
      for (int j = k; j < 4; j++) m[i][j] -= f * m[k][j];
    }
  }
  x[2] = m[2][3] / m[2][2];
  x[1] = (m[1][3] - m[1][2] * x[2]) / m[1][1];
  x[0] = (m[0][3] - m[0][2] * x[2] - m[0][1] * x[1]) / m[0][0];
}

struct T10Quad {
  double x[5], y[5], z[5], w[5];
};

__device__ __forceinline__ void t10_dN(double xi, double eta, double zeta, double dN[10][3]) {
  const double L[4] = {1.0 - xi - eta - zeta, xi, eta, zeta};
  const double dL[4][3] = {{-1.0, -1.0, -1.0}, {1.0, 0.0, 0.0}, {0.0, 1.0, 0.0}, {0.0, 0.0, 1.0}};
  const int ed[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const double f = 4.0 * L[i] - 1.0;
#pragma unroll
    for (int j = 0; j < 3; j++) dN[i][j] = f * dL[i][j];
  }
#pragma unroll
  for (int k = 0; k < 6; k++) {
#pragma unroll
    for (int d = 0; d < 3; d++) dN[k + 4][d] = 4.0 * (L[ed[k][0]] * dL[ed[k][1]][d] + L[ed[k][1]] * dL[ed[k][0]][d]);
  }
}

__global__ void t10_precompute_kernel(int n_elem, const int *__restrict__ ecoef, const double *__restrict__ x,
                                      const double *__restrict__ y, const double *__restrict__ z, T10Quad qd,
                                      double *gradN, double *detJ) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_elem * 5) return;
  const int e = (int)(tid / 5), q = (int)(tid % 5);
  double dN[10][3];
  t10_dN(qd.x[q], qd.y[q], qd.z[q], dN);
  double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
  for (int a = 0; a < 10; a++) {
    const int c = ecoef[(int64_t)e * 10 + a];
    const double X[3] = {x[c], y[c], z[c]};
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) J[i][j] += X[i] * dN[a][j];
  }
  detJ[tid] = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
              J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
  double JT[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) JT[i][j] = J[j][i];
  double *g = gradN + tid * 30;
#pragma unroll
  for (int a = 0; a < 10; a++) {
    double h[3];
    solve3_pp(JT, dN[a], h);
    g[a] = h[0];
    g[10 + a] = h[1];
    g[20 + a] = h[2];
  }
}

// Affine detection + compressed record: J at the 5 points; if it does not vary (1e-12 of its largest
// entry) the element is straight-sided and gradN[q][j][a] = sum_k Ginv[j][k] dN_a/dxi_k(q) with Ginv = J^-T.
__global__ void t10_affine_kernel(int n_elem, const int *__restrict__ ecoef, const double *__restrict__ x,
                                  const double *__restrict__ y, const double *__restrict__ z, T10Quad qd, double *aff,
                                  unsigned char *flag) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elem) return;
  double X[10][3];
#pragma unroll
  for (int a = 0; a < 10; a++) {
    const int c = ecoef[(int64_t)e * 10 + a];
    X[a][0] = x[c], X[a][1] = y[c], X[a][2] = z[c];
  }
  double J0[3][3], dmax = 0.0, jmax = 0.0;
  for (int q = 0; q < 5; q++) {
    double dN[10][3], J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    t10_dN(qd.x[q], qd.y[q], qd.z[q], dN);
#pragma unroll
    for (int a = 0; a < 10; a++)
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) J[i][j] += X[a][i] * dN[a][j];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) {
        if (q == 0) J0[i][j] = J[i][j], jmax = fmax(jmax, fabs(J[i][j]));
        dmax = fmax(dmax, fabs(J[i][j] - J0[i][j]));
      }
  }
  const double det = J0[0][0] * (J0[1][1] * J0[2][2] - J0[1][2] * J0[2][1]) -
                     J0[0][1] * (J0[1][0] * J0[2][2] - J0[1][2] * J0[2][0]) +
                     J0[0][2] * (J0[1][0] * J0[2][1] - J0[1][1] * J0[2][0]);
  double JT[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) JT[i][j] = J0[j][i];
  double *o = aff + (int64_t)e * 10;
  bool ok = (dmax <= 1e-12 * jmax) && jmax > 0.0;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const double ek[3] = {k == 0 ? 1.0 : 0.0, k == 1 ? 1.0 : 0.0, k == 2 ? 1.0 : 0.0};
    double hcol[3];
    solve3_pp(JT, ek, hcol);  // column k of J^-T (zero if a pivot vanishes: then the general path keeps the element)
    ok = ok && (hcol[0] != 0.0 || hcol[1] != 0.0 || hcol[2] != 0.0);
    o[0 * 3 + k] = hcol[0];
    o[1 * 3 + k] = hcol[1];
    o[2 * 3 + k] = hcol[2];
  }
  o[9] = det;
  flag[e] = ok ? 1 : 0;
}

// element mass matrix rho * sum_q N_a N_b detJ w  (mass_matrix_qp_kernel, FEAT10Data.cu:206-278)
__global__ void t10_elem_mass_kernel(int n_elem, const double *__restrict__ detJ, T10Quad qd, double rho,
                                     double *Me) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_elem * 100) return;
  const int e = (int)(tid / 100), r = (int)(tid % 100), il = r / 10, jl = r % 10;
  const int ed[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};
  double m = 0.0;
  for (int q = 0; q < 5; q++) {
    const double L[4] = {1.0 - qd.x[q] - qd.y[q] - qd.z[q], qd.x[q], qd.y[q], qd.z[q]};
    double N[10];
#pragma unroll
    for (int k = 0; k < 4; k++) N[k] = L[k] * (2.0 * L[k] - 1.0);
#pragma unroll
    for (int k = 0; k < 6; k++) N[k + 4] = 4.0 * L[ed[k][0]] * L[ed[k][1]];
    m += rho * N[il] * N[jl] * detJ[(int64_t)e * 5 + q] * qd.w[q];
  }
  Me[tid] = m;
}

// deterministic gather of element mass matrices into the coefficient-level CSR:
// one thread per row, ascending element order (replaces search + atomicAdd).
__global__ void mass_gather_kernel(int n_coef, int S, const int *__restrict__ off, const int *__restrict__ inc_ptr,
                                   const int *__restrict__ inc_ea, const uint16_t *__restrict__ slot,
                                   const double *__restrict__ Me, double *mass) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n_coef) return;
  const int rs = off[row], re = off[row + 1];
  for (int i = rs; i < re; i++) mass[i] = 0.0;
  for (int v = inc_ptr[row]; v < inc_ptr[row + 1]; v++) {
    const int ea = inc_ea[v];
    const int e = ea / S, a = ea - e * S;
    for (int b = 0; b < S; b++) mass[rs + slot[(int64_t)v * S + b]] += Me[((int64_t)e * S + a) * S + b];
  }
}

__global__ void fixed_constraint_kernel(int n_fixed, const int *__restrict__ fixed, const double *x, const double *y,
                                        const double *z, const double *xr, const double *yr, const double *zr,
                                        double *c) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_fixed) return;
  const int n = fixed[i];
  c[3 * i + 0] = x[n] - xr[n];
  c[3 * i + 1] = y[n] - yr[n];
  c[3 * i + 2] = z[n] - zr[n];
}

// c = J q - rhs with q the interleaved coefficient vector (ANCF3243Data.cuh:810-940 linear-CSR mode)
__global__ void linear_constraint_kernel(int n_c, const int *__restrict__ off, const int *__restrict__ cols,
                                         const double *__restrict__ vals, const double *__restrict__ rhs,
                                         const double *x, const double *y, const double *z, double *c) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_c) return;
  double s = 0.0;
  for (int k = off[i]; k < off[i + 1]; k++) {
    const int dof = cols[k], coef = dof / 3, d = dof - 3 * coef;
    s += vals[k] * (d == 0 ? x[coef] : (d == 1 ? y[coef] : z[coef]));
  }
  c[i] = s - rhs[i];
}

__global__ void t10_ecoef_kernel(int n_elem, const int *__restrict__ conn, int *ecoef) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_elem * 10) return;
  const int e = (int)(tid / 10), n = (int)(tid % 10);
  ecoef[tid] = conn[(int64_t)n * n_elem + e];
}

template <int S, int Q>
static int launch_gp_t(tlfea_elem *h, const GPArgs &a, bool damp) {
  using Sh = GPShape<S, Q>;
  const unsigned grid = blocks_for(h->n_elem, Sh::EPB);
  const size_t smem = Sh::SMEM;  // gradN staging tile (always), reused for the force partials
#define TL_GP(MODEL, DAMP)                                                                                     \
  do {                                                                                                         \
    if (smem > 48 * 1024)                                                                                      \
      TL_CUDA(cudaFuncSetAttribute(gp_kernel<S, Q, MODEL, DAMP>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                   (int)smem));                                                                \
    TL_LAUNCH((gp_kernel<S, Q, MODEL, DAMP>), grid, Sh::THREADS, smem, h->stream, a);                          \
  } while (0)
  if (h->model == TL_MODEL_SVK) {
    if (damp)
      TL_GP(TL_MODEL_SVK, true);
    else
      TL_GP(TL_MODEL_SVK, false);
  } else {
    if (damp)
      TL_GP(TL_MODEL_MR, true);
    else
      TL_GP(TL_MODEL_MR, false);
  }
#undef TL_GP
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

int launch_gp(tlfea_elem *h, const double *v, double *rec, double *fe, double *F_out, double *P_out) {
  GPArgs a;
  a.n_elem = h->n_elem;
  a.ecoef = h->d_ecoef;
  a.x = h->d_x;
  a.y = h->d_y;
  a.z = h->d_z;
  a.v = v;
  a.gradN = h->d_gradN;
  a.detJ = h->d_detJ;
  a.qw = h->qw;
  a.lambda = h->lambda;
  a.mu = h->mu;
  a.mu10 = h->mu10;
  a.mu01 = h->mu01;
  a.kappa = h->kappa;
  a.eta = h->eta;
  a.lambda_d = h->lambda_d;
  a.rec = rec;
  a.fe = fe;
  a.F_out = F_out;
  a.P_out = P_out;
  const bool damp = (v != nullptr) && (h->eta != 0.0 || h->lambda_d != 0.0);  // FEAT10DataFunc.cuh:137
  switch (h->type) {
    case TLFEA_TYPE_T10: return launch_gp_t<10, 5>(h, a, damp);
    case TLFEA_TYPE_3243: return launch_gp_t<8, 12>(h, a, damp);
    case TLFEA_TYPE_3443: return launch_gp_t<16, 48>(h, a, damp);
  }
  return fail(TLFEA_ERR_INVALID, "bad element type");
}

int ensure_pattern(tlfea_elem *h) {
  if (h->is_pattern) return TLFEA_OK;
  TL_CHECK(h->is_setup, "element must be set up before building the pattern");
  // General linear constraints couple the coefficients of a constraint row through h^2 rho J^T J: the reference
  // widens the Hessian pattern with those pairs (SyncedNewton.cu:556-801).  Here the H pattern is 9x the
  // coefficient pattern, so the pairs join the adjacency keys; the mass getter filters them out again.
  std::vector<unsigned long long> extra;
  if (h->cons_mode == TL_CONS_LINEAR_CSR) {
    std::vector<int> row;
    for (int c = 0; c < h->n_constraint; c++) {
      row.clear();
      for (int k = h->h_j_off[c]; k < h->h_j_off[c + 1]; k++) row.push_back(h->h_j_cols[k] / 3);
      std::sort(row.begin(), row.end());
      row.erase(std::unique(row.begin(), row.end()), row.end());
      for (int a : row)
        for (int b : row) extra.push_back(((unsigned long long)(unsigned)a << 32) | (unsigned)b);
    }
  }
  h->pat_extra = !extra.empty();
  TL_TRY(build_pattern(h->d_ecoef, h->n_elem, h->S, h->n_coef, h->stream, &h->pat, extra.data(),
                       (int64_t)extra.size()));
  TL_TRY(dmalloc(&h->d_mass, (size_t)h->pat.nnz));
  TL_CUDA(cudaMemsetAsync(h->d_mass, 0, (size_t)h->pat.nnz * sizeof(double), h->stream));
  h->is_pattern = true;
  h->pattern_version++;
  return TLFEA_OK;
}

}  // namespace tlfea

// f_ext = base + alpha * contact   (test_item_drop.cc:589-598: gravity copied in, contact forces added by cublasDaxpy)
__global__ void k_fext_sum(int n, const double *__restrict__ base, const double *__restrict__ contact, double alpha,
                           double *f_ext) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) f_ext[i] = contact ? base[i] + alpha * contact[i] : base[i];
}

using namespace tlfea;

extern "C" {

const char *tlfea_last_error(void) { return g_last_error.c_str(); }
int tlfea_version(void) { return 100; }
int64_t tlfea_launch_count(void) { return g_launches.load(); }
int tlfea_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int tlfea_elem_create(int type, int n_elem, int n_nodes, int device, tlfea_elem **out) {
  TL_CHECK(out, "null out");
  *out = nullptr;
  TL_CHECK(n_elem > 0 && n_nodes > 0, "n_elem and n_nodes must be positive");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(TLFEA_ERR_NO_DEVICE, "no CUDA device visible: libtlfea_b200 has no CPU fallback");
  TL_CHECK(device >= 0 && device < ndev, "device %d out of range", device);
  TL_CUDA(cudaSetDevice(device));
  tlfea_elem *h = new tlfea_elem();
  h->type = type;
  h->n_elem = n_elem;
  h->n_nodes = n_nodes;
  h->device = device;
  switch (type) {
    case TLFEA_TYPE_T10:
      h->S = 10, h->Q = 5, h->n_coef = n_nodes, h->nodes_per_elem = 10;
      break;
    case TLFEA_TYPE_3243:
      h->S = 8, h->Q = 12, h->n_coef = 4 * n_nodes, h->nodes_per_elem = 2;
      break;
    case TLFEA_TYPE_3443:
      h->S = 16, h->Q = 48, h->n_coef = 4 * n_nodes, h->nodes_per_elem = 4;
      break;
    default:
      delete h;
      return fail(TLFEA_ERR_INVALID, "unknown element type %d", type);
  }
  TL_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  const size_t nc = h->n_coef, ne = n_elem;
  TL_TRY(dmalloc(&h->d_conn, ne * h->nodes_per_elem));
  TL_TRY(dmalloc(&h->d_ecoef, ne * h->S));
  TL_TRY(dmalloc(&h->d_x, nc));
  TL_TRY(dmalloc(&h->d_y, nc));
  TL_TRY(dmalloc(&h->d_z, nc));
  TL_TRY(dmalloc(&h->d_xr, nc));
  TL_TRY(dmalloc(&h->d_yr, nc));
  TL_TRY(dmalloc(&h->d_zr, nc));
  TL_TRY(dmalloc(&h->d_gradN, ne * h->Q * 3 * h->S));
  TL_TRY(dmalloc(&h->d_detJ, ne * h->Q));
  TL_TRY(dmalloc(&h->d_f_int, 3 * nc));
  TL_TRY(dmalloc(&h->d_f_ext, 3 * nc));
  TL_TRY(dmalloc(&h->d_fe, ne * h->S * 3));
  TL_CUDA(cudaMemsetAsync(h->d_f_int, 0, 3 * nc * sizeof(double), h->stream));
  TL_CUDA(cudaMemsetAsync(h->d_f_ext, 0, 3 * nc * sizeof(double), h->stream));
  elem_set_live(h, true);
  *out = h;
  return TLFEA_OK;
}

int tlfea_elem_destroy(tlfea_elem *h) {
  if (!h) return TLFEA_OK;
  elem_set_live(h, false);
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  dfree(h->d_conn);
  dfree(h->d_ecoef);
  dfree(h->d_x);
  dfree(h->d_y);
  dfree(h->d_z);
  dfree(h->d_xr);
  dfree(h->d_yr);
  dfree(h->d_zr);
  dfree(h->d_gradN);
  dfree(h->d_aff);
  dfree(h->d_aff_flag);
  dfree(h->d_detJ);
  dfree(h->d_F);
  dfree(h->d_P);
  dfree(h->d_f_int);
  dfree(h->d_f_ext);
  dfree(h->d_fe);
  dfree(h->d_mass);
  dfree(h->d_LWH);
  dfree(h->d_Binv);
  dfree(h->d_fixed);
  dfree(h->d_constraint);
  dfree(h->d_j_off);
  dfree(h->d_j_cols);
  dfree(h->d_j_vals);
  dfree(h->d_j_rhs);
  dfree(h->d_jt_off);
  dfree(h->d_jt_cols);
  dfree(h->d_jt_vals);
  h->pat.release();
  if (h->own_stream) cudaStreamDestroy(h->stream);
  delete h;
  return TLFEA_OK;
}

int tlfea_elem_set_stream(tlfea_elem *h, void *stream) {
  TL_CHECK(h, "null handle");
  TL_CUDA(cudaSetDevice(h->device));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  if (h->own_stream) cudaStreamDestroy(h->stream);
  h->stream = (cudaStream_t)stream;
  h->own_stream = false;
  return TLFEA_OK;
}

int tlfea_elem_sync(tlfea_elem *h) {
  TL_CHECK(h, "null handle");
  TL_CUDA(cudaSetDevice(h->device));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  return TLFEA_OK;
}

static int upload_coords(tlfea_elem *h, const double *x, const double *y, const double *z, bool also_ref) {
  const size_t b = (size_t)h->n_coef * sizeof(double);
  TL_CUDA(cudaMemcpyAsync(h->d_x, x, b, cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(h->d_y, y, b, cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(h->d_z, z, b, cudaMemcpyHostToDevice, h->stream));
  if (also_ref) {
    TL_CUDA(cudaMemcpyAsync(h->d_xr, x, b, cudaMemcpyHostToDevice, h->stream));
    TL_CUDA(cudaMemcpyAsync(h->d_yr, y, b, cudaMemcpyHostToDevice, h->stream));
    TL_CUDA(cudaMemcpyAsync(h->d_zr, z, b, cudaMemcpyHostToDevice, h->stream));
  }
  TL_CUDA(cudaStreamSynchronize(h->stream));
  return TLFEA_OK;
}

// An out-of-range node id would only show up as an illegal address deep inside a kernel (the reference does not
// check either): refuse it here, where the caller can still be told which entry is wrong.
static int check_node_ids(const int *conn, int64_t n, int n_nodes) {
  for (int64_t i = 0; i < n; i++)
    TL_CHECK(conn[i] >= 0 && conn[i] < n_nodes, "connectivity entry %lld = %d is outside [0, %d)", (long long)i,
             conn[i], n_nodes);
  return TLFEA_OK;
}

int tlfea_elem_setup_t10(tlfea_elem *h, const double *qx, const double *qy, const double *qz, const double *qw,
                         const double *x, const double *y, const double *z, const int *conn) {
  TL_CHECK(h && h->type == TLFEA_TYPE_T10, "handle is not FEAT10");
  TL_CHECK(!h->is_setup, "GPU_FEAT10_Data is already set up.");
  TL_CHECK(qx && qy && qz && qw && x && y && z && conn, "null input");
  TL_TRY(check_node_ids(conn, (int64_t)h->n_elem * 10, h->n_nodes));
  TL_CUDA(cudaSetDevice(h->device));
  for (int q = 0; q < 5; q++) {
    h->t10_q[0][q] = qx[q];
    h->t10_q[1][q] = qy[q];
    h->t10_q[2][q] = qz[q];
    h->t10_q[3][q] = qw[q];
    h->qw.w[q] = qw[q];
  }
  TL_TRY(upload_coords(h, x, y, z, true));
  TL_CUDA(cudaMemcpyAsync(h->d_conn, conn, (size_t)h->n_elem * 10 * sizeof(int), cudaMemcpyHostToDevice, h->stream));
  TL_LAUNCH(t10_ecoef_kernel, blocks_for((int64_t)h->n_elem * 10, 256), 256, 0, h->stream, h->n_elem, h->d_conn,
            h->d_ecoef);
  TL_CUDA(cudaMemsetAsync(h->d_gradN, 0, (size_t)h->n_elem * 5 * 30 * sizeof(double), h->stream));
  TL_CUDA(cudaMemsetAsync(h->d_detJ, 0, (size_t)h->n_elem * 5 * sizeof(double), h->stream));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  h->is_setup = true;
  return TLFEA_OK;
}

int tlfea_elem_setup_3243(tlfea_elem *h, const double *L, const double *W, const double *H, const double *x,
                          const double *y, const double *z, const int *conn) {
  TL_CHECK(h && h->type == TLFEA_TYPE_3243, "handle is not ANCF3243");
  TL_CHECK(!h->is_setup, "GPU_ANCF3243_Data is already set up.");
  TL_CHECK(L && W && H && x && y && z && conn, "null input");
  TL_TRY(check_node_ids(conn, (int64_t)h->n_elem * 2, h->n_nodes));
  TL_CUDA(cudaSetDevice(h->device));
  TL_TRY(upload_coords(h, x, y, z, true));
  TL_TRY(ancf_setup(h, L, W, H, conn));
  h->is_setup = true;
  return TLFEA_OK;
}

int tlfea_elem_setup_3443(tlfea_elem *h, const double *L, const double *W, const double *H, const double *x,
                          const double *y, const double *z, const int *conn) {
  TL_CHECK(h && h->type == TLFEA_TYPE_3443, "handle is not ANCF3443");
  TL_CHECK(!h->is_setup, "GPU_ANCF3443_Data is already set up.");
  TL_CHECK(L && W && H && x && y && z && conn, "null input");
  TL_TRY(check_node_ids(conn, (int64_t)h->n_elem * 4, h->n_nodes));
  TL_CUDA(cudaSetDevice(h->device));
  TL_TRY(upload_coords(h, x, y, z, true));
  TL_TRY(ancf_setup(h, L, W, H, conn));
  h->is_setup = true;
  return TLFEA_OK;
}

int tlfea_elem_set_density(tlfea_elem *h, double rho0) {
  TL_CHECK(h && h->is_setup, "element must be set up before setting density.");
  h->rho0 = rho0;
  return TLFEA_OK;
}
int tlfea_elem_set_damping(tlfea_elem *h, double eta, double lambda_d) {
  TL_CHECK(h && h->is_setup, "element must be set up before setting damping.");
  h->eta = eta;
  h->lambda_d = lambda_d;
  return TLFEA_OK;
}
int tlfea_elem_set_svk(tlfea_elem *h, double E, double nu) {
  TL_CHECK(h && h->is_setup, "element must be set up before setting material.");
  h->E = E;
  h->nu = nu;
  h->mu = E / (2 * (1 + nu));                          // FEAT10Data.cuh:604-605
  h->lambda = (E * nu) / ((1 + nu) * (1 - 2 * nu));
  h->model = TL_MODEL_SVK;
  h->mu10 = h->mu01 = h->kappa = 0.0;
  return TLFEA_OK;
}
int tlfea_elem_set_mooney_rivlin(tlfea_elem *h, double mu10, double mu01, double kappa) {
  TL_CHECK(h && h->is_setup, "element must be set up before setting material.");
  h->model = TL_MODEL_MR;
  h->mu10 = mu10;
  h->mu01 = mu01;
  h->kappa = kappa;
  return TLFEA_OK;
}
int tlfea_elem_set_external_force(tlfea_elem *h, const double *f, int n) {
  TL_CHECK(h && f, "null input");
  TL_CHECK(n == 3 * h->n_coef, "External force vector size mismatch.");
  TL_CUDA(cudaSetDevice(h->device));
  TL_CUDA(cudaMemcpyAsync(h->d_f_ext, f, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  return TLFEA_OK;
}

int tlfea_elem_set_nodal_fixed(tlfea_elem *h, const int *fixed, int n_fixed) {
  TL_CHECK(h && (fixed || n_fixed == 0), "null input");
  TL_CHECK(h->cons_mode == TL_CONS_NONE, "CONSTRAINT is already set up.");
  TL_CHECK(n_fixed >= 0, "negative count");
  TL_CUDA(cudaSetDevice(h->device));
  // ANCF pins all 4 coefficients of a node? No: the reference pins what the caller lists
  // (coefficient indices for ANCF, node indices for FEAT10) -- 3 constraints per entry.
  for (int i = 0; i < n_fixed; i++) TL_CHECK(fixed[i] >= 0 && fixed[i] < h->n_coef, "fixed index out of range");
  {  // a repeated entry would map two constraint rows to the same H slots (the reference's atomicAdd adds both)
    std::vector<int> srt(fixed, fixed + n_fixed);
    std::sort(srt.begin(), srt.end());
    TL_CHECK(std::adjacent_find(srt.begin(), srt.end()) == srt.end(), "SetNodalFixed: an index is listed twice");
  }
  h->n_fixed = n_fixed;
  h->n_constraint = 3 * n_fixed;
  TL_TRY(dmalloc(&h->d_fixed, (size_t)n_fixed));
  TL_TRY(dmalloc(&h->d_constraint, (size_t)h->n_constraint));
  TL_CUDA(cudaMemcpyAsync(h->d_fixed, fixed, (size_t)n_fixed * sizeof(int), cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaMemsetAsync(h->d_constraint, 0, (size_t)h->n_constraint * sizeof(double), h->stream));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  h->cons_mode = TL_CONS_FIXED_NODES;
  return TLFEA_OK;
}

// UpdateNodalFixed (FEAT10Data.cu:751-839): replace the pinned set; buffers are re-made when the count changes.
// A solver bound to this element re-derives its constraint slots on the next call (same count required).
int tlfea_elem_update_nodal_fixed(tlfea_elem *h, const int *fixed, int n_fixed) {
  TL_CHECK(h, "null handle");
  if (h->cons_mode == TL_CONS_NONE) return tlfea_elem_set_nodal_fixed(h, fixed, n_fixed);
  TL_CHECK(h->cons_mode == TL_CONS_FIXED_NODES, "UpdateNodalFixed needs fixed-node constraints");
  TL_CUDA(cudaSetDevice(h->device));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  dfree(h->d_fixed);
  dfree(h->d_constraint);
  h->d_fixed = nullptr;
  h->d_constraint = nullptr;
  h->cons_mode = TL_CONS_NONE;
  h->cons_version++;
  return tlfea_elem_set_nodal_fixed(h, fixed, n_fixed);
}

int tlfea_elem_set_linear_constraints_csr(tlfea_elem *h, int n_c, const int *off, const int *cols,
                                          const double *vals, const double *rhs) {
  TL_CHECK(h && off && cols && vals && rhs, "null input");
  TL_CHECK(h->cons_mode == TL_CONS_NONE, "CONSTRAINT is already set up.");
  TL_CHECK(n_c > 0, "n_constraints must be positive");
  TL_CUDA(cudaSetDevice(h->device));
  const int nnz = off[n_c];
  const int n_dof = 3 * h->n_coef;
  for (int k = 0; k < nnz; k++) TL_CHECK(cols[k] >= 0 && cols[k] < n_dof, "constraint column out of range");
  // transpose on the host in ascending (dof, constraint) order
  std::vector<int> t_off(n_dof + 1, 0), t_cols(nnz);
  std::vector<double> t_vals(nnz);
  for (int k = 0; k < nnz; k++) t_off[cols[k] + 1]++;
  for (int i = 0; i < n_dof; i++) t_off[i + 1] += t_off[i];
  std::vector<int> cur(t_off.begin(), t_off.end() - 1);
  for (int c = 0; c < n_c; c++)
    for (int k = off[c]; k < off[c + 1]; k++) {
      int p = cur[cols[k]]++;
      t_cols[p] = c;
      t_vals[p] = vals[k];
    }
  h->n_constraint = n_c;
  h->j_nnz = nnz;
  TL_TRY(dmalloc(&h->d_j_off, (size_t)n_c + 1));
  TL_TRY(dmalloc(&h->d_j_cols, (size_t)nnz));
  TL_TRY(dmalloc(&h->d_j_vals, (size_t)nnz));
  TL_TRY(dmalloc(&h->d_j_rhs, (size_t)n_c));
  TL_TRY(dmalloc(&h->d_jt_off, (size_t)n_dof + 1));
  TL_TRY(dmalloc(&h->d_jt_cols, (size_t)nnz));
  TL_TRY(dmalloc(&h->d_jt_vals, (size_t)nnz));
  TL_TRY(dmalloc(&h->d_constraint, (size_t)n_c));
  TL_CUDA(cudaMemcpy(h->d_j_off, off, ((size_t)n_c + 1) * sizeof(int), cudaMemcpyHostToDevice));
  TL_CUDA(cudaMemcpy(h->d_j_cols, cols, (size_t)nnz * sizeof(int), cudaMemcpyHostToDevice));
  TL_CUDA(cudaMemcpy(h->d_j_vals, vals, (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice));
  TL_CUDA(cudaMemcpy(h->d_j_rhs, rhs, (size_t)n_c * sizeof(double), cudaMemcpyHostToDevice));
  TL_CUDA(cudaMemcpy(h->d_jt_off, t_off.data(), ((size_t)n_dof + 1) * sizeof(int), cudaMemcpyHostToDevice));
  TL_CUDA(cudaMemcpy(h->d_jt_cols, t_cols.data(), (size_t)nnz * sizeof(int), cudaMemcpyHostToDevice));
  TL_CUDA(cudaMemcpy(h->d_jt_vals, t_vals.data(), (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice));
  TL_CUDA(cudaDeviceSynchronize());  // pageable copies may return before their DMA lands; the handle's stream is non-blocking
  TL_CUDA(cudaMemset(h->d_constraint, 0, (size_t)n_c * sizeof(double)));
  h->cons_mode = TL_CONS_LINEAR_CSR;
  h->h_j_off.assign(off, off + n_c + 1);
  h->h_j_cols.assign(cols, cols + nnz);
  if (h->is_pattern) {  // pattern was built before the constraints arrived: rebuild it with the J^T J couplings
    const bool had_mass = h->is_mass;
    TL_CUDA(cudaStreamSynchronize(h->stream));
    h->pat.release();
    h->pat = Pattern();
    dfree(h->d_mass);
    h->d_mass = nullptr;
    h->is_pattern = h->is_mass = false;
    TL_TRY(ensure_pattern(h));
    if (had_mass) TL_TRY(tlfea_elem_calc_mass(h));
  }
  return TLFEA_OK;
}

int tlfea_elem_update_positions(tlfea_elem *h, const double *x, const double *y, const double *z, int n) {
  TL_CHECK(h && x && y && z, "null input");
  TL_CHECK(n == h->n_coef, "Position vector size mismatch.");
  TL_CUDA(cudaSetDevice(h->device));
  return upload_coords(h, x, y, z, false);
}

int tlfea_elem_update_constraint_targets(tlfea_elem *h, const double *x, const double *y, const double *z, int n) {
  TL_CHECK(h && x && y && z, "null input");
  TL_CHECK(n == h->n_coef, "Position vector size mismatch.");
  TL_CUDA(cudaSetDevice(h->device));
  const size_t b = (size_t)n * sizeof(double);
  TL_CUDA(cudaMemcpyAsync(h->d_xr, x, b, cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(h->d_yr, y, b, cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(h->d_zr, z, b, cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  return TLFEA_OK;
}

int tlfea_elem_precompute(tlfea_elem *h) {
  TL_CHECK(h && h->is_setup, "element must be set up before CalcDnDuPre/CalcDsDuPre");
  TL_CUDA(cudaSetDevice(h->device));
  if (h->type == TLFEA_TYPE_T10) {
    T10Quad qd;
    for (int q = 0; q < 5; q++) {
      qd.x[q] = h->t10_q[0][q];
      qd.y[q] = h->t10_q[1][q];
      qd.z[q] = h->t10_q[2][q];
      qd.w[q] = h->t10_q[3][q];
    }
    // reference gradients are taken at the CURRENT coordinates, as dn_du_pre_kernel reads x12
    TL_LAUNCH(t10_precompute_kernel, blocks_for((int64_t)h->n_elem * 5, 128), 128, 0, h->stream, h->n_elem,
              h->d_ecoef, h->d_x, h->d_y, h->d_z, qd, h->d_gradN, h->d_detJ);
    if (!h->d_aff) TL_TRY(dmalloc(&h->d_aff, (size_t)h->n_elem * 10));
    if (!h->d_aff_flag) TL_TRY(dmalloc(&h->d_aff_flag, (size_t)h->n_elem));
    TL_LAUNCH(t10_affine_kernel, blocks_for(h->n_elem, 128), 128, 0, h->stream, h->n_elem, h->d_ecoef, h->d_x, h->d_y,
              h->d_z, qd, h->d_aff, h->d_aff_flag);
    for (int q = 0; q < 5; q++) {  // the mesh-independent table dN_a/dxi_k at the caller's points (host copy)
      const double xi = qd.x[q], eta = qd.y[q], zeta = qd.z[q];
      const double L[4] = {1.0 - xi - eta - zeta, xi, eta, zeta};
      const double dL[4][3] = {{-1.0, -1.0, -1.0}, {1.0, 0.0, 0.0}, {0.0, 1.0, 0.0}, {0.0, 0.0, 1.0}};
      const int ed[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};
      for (int a = 0; a < 4; a++)
        for (int k = 0; k < 3; k++) h->t10_dn[(q * 10 + a) * 3 + k] = (4.0 * L[a] - 1.0) * dL[a][k];
      for (int m = 0; m < 6; m++)
        for (int k = 0; k < 3; k++)
          h->t10_dn[(q * 10 + 4 + m) * 3 + k] = 4.0 * (L[ed[m][0]] * dL[ed[m][1]][k] + L[ed[m][1]] * dL[ed[m][0]][k]);
    }
    h->aff_version++;
    TL_CUDA(cudaGetLastError());
  } else {
    TL_TRY(ancf_precompute(h));
  }
  h->is_precomputed = true;
  return TLFEA_OK;
}

int tlfea_elem_build_mass_pattern(tlfea_elem *h) {
  TL_CHECK(h, "null handle");
  TL_CUDA(cudaSetDevice(h->device));
  return ensure_pattern(h);
}

int tlfea_elem_calc_mass(tlfea_elem *h) {
  TL_CHECK(h && h->is_setup, "element must be set up");
  TL_CUDA(cudaSetDevice(h->device));
  TL_TRY(ensure_pattern(h));
  double *Me = nullptr;
  TL_TRY(dmalloc(&Me, (size_t)h->n_elem * h->S * h->S));
  if (h->type == TLFEA_TYPE_T10) {
    TL_CHECK(h->is_precomputed, "CalcDnDuPre must run before CalcMassMatrix (detJ is needed)");
    T10Quad qd;
    for (int q = 0; q < 5; q++) {
      qd.x[q] = h->t10_q[0][q];
      qd.y[q] = h->t10_q[1][q];
      qd.z[q] = h->t10_q[2][q];
      qd.w[q] = h->t10_q[3][q];
    }
    TL_LAUNCH(t10_elem_mass_kernel, blocks_for((int64_t)h->n_elem * 100, 128), 128, 0, h->stream, h->n_elem,
              h->d_detJ, qd, h->rho0, Me);
  } else {
    int s = ancf_elem_mass(h, Me);
    if (s != TLFEA_OK) {
      cudaFree(Me);
      return s;
    }
  }
  TL_LAUNCH(mass_gather_kernel, blocks_for(h->n_coef, 128), 128, 0, h->stream, h->n_coef, h->S, h->pat.d_off,
            h->pat.d_inc_ptr, h->pat.d_inc_ea, h->pat.d_slot, Me, h->d_mass);
  TL_CUDA(cudaGetLastError());
  TL_CUDA(cudaStreamSynchronize(h->stream));
  cudaFree(Me);
  h->is_mass = true;
  h->mass_version++;
  return TLFEA_OK;
}

int tlfea_elem_calc_p(tlfea_elem *h) {
  TL_CHECK(h && h->is_precomputed, "CalcDnDuPre must run before CalcP");
  TL_CUDA(cudaSetDevice(h->device));
  const size_t n = (size_t)h->n_elem * h->Q * 9;
  if (!h->d_F) {
    TL_TRY(dmalloc(&h->d_F, n));
    TL_TRY(dmalloc(&h->d_P, n));
  }
  // standalone CalcP passes v_guess = nullptr: elastic stress only (FEAT10Data.cu:302-304)
  TL_TRY(launch_gp(h, nullptr, nullptr, nullptr, h->d_F, h->d_P));
  return TLFEA_OK;
}

int tlfea_elem_calc_internal_force(tlfea_elem *h) {
  TL_CHECK(h && h->d_P, "CalcP must run before CalcInternalForce");
  TL_CUDA(cudaSetDevice(h->device));
  TL_TRY(ensure_pattern(h));
  const int64_t n = (int64_t)h->n_elem * h->S;
  switch (h->type) {
    case TLFEA_TYPE_T10:
      TL_LAUNCH((fe_from_P_kernel<10, 5>), blocks_for(n, 128), 128, 0, h->stream, h->n_elem, h->d_P, h->d_gradN,
                h->d_detJ, h->qw, h->d_fe);
      break;
    case TLFEA_TYPE_3243:
      TL_LAUNCH((fe_from_P_kernel<8, 12>), blocks_for(n, 128), 128, 0, h->stream, h->n_elem, h->d_P, h->d_gradN,
                h->d_detJ, h->qw, h->d_fe);
      break;
    default:
      TL_LAUNCH((fe_from_P_kernel<16, 48>), blocks_for(n, 128), 128, 0, h->stream, h->n_elem, h->d_P, h->d_gradN,
                h->d_detJ, h->qw, h->d_fe);
  }
  // the reference ACCUMULATES into f_int (clearing is the solver's job, SyncedNewton.cu:1049)
  TL_LAUNCH(force_gather_kernel, blocks_for((int64_t)h->n_coef * 3, 128), 128, 0, h->stream, h->n_coef,
            h->pat.d_inc_ptr, h->pat.d_inc_ea, h->d_fe, h->d_f_int, 1);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

int tlfea_elem_clear_internal_force(tlfea_elem *h) {
  TL_CHECK(h, "null handle");
  TL_CUDA(cudaSetDevice(h->device));
  TL_CUDA(cudaMemsetAsync(h->d_f_int, 0, (size_t)3 * h->n_coef * sizeof(double), h->stream));
  return TLFEA_OK;
}

int tlfea_elem_calc_constraint(tlfea_elem *h) {
  TL_CHECK(h, "null handle");
  TL_CHECK(h->cons_mode != TL_CONS_NONE, "constraint is not set up");
  if (h->n_constraint == 0) return TLFEA_OK;
  TL_CUDA(cudaSetDevice(h->device));
  if (h->cons_mode == TL_CONS_FIXED_NODES) {
    TL_LAUNCH(fixed_constraint_kernel, blocks_for(h->n_fixed, 128), 128, 0, h->stream, h->n_fixed, h->d_fixed,
              h->d_x, h->d_y, h->d_z, h->d_xr, h->d_yr, h->d_zr, h->d_constraint);
  } else {
    TL_LAUNCH(linear_constraint_kernel, blocks_for(h->n_constraint, 128), 128, 0, h->stream, h->n_constraint,
              h->d_j_off, h->d_j_cols, h->d_j_vals, h->d_j_rhs, h->d_x, h->d_y, h->d_z, h->d_constraint);
  }
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

int tlfea_elem_n_elem(const tlfea_elem *h) { return h ? h->n_elem : -1; }
int tlfea_elem_n_coef(const tlfea_elem *h) { return h ? h->n_coef : -1; }
int tlfea_elem_n_shape(const tlfea_elem *h) { return h ? h->S : -1; }
int tlfea_elem_n_qp(const tlfea_elem *h) { return h ? h->Q : -1; }
int tlfea_elem_n_constraint(const tlfea_elem *h) { return h ? h->n_constraint : -1; }
int64_t tlfea_elem_mass_nnz(const tlfea_elem *h) { return (h && h->is_pattern) ? h->pat.nnz : -1; }

static int d2h(tlfea_elem *h, void *dst, const void *src, size_t bytes) {
  TL_CHECK(src, "buffer has not been computed yet");
  TL_CUDA(cudaSetDevice(h->device));
  TL_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, h->stream));
  TL_CUDA(cudaStreamSynchronize(h->stream));
  return TLFEA_OK;
}

int tlfea_elem_get_internal_force(tlfea_elem *h, double *out, int n) {
  TL_CHECK(h && out && n == 3 * h->n_coef, "size mismatch");
  return d2h(h, out, h->d_f_int, (size_t)n * sizeof(double));
}
int tlfea_elem_get_positions(tlfea_elem *h, double *x, double *y, double *z, int n) {
  TL_CHECK(h && x && y && z && n == h->n_coef, "size mismatch");
  TL_TRY(d2h(h, x, h->d_x, (size_t)n * sizeof(double)));
  TL_TRY(d2h(h, y, h->d_y, (size_t)n * sizeof(double)));
  return d2h(h, z, h->d_z, (size_t)n * sizeof(double));
}
int tlfea_elem_get_F(tlfea_elem *h, double *out, int64_t n) {
  TL_CHECK(h && out && n == (int64_t)h->n_elem * h->Q * 9, "size mismatch");
  return d2h(h, out, h->d_F, (size_t)n * sizeof(double));
}
int tlfea_elem_get_P(tlfea_elem *h, double *out, int64_t n) {
  TL_CHECK(h && out && n == (int64_t)h->n_elem * h->Q * 9, "size mismatch");
  return d2h(h, out, h->d_P, (size_t)n * sizeof(double));
}
int tlfea_elem_get_grad_N(tlfea_elem *h, double *out, int64_t n) {
  TL_CHECK(h && out && n == (int64_t)h->n_elem * h->Q * 3 * h->S, "size mismatch");
  return d2h(h, out, h->d_gradN, (size_t)n * sizeof(double));
}
int tlfea_elem_get_detJ(tlfea_elem *h, double *out, int64_t n) {
  TL_CHECK(h && out && n == (int64_t)h->n_elem * h->Q, "size mismatch");
  return d2h(h, out, h->d_detJ, (size_t)n * sizeof(double));
}
int tlfea_elem_get_mass_csr(tlfea_elem *h, int *off, int *cols, double *vals) {
  TL_CHECK(h && h->is_pattern, "mass pattern has not been built");
  if (!h->pat_extra) {
    if (off) TL_TRY(d2h(h, off, h->pat.d_off, ((size_t)h->n_coef + 1) * sizeof(int)));
    if (cols) TL_TRY(d2h(h, cols, h->pat.d_cols, (size_t)h->pat.nnz * sizeof(int)));
    if (vals) TL_TRY(d2h(h, vals, h->d_mass, (size_t)h->pat.nnz * sizeof(double)));
    return TLFEA_OK;
  }
  // the stored pattern also holds constraint couplings: hand out the element adjacency only (BuildMassCSRPattern)
  std::vector<int> xo(h->n_coef + 1), xc(h->pat.nnz), ec((size_t)h->n_elem * h->S);
  std::vector<double> xv(h->pat.nnz, 0.0);
  TL_TRY(d2h(h, xo.data(), h->pat.d_off, xo.size() * sizeof(int)));
  TL_TRY(d2h(h, xc.data(), h->pat.d_cols, xc.size() * sizeof(int)));
  TL_TRY(d2h(h, xv.data(), h->d_mass, xv.size() * sizeof(double)));
  TL_TRY(d2h(h, ec.data(), h->d_ecoef, ec.size() * sizeof(int)));
  std::vector<char> keep(h->pat.nnz, 0);
  for (int e = 0; e < h->n_elem; e++)
    for (int a = 0; a < h->S; a++) {
      const int i = ec[(size_t)e * h->S + a];
      for (int b = 0; b < h->S; b++) {
        const int j = ec[(size_t)e * h->S + b];
        keep[std::lower_bound(xc.begin() + xo[i], xc.begin() + xo[i + 1], j) - xc.begin()] = 1;
      }
    }
  int n = 0;
  for (int i = 0; i < h->n_coef; i++) {
    if (off) off[i] = n;
    for (int k = xo[i]; k < xo[i + 1]; k++)
      if (keep[k]) {
        if (cols) cols[n] = xc[k];
        if (vals) vals[n] = xv[k];
        n++;
      }
  }
  if (off) off[h->n_coef] = n;
  return TLFEA_OK;
}
int tlfea_elem_get_constraint(tlfea_elem *h, double *out, int n) {
  TL_CHECK(h && out && n == h->n_constraint, "size mismatch");
  if (n == 0) return TLFEA_OK;
  return d2h(h, out, h->d_constraint, (size_t)n * sizeof(double));
}
// Constraint Jacobian J (rows = constraints, columns = DOFs) in CSR: unit selection rows for pinned nodes
// (FEAT10Data.cu:501-536), the caller's matrix for linear constraints (ANCF3243Data.cuh:810-940).
int64_t tlfea_elem_constraint_jacobian_nnz(tlfea_elem *h) {
  if (!h) return -1;
  if (h->cons_mode == TL_CONS_FIXED_NODES) return 3 * (int64_t)h->n_fixed;
  if (h->cons_mode == TL_CONS_LINEAR_CSR) return h->j_nnz;
  return 0;
}
int tlfea_elem_get_constraint_jacobian_csr(tlfea_elem *h, int *offsets, int *columns, double *values) {
  TL_CHECK(h && offsets && columns && values, "null input");
  TL_CHECK(h->cons_mode != TL_CONS_NONE, "no constraints are set up");
  TL_CUDA(cudaSetDevice(h->device));
  if (h->cons_mode == TL_CONS_FIXED_NODES) {
    std::vector<int> fx((size_t)h->n_fixed);
    if (h->n_fixed > 0) TL_TRY(d2h(h, fx.data(), h->d_fixed, fx.size() * sizeof(int)));
    for (int c = 0; c < 3 * h->n_fixed; c++) {
      offsets[c] = c;
      columns[c] = 3 * fx[(size_t)c / 3] + c % 3;
      values[c] = 1.0;
    }
    offsets[3 * h->n_fixed] = 3 * h->n_fixed;
    return TLFEA_OK;
  }
  TL_TRY(d2h(h, offsets, h->d_j_off, ((size_t)h->n_constraint + 1) * sizeof(int)));
  TL_TRY(d2h(h, columns, h->d_j_cols, (size_t)h->j_nnz * sizeof(int)));
  return d2h(h, values, h->d_j_vals, (size_t)h->j_nnz * sizeof(double));
}
int tlfea_elem_get_connectivity(tlfea_elem *h, int *out, int64_t n) {
  TL_CHECK(h && out && n == (int64_t)h->n_elem * h->nodes_per_elem, "size mismatch");
  return d2h(h, out, h->d_conn, (size_t)n * sizeof(int));
}

const double *tlfea_elem_devptr_x(tlfea_elem *h) { return h ? h->d_x : nullptr; }
const double *tlfea_elem_devptr_y(tlfea_elem *h) { return h ? h->d_y : nullptr; }
const double *tlfea_elem_devptr_z(tlfea_elem *h) { return h ? h->d_z : nullptr; }
double *tlfea_elem_devptr_f_ext(tlfea_elem *h) { return h ? h->d_f_ext : nullptr; }

// ---- contact coupling (CollisionSystemInput, lib_src/collision/CollisionSystemBase.h:17-29; the hand-off of
// lib_bin/collision_system/test_item_drop.cc:560-602), all on the handle's stream, no host copy
int tlfea_elem_pack_nodes_xyz(tlfea_elem *h, double *d_nodes_xyz, int n_nodes) {
  TL_CHECK(h && d_nodes_xyz, "null input");
  TL_CHECK(n_nodes == h->n_coef, "n_nodes must equal the element's coefficient count");
  TL_CUDA(cudaSetDevice(h->device));
  const size_t b = (size_t)n_nodes * sizeof(double);
  TL_CUDA(cudaMemcpyAsync(d_nodes_xyz, h->d_x, b, cudaMemcpyDeviceToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(d_nodes_xyz + n_nodes, h->d_y, b, cudaMemcpyDeviceToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(d_nodes_xyz + 2 * (size_t)n_nodes, h->d_z, b, cudaMemcpyDeviceToDevice, h->stream));
  return TLFEA_OK;
}

int tlfea_elem_set_external_force_device(tlfea_elem *h, const double *d_base, const double *d_contact, double alpha,
                                         int n) {
  TL_CHECK(h && d_base, "null input");
  TL_CHECK(n == 3 * h->n_coef, "External force vector size mismatch.");
  TL_CUDA(cudaSetDevice(h->device));
  TL_LAUNCH(k_fext_sum, blocks_for(n, 256), 256, 0, h->stream, n, d_base, d_contact, alpha, h->d_f_ext);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}
double *tlfea_elem_devptr_f_int(tlfea_elem *h) { return h ? h->d_f_int : nullptr; }

}  // extern "C"
