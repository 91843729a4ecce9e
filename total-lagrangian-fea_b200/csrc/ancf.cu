// ancf.cu -- ANCF3243 beam / ANCF3443 shell set-up: shape-coefficient matrices B^-T, reference-gradient
// precompute and element mass.  The force / tangent kernels are the generic ones in kernels.cuh.
//
// Reference behaviour replaced (file:line relative to the reference root):
//   ANCF3243_B12_matrix / ANCF3443_B12_matrix      lib_utils/cpu_utils.cc:125-209, 211-441  (Eigen inverse -> own LU)
//   precompute_reference_kernel                    ANCF3243Data.cu:102-198, ANCF3443Data.cu:96-182
//   mass_matrix_qp_kernel                          ANCF3243Data.cu:200-288, ANCF3443Data.cu:184-254
//   basis b(u,v,w) and its derivatives             ANCF3243DataFunc.cuh:115-125,132-152, ANCF3443DataFunc.cuh:114-238
//   Gauss-Legendre tables                          lib_utils/quadrature_utils.h:26-134
// Shape functions s = B^-T b with b the monomial basis; row 4p+0..3 of B holds b, db/du, db/dv, db/dw at the
// element's p-th node (+-L/2 on the beam axis; the four mid-surface corners of the shell).
#include <cmath>
#include <cstring>
#include <vector>

#include "ancf.cuh"

namespace tlfea {

// ---- monomial basis, shared by host (B matrix) and device (quadrature) --------------------------------------
template <int S>
__host__ __device__ inline void ancf_basis(double u, double v, double w, double *b, double *bu, double *bv, double *bw) {
  if (S == 8) {  // 1 u v w uv uw u^2 u^3
    const double bb[8] = {1.0, u, v, w, u * v, u * w, u * u, u * u * u};
    const double du[8] = {0.0, 1.0, 0.0, 0.0, v, w, 2.0 * u, 3.0 * u * u};
    const double dv[8] = {0.0, 0.0, 1.0, 0.0, u, 0.0, 0.0, 0.0};
    const double dw[8] = {0.0, 0.0, 0.0, 1.0, 0.0, u, 0.0, 0.0};
    for (int i = 0; i < 8; i++) b[i] = bb[i], bu[i] = du[i], bv[i] = dv[i], bw[i] = dw[i];
  } else {  // 1 u v w uv uw vw uvw u^2 v^2 u^2v uv^2 u^3 v^3 u^3v uv^3
    const double bb[16] = {1.0, u, v, w, u * v, u * w, v * w, u * v * w, u * u, v * v, (u * u) * v, u * (v * v),
                           u * u * u, v * v * v, (u * u * u) * v, u * (v * v * v)};
    const double du[16] = {0.0, 1.0, 0.0, 0.0, v, w, 0.0, v * w, 2.0 * u, 0.0, 2.0 * u * v, v * v,
                           3.0 * u * u, 0.0, 3.0 * u * u * v, v * v * v};
    const double dv[16] = {0.0, 0.0, 1.0, 0.0, u, 0.0, w, u * w, 0.0, 2.0 * v, u * u, 2.0 * u * v,
                           0.0, 3.0 * v * v, u * u * u, 3.0 * u * v * v};
    const double dw[16] = {0.0, 0.0, 0.0, 1.0, 0.0, u, v, u * v, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    for (int i = 0; i < 16; i++) b[i] = bb[i], bu[i] = du[i], bv[i] = dv[i], bw[i] = dw[i];
  }
}

// inverse of an n x n matrix by Gauss-Jordan with partial pivoting (what Eigen's inverse() does for n > 4)
static bool invert(std::vector<double> &a, int n) {
  std::vector<double> inv((size_t)n * n, 0.0);
  for (int i = 0; i < n; i++) inv[(size_t)i * n + i] = 1.0;
  for (int c = 0; c < n; c++) {
    int piv = c;
    for (int r = c + 1; r < n; r++)
      if (std::fabs(a[(size_t)r * n + c]) > std::fabs(a[(size_t)piv * n + c])) piv = r;
    if (a[(size_t)piv * n + c] == 0.0) return false;
    if (piv != c)
      for (int k = 0; k < n; k++) {
        std::swap(a[(size_t)c * n + k], a[(size_t)piv * n + k]);
        std::swap(inv[(size_t)c * n + k], inv[(size_t)piv * n + k]);
      }
    const double d = 1.0 / a[(size_t)c * n + c];
    for (int k = 0; k < n; k++) a[(size_t)c * n + k] *= d, inv[(size_t)c * n + k] *= d;
    for (int r = 0; r < n; r++) {
      if (r == c) continue;
      const double f = a[(size_t)r * n + c];
      if (f == 0.0) continue;
      for (int k = 0; k < n; k++) {
        a[(size_t)r * n + k] -= f * a[(size_t)c * n + k];
        inv[(size_t)r * n + k] -= f * inv[(size_t)c * n + k];
      }
    }
  }
  a.swap(inv);
  return true;
}

// B^-T of one element, column-major like the reference's Eigen::MatrixXd buffer: out[j*S+i] = Binv(i,j)
template <int S>
static bool b_inv_T(double L, double W, double /*H*/, double *out) {
  double pu[4], pv[4];
  int np;
  if (S == 8) {
    np = 2;
    pu[0] = -L / 2.0, pu[1] = L / 2.0, pv[0] = pv[1] = 0.0;
  } else {
    np = 4;
    pu[0] = -L / 2.0, pv[0] = -W / 2.0;
    pu[1] = L / 2.0, pv[1] = -W / 2.0;
    pu[2] = L / 2.0, pv[2] = W / 2.0;
    pu[3] = -L / 2.0, pv[3] = W / 2.0;
  }
  std::vector<double> Bt((size_t)S * S);  // row-major B^T: Bt[c*S + r] = B(r,c)
  for (int p = 0; p < np; p++) {
    double b[16], bu[16], bv[16], bw[16];
    ancf_basis<S>(pu[p], pv[p], 0.0, b, bu, bv, bw);
    for (int c = 0; c < S; c++) {
      Bt[(size_t)c * S + 4 * p + 0] = b[c];
      Bt[(size_t)c * S + 4 * p + 1] = bu[c];
      Bt[(size_t)c * S + 4 * p + 2] = bv[c];
      Bt[(size_t)c * S + 4 * p + 3] = bw[c];
    }
  }
  if (!invert(Bt, S)) return false;
  for (int i = 0; i < S; i++)
    for (int j = 0; j < S; j++) out[(size_t)j * S + i] = Bt[(size_t)i * S + j];
  return true;
}

__global__ void ancf_ecoef_kernel(int n_elem, int type, const int *__restrict__ conn, int S, int *ecoef) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_elem * S) return;
  const int e = (int)(tid / S), a = (int)(tid % S), ln = a >> 2, dof = a & 3;
  // ANCF3243: row-major [e][2] (ANCF3243Data.cuh:310-312); ANCF3443: column-major [4][n_elem] (ANCF3443Data.cuh:212-213)
  const int node = (type == TLFEA_TYPE_3243) ? conn[(int64_t)e * 2 + ln] : conn[(int64_t)ln * n_elem + e];
  ecoef[tid] = node * 4 + dof;
}

int ancf_setup(tlfea_elem *h, const double *L, const double *W, const double *H, const int *conn) {
  const int S = h->S, ne = h->n_elem;
  TL_CUDA(cudaMemcpyAsync(h->d_conn, conn, (size_t)ne * h->nodes_per_elem * sizeof(int), cudaMemcpyHostToDevice,
                          h->stream));
  TL_LAUNCH(ancf_ecoef_kernel, blocks_for((int64_t)ne * S, 256), 256, 0, h->stream, ne, h->type, h->d_conn, S,
            h->d_ecoef);
  std::vector<double> lwh((size_t)ne * 3), binv((size_t)ne * S * S);
  for (int e = 0; e < ne; e++) {
    lwh[(size_t)e * 3] = L[e], lwh[(size_t)e * 3 + 1] = W[e], lwh[(size_t)e * 3 + 2] = H[e];
    double *dst = binv.data() + (size_t)e * S * S;
    if (e > 0 && L[e] == L[e - 1] && W[e] == W[e - 1] && H[e] == H[e - 1]) {
      std::memcpy(dst, dst - (size_t)S * S, (size_t)S * S * sizeof(double));  // uniform meshes: invert once
      continue;
    }
    const bool ok = (S == 8) ? b_inv_T<8>(L[e], W[e], H[e], dst) : b_inv_T<16>(L[e], W[e], H[e], dst);
    TL_CHECK(ok, "element %d: singular shape-coefficient matrix (L=%g W=%g H=%g)", e, L[e], W[e], H[e]);
  }
  TL_TRY(dmalloc(&h->d_LWH, lwh.size()));
  TL_TRY(dmalloc(&h->d_Binv, binv.size()));
  TL_CUDA(cudaMemcpyAsync(h->d_LWH, lwh.data(), lwh.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaMemcpyAsync(h->d_Binv, binv.data(), binv.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  TL_CUDA(cudaGetLastError());
  TL_CUDA(cudaStreamSynchronize(h->stream));
  h->is_setup = true;
  return TLFEA_OK;
}

// ---- Gauss-Legendre tables (lib_utils/quadrature_utils.h) ---------------------------------------------------
struct AncfRule {
  int n[3];
  double p[3][7], w[3][7];
};
static void set_axis(AncfRule &r, int ax, int n) {
  static const double p2[2] = {-0.57735026918962576, 0.57735026918962576}, w2[2] = {1.0, 1.0};
  static const double p3[3] = {-0.77459666924148340, 0.0, 0.77459666924148340};
  static const double w3[3] = {0.55555555555555556, 0.88888888888888889, 0.55555555555555556};
  static const double p4[4] = {-0.8611363115940526, -0.3399810435848563, 0.3399810435848563, 0.8611363115940526};
  static const double w4[4] = {0.3478548451374538, 0.6521451548625461, 0.6521451548625461, 0.3478548451374538};
  static const double p6[6] = {-0.93246951420315202, -0.66120938646626451, -0.23861918608319691,
                               0.23861918608319691,  0.66120938646626451,  0.93246951420315202};
  static const double w6[6] = {0.17132449237917034, 0.36076157304813861, 0.46791393457269104,
                               0.46791393457269104, 0.36076157304813861, 0.17132449237917034};
  static const double p7[7] = {-0.949107912342759, -0.741531185599394, -0.405845151377397, 0.0,
                               0.405845151377397,  0.741531185599394,  0.949107912342759};
  static const double w7[7] = {0.129484966168870, 0.279705391489277, 0.381830050505119, 0.417959183673469,
                               0.381830050505119, 0.279705391489277, 0.129484966168870};
  // the mass rule's zeta axis of the shell uses the 16-digit 3-point table (gauss_zeta_m_3)
  static const double p3m[3] = {-0.7745966692414834, 0.0, 0.7745966692414834};
  static const double w3m[3] = {0.5555555555555556, 0.8888888888888888, 0.5555555555555556};
  const double *p = nullptr, *w = nullptr;
  switch (n) {
    case 2: p = p2, w = w2; break;
    case 3: p = p3, w = w3; break;
    case -3: p = p3m, w = w3m, n = 3; break;
    case 4: p = p4, w = w4; break;
    case 6: p = p6, w = w6; break;
    default: p = p7, w = w7, n = 7;
  }
  r.n[ax] = n;
  for (int i = 0; i < n; i++) r.p[ax][i] = p[i], r.w[ax][i] = w[i];
}
static AncfRule stiffness_rule(int type) {
  AncfRule r{};
  if (type == TLFEA_TYPE_3243) {  // 3 x 2 x 2   (ANCF3243Data.cu:113-119)
    set_axis(r, 0, 3), set_axis(r, 1, 2), set_axis(r, 2, 2);
  } else {  // 4 x 4 x 3   (ANCF3443Data.cu:107-113)
    set_axis(r, 0, 4), set_axis(r, 1, 4), set_axis(r, 2, 3);
  }
  return r;
}
static AncfRule mass_rule(int type) {
  AncfRule r{};
  if (type == TLFEA_TYPE_3243) {  // 6 x 2 x 2   (ANCF3243Data.cu:201-202,249-258)
    set_axis(r, 0, 6), set_axis(r, 1, 2), set_axis(r, 2, 2);
  } else {  // 7 x 7 x 3   (ANCF3443Data.cu:185-205)
    set_axis(r, 0, 7), set_axis(r, 1, 7), set_axis(r, 2, -3);
  }
  return r;
}

// shape functions and their xi/eta/zeta derivatives at a normalised point: s = B^-T b, ds = B^-T db
template <int S>
__device__ __forceinline__ void ancf_shape(const double *__restrict__ Binv, double L, double W, double H, double xi,
                                           double eta, double zeta, double *s, double *dxi, double *deta,
                                           double *dzeta) {
  double b[S], bu[S], bv[S], bw[S];
  ancf_basis<S>(0.5 * L * xi, 0.5 * W * eta, 0.5 * H * zeta, b, bu, bv, bw);
#pragma unroll
  for (int i = 0; i < S; i++) bu[i] *= 0.5 * L, bv[i] *= 0.5 * W, bw[i] *= 0.5 * H;  // chain rule
  for (int i = 0; i < S; i++) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    for (int j = 0; j < S; j++) {
      const double m = Binv[j * S + i];  // column-major B^-T(i,j)
      a0 += m * b[j];
      a1 += m * bu[j];
      a2 += m * bv[j];
      a3 += m * bw[j];
    }
    s[i] = a0, dxi[i] = a1, deta[i] = a2, dzeta[i] = a3;
  }
}

template <int S>
__device__ __forceinline__ void ancf_jacobian(const int *__restrict__ ec, const double *__restrict__ x,
                                              const double *__restrict__ y, const double *__restrict__ z,
                                              const double *dxi, const double *deta, const double *dzeta,
                                              double J[3][3]) {
#pragma unroll
  for (int i = 0; i < 3; i++) J[i][0] = J[i][1] = J[i][2] = 0.0;
  for (int a = 0; a < S; a++) {
    const int c = ec[a];
    const double X[3] = {x[c], y[c], z[c]};
#pragma unroll
    for (int i = 0; i < 3; i++) {
      J[i][0] += X[i] * dxi[a];
      J[i][1] += X[i] * deta[a];
      J[i][2] += X[i] * dzeta[a];
    }
  }
}

__device__ __forceinline__ double det33(const double J[3][3]) {
  return J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
         J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
}

// Gaussian elimination with partial pivoting, |pivot| < 1e-14 => zero (ancf3243_solve_3x3_system)
__device__ __forceinline__ void solve33(const double A[3][3], const double b[3], double x[3]) {
  double m[3][4];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) m[i][j] = A[i][j];
    m[i][3] = b[i];
  }
  for (int k = 0; k < 3; k++) {
    int piv = k;
    for (int i = k + 1; i < 3; i++)
      if (fabs(m[i][k]) > fabs(m[piv][k])) piv = i;
    if (piv != k)
      for (int j = 0; j < 4; j++) {
        const double t = m[k][j];
        m[k][j] = m[piv][j];
        m[piv][j] = t;
      }
    if (fabs(m[k][k]) < 1e-14) {
      x[0] = x[1] = x[2] = 0.0;
      return;
    }
    for (int i = k + 1; i < 3; i++) {
      const double f = m[i][k] / m[k][k];
      for (int j = k; j < 4; j++) m[i][j] -= f * m[k][j];
    }
  }
  x[2] = m[2][3] / m[2][2];
  x[1] = (m[1][3] - m[1][2] * x[2]) / m[1][1];
  x[0] = (m[0][3] - m[0][2] * x[2] - m[0][1] * x[1]) / m[0][0];
}

// precompute_reference_kernel: detJ and grad_s = J^-T [ds/dxi ds/deta ds/dzeta] at every stiffness Gauss point
template <int S>
__global__ void ancf_precompute_kernel(int n_elem, int Q, AncfRule r, const int *__restrict__ ecoef,
                                       const double *__restrict__ x, const double *__restrict__ y,
                                       const double *__restrict__ z, const double *__restrict__ LWH,
                                       const double *__restrict__ Binv, double *gradN, double *detJ) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_elem * Q) return;
  const int e = (int)(tid / Q), q = (int)(tid - (int64_t)e * Q);
  const int ix = q / (r.n[1] * r.n[2]), ie = (q / r.n[2]) % r.n[1], iz = q % r.n[2];
  const double L = LWH[(int64_t)e * 3], W = LWH[(int64_t)e * 3 + 1], H = LWH[(int64_t)e * 3 + 2];
  double s[S], dxi[S], deta[S], dzeta[S];
  ancf_shape<S>(Binv + (int64_t)e * S * S, L, W, H, r.p[0][ix], r.p[1][ie], r.p[2][iz], s, dxi, deta, dzeta);
  double J[3][3];
  ancf_jacobian<S>(ecoef + (int64_t)e * S, x, y, z, dxi, deta, dzeta, J);
  detJ[tid] = det33(J);
  double JT[3][3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) JT[i][j] = J[j][i];
  double *g = gradN + tid * 3 * S;
  for (int a = 0; a < S; a++) {
    const double rhs[3] = {dxi[a], deta[a], dzeta[a]};
    double gr[3];
    solve33(JT, rhs, gr);
    g[a] = gr[0];
    g[S + a] = gr[1];
    g[2 * S + a] = gr[2];
  }
}

// element mass rho * sum_q s_i s_j w detJ on the mass rule (shape functions and detJ evaluated on the fly)
template <int S>
__global__ void ancf_mass_kernel(int n_elem, AncfRule r, const int *__restrict__ ecoef, const double *__restrict__ x,
                                 const double *__restrict__ y, const double *__restrict__ z,
                                 const double *__restrict__ LWH, const double *__restrict__ Binv, double rho,
                                 double *Me) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elem) return;
  const double L = LWH[(int64_t)e * 3], W = LWH[(int64_t)e * 3 + 1], H = LWH[(int64_t)e * 3 + 2];
  double *M = Me + (int64_t)e * S * S;
  for (int i = 0; i < S * S; i++) M[i] = 0.0;
  const int nq = r.n[0] * r.n[1] * r.n[2];
  for (int q = 0; q < nq; q++) {
    const int ix = q / (r.n[1] * r.n[2]), ie = (q / r.n[2]) % r.n[1], iz = q % r.n[2];
    const double weight = r.w[0][ix] * r.w[1][ie] * r.w[2][iz];
    double s[S], dxi[S], deta[S], dzeta[S];
    ancf_shape<S>(Binv + (int64_t)e * S * S, L, W, H, r.p[0][ix], r.p[1][ie], r.p[2][iz], s, dxi, deta, dzeta);
    double J[3][3];
    ancf_jacobian<S>(ecoef + (int64_t)e * S, x, y, z, dxi, deta, dzeta, J);
    const double dj = det33(J);
    for (int i = 0; i < S; i++)
      for (int j = 0; j < S; j++) M[i * S + j] += rho * s[i] * s[j] * weight * dj;  // reference operand order
  }
}

int ancf_precompute(tlfea_elem *h) {
  const AncfRule r = stiffness_rule(h->type);
  for (int q = 0; q < h->Q; q++) {
    const int ix = q / (r.n[1] * r.n[2]), ie = (q / r.n[2]) % r.n[1], iz = q % r.n[2];
    h->qw.w[q] = r.w[0][ix] * r.w[1][ie] * r.w[2][iz];  // ANCF3243DataFunc.cuh:432-435, ANCF3443DataFunc.cuh:514-516
  }
  const unsigned grid = blocks_for((int64_t)h->n_elem * h->Q, 64);
  // reference gradients are taken at the REFERENCE coordinates (x12_jac), as the reference kernel does
  if (h->type == TLFEA_TYPE_3243)
    TL_LAUNCH(ancf_precompute_kernel<8>, grid, 64, 0, h->stream, h->n_elem, h->Q, r, h->d_ecoef, h->d_xr, h->d_yr,
              h->d_zr, h->d_LWH, h->d_Binv, h->d_gradN, h->d_detJ);
  else
    TL_LAUNCH(ancf_precompute_kernel<16>, grid, 64, 0, h->stream, h->n_elem, h->Q, r, h->d_ecoef, h->d_xr, h->d_yr,
              h->d_zr, h->d_LWH, h->d_Binv, h->d_gradN, h->d_detJ);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

int ancf_elem_mass(tlfea_elem *h, double *d_Me) {
  const AncfRule r = mass_rule(h->type);
  const unsigned grid = blocks_for(h->n_elem, 64);
  if (h->type == TLFEA_TYPE_3243)
    TL_LAUNCH(ancf_mass_kernel<8>, grid, 64, 0, h->stream, h->n_elem, r, h->d_ecoef, h->d_xr, h->d_yr, h->d_zr,
              h->d_LWH, h->d_Binv, h->rho0, d_Me);
  else
    TL_LAUNCH(ancf_mass_kernel<16>, grid, 64, 0, h->stream, h->n_elem, r, h->d_ecoef, h->d_xr, h->d_yr, h->d_zr,
              h->d_LWH, h->d_Binv, h->rho0, d_Me);
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

}  // namespace tlfea
