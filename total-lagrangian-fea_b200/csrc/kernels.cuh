// kernels.cuh -- the hot path: fused Gauss-point kernel + deterministic row-gather
// assembly kernel, generic in <S shape functions, Q Gauss points> so FEAT10 (10,5),
// ANCF3243 (8,12) and ANCF3443 (16,48) share one code path.
//
// Reference behaviour replaced (file:line relative to the reference root):
//   compute_p                     FEAT10DataFunc.cuh:85-293, ANCF3243DataFunc.cuh:189-412, ANCF3443DataFunc.cuh:281-495
//   compute_internal_force        FEAT10DataFunc.cuh:397-458 (+ANCF twins)
//   compute_hessian_assemble_csr  FEAT10DataFunc.cuh:512-791 (+ANCF twins)
//   assemble_sparse_hessian_mass  SyncedNewton.cu:213-259
//   svk_* / mr_*                  SVK.cuh:14-55, MooneyRivlin.cuh:17-225
//
// Design: the reference gives one thread a whole 3S x 3S K_elem in
// local memory and scatters it with binary searches + FP64 atomics.  Here
//   (1) gp_kernel: one thread per Gauss point computes F, the material state and the
//       element force; it leaves a 128-byte record per point (F, dV*FF^T, dV*(lambda trE - mu)),
//   (2) row_kernel: one warp per coefficient ROW of H.  Lanes are (visit k, column b):
//       32/S incident elements are processed at once, each lane accumulating one 3x3
//       block over the Q points in registers with a 16-accumulator factorisation
//           K_ab = c1*A + c2*A^T + c3*G + diag,  A = sum_q dV u_a u_b^T,
//           G = sum_q (h_a.h_b) dV FF^T,  u = F h
//       (20 FMA per block per point instead of the reference's ~170 flop), then the
//       blocks are added to the warp's shared-memory row in ascending element order and
//       the finished row (mass/h included) is streamed to H exactly once.
// No atomics, no search, fixed summation order => bit-reproducible.
#pragma once
#include "common.cuh"

namespace tlfea {

constexpr int REC_SVK = 16;  // doubles per Gauss-point record, SVK
constexpr int REC_MR = 96;   // F(9) | dV*A(81) | dV*FF^T(6)

struct GPArgs {
  int n_elem;
  const int *ecoef;
  const double *x, *y, *z;
  const double *v;  // interleaved velocity or nullptr
  const double *gradN, *detJ;
  QuadW qw;
  double lambda, mu, mu10, mu01, kappa, eta, lambda_d;
  double *rec;    // nullptr: skip
  double *fe;     // nullptr: skip
  double *F_out;  // nullptr: skip (column-major 3x3, reference layout)
  double *P_out;
};

struct RowArgs {
  int n_coef;
  const int *off;
  const int *inc_ptr, *inc_ea;
  const uint16_t *slot;
  const double *mass;
  const double *rec, *gradN, *detJ, *fe;
  QuadW qw;
  double h, inv_h;
  double c1, c2, c3, c4, mu;  // c1=h*lambda+lambda_d, c2=c3=h*mu+eta, c4=h ; MR: see row_kernel
  double lambda_d, eta;
  double *H;
  double *f_int;
  int row_stride;  // doubles of shared memory per warp
  const int *perm;  // optional row order (Morton tiles): rows that share elements run together
  int chunk;        // consecutive rows (in that order) handled by one block
};

// ---------------------------------------------------------------------------
// materials
// ---------------------------------------------------------------------------
__device__ __forceinline__ double det3(const double F[3][3]) {
  return F[0][0] * (F[1][1] * F[2][2] - F[1][2] * F[2][1]) - F[0][1] * (F[1][0] * F[2][2] - F[1][2] * F[2][0]) +
         F[0][2] * (F[1][0] * F[2][1] - F[1][1] * F[2][0]);
}

struct MRState {
  double C[3][3], FC[3][3], FFT[3][3], Fit[3][3];
  double I1, I2, J, t1, t2, t3;
};

__device__ __forceinline__ void mr_state(const double F[3][3], double mu10, double mu01, double kappa, MRState &m) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double c = 0.0, f = 0.0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        c += F[k][i] * F[k][j];
        f += F[i][k] * F[j][k];
      }
      m.C[i][j] = c;
      m.FFT[i][j] = f;
    }
  m.I1 = m.C[0][0] + m.C[1][1] + m.C[2][2];
  double trC2 = 0.0;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int k = 0; k < 3; k++) trC2 += m.C[i][k] * m.C[k][i];
  m.I2 = 0.5 * (m.I1 * m.I1 - trC2);
  m.J = det3(F);
  double sd = m.J;
  if (fabs(sd) < 1e-12) sd = (sd >= 0.0) ? 1e-12 : -1e-12;  // MooneyRivlin.cuh:25-31
  double id = 1.0 / sd;
  m.Fit[0][0] = (F[1][1] * F[2][2] - F[1][2] * F[2][1]) * id;
  m.Fit[0][1] = (F[1][2] * F[2][0] - F[1][0] * F[2][2]) * id;
  m.Fit[0][2] = (F[1][0] * F[2][1] - F[1][1] * F[2][0]) * id;
  m.Fit[1][0] = (F[0][2] * F[2][1] - F[0][1] * F[2][2]) * id;
  m.Fit[1][1] = (F[0][0] * F[2][2] - F[0][2] * F[2][0]) * id;
  m.Fit[1][2] = (F[0][1] * F[2][0] - F[0][0] * F[2][1]) * id;
  m.Fit[2][0] = (F[0][1] * F[1][2] - F[0][2] * F[1][1]) * id;
  m.Fit[2][1] = (F[0][2] * F[1][0] - F[0][0] * F[1][2]) * id;
  m.Fit[2][2] = (F[0][0] * F[1][1] - F[0][1] * F[1][0]) * id;
  double J13 = cbrt(m.J);
  double Jm23 = 1.0 / (J13 * J13);
  double Jm43 = Jm23 * Jm23;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 3; k++) s += F[i][k] * m.C[k][j];
      m.FC[i][j] = s;
    }
  m.t1 = 2.0 * mu10 * Jm23;
  m.t2 = 2.0 * mu01 * Jm43;
  m.t3 = kappa * (m.J - 1.0) * m.J;
}

// ---------------------------------------------------------------------------
// (1) Gauss-point kernel
// ---------------------------------------------------------------------------
template <int S, int Q>
struct GPShape {
  static constexpr int EPB = (S == 10) ? 32 : (S == 8 ? 16 : 4);  // elements per block
  static constexpr int THREADS = EPB * Q;
  static constexpr size_t SMEM = (size_t)EPB * Q * 3 * S * sizeof(double);  // gradN tile, then force partials
};

template <int S, int Q, int MODEL, bool DAMP>
__global__ void __launch_bounds__(GPShape<S, Q>::THREADS) gp_kernel(GPArgs p) {
  constexpr int EPB = GPShape<S, Q>::EPB;
  extern __shared__ double part[];  // [EPB][Q][3S]: gradN of the block's elements, later the force partials
  const int el = threadIdx.x / Q, q = threadIdx.x % Q;
  const int e = blockIdx.x * EPB + el;
  const bool active = e < p.n_elem;
  {
    // coalesced copy of the block's contiguous gradN range into shared memory
    const int64_t base = (int64_t)blockIdx.x * EPB * Q * 3 * S;
    const int64_t n_here = min((int64_t)EPB, (int64_t)p.n_elem - (int64_t)blockIdx.x * EPB) * Q * 3 * S;
    const double2 *src = reinterpret_cast<const double2 *>(p.gradN + base);
    double2 *dst = reinterpret_cast<double2 *>(part);
    for (int64_t i = threadIdx.x; i < n_here / 2; i += GPShape<S, Q>::THREADS) dst[i] = src[i];
  }
  __syncthreads();
  double hreg[3 * S];
  if (active) {
    const double *gs = part + (size_t)(el * Q + q) * 3 * S;
#pragma unroll
    for (int i = 0; i < 3 * S; i++) hreg[i] = gs[i];
  }
  __syncthreads();  // part is reused for the force partials below
  if (active) {
    const int64_t t = (int64_t)e * Q + q;
    const int *ec = p.ecoef + (int64_t)e * S;
    const double *g = hreg;
    double F[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
    double Fd[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
#pragma unroll
    for (int a = 0; a < S; a++) {
      const int c = ec[a];
      const double xa[3] = {p.x[c], p.y[c], p.z[c]};
      const double ha[3] = {g[a], g[S + a], g[2 * S + a]};
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) F[i][j] += xa[i] * ha[j];
      if (DAMP) {
        const double va[3] = {p.v[3 * (int64_t)c], p.v[3 * (int64_t)c + 1], p.v[3 * (int64_t)c + 2]};
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) Fd[i][j] += va[i] * ha[j];
      }
    }
    const double dV = p.detJ[t] * p.qw.w[q];

    double FFT[3][3], FtF_tr = 0.0;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) {
        FFT[i][j] = F[i][0] * F[j][0] + F[i][1] * F[j][1] + F[i][2] * F[j][2];
        FtF_tr += F[i][j] * F[i][j];
      }
    const double trE = 0.5 * (FtF_tr - 3.0);

    double P[3][3];
    if (MODEL == TL_MODEL_SVK) {
      const double lf = p.lambda * (0.5 * FtF_tr - 1.5);
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          double FFtF = FFT[i][0] * F[0][j] + FFT[i][1] * F[1][j] + FFT[i][2] * F[2][j];
          P[i][j] = lf * F[i][j] + p.mu * (FFtF - F[i][j]);
        }
      if (p.rec) {
        double *r = p.rec + t * REC_SVK;
        double2 *r2 = reinterpret_cast<double2 *>(r);
        r2[0] = make_double2(F[0][0], F[0][1]);
        r2[1] = make_double2(F[0][2], F[1][0]);
        r2[2] = make_double2(F[1][1], F[1][2]);
        r2[3] = make_double2(F[2][0], F[2][1]);
        r2[4] = make_double2(F[2][2], dV * FFT[0][0]);
        r2[5] = make_double2(dV * FFT[0][1], dV * FFT[0][2]);
        r2[6] = make_double2(dV * FFT[1][1], dV * FFT[1][2]);
        r2[7] = make_double2(dV * FFT[2][2], dV * (p.lambda * trE - p.mu));
      }
    } else {
      MRState m;
      mr_state(F, p.mu10, p.mu01, p.kappa, m);
      double term1[3][3], term2[3][3];
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          term1[i][j] = F[i][j] - (m.I1 / 3.0) * m.Fit[i][j];
          term2[i][j] = m.I1 * F[i][j] - m.FC[i][j] - (2.0 * m.I2 / 3.0) * m.Fit[i][j];
          P[i][j] = m.t1 * term1[i][j] + m.t2 * term2[i][j] + m.t3 * m.Fit[i][j];
        }
      if (p.rec) {
        double *r = p.rec + t * REC_MR;
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) r[i * 3 + j] = F[i][j];
        // A_{iJkL} = dP_iJ/dF_kL  (MooneyRivlin.cuh:193-223), stored dV-scaled as [i][J][k][L]
        for (int i = 0; i < 3; i++)
          for (int j = 0; j < 3; j++)
#pragma unroll
            for (int k = 0; k < 3; k++)
#pragma unroll
              for (int l = 0; l < 3; l++) {
                const double dik = (i == k) ? 1.0 : 0.0, djl = (j == l) ? 1.0 : 0.0;
                const double dFit = -m.Fit[i][l] * m.Fit[k][j];
                const double dt1 = (-2.0 / 3.0) * m.t1 * m.Fit[k][l];
                const double dt2 = (-4.0 / 3.0) * m.t2 * m.Fit[k][l];
                const double dt3 = (p.kappa * (2.0 * m.J - 1.0) * m.J) * m.Fit[k][l];
                const double dT1 = dik * djl - (2.0 / 3.0) * F[k][l] * m.Fit[i][j] +
                                   (m.I1 / 3.0) * m.Fit[i][l] * m.Fit[k][j];
                const double dT2 = 2.0 * F[k][l] * F[i][j] + m.I1 * dik * djl -
                                   (dik * m.C[l][j] + F[i][l] * F[k][j] + djl * m.FFT[i][k]) -
                                   (4.0 / 3.0) * (m.I1 * F[k][l] - m.FC[k][l]) * m.Fit[i][j] +
                                   (2.0 * m.I2 / 3.0) * m.Fit[i][l] * m.Fit[k][j];
                r[9 + ((i * 3 + j) * 3 + k) * 3 + l] =
                    dV * (dt1 * term1[i][j] + m.t1 * dT1 + dt2 * term2[i][j] + m.t2 * dT2 + dt3 * m.Fit[i][j] +
                          m.t3 * dFit);
              }
        r[90] = dV * FFT[0][0];
        r[91] = dV * FFT[0][1];
        r[92] = dV * FFT[0][2];
        r[93] = dV * FFT[1][1];
        r[94] = dV * FFT[1][2];
        r[95] = dV * FFT[2][2];
      }
    }
    if (DAMP) {
      // P_vis = F (2 eta Edot + lambda_d tr(Edot) I), Edot = sym(Fdot^T F)   FEAT10DataFunc.cuh:139-223
      double Ed[3][3];
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          double a1 = 0.0, a2 = 0.0;
#pragma unroll
          for (int k = 0; k < 3; k++) {
            a1 += Fd[k][i] * F[k][j];
            a2 += F[k][i] * Fd[k][j];
          }
          Ed[i][j] = 0.5 * (a1 + a2);
        }
      const double trEd = Ed[0][0] + Ed[1][1] + Ed[2][2];
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          double s = 0.0;
#pragma unroll
          for (int k = 0; k < 3; k++)
            s += F[i][k] * (2.0 * p.eta * Ed[k][j] + ((k == j) ? p.lambda_d * trEd : 0.0));
          P[i][j] += s;
        }
    }
    if (p.F_out) {
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          p.F_out[t * 9 + j * 3 + i] = F[i][j];
          p.P_out[t * 9 + j * 3 + i] = P[i][j];
        }
    }
    if (p.fe) {
      double *pt = part + (size_t)(el * Q + q) * 3 * S;
#pragma unroll
      for (int a = 0; a < S; a++) {
        const double ha[3] = {g[a], g[S + a], g[2 * S + a]};
#pragma unroll
        for (int i = 0; i < 3; i++) pt[a * 3 + i] = (P[i][0] * ha[0] + P[i][1] * ha[1] + P[i][2] * ha[2]) * dV;
      }
    }
  }
  if (p.fe) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < EPB * 3 * S; idx += GPShape<S, Q>::THREADS) {
      const int l = idx / (3 * S), c = idx % (3 * S);
      const int ee = blockIdx.x * EPB + l;
      if (ee >= p.n_elem) continue;
      double s = 0.0;
      for (int qq = 0; qq < Q; qq++) s += part[(size_t)(l * Q + qq) * 3 * S + c];  // fixed q order
      p.fe[(int64_t)ee * 3 * S + c] = s;
    }
  }
}

// ---------------------------------------------------------------------------
// (2) row-gather assembly kernel
// ---------------------------------------------------------------------------
constexpr int ROW_WARPS = 8;

template <int S, int Q, int MODEL, bool DAMP>
__global__ void __launch_bounds__(ROW_WARPS * 32) row_kernel(RowArgs p) {
  constexpr int VPB = 32 / S;  // visits handled at once by one warp
  extern __shared__ double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *rowbuf = smem + (size_t)warp * p.row_stride;
  const int k = lane / S, b = lane % S;
  const bool lane_on = lane < VPB * S;

  const int n_chunks = (p.n_coef + p.chunk - 1) / p.chunk;
  for (int ch = blockIdx.x; ch < n_chunks; ch += gridDim.x)
  for (int t8 = 0; t8 < p.chunk; t8 += ROW_WARPS) {
    // rotate the warp<->row assignment: in Morton order every 8th row is a hub (vertex) row
    const int ri = ch * p.chunk + t8 + ((warp + ROW_WARPS - (t8 / ROW_WARPS) % ROW_WARPS) % ROW_WARPS);
    if (ri >= p.n_coef || ri >= (ch + 1) * p.chunk) continue;
    const int row = p.perm ? p.perm[ri] : ri;
    const int rs = p.off[row];
    const int deg = p.off[row + 1] - rs;
    // H row starts from the mass term: H(3i+d,3j+d) = M_ij / h   (SyncedNewton.cu:213-259)
    for (int idx = lane; idx < deg * 9; idx += 32) {
      const int s = idx / 9, c = idx - s * 9;
      rowbuf[idx] = (c == 0 || c == 4 || c == 8) ? p.mass[rs + s] * p.inv_h : 0.0;
    }
    const int vb = p.inc_ptr[row], ve = p.inc_ptr[row + 1];
    // internal force of this coefficient: fixed (ascending element) order
    if (lane < 3) {
      double f = 0.0;
      for (int v = vb; v < ve; v++) f += p.fe[(int64_t)p.inc_ea[v] * 3 + lane];
      p.f_int[(int64_t)row * 3 + lane] = f;
    }
    __syncwarp();

    for (int v0 = vb; v0 < ve; v0 += VPB) {
      const int v = v0 + k;
      const bool valid = lane_on && (v < ve);
      double val[9];
      int slot = 0;
      if (valid) {
        const int ea = p.inc_ea[v];
        const int e = ea / S, a = ea - e * S;
        slot = p.slot[(int64_t)v * S + b];
        double A[9], G[6], sacc = 0.0;
#pragma unroll
        for (int i = 0; i < 9; i++) A[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; i++) G[i] = 0.0;
        double Km[9];
        if (MODEL == TL_MODEL_MR) {
#pragma unroll
          for (int i = 0; i < 9; i++) Km[i] = 0.0;
        }
        const int64_t t0 = (int64_t)e * Q;
#pragma unroll(Q <= 12 ? Q : 4)
        for (int q = 0; q < Q; q++) {
          const int64_t t = t0 + q;
          const double *g = p.gradN + t * 3 * S;
          const double hb0 = g[b], hb1 = g[S + b], hb2 = g[2 * S + b];
          const double ha0 = g[a], ha1 = g[S + a], ha2 = g[2 * S + a];
          const double hab = ha0 * hb0 + ha1 * hb1 + ha2 * hb2;
          if (MODEL == TL_MODEL_SVK) {
            const double2 *r2 = reinterpret_cast<const double2 *>(p.rec + t * REC_SVK);
            const double2 r0 = r2[0], r1 = r2[1], r2_ = r2[2], r3 = r2[3], r4 = r2[4], r5 = r2[5], r6 = r2[6],
                          r7 = r2[7];
            const double dV = p.detJ[t] * p.qw.w[q];
            const double ub0 = r0.x * hb0 + r0.y * hb1 + r1.x * hb2;
            const double ub1 = r1.y * hb0 + r2_.x * hb1 + r2_.y * hb2;
            const double ub2 = r3.x * hb0 + r3.y * hb1 + r4.x * hb2;
            const double ua0 = (r0.x * ha0 + r0.y * ha1 + r1.x * ha2) * dV;
            const double ua1 = (r1.y * ha0 + r2_.x * ha1 + r2_.y * ha2) * dV;
            const double ua2 = (r3.x * ha0 + r3.y * ha1 + r4.x * ha2) * dV;
            A[0] += ua0 * ub0; A[1] += ua0 * ub1; A[2] += ua0 * ub2;
            A[3] += ua1 * ub0; A[4] += ua1 * ub1; A[5] += ua1 * ub2;
            A[6] += ua2 * ub0; A[7] += ua2 * ub1; A[8] += ua2 * ub2;
            G[0] += hab * r4.y; G[1] += hab * r5.x; G[2] += hab * r5.y;
            G[3] += hab * r6.x; G[4] += hab * r6.y; G[5] += hab * r7.x;
            sacc += hab * r7.y;
          } else {
            const double *r = p.rec + t * REC_MR;
            const double w[9] = {ha0 * hb0, ha0 * hb1, ha0 * hb2, ha1 * hb0, ha1 * hb1,
                                 ha1 * hb2, ha2 * hb0, ha2 * hb1, ha2 * hb2};  // w[J*3+L]
            // K[d][e] += sum_JL (dV A[d][J][e][L]) h_a[J] h_b[L]   FEAT10DataFunc.cuh:628-643
#pragma unroll
            for (int d = 0; d < 3; d++)
#pragma unroll
              for (int e2 = 0; e2 < 3; e2++) {
                double s = Km[d * 3 + e2];
#pragma unroll
                for (int J = 0; J < 3; J++)
#pragma unroll
                  for (int L = 0; L < 3; L++) s += r[9 + ((d * 3 + J) * 3 + e2) * 3 + L] * w[J * 3 + L];
                Km[d * 3 + e2] = s;
              }
            if (DAMP) {
              const double dV = p.detJ[t] * p.qw.w[q];
              const double ub0 = r[0] * hb0 + r[1] * hb1 + r[2] * hb2;
              const double ub1 = r[3] * hb0 + r[4] * hb1 + r[5] * hb2;
              const double ub2 = r[6] * hb0 + r[7] * hb1 + r[8] * hb2;
              const double ua0 = (r[0] * ha0 + r[1] * ha1 + r[2] * ha2) * dV;
              const double ua1 = (r[3] * ha0 + r[4] * ha1 + r[5] * ha2) * dV;
              const double ua2 = (r[6] * ha0 + r[7] * ha1 + r[8] * ha2) * dV;
              A[0] += ua0 * ub0; A[1] += ua0 * ub1; A[2] += ua0 * ub2;
              A[3] += ua1 * ub0; A[4] += ua1 * ub1; A[5] += ua1 * ub2;
              A[6] += ua2 * ub0; A[7] += ua2 * ub1; A[8] += ua2 * ub2;
              G[0] += hab * r[90]; G[1] += hab * r[91]; G[2] += hab * r[92];
              G[3] += hab * r[93]; G[4] += hab * r[94]; G[5] += hab * r[95];
            }
          }
        }
        if (MODEL == TL_MODEL_SVK) {
          // h*K_ab + C_ab = c1*A + c2*A^T + c3*G + c4*(sacc + mu*tr A) I
          //   SVK.cuh:35-55 terms A..F and the Kelvin-Voigt block FEAT10DataFunc.cuh:695-762
          const double dg = p.c4 * (sacc + p.mu * (A[0] + A[4] + A[8]));
          val[0] = p.c1 * A[0] + p.c2 * A[0] + p.c3 * G[0] + dg;
          val[1] = p.c1 * A[1] + p.c2 * A[3] + p.c3 * G[1];
          val[2] = p.c1 * A[2] + p.c2 * A[6] + p.c3 * G[2];
          val[3] = p.c1 * A[3] + p.c2 * A[1] + p.c3 * G[1];
          val[4] = p.c1 * A[4] + p.c2 * A[4] + p.c3 * G[3] + dg;
          val[5] = p.c1 * A[5] + p.c2 * A[7] + p.c3 * G[4];
          val[6] = p.c1 * A[6] + p.c2 * A[2] + p.c3 * G[2];
          val[7] = p.c1 * A[7] + p.c2 * A[5] + p.c3 * G[4];
          val[8] = p.c1 * A[8] + p.c2 * A[8] + p.c3 * G[5] + dg;
        } else {
#pragma unroll
          for (int i = 0; i < 9; i++) val[i] = p.h * Km[i];
          if (DAMP) {
            val[0] += p.lambda_d * A[0] + p.eta * A[0] + p.eta * G[0];
            val[1] += p.lambda_d * A[1] + p.eta * A[3] + p.eta * G[1];
            val[2] += p.lambda_d * A[2] + p.eta * A[6] + p.eta * G[2];
            val[3] += p.lambda_d * A[3] + p.eta * A[1] + p.eta * G[1];
            val[4] += p.lambda_d * A[4] + p.eta * A[4] + p.eta * G[3];
            val[5] += p.lambda_d * A[5] + p.eta * A[7] + p.eta * G[4];
            val[6] += p.lambda_d * A[6] + p.eta * A[2] + p.eta * G[2];
            val[7] += p.lambda_d * A[7] + p.eta * A[5] + p.eta * G[4];
            val[8] += p.lambda_d * A[8] + p.eta * A[8] + p.eta * G[5];
          }
        }
      }
      // ordered accumulation: visit k=0 first, then 1, ...  Lanes of one visit hit distinct slots.
#pragma unroll
      for (int kk = 0; kk < VPB; kk++) {
        if (valid && k == kk) {
          double *dst = rowbuf + slot * 9;
#pragma unroll
          for (int i = 0; i < 9; i++) dst[i] += val[i];
        }
        __syncwarp();
      }
    }
    // stream the finished block row out as three scalar CSR rows of length 3*deg
    const int64_t base = 9 * (int64_t)rs;
    const int rowlen = 3 * deg;
    for (int idx = lane; idx < 3 * rowlen; idx += 32) {
      const int d = idx / rowlen, c = idx - d * rowlen;  // c = 3*s + e
      const int s = c / 3, e2 = c - 3 * s;
      p.H[base + idx] = rowbuf[s * 9 + d * 3 + e2];
    }
    __syncwarp();
  }
}

// force-only gather (what the gradient-only solvers need)
static __global__ void force_gather_kernel(int n_coef, const int *__restrict__ inc_ptr, const int *__restrict__ inc_ea,
                                    const double *__restrict__ fe, double *f_int, int accumulate) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_coef * 3) return;
  const int row = (int)(tid / 3), d = (int)(tid % 3);
  double f = accumulate ? f_int[tid] : 0.0;
  for (int v = inc_ptr[row]; v < inc_ptr[row + 1]; v++) f += fe[(int64_t)inc_ea[v] * 3 + d];
  f_int[tid] = f;
}

// element force from STORED P (CalcInternalForce semantics, FEAT10Data.cu:314-333)
template <int S, int Q>
__global__ void fe_from_P_kernel(int n_elem, const double *__restrict__ P, const double *__restrict__ gradN,
                                 const double *__restrict__ detJ, QuadW qw, double *fe) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n_elem * S) return;
  const int e = (int)(tid / S), a = (int)(tid % S);
  double f[3] = {0.0, 0.0, 0.0};
  for (int q = 0; q < Q; q++) {
    const int64_t t = (int64_t)e * Q + q;
    const double *Pq = P + t * 9;
    const double *g = gradN + t * 3 * S;
    const double dV = detJ[t] * qw.w[q];
#pragma unroll
    for (int i = 0; i < 3; i++) f[i] += (Pq[i] * g[a] + Pq[3 + i] * g[S + a] + Pq[6 + i] * g[2 * S + a]) * dV;
  }
#pragma unroll
  for (int i = 0; i < 3; i++) fe[tid * 3 + i] = f[i];
}

}  // namespace tlfea
