// ancf_ek.cuh -- phase E of the two-phase deterministic assembly for the ANCF elements (3243: S = 8, Q = 12;
// 3443: S = 16, Q = 48), St. Venant-Kirchhoff with optional Kelvin-Voigt damping.
//
// Reference behaviour replaced (file:line relative to the reference root):
//   compute_p                 ANCF3243DataFunc.cuh:189-412   ANCF3443DataFunc.cuh:281-495
//   compute_internal_force    ANCF3243DataFunc.cuh:414-454   ANCF3443DataFunc.cuh:498-535   (atomicAdd per component)
//   compute_hessian_assemble  ANCF3243DataFunc.cuh:642-940   ANCF3443DataFunc.cuh:724-1033  (K_elem[3S][3S] in local
//                             memory per Gauss-point thread, S*S*9 searches + atomics per Gauss point)
//
// Same algebra as the FEAT10 element kernel (ek_kernel.cuh): with u_a = F h_a per Gauss point,
//   h K_ab + C_ab = c1 A + c2 A^T + c4 mu tr(A) I + G,  A = sum_q dV u_a u_b^T,  G = sum_q (h_a.h_b) M_q,
//   M_q = dV[(h mu + eta) F F^T + h (lambda trE - mu) I],
// i.e. 18 DFMA per (pair, point) and only the S(S+1)/2 pairs a <= b are evaluated (K_ba = K_ab^T).
// One CTA works on EPB elements at a time: the elements' gradN rows are staged in shared memory once and reused by
// the Gauss-point step (thread = (element, q)), the force step (thread = (element, a, i)) and the block step
// (thread = (element, pair)); finished blocks are staged as [a][b][9] and leave as one contiguous burst per element
// -- the visit-major record the row-reduce kernel (rk_kernel) consumes.  No atomics, fixed summation order.
#pragma once
#include "common.cuh"

namespace tlfea {

struct AEKArgs {
  int n_elem;
  const int *ecoef;              // [n_elem][S]
  const double *x, *y, *z;       // coefficient SoA
  const double *v;               // interleaved velocity or nullptr
  const double *gradN, *detJ;    // [n_elem][Q][3][S], [n_elem][Q]
  QuadW qw;
  double lambda, mu, eta, lambda_d;
  double c1, c2, c4;             // c1 = h lambda + lambda_d, c2 = h mu + eta, c4 = h
  double *fe;                    // [n_elem][S][3]
  double *stage;                 // [n_elem][S][S][9] or nullptr (force only)
};

template <int S, int Q, int EPB, bool DAMP>
struct AEKShape {
  static constexpr int NPAIR = S * (S + 1) / 2;
  static constexpr int HS = 3 * S + 1;           // padded stride of a Gauss point's gradN / u row (odd: lanes = q)
  static constexpr int RS = 9 * S + 1;           // padded stride of a staged block row (odd: lanes = b)
  static constexpr int WORK = (EPB * NPAIR > EPB * 3 * S ? EPB * NPAIR : EPB * 3 * S);
  static constexpr int THREADS = ((WORK > EPB * Q ? WORK : EPB * Q) + 31) / 32 * 32;
  static constexpr int H_D = Q * HS;             // per element
  static constexpr int STG_D = S * RS;           // per element; aliases gradN + u after the block step
  static_assert(STG_D <= 2 * H_D, "block staging must fit in the gradN + u region");
  static constexpr int M_D = Q * 8, P_D = Q * 9, X_D = S * 3;  // M_q padded to 8: four 16-byte loads in the block step
  static constexpr int H2_D = (2 * H_D + 1) / 2 * 2;            // keeps M 16-byte aligned
  // odd element stride: the lanes of a warp that belong to different elements must not share banks
  static constexpr int ELEM_D = (EPB > 1) ? ((H2_D + M_D + P_D + X_D * (DAMP ? 2 : 1) + 1) | 3) - 1 : H2_D + M_D + P_D + X_D * (DAMP ? 2 : 1);
  static constexpr size_t SMEM = (size_t)EPB * ELEM_D * sizeof(double);
};

template <int S, int Q, int EPB, bool DAMP, bool TANGENT>
__global__ void __launch_bounds__(AEKShape<S, Q, EPB, DAMP>::THREADS, S == 16 ? 5 : 4) aek_kernel(AEKArgs p) {
  using Sh = AEKShape<S, Q, EPB, DAMP>;
  constexpr int T = Sh::THREADS, HS = Sh::HS, RS = Sh::RS, NP = Sh::NPAIR;
  extern __shared__ __align__(16) double aek_smem[];
  const int tid = threadIdx.x;
  // pair index -> (a, b), a <= b, row-major upper triangle (fixed per thread)
  int pa = 0, pb = 0, pel = tid / NP;
  {
    int r = tid - pel * NP, a = 0;
    while (r >= S - a) {
      r -= S - a;
      a++;
    }
    pa = a;
    pb = a + r;
  }
  const int n_groups = (p.n_elem + EPB - 1) / EPB;
  for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
    const int e0 = grp * EPB;
    const int ne = min(EPB, p.n_elem - e0);
    __syncthreads();  // the previous group's staging (aliasing gradN / u) has been copied out
    // ---- stage gradN (padded rows) and the gathered coefficients
    {
      const double *src = p.gradN + (int64_t)e0 * Q * 3 * S;
      for (int i = tid; i < ne * Q * 3 * S; i += T) {
        const int el = i / (Q * 3 * S), r = i - el * (Q * 3 * S), q = r / (3 * S), c = r - q * 3 * S;
        aek_smem[el * Sh::ELEM_D + q * HS + c] = src[i];
      }
      for (int i = tid; i < ne * S; i += T) {
        const int el = i / S, a = i - el * S;
        const int c = p.ecoef[(int64_t)(e0 + el) * S + a];
        double *sx = aek_smem + el * Sh::ELEM_D + Sh::H2_D + Sh::M_D + Sh::P_D;
        sx[a * 3 + 0] = p.x[c];
        sx[a * 3 + 1] = p.y[c];
        sx[a * 3 + 2] = p.z[c];
        if (DAMP) {
          sx[Sh::X_D + a * 3 + 0] = p.v[3 * (int64_t)c + 0];
          sx[Sh::X_D + a * 3 + 1] = p.v[3 * (int64_t)c + 1];
          sx[Sh::X_D + a * 3 + 2] = p.v[3 * (int64_t)c + 2];
        }
      }
    }
    __syncthreads();
    // ---- Gauss-point step: thread = (element, q)
    if (tid < ne * Q) {
      const int el = tid / Q, q = tid - el * Q;
      double *se = aek_smem + el * Sh::ELEM_D;
      const double *hq = se + q * HS;
      double *uq = se + Sh::H_D + q * HS;
      const double *sx = se + Sh::H2_D + Sh::M_D + Sh::P_D;
      double F[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
      double Fd[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
#pragma unroll 4
      for (int a = 0; a < S; a++) {
        const double ha[3] = {hq[a], hq[S + a], hq[2 * S + a]};
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) F[i][j] += sx[a * 3 + i] * ha[j];
        if (DAMP) {
#pragma unroll
          for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) Fd[i][j] += sx[Sh::X_D + a * 3 + i] * ha[j];
        }
      }
      const double dV = p.detJ[(int64_t)(e0 + el) * Q + q] * p.qw.w[q];
      double FFT[3][3], trFtF = 0.0;
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          FFT[i][j] = F[i][0] * F[j][0] + F[i][1] * F[j][1] + F[i][2] * F[j][2];
          trFtF += F[i][j] * F[i][j];
        }
      const double trE = 0.5 * (trFtF - 3.0);
      const double lf = p.lambda * (0.5 * trFtF - 1.5);
      double P[3][3];
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          const double FFtF = FFT[i][0] * F[0][j] + FFT[i][1] * F[1][j] + FFT[i][2] * F[2][j];
          P[i][j] = lf * F[i][j] + p.mu * (FFtF - F[i][j]);
        }
      if (DAMP) {  // P_vis = F (2 eta Edot + lambda_d tr(Edot) I), Edot = sym(Fdot^T F)
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
            double sv = 0.0;
#pragma unroll
            for (int k = 0; k < 3; k++)
              sv += F[i][k] * (2.0 * p.eta * Ed[k][j] + ((k == j) ? p.lambda_d * trEd : 0.0));
            P[i][j] += sv;
          }
      }
      double *sP = se + Sh::H2_D + Sh::M_D + q * 9;
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) sP[i * 3 + j] = P[i][j] * dV;
      if (TANGENT) {
#pragma unroll 4
        for (int a = 0; a < S; a++) {
          const double ha[3] = {hq[a], hq[S + a], hq[2 * S + a]};
#pragma unroll
          for (int i = 0; i < 3; i++) uq[i * S + a] = F[i][0] * ha[0] + F[i][1] * ha[1] + F[i][2] * ha[2];
        }
        double *sM = se + Sh::H2_D + q * 8;
        const double sI = p.c4 * dV * (p.lambda * trE - p.mu);
        const double cf = p.c2 * dV;
        sM[0] = cf * FFT[0][0] + sI;
        sM[1] = cf * FFT[0][1];
        sM[2] = cf * FFT[0][2];
        sM[3] = cf * FFT[1][1] + sI;
        sM[4] = cf * FFT[1][2];
        sM[5] = cf * FFT[2][2] + sI;
        sM[6] = dV;
      }
    }
    __syncthreads();
    // ---- element forces: thread = (element, a, i), points in ascending order
    if (tid < ne * 3 * S) {
      const int el = tid / (3 * S), r = tid - el * 3 * S, a = r / 3, i = r - 3 * a;
      const double *se = aek_smem + el * Sh::ELEM_D;
      const double *sP = se + Sh::H2_D + Sh::M_D;
      double f = 0.0;
      for (int q = 0; q < Q; q++) {
        const double *hq = se + q * HS;
        f += sP[q * 9 + i * 3] * hq[a] + sP[q * 9 + i * 3 + 1] * hq[S + a] + sP[q * 9 + i * 3 + 2] * hq[2 * S + a];
      }
      p.fe[((int64_t)(e0 + el) * S + a) * 3 + i] = f;
    }
    if (TANGENT) {
      // ---- block step: thread = (element, pair a <= b)
      double k[9];
      const bool on = pel < ne;
      if (on) {
        const double *se = aek_smem + pel * Sh::ELEM_D;
        double A[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, G[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll 2
        for (int q = 0; q < Q; q++) {
          const double *hq = se + q * HS, *uq = se + Sh::H_D + q * HS;
          const double2 *m2 = reinterpret_cast<const double2 *>(se + Sh::H2_D + q * 8);  // same address in every lane
          const double2 m01 = m2[0], m23 = m2[1], m45 = m2[2], m67 = m2[3];
          const double mq[6] = {m01.x, m01.y, m23.x, m23.y, m45.x, m45.y};
          const double dV = m67.x;
          const double hab = hq[pa] * hq[pb] + hq[S + pa] * hq[S + pb] + hq[2 * S + pa] * hq[2 * S + pb];
          const double ua[3] = {uq[pa] * dV, uq[S + pa] * dV, uq[2 * S + pa] * dV};
          const double ub[3] = {uq[pb], uq[S + pb], uq[2 * S + pb]};
#pragma unroll
          for (int i = 0; i < 3; i++)
#pragma unroll
            for (int c = 0; c < 3; c++) A[i * 3 + c] += ua[i] * ub[c];
#pragma unroll
          for (int i = 0; i < 6; i++) G[i] += hab * mq[i];
        }
        const double dg = p.c4 * p.mu * (A[0] + A[4] + A[8]);
        k[0] = p.c1 * A[0] + p.c2 * A[0] + G[0] + dg;
        k[1] = p.c1 * A[1] + p.c2 * A[3] + G[1];
        k[2] = p.c1 * A[2] + p.c2 * A[6] + G[2];
        k[3] = p.c1 * A[3] + p.c2 * A[1] + G[1];
        k[4] = p.c1 * A[4] + p.c2 * A[4] + G[3] + dg;
        k[5] = p.c1 * A[5] + p.c2 * A[7] + G[4];
        k[6] = p.c1 * A[6] + p.c2 * A[2] + G[2];
        k[7] = p.c1 * A[7] + p.c2 * A[5] + G[4];
        k[8] = p.c1 * A[8] + p.c2 * A[8] + G[5] + dg;
      }
      __syncthreads();  // every thread is done with gradN / u: the staging may overwrite them
      if (on) {
        double *st = aek_smem + pel * Sh::ELEM_D;
        double *d1 = st + pa * RS + pb * 9;
#pragma unroll
        for (int i = 0; i < 9; i++) d1[i] = k[i];
        if (pa != pb) {
          double *d2 = st + pb * RS + pa * 9;
#pragma unroll
          for (int i = 0; i < 3; i++)
#pragma unroll
            for (int c = 0; c < 3; c++) d2[c * 3 + i] = k[i * 3 + c];
        }
      }
      __syncthreads();
      // ---- records [e][a][b][9] leave as one contiguous burst per group
      double *dst = p.stage + (int64_t)e0 * S * S * 9;
      for (int i = tid; i < ne * S * S * 9; i += T) {
        const int el = i / (S * S * 9), r = i - el * (S * S * 9), a = r / (S * 9), c = r - a * (S * 9);
        dst[i] = aek_smem[el * Sh::ELEM_D + a * RS + c];
      }
    }
  }
}

}  // namespace tlfea
