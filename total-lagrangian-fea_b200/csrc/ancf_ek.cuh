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
//
// Mooney-Rivlin (MODEL = TL_MODEL_MR; MooneyRivlin.cuh:45-225): the Gauss-point step runs on threads (element, q, row i)
// and leaves, per point, P dV and the 81-entry tensor T[i][k][J][L] = h dV dP_iJ/dF_kL with the Kelvin-Voigt block
// folded in (same form as the FEAT10 kernel, ek_kernel.cuh); the block step is the contraction
// K_ab[i][k] = sum_J h_a[J] W_b[i][k][J] with W_b = sum_L T[i][k][J][L] h_b[L] formed once per (point, column b) and
// shared by all rows a: 27 DFMA per (pair, point).
#pragma once
#include "common.cuh"
#include "kernels.cuh"  // MRState / mr_state

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
  double mu10, mu01, kappa;      // Mooney-Rivlin (MODEL == TL_MODEL_MR)
  double *fe;                    // [n_elem][S][3]
  double *stage;                 // [n_elem][S][S][9] or nullptr (force only)
};

template <int S, int Q, int EPB, bool DAMP, int MODEL = TL_MODEL_SVK>
struct AEKShape {
  static constexpr bool MR = (MODEL == TL_MODEL_MR);
  static constexpr int NPAIR = S * (S + 1) / 2;
  static constexpr int HS = 3 * S + 1;           // padded stride of a Gauss point's gradN / u row (odd: lanes = q)
  static constexpr int RS = 9 * S + 1;           // padded stride of a staged block row (odd: lanes = b)
  static constexpr int TS = 90;                  // Mooney-Rivlin: T of one point, 9 groups (i,k) of [J][L] padded to 10
  static constexpr int GPT = MR ? EPB * Q * 3 : EPB * Q;  // threads of the Gauss-point step
  static constexpr int NT2 = (S / 2) * (S / 2 + 1);       // Mooney-Rivlin block step: tiles of two rows (2t, 2t+1) x column b >= 2t
  static constexpr int BLK = MR ? EPB * NT2 : EPB * NPAIR;
  static constexpr int WORK = (BLK > EPB * 3 * S ? BLK : EPB * 3 * S);
  static constexpr int THREADS = ((WORK > GPT ? WORK : GPT) + 31) / 32 * 32;
  static constexpr int H_D = Q * HS;             // per element
  static constexpr int U_D = MR ? Q * TS : Q * HS;  // u rows (SVK) or the tangent tensors (Mooney-Rivlin)
  static constexpr int STG_D = S * RS;           // per element; aliases gradN + u / T after the block step
  static_assert(STG_D <= H_D + U_D, "block staging must fit in the gradN + u region");
  static_assert(H_D % 2 == 0, "the region after gradN must stay 16-byte aligned");
  static constexpr int M_D = Q * 8, P_D = Q * 9, X_D = S * 3;  // M_q padded to 8: four 16-byte loads in the block step
  static constexpr int H2_D = (H_D + U_D + 1) / 2 * 2;          // keeps M 16-byte aligned
  // Mooney-Rivlin block step: W_b = T h_b of QB points at a time, [point][27][b] (b fastest: the pairs of a warp read
  // consecutive addresses), shared by every row a of column b
  static constexpr int QB = 3;
  static_assert(Q % QB == 0, "point batches");
  static constexpr int OFF_W = H2_D + M_D + P_D + X_D * (DAMP ? 2 : 1);
  static constexpr int W_D = MR ? QB * 27 * S : 0;
  // odd element stride: the lanes of a warp that belong to different elements must not share banks
  static constexpr int ELEM_D = (EPB > 1) ? ((OFF_W + W_D + 1) | 3) - 1 : OFF_W + W_D;
  static constexpr size_t SMEM = (size_t)EPB * ELEM_D * sizeof(double);
  static constexpr int MINB = MR ? (S == 16 ? 3 : 4) : (S == 16 ? 5 : 4);
};

template <int S, int Q, int EPB, bool DAMP, bool TANGENT, int MODEL = TL_MODEL_SVK>
__global__ void __launch_bounds__(AEKShape<S, Q, EPB, DAMP, MODEL>::THREADS, AEKShape<S, Q, EPB, DAMP, MODEL>::MINB)
    aek_kernel(AEKArgs p) {
  using Sh = AEKShape<S, Q, EPB, DAMP, MODEL>;
  constexpr bool MR = Sh::MR;
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
  if (MR) {  // thread = (element, row pair t, column b >= 2t), b fastest: the lanes of a warp read consecutive W_b
    pel = tid / Sh::NT2;
    int r = tid - pel * Sh::NT2, t = 0;
    while (r >= S - 2 * t) {
      r -= S - 2 * t;
      t++;
    }
    pa = 2 * t;      // rows pa and pa + 1 (the second one only where pa + 1 <= pb)
    pb = 2 * t + r;
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
    // ---- Gauss-point step, Mooney-Rivlin: thread = (element, q, row i)
    if (MR && tid < ne * Q * 3) {
      const int idx = tid / 3, part = tid - 3 * idx;
      const int el = idx / Q, q = idx - el * Q;
      double *se = aek_smem + el * Sh::ELEM_D;
      const double *hq = se + q * HS;
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
      MRState m;
      mr_state(F, p.mu10, p.mu01, p.kappa, m);
      double term1[3][3], term2[3][3], P[3][3];
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          term1[i][j] = F[i][j] - (m.I1 / 3.0) * m.Fit[i][j];
          term2[i][j] = m.I1 * F[i][j] - m.FC[i][j] - (2.0 * m.I2 / 3.0) * m.Fit[i][j];
          P[i][j] = m.t1 * term1[i][j] + m.t2 * term2[i][j] + m.t3 * m.Fit[i][j];
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
      // this thread's row i = part of everything indexed by i
      double Fi[3], Fiti[3], FFTi[3], t1i[3], t2i[3], dk[3];
      double *sP = se + Sh::H2_D + Sh::M_D + q * 9;
#pragma unroll
      for (int j = 0; j < 3; j++) {
        sP[part * 3 + j] = dV * (part == 0 ? P[0][j] : (part == 1 ? P[1][j] : P[2][j]));
        Fi[j] = part == 0 ? F[0][j] : (part == 1 ? F[1][j] : F[2][j]);
        Fiti[j] = part == 0 ? m.Fit[0][j] : (part == 1 ? m.Fit[1][j] : m.Fit[2][j]);
        FFTi[j] = part == 0 ? m.FFT[0][j] : (part == 1 ? m.FFT[1][j] : m.FFT[2][j]);
        t1i[j] = part == 0 ? term1[0][j] : (part == 1 ? term1[1][j] : term1[2][j]);
        t2i[j] = part == 0 ? term2[0][j] : (part == 1 ? term2[1][j] : term2[2][j]);
        dk[j] = (part == j) ? 1.0 : 0.0;
      }
      if (TANGENT) {
        // T[i][k][J][L] = h dV dP_iJ/dF_kL (MooneyRivlin.cuh:193-223) + dV (lambda_d F_iJ F_kL + eta F_kJ F_iL
        // + eta d_JL (FF^T)_ik)  (Kelvin-Voigt, ANCF3243DataFunc.cuh:843-917), groups (i,k) of 9 padded to 10
        double *sT = se + Sh::H_D + q * Sh::TS + part * 30;
        const double hdV = p.c4 * dV;
        const double kt3 = p.kappa * (2.0 * m.J - 1.0) * m.J;
#pragma unroll
        for (int j = 0; j < 3; j++)
#pragma unroll
          for (int k = 0; k < 3; k++)
#pragma unroll
            for (int l = 0; l < 3; l++) {
              const double djl = (j == l) ? 1.0 : 0.0;
              const double dFit = -Fiti[l] * m.Fit[k][j];
              const double dt1 = (-2.0 / 3.0) * m.t1 * m.Fit[k][l];
              const double dt2 = (-4.0 / 3.0) * m.t2 * m.Fit[k][l];
              const double dt3 = kt3 * m.Fit[k][l];
              const double dT1 = dk[k] * djl - (2.0 / 3.0) * F[k][l] * Fiti[j] + (m.I1 / 3.0) * Fiti[l] * m.Fit[k][j];
              const double dT2 = 2.0 * F[k][l] * Fi[j] + m.I1 * dk[k] * djl -
                                 (dk[k] * m.C[l][j] + Fi[l] * F[k][j] + djl * FFTi[k]) -
                                 (4.0 / 3.0) * (m.I1 * F[k][l] - m.FC[k][l]) * Fiti[j] +
                                 (2.0 * m.I2 / 3.0) * Fiti[l] * m.Fit[k][j];
              double t = hdV * (dt1 * t1i[j] + m.t1 * dT1 + dt2 * t2i[j] + m.t2 * dT2 + dt3 * Fiti[j] + m.t3 * dFit);
              if (DAMP) t += dV * (p.lambda_d * Fi[j] * F[k][l] + p.eta * (F[k][j] * Fi[l] + djl * FFTi[k]));
              sT[k * 10 + j * 3 + l] = t;
            }
      }
    }
    // ---- Gauss-point step, SVK: thread = (element, q)
    if (!MR && tid < ne * Q) {
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
      double k[9], k1[9];  // k1: second row of a Mooney-Rivlin tile
      const bool on = pel < ne;
      const bool on1 = MR && on && pa + 1 <= pb;
      if (MR) {
        // K_ab[i][k] = sum_q sum_J h_a[J] W_b[i][k][J],  W_b[i][k][J] = sum_L T[i][k][J][L] h_b[L]: W_b does not depend
        // on the row a, so it is formed once per (point, b) for QB points at a time (phase 1, thread = (point, group
        // (i,k), b)) and every pair of column b reads it (phase 2, thread = two rows x column b): 27 DFMA per (pair, point)
        // instead of 108
#pragma unroll
        for (int i = 0; i < 9; i++) k[i] = 0.0, k1[i] = 0.0;
        constexpr int QB = Sh::QB, ITEMS = QB * 9 * S;
        const int pa1 = min(pa + 1, S - 1);
#pragma unroll 1
        for (int q0 = 0; q0 < Q; q0 += QB) {
          for (int it = tid; it < ne * ITEMS; it += T) {
            const int el = it / ITEMS, r = it - el * ITEMS, qq = r / (9 * S), r2 = r - qq * 9 * S, g = r2 / S, b = r2 - g * S;
            const double *se = aek_smem + el * Sh::ELEM_D;
            const double *hq = se + (q0 + qq) * HS;
            const double2 *t2 = reinterpret_cast<const double2 *>(se + Sh::H_D + (q0 + qq) * Sh::TS + g * 10);
            const double2 v0 = t2[0], v1 = t2[1], v2 = t2[2], v3 = t2[3], v4 = t2[4];
            const double h0 = hq[b], h1 = hq[S + b], h2 = hq[2 * S + b];
            double *w = aek_smem + el * Sh::ELEM_D + Sh::OFF_W + (qq * 27 + g * 3) * S + b;
            w[0] = v0.x * h0 + v0.y * h1 + v1.x * h2;
            w[S] = v1.y * h0 + v2.x * h1 + v2.y * h2;
            w[2 * S] = v3.x * h0 + v3.y * h1 + v4.x * h2;
          }
          __syncthreads();
          if (on) {
            const double *se = aek_smem + pel * Sh::ELEM_D;
#pragma unroll
            for (int qq = 0; qq < QB; qq++) {
              const double *hq = se + (q0 + qq) * HS;
              const double *w = se + Sh::OFF_W + qq * 27 * S + pb;
              const double ha0 = hq[pa], ha1 = hq[S + pa], ha2 = hq[2 * S + pa];
              const double hc0 = hq[pa1], hc1 = hq[S + pa1], hc2 = hq[2 * S + pa1];
#pragma unroll
              for (int g = 0; g < 9; g++) {  // one read of W_b serves both rows of the tile
                const double w0 = w[(g * 3) * S], w1 = w[(g * 3 + 1) * S], w2 = w[(g * 3 + 2) * S];
                k[g] += ha0 * w0 + ha1 * w1 + ha2 * w2;
                k1[g] += hc0 * w0 + hc1 * w1 + hc2 * w2;
              }
            }
          }
          __syncthreads();  // the next batch overwrites W
        }
      }
      if (on && !MR) {
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
        if (on1) {
          double *e1 = st + (pa + 1) * RS + pb * 9;
#pragma unroll
          for (int i = 0; i < 9; i++) e1[i] = k1[i];
          if (pa + 1 != pb) {
            double *e2 = st + pb * RS + (pa + 1) * 9;
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
              for (int c = 0; c < 3; c++) e2[c * 3 + i] = k1[i * 3 + c];
          }
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
