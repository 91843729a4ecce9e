// ek_kernel.cuh -- two-phase deterministic assembly for FEAT10 (S = 10, Q = 5): SVK or Mooney-Rivlin, each with
// optional Kelvin-Voigt damping.  The row-reduce kernel at the end of the file also serves the ANCF elements.
//
// Reference behaviour replaced (file:line relative to the reference root):
//   compute_p                     FEAT10DataFunc.cuh:85-293
//   compute_internal_force        FEAT10DataFunc.cuh:397-458   (atomicAdd per nodal component)
//   compute_hessian_assemble_csr  FEAT10DataFunc.cuh:512-791   (900 binary searches + 900 FP64 atomics per thread)
//   assemble_sparse_hessian_mass  SyncedNewton.cu:213-259
//   svk_compute_tangent_block     SVK.cuh:35-55
//   mr_compute_P / mr_compute_tangent_tensor  MooneyRivlin.cuh:45-225
//
// Phase E (ek_kernel, element-parallel, FP64-pipe + shared-memory work): ONE WARP PER GROUP OF G ELEMENTS.
//   * gradN (one TMA bulk copy per group, per-warp mbarrier) and detJ / connectivity / gathered nodal
//     coordinates (cp.async) of the NEXT group are in flight into a per-warp double buffer while the current
//     one computes;
//   * Gauss-point step: lane = (element-in-group, q) builds F, P (SVK + Kelvin-Voigt), u_a = F h_a, the
//     force partials P dV h_a and the 6 numbers of
//       M_q = dV [ (h mu + eta) F F^T + h (lambda trE - mu) I ]          (symmetric 3x3)
//     and leaves them in shared memory; the element force is the fixed-order sum of its 5 point partials;
//   * block step: two elements at a time, lane = (element, 2x2 tile (I,J), I <= J, of the symmetric 5x5
//     tile grid of shape-function pairs): per Gauss point the lane loads h, u of its 2+2 shape functions
//     and M_q once and updates 4 blocks
//       A  += (dV u_a) u_b^T   (9 DFMA)      G += (h_a.h_b) M_q   (3 + 6 DFMA)
//     (18 DFMA per (pair, point) instead of the reference's ~170 flop), then
//       K_ab = c1 A + c2 A^T + h mu tr(A) I + G,   K_ba = K_ab^T;
//   * the 100 blocks of an element go to its visit-major record [a][b][9] (7 200 B) through a chunked
//     shared-memory staging area laid out so that tile stores and row reads are bank-conflict free (see
//     ek_chunk_off), so the global write is a sequence of coalesced 720-byte rows and every row visit of
//     phase R reads 720 contiguous bytes; element forces go to fe[e][a][3];
//   * Mooney-Rivlin: the Gauss-point step runs on lanes (element, q, row i) and leaves the 81-entry tensor
//     T = h dV dP/dF (+ the Kelvin-Voigt block in the same form) per point; the block step is the contraction
//     K_ab[i][k] = sum_J h_a[J] (sum_L T[i][J][k][L] h_b[L]).
// Phase R (rk_kernel, row-parallel, HBM work): ONE WARP PER COEFFICIENT ROW of H.  The warp zero-fills the
//   block row in shared memory with M/h on the 3x3 diagonals, adds the row's blocks of every incident element
//   in ASCENDING ELEMENT ORDER (ceil(9S/32) coalesced loads per visit record), sums the nodal force in the
//   same order, and streams the finished 3 scalar CSR rows to H exactly once.
// No atomics, no searches, fixed summation order => bit-reproducible results.
#pragma once
#include "common.cuh"
#include "kernels.cuh"  // MRState / mr_state

namespace tlfea {

// Cache hints on the streams that are touched exactly once (A/B switch: -DTLFEA_HINTS=mask at build time).
//   bit 0: visit records are read by the row-reduce kernel with ld.global.cs (evict-first)
//   bit 1: finished H rows leave the row-reduce kernel with st.global.cs
//   bit 2: the element kernel writes its records with st.global.cs
#ifndef TLFEA_HINTS
#define TLFEA_HINTS 0
#endif
__device__ __forceinline__ double rec_load(const double *p) {
#if TLFEA_HINTS & 1
  return __ldcs(p);
#else
  return *p;
#endif
}
__device__ __forceinline__ void row_store(double *p, double v) {
#if TLFEA_HINTS & 2
  __stcs(p, v);
#else
  *p = v;
#endif
}
__device__ __forceinline__ void rec_store(double *p, double v) {
#if TLFEA_HINTS & 4
  __stcs(p, v);
#else
  *p = v;
#endif
}

template <int S>
struct SymLayout {
  static constexpr int D = S / 2;               // diagonals d = 0..D-1 are stored for every a
  static constexpr int NPAIR = S * D + S / 2;   // + the half diagonal d = D for a < D   (= S(S+1)/2)
  static constexpr int ELEM_DOUBLES = NPAIR * 9;
};

struct EKArgs {
  int n_elem;
  const int *ecoef;
  const double *x, *y, *z;
  const double *v;  // interleaved velocity (Kelvin-Voigt) or nullptr
  const double *gradN, *detJ;
  QuadW qw;
  double lambda, mu, eta, lambda_d;
  double c1, c2, c3, c4;  // c1 = h lambda + lambda_d, c2 = c3 = h mu + eta, c4 = h
  double mu10, mu01, kappa;  // Mooney-Rivlin (MODEL == TL_MODEL_MR)
  double *fe;             // [n_elem][10][3]
  double *stage;          // [n_elem][10][10][9] visit-major block records, or nullptr (force only)
};

#ifndef TL_HAVE_CP_ASYNC
#define TL_HAVE_CP_ASYNC
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }
#endif

// TMA bulk copy (cp.async.bulk) of a contiguous global range into shared memory, completion on an mbarrier.
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(b), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void *smem, const void *gmem, unsigned bytes, uint64_t *bar) {
  const unsigned b = (unsigned)__cvta_generic_to_shared(bar), s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(s),
               "l"(gmem), "r"(bytes), "r"(b)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
  const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(b),
      "r"(parity)
      : "memory");
}

constexpr int EK_REC = 9;     // per Gauss point: M(6) | dV | pad(2); odd stride => conflict-free lane writes
constexpr int EK_GQ = 33;     // stride of the per-(element, point) u / force-partial rows (30 used); = 1 (mod 16)
constexpr int EK_VIS = 90;    // doubles of one visit record: K_ab, b = 0..9
constexpr int EK_ELEM = 900;  // doubles of one element record: [a][b][9]
// Staging of one element's block records on the way to global memory.  A chunk (I, p) holds rows a = 2I, 2I+1 of
// block columns b = 2p, 2p+1 as [a & 1][18 doubles].  Chunk bases are chosen so that, per half warp (the unit the
// shared-memory pipe serves for 64-bit accesses), the 15 tiles of an element store to 15 distinct banks -- direct
// and transposed -- and a row of 90 doubles reads back linearly without conflicts: base(I, p) = 3 I + 2 p (mod 16),
// except chunk (4, 4) which sits one bank lower (found by exhaustive search; 134 wavefronts per element against
// 132 ideal).
constexpr int EK_CRS = 195;                 // stride of a row-pair region (186 used), = 3 (mod 16)
constexpr int EK_CX = 193;                  // offset of the exception chunk (4, 4) in its region
constexpr int EK_ES = 4 * EK_CRS + EK_CX + 36;  // staged element
__host__ __device__ constexpr int ek_chunk_off(int I, int p) {
  return I * EK_CRS + ((I == 4 && p == 4) ? EK_CX : 18 * p + 96 * (p & 1));
}

constexpr int EK_GQM = 35;  // Mooney-Rivlin: stride of the force-partial rows, lane = (element, point, row i); = 3 (mod 16)
constexpr int EK_TS = 83;   // Mooney-Rivlin: stride of the per-point tangent tensor T[j][k][l][i] (81 used); = 3 (mod 16)

template <int G, bool DAMP, int MODEL = TL_MODEL_SVK>
struct EKShape {
  static constexpr int H_D = G * 150;               // gradN of a group
  static constexpr int DJ_D = (G * 5 + 1) / 2 * 2;  // detJ of a group
  static constexpr int X_D = G * 30;                // nodal coordinates [el][a][3]
  static constexpr int EC_D = (G * 5 + 1) / 2 * 2;  // G*10 ints
  static constexpr bool MR = (MODEL == TL_MODEL_MR);
  static constexpr int U_D = MR ? G * 5 * EK_TS : G * 5 * EK_GQ;  // u = F h per (element, point); MR: tensor T
  static constexpr int REC_D = MR ? 0 : G * 5 * EK_REC;
  static constexpr int K_D = 2 * EK_ES;             // chunked staging of the element pair in flight; also the force partials
  static constexpr int OFF_H = 0;
  static constexpr int OFF_DJ = OFF_H + 2 * H_D;
  static constexpr int OFF_X = OFF_DJ + 2 * DJ_D;
  static constexpr int OFF_V = OFF_X + 2 * X_D;
  static constexpr int OFF_EC = OFF_V + (DAMP ? 2 * X_D : 0);
  static constexpr int OFF_U = OFF_EC + 2 * EC_D;
  static constexpr int OFF_REC = OFF_U + (U_D + 1) / 2 * 2;
  static constexpr int OFF_K = OFF_REC + (REC_D + 1) / 2 * 2;
  static constexpr int WARP_D = OFF_K + K_D;
  static_assert(G * 5 * EK_GQM <= K_D, "force partials alias the record staging");
  static constexpr size_t smem_bytes(int warps) { return (size_t)warps * WARP_D * sizeof(double); }
};

template <int G, int W, bool DAMP, bool TANGENT, int MODEL = TL_MODEL_SVK>
__global__ void __launch_bounds__(W * 32, 1) ek_kernel(EKArgs p) {
  using Sh = EKShape<G, DAMP, MODEL>;
  constexpr bool MR = (MODEL == TL_MODEL_MR);
  static_assert(!MR || G == 2, "the Mooney-Rivlin Gauss-point step maps lane = (element, point, row): G * 15 <= 32");
  static_assert(G * 5 <= 32 && G % 2 == 0, "a group is G*5 Gauss-point lanes and G/2 element pairs");
  extern __shared__ __align__(16) double ek_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *wb = ek_smem + (size_t)warp * Sh::WARP_D;
  double *s_u = wb + Sh::OFF_U, *s_rec = wb + Sh::OFF_REC, *s_k = wb + Sh::OFF_K;
  double *s_fp = s_k;  // force partials live in the record staging before the block step needs it
  __shared__ uint64_t s_bar[W][2];  // per warp, per buffer: the bulk copy of gradN has landed
  if (lane == 0) {
    mbar_init(&s_bar[warp][0], 1);
    mbar_init(&s_bar[warp][1], 1);
  }
  __syncwarp();
  const int n_groups = (p.n_elem + G - 1) / G;
  const int gstride = gridDim.x * W;
  const int g_first = blockIdx.x * W + warp;

  // ---- prefetch helpers (per warp; every lane issues its share and commits one group)
  auto issue_ec = [&](int grp, int buf) {
    if (grp >= n_groups) return;
    const int e0 = grp * G;
    const int n_int = min(G, p.n_elem - e0) * 10;
    int *dst = reinterpret_cast<int *>(wb + Sh::OFF_EC + buf * Sh::EC_D);
    const int *src = p.ecoef + (int64_t)e0 * 10;
    for (int i = lane; i < n_int / 2; i += 32) cp_async8(dst + 2 * i, src + 2 * i);
  };
  auto issue_data = [&](int grp, int buf) {
    if (grp >= n_groups) return;
    const int e0 = grp * G;
    const int ne = min(G, p.n_elem - e0);
    double *dh = wb + Sh::OFF_H + buf * Sh::H_D;
    const double *sh = p.gradN + (int64_t)e0 * 150;
    if (lane == 0) bulk_load(dh, sh, ne * 1200u, &s_bar[warp][buf]);  // one TMA bulk copy: gradN of the group
    double *dj = wb + Sh::OFF_DJ + buf * Sh::DJ_D;
    const double *sj = p.detJ + (int64_t)e0 * 5;
    for (int i = lane; i < ne * 5; i += 32) cp_async8(dj + i, sj + i);
    const int *ec = reinterpret_cast<const int *>(wb + Sh::OFF_EC + buf * Sh::EC_D);
    double *dx = wb + Sh::OFF_X + buf * Sh::X_D;
    for (int i = lane; i < ne * 10; i += 32) {
      const int c = ec[i];
      cp_async8(dx + 3 * i + 0, p.x + c);
      cp_async8(dx + 3 * i + 1, p.y + c);
      cp_async8(dx + 3 * i + 2, p.z + c);
      if (DAMP) {
        double *dv = wb + Sh::OFF_V + buf * Sh::X_D;
        cp_async8(dv + 3 * i + 0, p.v + 3 * (int64_t)c + 0);
        cp_async8(dv + 3 * i + 1, p.v + 3 * (int64_t)c + 1);
        cp_async8(dv + 3 * i + 2, p.v + 3 * (int64_t)c + 2);
      }
    }
  };

  // prologue: connectivity of the first group, then its data + connectivity of the second
  issue_ec(g_first, 0);
  cp_async_commit();
  cp_async_wait_all();
  __syncwarp();
  issue_ec(g_first + gstride, 1);
  issue_data(g_first, 0);
  cp_async_commit();

  // block step: lane = (element of the pair, 2x2 tile (I, J), I <= J, of the 5x5 tile grid)
  const int el2 = lane >> 4;  // one element per half warp
  int tile = lane & 15;
  const bool lane_on = tile < 15;
  if (tile > 14) tile = 14;  // lanes 15, 31 shadow the last tile and never store
  int tI, tJ;
  if (tile < 5) {
    tI = 0; tJ = tile;
  } else if (tile < 9) {
    tI = 1; tJ = tile - 4;
  } else if (tile < 12) {
    tI = 2; tJ = tile - 7;
  } else if (tile < 14) {
    tI = 3; tJ = tile - 9;
  } else {
    tI = 4; tJ = 4;
  }
  const int a1 = 2 * tI, b1 = 2 * tJ;
  const int st_d = el2 * EK_ES + ek_chunk_off(tI, tJ);  // staging chunk of the tile, and of its transpose
  const int st_t = el2 * EK_ES + ek_chunk_off(tJ, tI);
  // copy-out: lane reads positions lane, lane + 32, lane + 64 of a 90-double record row
  int cp_off[3], cp_x2;
  {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int pos = lane + 32 * j, pp = pos / 18;
      cp_off[j] = 18 * pp + 96 * (pp & 1) + (pos - 18 * pp);
    }
    cp_x2 = (lane + 64) / 18 == 4 ? EK_CX + (lane + 64 - 72) : cp_off[2];
  }

  int it = 0;
  for (int grp = g_first; grp < n_groups; grp += gstride, it++) {
    const int buf = it & 1;
    cp_async_wait_all();
    mbar_wait(&s_bar[warp][buf], (it >> 1) & 1);
    __syncwarp();
    // next group's data (its connectivity landed with the wait above) and the one after's connectivity
    issue_ec(grp + 2 * gstride, buf);
    issue_data(grp + gstride, buf ^ 1);
    cp_async_commit();

    const int e0 = grp * G;
    const int ne = min(G, p.n_elem - e0);
    const double *s_h = wb + Sh::OFF_H + buf * Sh::H_D;
    const double *s_dj = wb + Sh::OFF_DJ + buf * Sh::DJ_D;
    const double *s_x = wb + Sh::OFF_X + buf * Sh::X_D;

    // ------------------------------------------------------------ Gauss-point step, Mooney-Rivlin
    // lane = (element, point, row i): F and the invariants are computed by the three lanes of a point, the force
    // partials and the tangent tensor are split by row.  T[i][J][k][L] = h dV dP_iJ/dF_kL (MooneyRivlin.cuh:193-223)
    // + the Kelvin-Voigt block (FEAT10DataFunc.cuh:695-762) written in the same form:
    //   lambda_d u_a u_b^T + eta u_b u_a^T + eta (h_a.h_b) F F^T = sum_JL h_a[J] dV (lambda_d F_iJ F_kL + eta F_kJ F_iL
    //   + eta d_JL (FF^T)_ik) h_b[L], so the block step is one contraction K_ab[i][k] = sum_JL h_a[J] T[i][J][k][L] h_b[L].
    if (MR && lane < ne * 15) {
      const int idx = lane / 3, part = lane - 3 * idx;
      const int el = idx / 5, q = idx - el * 5;
      const double *hq = s_h + el * 150 + q * 30;
      const double *xe = s_x + el * 30;
      double F[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
      double Fd[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
#pragma unroll
      for (int aa = 0; aa < 10; aa++) {
        const double xa[3] = {xe[aa * 3], xe[aa * 3 + 1], xe[aa * 3 + 2]};
        const double ha[3] = {hq[aa], hq[10 + aa], hq[20 + aa]};
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) F[i][j] += xa[i] * ha[j];
        if (DAMP) {
          const double *ve = wb + Sh::OFF_V + buf * Sh::X_D + el * 30;
          const double va[3] = {ve[aa * 3], ve[aa * 3 + 1], ve[aa * 3 + 2]};
#pragma unroll
          for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) Fd[i][j] += va[i] * ha[j];
        }
      }
      const double dV = s_dj[idx] * p.qw.w[q];
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
      if (DAMP) {
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
      // this lane's row i = part of everything indexed by i
      double Pi[3], Fi[3], Fiti[3], FFTi[3], t1i[3], t2i[3], dk[3];
#pragma unroll
      for (int j = 0; j < 3; j++) {
        Pi[j] = dV * (part == 0 ? P[0][j] : (part == 1 ? P[1][j] : P[2][j]));
        Fi[j] = part == 0 ? F[0][j] : (part == 1 ? F[1][j] : F[2][j]);
        Fiti[j] = part == 0 ? m.Fit[0][j] : (part == 1 ? m.Fit[1][j] : m.Fit[2][j]);
        FFTi[j] = part == 0 ? m.FFT[0][j] : (part == 1 ? m.FFT[1][j] : m.FFT[2][j]);
        t1i[j] = part == 0 ? term1[0][j] : (part == 1 ? term1[1][j] : term1[2][j]);
        t2i[j] = part == 0 ? term2[0][j] : (part == 1 ? term2[1][j] : term2[2][j]);
        dk[j] = (part == j) ? 1.0 : 0.0;
      }
      double *fpq = s_fp + idx * EK_GQM;
#pragma unroll
      for (int aa = 0; aa < 10; aa++)
        fpq[aa * 3 + part] = Pi[0] * hq[aa] + Pi[1] * hq[10 + aa] + Pi[2] * hq[20 + aa];
      if (TANGENT) {
        double *sT = s_u + idx * EK_TS + part;
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
              sT[((j * 3 + k) * 3 + l) * 3] = t;
            }
      }
    }
    // ------------------------------------------------------------ Gauss-point step, SVK
    if (!MR && lane < ne * 5) {
      const int el = lane / 5, q = lane - el * 5;
      const int rot = lane >> 3;  // lanes 8 apart would hit the same banks: rotate their a order
      const double *hq = s_h + el * 150 + q * 30;
      const double *xe = s_x + el * 30;
      double F[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
      double Fd[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
#pragma unroll
      for (int aa = 0; aa < 10; aa++) {
        int ar = aa + rot;
        if (ar >= 10) ar -= 10;
        const double xa[3] = {xe[ar * 3], xe[ar * 3 + 1], xe[ar * 3 + 2]};
        const double ha[3] = {hq[ar], hq[10 + ar], hq[20 + ar]};
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) F[i][j] += xa[i] * ha[j];
        if (DAMP) {
          const double *ve = wb + Sh::OFF_V + buf * Sh::X_D + el * 30;
          const double va[3] = {ve[ar * 3], ve[ar * 3 + 1], ve[ar * 3 + 2]};
#pragma unroll
          for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) Fd[i][j] += va[i] * ha[j];
        }
      }
      const double dV = s_dj[lane] * p.qw.w[q];
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
      if (DAMP) {
        // P_vis = F (2 eta Edot + lambda_d tr(Edot) I), Edot = sym(Fdot^T F)   FEAT10DataFunc.cuh:139-223
        double Ed[3][3];
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) {
            double s1 = 0.0, s2 = 0.0;
#pragma unroll
            for (int k = 0; k < 3; k++) {
              s1 += Fd[k][i] * F[k][j];
              s2 += F[k][i] * Fd[k][j];
            }
            Ed[i][j] = 0.5 * (s1 + s2);
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
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) P[i][j] *= dV;
      // element-force partials of this point and u_a = F h_a for the block step
      double *fpq = s_fp + lane * EK_GQ;
      double *uq = s_u + lane * EK_GQ;
#pragma unroll
      for (int aa = 0; aa < 10; aa++) {
        int ar = aa + rot;  // same rotation as above keeps the gradN reads and these stores conflict-free
        if (ar >= 10) ar -= 10;
        const double ha[3] = {hq[ar], hq[10 + ar], hq[20 + ar]};
#pragma unroll
        for (int i = 0; i < 3; i++) {
          fpq[ar * 3 + i] = P[i][0] * ha[0] + P[i][1] * ha[1] + P[i][2] * ha[2];
          if (TANGENT) uq[i * 10 + ar] = F[i][0] * ha[0] + F[i][1] * ha[1] + F[i][2] * ha[2];
        }
      }
      if (TANGENT) {
        double *r = s_rec + lane * EK_REC;
        const double sI = p.c4 * dV * (p.lambda * trE - p.mu);
        const double cf = p.c3 * dV;
        r[0] = cf * FFT[0][0] + sI;
        r[1] = cf * FFT[0][1];
        r[2] = cf * FFT[0][2];
        r[3] = cf * FFT[1][1] + sI;
        r[4] = cf * FFT[1][2];
        r[5] = cf * FFT[2][2] + sI;
        r[6] = dV;
      }
    }
    __syncwarp();
    // element forces: sum of the 5 point partials in fixed order, one coalesced 240-byte row per element
    for (int el = 0; el < ne; el++) {
      if (lane < 30) {
        constexpr int FS = MR ? EK_GQM : EK_GQ;
        const double *fq = s_fp + el * 5 * FS + lane;
        p.fe[(int64_t)(e0 + el) * 30 + lane] = (((fq[0] + fq[FS]) + fq[2 * FS]) + fq[3 * FS]) + fq[4 * FS];
      }
    }
    __syncwarp();

    // ------------------------------------------------------------ block step, two elements at a time
    if (TANGENT) {
      for (int pr = 0; pr < ne; pr += 2) {
        const int el = min(pr + el2, ne - 1);  // odd tail: the second half shadows the first and never stores
        const bool on = lane_on && (pr + el2 < ne);
        const double *he = s_h + el * 150;
        const double *ue = s_u + el * 5 * EK_GQ;
        const double *re = s_rec + el * 5 * EK_REC;
        double A[4][9], Gs[4][6];  // SVK accumulators; Mooney-Rivlin accumulates the finished blocks in A
#pragma unroll
        for (int t = 0; t < 4; t++) {
#pragma unroll
          for (int i = 0; i < 9; i++) A[t][i] = 0.0;
#pragma unroll
          for (int i = 0; i < 6; i++) Gs[t][i] = 0.0;
        }
        if (MR) {
          // K_ab[i][k] += sum_J h_a[J] (sum_L T[i][J][k][L] h_b[L]): 3 loads + 10 DFMA per (i, J, k) for the 2x2 tile
          const double *Te = s_u + el * 5 * EK_TS;
#pragma unroll 1
          for (int q = 0; q < 5; q++) {
            const double *hq = he + q * 30, *Tq = Te + q * EK_TS;
            double ha[2][3], hb[2][3];
#pragma unroll
            for (int j = 0; j < 3; j++) {
              ha[0][j] = hq[j * 10 + a1];
              ha[1][j] = hq[j * 10 + a1 + 1];
              hb[0][j] = hq[j * 10 + b1];
              hb[1][j] = hq[j * 10 + b1 + 1];
            }
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
              for (int j = 0; j < 3; j++)
#pragma unroll
                for (int k = 0; k < 3; k++) {
                  const double t0 = Tq[((j * 3 + k) * 3 + 0) * 3 + i], t1 = Tq[((j * 3 + k) * 3 + 1) * 3 + i],
                               t2 = Tq[((j * 3 + k) * 3 + 2) * 3 + i];
                  const double s0 = t0 * hb[0][0] + t1 * hb[0][1] + t2 * hb[0][2];
                  const double s1 = t0 * hb[1][0] + t1 * hb[1][1] + t2 * hb[1][2];
                  A[0][i * 3 + k] += ha[0][j] * s0;
                  A[1][i * 3 + k] += ha[0][j] * s1;
                  A[2][i * 3 + k] += ha[1][j] * s0;
                  A[3][i * 3 + k] += ha[1][j] * s1;
                }
          }
        }
#pragma unroll
        for (int q = 0; q < (MR ? 0 : 5); q++) {
          const double *hq = he + q * 30, *uq = ue + q * EK_GQ, *rq = re + q * EK_REC;
          const double dV = rq[6];
          double ha[2][3], hb[2][3], ua[2][3], ub[2][3], M[6];
#pragma unroll
          for (int j = 0; j < 3; j++) {
            ha[0][j] = hq[j * 10 + a1];
            ha[1][j] = hq[j * 10 + a1 + 1];
            hb[0][j] = hq[j * 10 + b1];
            hb[1][j] = hq[j * 10 + b1 + 1];
            ua[0][j] = uq[j * 10 + a1] * dV;
            ua[1][j] = uq[j * 10 + a1 + 1] * dV;
            ub[0][j] = uq[j * 10 + b1];
            ub[1][j] = uq[j * 10 + b1 + 1];
          }
#pragma unroll
          for (int i = 0; i < 6; i++) M[i] = rq[i];
#pragma unroll
          for (int ta = 0; ta < 2; ta++)
#pragma unroll
            for (int tb = 0; tb < 2; tb++) {
              const int t = ta * 2 + tb;
              const double hab = ha[ta][0] * hb[tb][0] + ha[ta][1] * hb[tb][1] + ha[ta][2] * hb[tb][2];
#pragma unroll
              for (int i = 0; i < 3; i++)
#pragma unroll
                for (int c = 0; c < 3; c++) A[t][i * 3 + c] += ua[ta][i] * ub[tb][c];
#pragma unroll
              for (int i = 0; i < 6; i++) Gs[t][i] += hab * M[i];
            }
        }
        // h K_ab + C_ab = c1 A + c2 A^T + c4 mu tr(A) I + G      (SVK.cuh:35-55, FEAT10DataFunc.cuh:695-762)
        if (on) {
#pragma unroll
          for (int ta = 0; ta < 2; ta++)
#pragma unroll
            for (int tb = 0; tb < 2; tb++) {
              const int t = ta * 2 + tb;
              const double dg = p.c4 * p.mu * (A[t][0] + A[t][4] + A[t][8]);
              double k[9];
              if (MR) {
#pragma unroll
                for (int i = 0; i < 9; i++) k[i] = A[t][i];
              } else {
              k[0] = p.c1 * A[t][0] + p.c2 * A[t][0] + Gs[t][0] + dg;
              k[1] = p.c1 * A[t][1] + p.c2 * A[t][3] + Gs[t][1];
              k[2] = p.c1 * A[t][2] + p.c2 * A[t][6] + Gs[t][2];
              k[3] = p.c1 * A[t][3] + p.c2 * A[t][1] + Gs[t][1];
              k[4] = p.c1 * A[t][4] + p.c2 * A[t][4] + Gs[t][3] + dg;
              k[5] = p.c1 * A[t][5] + p.c2 * A[t][7] + Gs[t][4];
              k[6] = p.c1 * A[t][6] + p.c2 * A[t][2] + Gs[t][2];
              k[7] = p.c1 * A[t][7] + p.c2 * A[t][5] + Gs[t][4];
              k[8] = p.c1 * A[t][8] + p.c2 * A[t][8] + Gs[t][5] + dg;
              }
              double *d1 = s_k + st_d + ta * 18 + tb * 9;
#pragma unroll
              for (int i = 0; i < 9; i++) d1[i] = k[i];
              if (tI != tJ) {  // K_ba = K_ab^T
                double *d2 = s_k + st_t + tb * 18 + ta * 9;
#pragma unroll
                for (int i = 0; i < 3; i++)
#pragma unroll
                  for (int c = 0; c < 3; c++) d2[c * 3 + i] = k[i * 3 + c];
              }
            }
        }
        __syncwarp();
        // the pair's records are contiguous in global memory: one coalesced burst
        const int npair = min(2, ne - pr);
        double *dst = p.stage + (int64_t)(e0 + pr) * EK_ELEM;
        for (int e = 0; e < npair; e++) {
          const double *se = s_k + e * EK_ES;
          double *de = dst + e * EK_ELEM + lane;
#pragma unroll
          for (int I = 0; I < 5; I++)
#pragma unroll
            for (int ta = 0; ta < 2; ta++) {
              const double *sr = se + I * EK_CRS + ta * 18;
              double *dr = de + (2 * I + ta) * EK_VIS;
              rec_store(dr, sr[cp_off[0]]);
              rec_store(dr + 32, sr[cp_off[1]]);
              if (lane < 26) rec_store(dr + 64, sr[I == 4 ? cp_x2 : cp_off[2]]);
            }
        }
        __syncwarp();
      }
    }
  }
  cp_async_wait_all();
}

// ---------------------------------------------------------------------------------------------
// Phase R
// ---------------------------------------------------------------------------------------------
struct RowDesc {  // one per row, stored in PROCESSING order (Morton): a single 32-byte load per row
  int row, rs, deg, vb, nv, pad0, pad1, pad2;
};

struct RKArgs {
  int n_coef;
  const RowDesc *desc;
  const int *inc_ea;
  const uint16_t *slot;
  const double *mass;
  const double *stage, *fe;
  double inv_h;
  double *H;
  double *f_int;
  int row_stride;  // doubles of shared memory per warp
  int *counter;    // ticket counter, zeroed before the launch
};

constexpr int RK_WARPS = 8;
constexpr int RK_T = 4;  // rows per ticket

static __global__ void k_row_desc(int n, const int *__restrict__ order, const int *__restrict__ off,
                           const int *__restrict__ inc_ptr, RowDesc *desc) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int row = order ? order[i] : i;
  RowDesc d;
  d.row = row;
  d.rs = off[row];
  d.deg = off[row + 1] - d.rs;
  d.vb = inc_ptr[row];
  d.nv = inc_ptr[row + 1] - d.vb;
  d.pad0 = d.pad1 = d.pad2 = 0;
  desc[i] = d;
}

// RK_U = visits in flight per warp, MINB = resident blocks per SM the register budget is set for
template <int S, int RK_U, int MINB>
__global__ void __launch_bounds__(RK_WARPS * 32, MINB) rk_kernel(RKArgs p) {
  constexpr int VIS = S * 9;
  constexpr int NK = (VIS + 31) / 32;  // a visit record is read as NK coalesced 32-double rows (3 for S <= 10)
  extern __shared__ double rk_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *rowbuf = rk_smem + (size_t)warp * p.row_stride;
  // lane l adds doubles l, l+32, l+64 of every visit record: block b = idx / 9, entry idx % 9
  int lb[NK], lk[NK];
#pragma unroll
  for (int k = 0; k < NK; k++) {
    const int idx = min(lane + 32 * k, VIS - 1);
    lb[k] = idx / 9;
    lk[k] = idx - 9 * lb[k];
  }
  const bool on2 = lane + 32 * (NK - 1) < VIS;  // the last row of a record may be partial

  // Dynamic schedule: a warp draws tickets of RK_T consecutive rows (processing order) from a global counter,
  // so every resident warp works at the same frontier of the Morton order and the element records the
  // rows share are still in L2 when the next row needs them (a static grid-stride schedule drifts apart
  // on long runs).  The next ticket and the current ticket's row descriptors / visit lists are fetched early.
  const int n_tickets = (p.n_coef + RK_T - 1) / RK_T;
  int t_next = 0;
  if (lane == 0) t_next = atomicAdd(p.counter, 1);
  while (true) {
    const int tk = __shfl_sync(0xffffffffu, t_next, 0);
    if (tk >= n_tickets) break;
    if (lane == 0) t_next = atomicAdd(p.counter, 1);
    // RK_T descriptors = RK_T*8 ints: one coalesced load
    int dword = 0;
    {
      const int64_t w = (int64_t)tk * RK_T * 8 + lane;
      if (lane < RK_T * 8 && w < (int64_t)p.n_coef * 8) dword = reinterpret_cast<const int *>(p.desc)[w];
    }
    int ea_r[RK_T];
#pragma unroll
    for (int r = 0; r < RK_T; r++) {
      const int vb_r = __shfl_sync(0xffffffffu, dword, r * 8 + 3), nv_r = __shfl_sync(0xffffffffu, dword, r * 8 + 4);
      ea_r[r] = (lane < nv_r) ? p.inc_ea[vb_r + lane] : 0;
    }
#pragma unroll 1  // one copy of the row body: four unrolled copies overflow the instruction cache
  for (int r = 0; r < RK_T; r++) {
    if (tk * RK_T + r >= p.n_coef) break;
    const int row = __shfl_sync(0xffffffffu, dword, r * 8 + 0), rs = __shfl_sync(0xffffffffu, dword, r * 8 + 1);
    const int deg = __shfl_sync(0xffffffffu, dword, r * 8 + 2), vb = __shfl_sync(0xffffffffu, dword, r * 8 + 3);
    const int nvis = __shfl_sync(0xffffffffu, dword, r * 8 + 4);
    static_assert(RK_T == 4, "select chain below");
    const int ea_cur = (r == 0) ? ea_r[0] : (r == 1) ? ea_r[1] : (r == 2) ? ea_r[2] : ea_r[3];
    // Every global load of the row's first phase is issued before anything waits on one:
    // the first batch of block values + slots, the first 10 element forces, the mass row.
    double *fbuf = rowbuf + p.row_stride - 32;
    double val[RK_U][NK];
    unsigned sl[RK_U][NK];
    auto load_batch = [&](int v0) {
#pragma unroll
      for (int u = 0; u < RK_U; u++) {
        const int v = v0 + u;
        if (v < nvis) {
          const int ea_s = __shfl_sync(0xffffffffu, ea_cur, v & 31);
          const int ea = (v < 32) ? ea_s : p.inc_ea[vb + v];
          const double *src = p.stage + (int64_t)ea * VIS + lane;  // record of visit (e, a): [b][9], contiguous
          const uint16_t *ss = p.slot + (int64_t)(vb + v) * S;
#pragma unroll
          for (int k = 0; k < NK - 1; k++) {
            val[u][k] = rec_load(src + 32 * k);
            sl[u][k] = ss[lb[k]];
          }
          if (on2) {
            val[u][NK - 1] = rec_load(src + 32 * (NK - 1));
            sl[u][NK - 1] = ss[lb[NK - 1]];
          }
        }
      }
    };
    const int fv = lane / 3, fc = lane - 3 * fv;
    auto load_force = [&](int v0) {
      const int v = v0 + fv;
      const int ea_s = __shfl_sync(0xffffffffu, ea_cur, v & 31);
      const int ea = (v < 32) ? ea_s : ((v < nvis) ? p.inc_ea[vb + v] : 0);
      return (lane < 30 && v < nvis) ? p.fe[(int64_t)ea * 3 + fc] : 0.0;
    };
    load_batch(0);
    double fe0 = load_force(0);
    double m[3];
#pragma unroll
    for (int k = 0; k < 3; k++) m[k] = (lane + 32 * k < deg) ? p.mass[rs + lane + 32 * k] : 0.0;
    for (int idx = lane; idx < deg * 9; idx += 32) rowbuf[idx] = 0.0;
    __syncwarp();
    // internal force of this coefficient: lanes (visit, component) park 10 element contributions at a time in
    // shared memory, lanes 0..2 add them in ascending element order
    double f = 0.0;
    for (int v0 = 0; v0 < nvis; v0 += 10) {
      if (v0 > 0) fe0 = load_force(v0);
      if (lane < 30) fbuf[lane] = fe0;
      __syncwarp();
      if (lane < 3) {
        const int mm = min(10, nvis - v0);
        for (int k = 0; k < mm; k++) f += fbuf[3 * k + lane];
      }
      __syncwarp();
    }
    // H row starts from the mass term: H(3i+d,3j+d) = M_ij / h   (SyncedNewton.cu:213-259)
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const int s2 = lane + 32 * k;
      if (s2 < deg) {
        const double mm = m[k] * p.inv_h;
        rowbuf[s2 * 9] = mm;
        rowbuf[s2 * 9 + 4] = mm;
        rowbuf[s2 * 9 + 8] = mm;
      }
    }
    for (int s2 = lane + 96; s2 < deg; s2 += 32) {  // very long rows
      const double mm = p.mass[rs + s2] * p.inv_h;
      rowbuf[s2 * 9] = mm;
      rowbuf[s2 * 9 + 4] = mm;
      rowbuf[s2 * 9 + 8] = mm;
    }
    __syncwarp();
    for (int v0 = 0; v0 < nvis; v0 += RK_U) {
      if (v0 > 0) load_batch(v0);
      // ordered accumulation: one visit at a time; the lanes of one visit hit distinct entries
#pragma unroll
      for (int u = 0; u < RK_U; u++) {
        if (v0 + u < nvis) {
#pragma unroll
          for (int k = 0; k < NK - 1; k++) rowbuf[sl[u][k] * 9 + lk[k]] += val[u][k];
          if (on2) rowbuf[sl[u][NK - 1] * 9 + lk[NK - 1]] += val[u][NK - 1];
        }
        __syncwarp();
      }
    }
    if (lane < 3 && p.f_int) p.f_int[(int64_t)row * 3 + lane] = f;  // nullptr: a separate force gather owns f_int
    // stream the finished block row out as three scalar CSR rows of length 3*deg
    const int64_t base = 9 * (int64_t)rs;
    const int rowlen = 3 * deg;
#pragma unroll
    for (int d = 0; d < 3; d++) {
      double *out = p.H + base + (int64_t)d * rowlen;
      for (int c = lane; c < rowlen; c += 32) {  // c = 3*s + e
        const int sl = c / 3, e2 = c - 3 * sl;
        row_store(out + c, rowbuf[sl * 9 + d * 3 + e2]);
      }
    }
    __syncwarp();
  }
  }
}

}  // namespace tlfea
