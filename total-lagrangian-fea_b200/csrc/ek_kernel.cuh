// ek_kernel.cuh -- two-phase deterministic assembly for FEAT10 (S = 10, Q = 5), SVK (+ Kelvin-Voigt).
//
// Reference behaviour replaced (file:line relative to the reference root):
//   compute_p                     FEAT10DataFunc.cuh:85-293
//   compute_internal_force        FEAT10DataFunc.cuh:397-458   (atomicAdd per nodal component)
//   compute_hessian_assemble_csr  FEAT10DataFunc.cuh:512-791   (900 binary searches + 900 FP64 atomics per thread)
//   assemble_sparse_hessian_mass  SyncedNewton.cu:213-259
//   svk_compute_tangent_block     SVK.cuh:35-55
//
// Phase E (ek_kernel, element-parallel, FP64-pipe work): ONE WARP PER ELEMENT.
//   * a warp pulls groups of G consecutive elements; gradN / detJ / connectivity / nodal coordinates of
//     the NEXT group are in flight (cp.async into a per-warp double buffer) while the current one computes;
//   * Gauss-point step: lane = (element-in-group, q) builds F, P, and the 6 numbers of
//       M_q = dV [ (h mu + eta) F F^T + h (lambda trE - mu) I ]          (symmetric 3x3)
//     and leaves them in shared memory;
//   * block step: lane = (a, g): shape function a and one third of the cyclic diagonals d of the
//     SYMMETRIC element matrix.  K_ab for b = (a + d) mod 10, d = 0..5 covers all 55 unordered pairs
//     exactly once (d = 5 only for a < 5).  h_a, u_a = F h_a stay in registers for all 5 points, the
//     b-side operands are shared-memory reads with <= 10 distinct addresses per warp (one wavefront):
//       A  = sum_q dV u_a u_b^T   (9 DFMA)      G = sum_q (h_a.h_b) M_q   (3 + 6 DFMA)
//       K_ab = c1 A + c2 A^T + h mu tr(A) I + G
//     18 DFMA per (pair, point) instead of the reference's ~170 flop;
//   * the 55 blocks (495 doubles) go to a global staging record of the element through shared memory so the
//     write is one contiguous 3 960-byte burst; element forces go to fe[e][a][3].
// Phase R (rk_kernel, row-parallel, HBM work): ONE WARP PER COEFFICIENT ROW of H.  The warp zero-fills the
//   block row in shared memory with M/h on the 3x3 diagonals, adds the row's blocks of every incident element
//   in ASCENDING ELEMENT ORDER (direct or transposed from the symmetric record), sums the nodal force in the
//   same order, and streams the finished 3 scalar CSR rows to H exactly once.
// No atomics, no searches, fixed summation order => bit-reproducible results.
#pragma once
#include "common.cuh"

namespace tlfea {

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
  double *fe;             // [n_elem][10][3]
  double *stage;          // [n_elem][55][9] or nullptr (force only)
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

constexpr int EK_REC = 26;  // per Gauss point: M(6) | F(9) | dV | P*dV(9) | pad

template <int G, bool DAMP>
struct EKShape {
  static constexpr int H_D = G * 150;                  // gradN of a group
  static constexpr int DJ_D = (G * 5 + 1) / 2 * 2;     // detJ of a group (even => 16-byte aligned sections)
  static constexpr int X_D = G * 30;                   // nodal coordinates [el][a][3]
  static constexpr int EC_D = G * 5;                   // G*10 ints
  static constexpr int U_D = 150;                      // u = F h of the element in flight
  static constexpr int FP_D = 90;                      // force partials [3][30]
  static constexpr int REC_D = G * 5 * EK_REC;
  static constexpr int K_D = 496;                      // symmetric block record
  static constexpr int OFF_H = 0;
  static constexpr int OFF_DJ = OFF_H + 2 * H_D;
  static constexpr int OFF_X = OFF_DJ + 2 * DJ_D;
  static constexpr int OFF_V = OFF_X + 2 * X_D;
  static constexpr int OFF_EC = OFF_V + (DAMP ? 2 * X_D : 0);
  static constexpr int OFF_U = OFF_EC + 2 * ((EC_D + 1) / 2 * 2);
  static constexpr int OFF_FP = OFF_U + U_D;
  static constexpr int OFF_REC = OFF_FP + FP_D;
  static constexpr int OFF_K = OFF_REC + REC_D;
  static constexpr int WARP_D = (OFF_K + K_D + 1) / 2 * 2;
  static constexpr size_t smem_bytes(int warps) { return (size_t)warps * WARP_D * sizeof(double); }
};

template <int G, int W, bool DAMP, bool TANGENT>
__global__ void __launch_bounds__(W * 32, 1) ek_kernel(EKArgs p) {
  using Sh = EKShape<G, DAMP>;
  constexpr int ELEM_D = SymLayout<10>::ELEM_DOUBLES;  // 495
  extern __shared__ __align__(16) double ek_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *wb = ek_smem + (size_t)warp * Sh::WARP_D;
  double *s_u = wb + Sh::OFF_U, *s_fp = wb + Sh::OFF_FP, *s_rec = wb + Sh::OFF_REC, *s_k = wb + Sh::OFF_K;

  const int n_groups = (p.n_elem + G - 1) / G;
  const int gstride = gridDim.x * W;
  const int g_first = blockIdx.x * W + warp;

  // ---- prefetch helpers (per warp; every lane issues its share and commits one group)
  auto issue_ec = [&](int grp, int buf) {
    if (grp >= n_groups) return;
    const int e0 = grp * G;
    const int n_int = min(G, p.n_elem - e0) * 10;
    int *dst = reinterpret_cast<int *>(wb + Sh::OFF_EC + buf * ((Sh::EC_D + 1) / 2 * 2));
    const int *src = p.ecoef + (int64_t)e0 * 10;
    for (int i = lane; i < n_int / 2; i += 32) cp_async8(dst + 2 * i, src + 2 * i);
  };
  auto issue_data = [&](int grp, int buf) {
    if (grp >= n_groups) return;
    const int e0 = grp * G;
    const int ne = min(G, p.n_elem - e0);
    double *dh = wb + Sh::OFF_H + buf * Sh::H_D;
    const double *sh = p.gradN + (int64_t)e0 * 150;
    for (int i = lane; i < ne * 75; i += 32) cp_async16(dh + 2 * i, sh + 2 * i);
    double *dj = wb + Sh::OFF_DJ + buf * Sh::DJ_D;
    const double *sj = p.detJ + (int64_t)e0 * 5;
    for (int i = lane; i < ne * 5; i += 32) cp_async8(dj + i, sj + i);
    const int *ec = reinterpret_cast<const int *>(wb + Sh::OFF_EC + buf * ((Sh::EC_D + 1) / 2 * 2));
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

  const int a = (lane < 30) ? lane % 10 : lane - 30;
  const int g = lane / 10;  // 0..2 work, 3 = the two spare lanes

  int it = 0;
  for (int grp = g_first; grp < n_groups; grp += gstride, it++) {
    const int buf = it & 1;
    cp_async_wait_all();
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

    // ------------------------------------------------------------ Gauss-point step
    if (lane < ne * 5) {
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
      double *r = s_rec + lane * EK_REC;
      const double sI = p.c4 * dV * (p.lambda * trE - p.mu);
      const double cf = p.c3 * dV;
      r[0] = cf * FFT[0][0] + sI;
      r[1] = cf * FFT[0][1];
      r[2] = cf * FFT[0][2];
      r[3] = cf * FFT[1][1] + sI;
      r[4] = cf * FFT[1][2];
      r[5] = cf * FFT[2][2] + sI;
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          r[6 + i * 3 + j] = F[i][j];
          r[16 + i * 3 + j] = P[i][j] * dV;
        }
      r[15] = dV;
    }
    __syncwarp();

    // ------------------------------------------------------------ block step, element by element
    for (int el = 0; el < ne; el++) {
      const int e = e0 + el;
      const double *he = s_h + el * 150;
      const double *re = s_rec + el * 5 * EK_REC;
      double ha[5][3];
#pragma unroll
      for (int q = 0; q < 5; q++)
#pragma unroll
        for (int j = 0; j < 3; j++) ha[q][j] = he[q * 30 + j * 10 + a];
      // u_a(q) = F_q h_a(q) and the force partials: point q is done by lane group q/2
      double fp[3] = {0.0, 0.0, 0.0};
#pragma unroll
      for (int q = 0; q < 5; q++) {
        if (q / 2 == g) {
          const double *rq = re + q * EK_REC;
#pragma unroll
          for (int i = 0; i < 3; i++) {
            s_u[q * 30 + i * 10 + a] = rq[6 + i * 3] * ha[q][0] + rq[7 + i * 3] * ha[q][1] + rq[8 + i * 3] * ha[q][2];
            fp[i] += rq[16 + i * 3] * ha[q][0] + rq[17 + i * 3] * ha[q][1] + rq[18 + i * 3] * ha[q][2];
          }
        }
      }
      if (g < 3) {
#pragma unroll
        for (int i = 0; i < 3; i++) s_fp[g * 30 + a * 3 + i] = fp[i];
      }
      __syncwarp();
      if (lane < 30) p.fe[(int64_t)e * 30 + lane] = (s_fp[lane] + s_fp[30 + lane]) + s_fp[60 + lane];
      if (TANGENT) {
        double ua[5][3];
#pragma unroll
        for (int q = 0; q < 5; q++) {
          const double dV = re[q * EK_REC + 15];
#pragma unroll
          for (int i = 0; i < 3; i++) ua[q][i] = s_u[q * 30 + i * 10 + a] * dV;
        }
#pragma unroll
        for (int s = 0; s < 2; s++) {
          const int d = 2 * g + s;
          int b = a + d;
          if (b >= 10) b -= 10;
          double A[9], Gs[6];
#pragma unroll
          for (int i = 0; i < 9; i++) A[i] = 0.0;
#pragma unroll
          for (int i = 0; i < 6; i++) Gs[i] = 0.0;
#pragma unroll
          for (int q = 0; q < 5; q++) {
            const double hb0 = he[q * 30 + b], hb1 = he[q * 30 + 10 + b], hb2 = he[q * 30 + 20 + b];
            const double ub0 = s_u[q * 30 + b], ub1 = s_u[q * 30 + 10 + b], ub2 = s_u[q * 30 + 20 + b];
            const double2 *m2 = reinterpret_cast<const double2 *>(re + q * EK_REC);
            const double2 m01 = m2[0], m23 = m2[1], m45 = m2[2];
            const double hab = ha[q][0] * hb0 + ha[q][1] * hb1 + ha[q][2] * hb2;
            A[0] += ua[q][0] * ub0; A[1] += ua[q][0] * ub1; A[2] += ua[q][0] * ub2;
            A[3] += ua[q][1] * ub0; A[4] += ua[q][1] * ub1; A[5] += ua[q][1] * ub2;
            A[6] += ua[q][2] * ub0; A[7] += ua[q][2] * ub1; A[8] += ua[q][2] * ub2;
            Gs[0] += hab * m01.x; Gs[1] += hab * m01.y; Gs[2] += hab * m23.x;
            Gs[3] += hab * m23.y; Gs[4] += hab * m45.x; Gs[5] += hab * m45.y;
          }
          // h K_ab + C_ab = c1 A + c2 A^T + c4 mu tr(A) I + G      (SVK.cuh:35-55, FEAT10DataFunc.cuh:695-762)
          const double dg = p.c4 * p.mu * (A[0] + A[4] + A[8]);
          const bool on = (g < 3) && (d < 5 || a < 5);
          if (on) {
            double *k = s_k + ((d < 5) ? (a * 5 + d) : (50 + a)) * 9;
            k[0] = p.c1 * A[0] + p.c2 * A[0] + Gs[0] + dg;
            k[1] = p.c1 * A[1] + p.c2 * A[3] + Gs[1];
            k[2] = p.c1 * A[2] + p.c2 * A[6] + Gs[2];
            k[3] = p.c1 * A[3] + p.c2 * A[1] + Gs[1];
            k[4] = p.c1 * A[4] + p.c2 * A[4] + Gs[3] + dg;
            k[5] = p.c1 * A[5] + p.c2 * A[7] + Gs[4];
            k[6] = p.c1 * A[6] + p.c2 * A[2] + Gs[2];
            k[7] = p.c1 * A[7] + p.c2 * A[5] + Gs[4];
            k[8] = p.c1 * A[8] + p.c2 * A[8] + Gs[5] + dg;
          }
        }
        __syncwarp();
        double *dst = p.stage + (int64_t)e * ELEM_D;
        for (int i = lane; i < ELEM_D; i += 32) dst[i] = s_k[i];
      }
      __syncwarp();
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
};

constexpr int RK_WARPS = 8;
constexpr int RK_U = 6;  // visits in flight per warp

__global__ void k_row_desc(int n, const int *__restrict__ order, const int *__restrict__ off,
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

// source of the three values lane (block bi, matrix row i) adds for a visit with local shape function a:
// packed (offset in the element record) | stride << 12 | local column node b << 16
template <int S>
__device__ __forceinline__ unsigned rk_source(int a, int bi, int i) {
  constexpr int D = SymLayout<S>::D;
  int b, pr;
  bool tr;
  if (bi < D) {  // own diagonal d = bi
    b = a + bi;
    if (b >= S) b -= S;
    pr = a * D + bi;
    tr = false;
  } else if (bi == D) {  // the half diagonal
    b = a + D;
    if (b >= S) b -= S;
    tr = (a >= D);
    pr = S * D + (tr ? a - D : a);
  } else {  // stored by the partner: K_ab = K_ba^T
    const int d = bi - D;
    b = a - d;
    if (b < 0) b += S;
    pr = b * D + d;
    tr = true;
  }
  const unsigned off = (unsigned)(pr * 9 + (tr ? i : 3 * i));
  return off | ((tr ? 3u : 1u) << 12) | ((unsigned)b << 16);
}

template <int S>
__global__ void __launch_bounds__(RK_WARPS * 32, 3) rk_kernel(RKArgs p) {
  static_assert(3 * S <= 32, "one lane per (block, matrix row)");
  constexpr int ELEM_D = SymLayout<S>::ELEM_DOUBLES;
  extern __shared__ double rk_smem[];
  __shared__ unsigned s_tab[S * 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *rowbuf = rk_smem + (size_t)warp * p.row_stride;
  for (int t = threadIdx.x; t < S * 32; t += RK_WARPS * 32) {
    const int a = t >> 5, l = t & 31;
    s_tab[t] = (l < 3 * S) ? rk_source<S>(a, l / 3, l % 3) : 0u;
  }
  __syncthreads();
  const bool lane_on = lane < 3 * S;
  const int li3 = 3 * (lane % 3);

  const int n_chunks = (p.n_coef + RK_WARPS - 1) / RK_WARPS;
  // chunk = 16 consecutive rows in processing order, one per warp; the warp<->row assignment rotates with
  // the chunk so the periodic hub (vertex) rows of a Morton order spread over all warps.  All resident
  // blocks work on neighbouring chunks => the element records they share are still in L2.
  auto row_index = [&](int ch, int it) { return ch * RK_WARPS + ((warp + it) & (RK_WARPS - 1)); };
  auto load_desc = [&](int ch, int it, int4 &d0, int &nv) {
    const int ri = row_index(ch, it);
    if (ch < n_chunks && ri < p.n_coef) {
      const int4 *q = reinterpret_cast<const int4 *>(p.desc + ri);
      d0 = q[0];
      nv = reinterpret_cast<const int *>(q + 1)[0];
    } else {
      d0 = make_int4(0, 0, 0, 0);
      nv = -1;
    }
  };
  int ch = blockIdx.x;
  int4 d0;
  int nv;
  int it = 0;
  load_desc(ch, it, d0, nv);
  int ea_l = (nv > lane) ? p.inc_ea[d0.w + lane] : 0;  // visits lane, (lane + 32 handled in the loop)
  for (; ch < n_chunks; ch += gridDim.x, it++) {
    const int row = d0.x, rs = d0.y, deg = d0.z, vb = d0.w, nvis = nv;
    const int ea_cur = ea_l;
    // prefetch the next row's descriptor and its visit list
    load_desc(ch + gridDim.x, it + 1, d0, nv);
    if (nvis < 0) continue;
    // internal force of this coefficient: element contributions fetched 10 at a time by lanes (visit, component),
    // parked in shared memory and summed by lanes 0..2 in ascending element order
    double *fbuf = rowbuf + p.row_stride - 32;
    double f = 0.0;
    {
      const int fv = lane / 3, fc = lane - 3 * fv;
      for (int v0 = 0; v0 < nvis; v0 += 10) {
        const int v = v0 + fv;
        const int ea_s = __shfl_sync(0xffffffffu, ea_cur, v & 31);
        const int ea = (v < 32) ? ea_s : ((v < nvis) ? p.inc_ea[vb + v] : 0);
        if (lane < 30 && v < nvis) fbuf[lane] = p.fe[(int64_t)ea * 3 + fc];
        __syncwarp();
        if (lane < 3) {
          const int m = min(10, nvis - v0);
          for (int k = 0; k < m; k++) f += fbuf[3 * k + lane];
        }
        __syncwarp();
      }
    }
    // H row starts from the mass term: H(3i+d,3j+d) = M_ij / h   (SyncedNewton.cu:213-259)
    {
      double m[3];
#pragma unroll
      for (int k = 0; k < 3; k++) m[k] = (lane + 32 * k < deg) ? p.mass[rs + lane + 32 * k] * p.inv_h : 0.0;
      for (int idx = lane; idx < deg * 9; idx += 32) rowbuf[idx] = 0.0;
      __syncwarp();
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const int sl = lane + 32 * k;
        if (sl < deg) {
          rowbuf[sl * 9] = m[k];
          rowbuf[sl * 9 + 4] = m[k];
          rowbuf[sl * 9 + 8] = m[k];
        }
      }
      for (int sl = lane + 96; sl < deg; sl += 32) {  // very long rows
        const double mm = p.mass[rs + sl] * p.inv_h;
        rowbuf[sl * 9] = mm;
        rowbuf[sl * 9 + 4] = mm;
        rowbuf[sl * 9 + 8] = mm;
      }
    }
    __syncwarp();
    for (int v0 = 0; v0 < nvis; v0 += RK_U) {
      double val[RK_U][3];
      int dst[RK_U];
#pragma unroll
      for (int u = 0; u < RK_U; u++) {
        const int v = v0 + u;
        dst[u] = -1;
        if (v < nvis) {
          const int ea = (v < 32) ? __shfl_sync(0xffffffffu, ea_cur, v & 31) : p.inc_ea[vb + v];
          if (lane_on) {
            const int e = ea / S, a = ea - e * S;
            const unsigned t = s_tab[a * 32 + lane];
            const double *src = p.stage + (int64_t)e * ELEM_D + (t & 0xfffu);
            const int st = (t >> 12) & 3;
            val[u][0] = src[0];
            val[u][1] = src[st];
            val[u][2] = src[2 * st];
            dst[u] = (int)p.slot[(int64_t)(vb + v) * S + (t >> 16)] * 9 + li3;
          }
        }
      }
      if (v0 == 0) ea_l = (nv > lane) ? p.inc_ea[d0.w + lane] : 0;  // next row's visits, in flight early
      // ordered accumulation: one visit at a time; the lanes of one visit hit distinct entries
#pragma unroll
      for (int u = 0; u < RK_U; u++) {
        if (dst[u] >= 0) {
          double *q = rowbuf + dst[u];
          q[0] += val[u][0];
          q[1] += val[u][1];
          q[2] += val[u][2];
        }
        __syncwarp();
      }
    }
    if (nvis == 0) ea_l = (nv > lane) ? p.inc_ea[d0.w + lane] : 0;
    if (lane < 3) p.f_int[(int64_t)row * 3 + lane] = f;
    // stream the finished block row out as three scalar CSR rows of length 3*deg
    const int64_t base = 9 * (int64_t)rs;
    const int rowlen = 3 * deg;
#pragma unroll
    for (int d = 0; d < 3; d++) {
      double *out = p.H + base + (int64_t)d * rowlen;
      for (int c = lane; c < rowlen; c += 32) {  // c = 3*s + e
        const int sl = c / 3, e2 = c - 3 * sl;
        out[c] = rowbuf[sl * 9 + d * 3 + e2];
      }
    }
    __syncwarp();
  }
}

}  // namespace tlfea
