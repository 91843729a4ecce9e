// pk_kernel.cuh -- single-pass deterministic FEAT10 assembly: element quadrature, on-chip accumulation per patch,
// in-place update of the CSR values of H.  No block records travel through HBM (the two-phase design of
// ek_kernel.cuh wrote and re-read 7.2 KB per element), no atomics, fixed summation order.
//
// Reference behaviour replaced (file:line relative to the reference root):
//   compute_p                     FEAT10DataFunc.cuh:85-293
//   compute_internal_force        FEAT10DataFunc.cuh:397-458   (element part; the nodal sum stays an ordered gather)
//   compute_hessian_assemble_csr  FEAT10DataFunc.cuh:512-791   (900 binary searches + 900 FP64 atomics per thread)
//   assemble_sparse_hessian_mass  SyncedNewton.cu:213-259
//   svk_compute_tangent_block     SVK.cuh:35-55
//   mr_compute_P / mr_compute_tangent_tensor  MooneyRivlin.cuh:45-225
//
// One CTA of W warps = one patch (patch_plan.cuh).  Warp w walks the elements of sub-patch w two at a time:
//   * the elements' plan records (128 B), gradN (one TMA bulk copy per element), detJ and gathered nodal
//     coordinates of the NEXT pair are in flight into a per-warp double buffer while the current pair computes;
//   * Gauss-point step and block step are those of ek_kernel.cuh (lane = (element, point) / lane = (element,
//     2x2 tile of the symmetric 5x5 tile grid); 18 DFMA per (pair, point) for SVK, one 81-entry contraction
//     for Mooney-Rivlin);
//   * instead of writing 100 block records, each lane adds its blocks (upper triangle by global id; transposed
//     where the local order is reversed) into the warp's PRIVATE accumulator in shared memory: first the first
//     element of the pair, then the second -- program order is the summation order.
// After a CTA barrier the patch's write-out lists arrive by one TMA bulk copy into the dead staging area and
// every thread walks (entry, column) positions: sum the entry's sources in warp order, start from M_ij/h (first
// touch of the H block in the colour sequence) or from the value in H (later touch), store to the three scalar
// CSR rows.  Loads of old values are issued before the first-touch entries are processed, so their latency
// hides under that work.  Patches of one launch share no coefficient (graph colouring): plain loads and
// stores, order per H entry = (colour, warp, element).
#pragma once
#include "ek_kernel.cuh"
#include "patch_plan.cuh"

namespace tlfea {

struct PKArgs {
  const PatchDesc *desc;  // first patch of this launch
  const PElem *elems;
  const unsigned char *blob;
  const double *x, *y, *z;
  const double *v;  // interleaved velocity (Kelvin-Voigt) or nullptr
  const double *gradN, *detJ;
  QuadW qw;
  double lambda, mu, eta, lambda_d;
  double c1, c2, c3, c4;  // c1 = h lambda + lambda_d, c2 = c3 = h mu + eta, c4 = h
  double mu10, mu01, kappa;
  double inv_h;
  double *fe;  // [n_elem][10][3]
  double *H;
};

template <bool DAMP, int MODEL>
struct PKShape {
  static constexpr bool MR = (MODEL == TL_MODEL_MR);
  static constexpr int ACC_D = PK_CAPU * 9;
  static constexpr int H_D = 300;   // gradN of an element pair
  static constexpr int DJ_D = 10;
  static constexpr int X_D = 60;
  static constexpr int PE_D = 32;   // two plan records
  static constexpr int U_D = MR ? 2 * 5 * EK_TS : 2 * 5 * EK_GQ;
  static constexpr int REC_D = MR ? 0 : 2 * 5 * EK_REC;
  static constexpr int FP_D = MR ? 2 * 5 * EK_GQM : 2 * 5 * EK_GQ;
  static constexpr int OFF_H = 0;
  static constexpr int OFF_DJ = OFF_H + 2 * H_D;
  static constexpr int OFF_X = OFF_DJ + 2 * DJ_D;
  static constexpr int OFF_V = OFF_X + 2 * X_D;
  static constexpr int OFF_PE = OFF_V + (DAMP ? 2 * X_D : 0);
  static constexpr int OFF_U = OFF_PE + 2 * PE_D;
  static constexpr int OFF_REC = OFF_U + (U_D + 1) / 2 * 2;
  static constexpr int OFF_FP = OFF_REC + (REC_D + 1) / 2 * 2;
  static constexpr int STG_D = OFF_FP + (FP_D + 1) / 2 * 2;  // staging doubles per warp (even)
  static constexpr size_t smem_bytes(int warps) { return (size_t)warps * (ACC_D + STG_D) * sizeof(double); }
  static constexpr int stage_bytes(int warps) { return warps * STG_D * (int)sizeof(double); }
};

__device__ __forceinline__ void bulk_prefetch_l2(const void *gmem, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(gmem), "r"(bytes) : "memory");
}

constexpr int PK_NB = 10;  // later-touch positions a thread keeps in flight

template <int W, bool DAMP, bool TANGENT, int MODEL>
__global__ void __launch_bounds__(W * 32, (W <= 4) ? 2 : 1) pk_kernel(PKArgs p) {
  using Sh = PKShape<DAMP, MODEL>;
  constexpr bool MR = (MODEL == TL_MODEL_MR);
  constexpr int NT = W * 32;
  extern __shared__ __align__(16) double pk_smem[];
  __shared__ uint64_t s_bar[W][2];
  __shared__ uint64_t s_bar_blob;
  __shared__ PatchDesc s_desc;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = pk_smem + (size_t)warp * Sh::ACC_D;
  double *stage0 = pk_smem + (size_t)W * Sh::ACC_D;
  double *wb = stage0 + (size_t)warp * Sh::STG_D;
  double *s_u = wb + Sh::OFF_U, *s_rec = wb + Sh::OFF_REC, *s_fp = wb + Sh::OFF_FP;

  if (threadIdx.x < 16)
    reinterpret_cast<int *>(&s_desc)[threadIdx.x] = reinterpret_cast<const int *>(p.desc + blockIdx.x)[threadIdx.x];
  if (lane == 0) {
    mbar_init(&s_bar[warp][0], 1);
    mbar_init(&s_bar[warp][1], 1);
  }
  if (threadIdx.x == 0) mbar_init(&s_bar_blob, 1);
  if (TANGENT)
    for (int i = lane; i < Sh::ACC_D; i += 32) acc[i] = 0.0;
  __syncthreads();
  if (TANGENT && threadIdx.x == 0)  // the write-out lists are needed last: pull them into L2 now
    bulk_prefetch_l2(p.blob + (size_t)s_desc.blob_off * 16,
                     (unsigned)BlobLayout(s_desc.n_ent, s_desc.n_first, s_desc.n_merge).total);

  const int k0 = s_desc.sub_ptr[warp], k1 = s_desc.sub_ptr[warp + 1];
  const int n_it = (k1 - k0 + 1) >> 1;

  // ---- prefetch helpers (per warp)
  auto issue_pe = [&](int it, int buf) {
    if (it >= n_it) return;
    const int kk = k0 + 2 * it, ne = min(2, k1 - kk);
    double *dst = wb + Sh::OFF_PE + buf * Sh::PE_D;
    const char *src = reinterpret_cast<const char *>(p.elems + kk);
    if (lane < ne * 8) cp_async16(reinterpret_cast<char *>(dst) + lane * 16, src + lane * 16);
  };
  auto issue_data = [&](int it, int buf) {
    if (it >= n_it) return;
    const int kk = k0 + 2 * it, ne = min(2, k1 - kk);
    const PElem *pe = reinterpret_cast<const PElem *>(wb + Sh::OFF_PE + buf * Sh::PE_D);
    double *dh = wb + Sh::OFF_H + buf * Sh::H_D;
    if (lane == 0) {
      const unsigned b = (unsigned)__cvta_generic_to_shared(&s_bar[warp][buf]);
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // the buffer was last read by ordinary loads
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(ne * 1200u) : "memory");
      for (int el = 0; el < ne; el++) {
        const unsigned s = (unsigned)__cvta_generic_to_shared(dh + el * 150);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(s),
                     "l"(p.gradN + (int64_t)pe[el].elem * 150), "r"(1200u), "r"(b)
                     : "memory");
      }
    }
    double *dj = wb + Sh::OFF_DJ + buf * Sh::DJ_D;
    if (lane < ne * 5) {
      const int el = lane / 5, q = lane - 5 * el;
      cp_async8(dj + lane, p.detJ + (int64_t)pe[el].elem * 5 + q);
    }
    double *dx = wb + Sh::OFF_X + buf * Sh::X_D;
    if (lane < ne * 10) {
      const int el = lane / 10, a = lane - 10 * el;
      const int c = pe[el].node[a];
      cp_async8(dx + 3 * lane + 0, p.x + c);
      cp_async8(dx + 3 * lane + 1, p.y + c);
      cp_async8(dx + 3 * lane + 2, p.z + c);
      if (DAMP) {
        double *dv = wb + Sh::OFF_V + buf * Sh::X_D;
        cp_async8(dv + 3 * lane + 0, p.v + 3 * (int64_t)c + 0);
        cp_async8(dv + 3 * lane + 1, p.v + 3 * (int64_t)c + 1);
        cp_async8(dv + 3 * lane + 2, p.v + 3 * (int64_t)c + 2);
      }
    }
  };

  // block step: lane = (element of the pair, 2x2 tile (I, J), I <= J, of the 5x5 tile grid)
  const int el2 = lane >> 4;
  int tile = lane & 15;
  const bool lane_on = tile < 15;
  if (tile > 14) tile = 14;
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

  if (n_it > 0) {
    issue_pe(0, 0);
    cp_async_commit();
    cp_async_wait_all();
    __syncwarp();
    issue_pe(1, 1);
    issue_data(0, 0);
    cp_async_commit();
  }

  for (int it = 0; it < n_it; it++) {
    const int buf = it & 1;
    cp_async_wait_all();
    mbar_wait(&s_bar[warp][buf], (it >> 1) & 1);
    __syncwarp();
    const int kk = k0 + 2 * it, ne = min(2, k1 - kk);
    // what this lane needs of the plan records, before the buffer is recycled
    unsigned blkw = 0, trb = 0;
    int eid[2];
    {
      const PElem *pe = reinterpret_cast<const PElem *>(wb + Sh::OFF_PE + buf * Sh::PE_D);
      const int el = min(el2, ne - 1);
      blkw = pe[el].blk4[tile];
      trb = (pe[el].tr[(tile * 4) >> 5] >> ((tile * 4) & 31)) & 15u;
      eid[0] = pe[0].elem;
      eid[1] = pe[ne - 1].elem;
    }
    __syncwarp();
    issue_data(it + 1, buf ^ 1);  // its plan records landed with the wait above
    issue_pe(it + 2, buf);
    cp_async_commit();

    const double *s_h = wb + Sh::OFF_H + buf * Sh::H_D;
    const double *s_dj = wb + Sh::OFF_DJ + buf * Sh::DJ_D;
    const double *s_x = wb + Sh::OFF_X + buf * Sh::X_D;

    // ------------------------------------------------------------ Gauss-point step, Mooney-Rivlin
    // (ek_kernel.cuh: lane = (element, point, row i); T[i][J][k][L] = h dV dP_iJ/dF_kL + Kelvin-Voigt block)
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
      const int rot = lane >> 3;
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
      double *fpq = s_fp + lane * EK_GQ;
      double *uq = s_u + lane * EK_GQ;
#pragma unroll
      for (int aa = 0; aa < 10; aa++) {
        int ar = aa + rot;
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
        p.fe[(int64_t)eid[el] * 30 + lane] = (((fq[0] + fq[FS]) + fq[2 * FS]) + fq[3 * FS]) + fq[4 * FS];
      }
    }

    // ------------------------------------------------------------ block step
    if (TANGENT) {
      const int el = min(el2, ne - 1);  // odd tail: the second half shadows the first and never adds
      const bool on = lane_on && (el2 < ne);
      const double *he = s_h + el * 150;
      const double *ue = s_u + el * 5 * EK_GQ;
      const double *re = s_rec + el * 5 * EK_REC;
      double A[4][9], Gs[4][6];
#pragma unroll
      for (int t = 0; t < 4; t++) {
#pragma unroll
        for (int i = 0; i < 9; i++) A[t][i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; i++) Gs[t][i] = 0.0;
      }
      if (MR) {
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
      // h K_ab + C_ab = c1 A + c2 A^T + c4 mu tr(A) I + G, finished in place   (SVK.cuh:35-55, FEAT10DataFunc.cuh:695-762)
      if (!MR) {
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const double dg = p.c4 * p.mu * (A[t][0] + A[t][4] + A[t][8]);
          double k[9];
          k[0] = p.c1 * A[t][0] + p.c2 * A[t][0] + Gs[t][0] + dg;
          k[1] = p.c1 * A[t][1] + p.c2 * A[t][3] + Gs[t][1];
          k[2] = p.c1 * A[t][2] + p.c2 * A[t][6] + Gs[t][2];
          k[3] = p.c1 * A[t][3] + p.c2 * A[t][1] + Gs[t][1];
          k[4] = p.c1 * A[t][4] + p.c2 * A[t][4] + Gs[t][3] + dg;
          k[5] = p.c1 * A[t][5] + p.c2 * A[t][7] + Gs[t][4];
          k[6] = p.c1 * A[t][6] + p.c2 * A[t][2] + Gs[t][2];
          k[7] = p.c1 * A[t][7] + p.c2 * A[t][5] + Gs[t][4];
          k[8] = p.c1 * A[t][8] + p.c2 * A[t][8] + Gs[t][5] + dg;
#pragma unroll
          for (int i = 0; i < 9; i++) A[t][i] = k[i];
        }
      }
      // ordered accumulation into the warp's private accumulator: first element of the pair, then the second.
      // Every lane walks the 9 MEMORY positions of its block in the same order (a transposed block picks the other
      // register), so the 15 lanes of a half warp hit 15 different banks when their slots differ modulo 16 -- which
      // the plan arranges (assign_slots, patch_plan.cu).
#pragma unroll 1
      for (int e = 0; e < 2; e++) {
        if (on && el2 == e) {
#pragma unroll
          for (int t = 0; t < 4; t++) {
            if (t == 2 && tI == tJ) continue;  // (a1+1, a1) = transpose of (a1, a1+1): the upper block carries it
            double *d = acc + ((blkw >> (8 * t)) & 255u) * 9;
            const bool tr = (trb >> t) & 1u;
#pragma unroll
            for (int m = 0; m < 9; m++) d[m] += tr ? A[t][(m % 3) * 3 + m / 3] : A[t][m];
          }
        }
        __syncwarp();
      }
    }
    __syncwarp();
  }
  cp_async_wait_all();
  if (!TANGENT) return;

  // ------------------------------------------------------------ write-out
  __syncthreads();  // every accumulator is final and the staging area is dead
  const int n_ent = s_desc.n_ent, n_first = s_desc.n_first, n_merge = s_desc.n_merge;
  const BlobLayout L(n_ent, n_first, n_merge);
  unsigned char *sb = reinterpret_cast<unsigned char *>(stage0);
  if (threadIdx.x == 0) {
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    bulk_load(sb, p.blob + (size_t)s_desc.blob_off * 16, (unsigned)L.total, &s_bar_blob);
  }
  mbar_wait(&s_bar_blob, 0);
  const int *round_end = reinterpret_cast<const int *>(sb);
  const uint32_t *ops = reinterpret_cast<const uint32_t *>(sb + L.ops);
  const uint2 *ent = reinterpret_cast<const uint2 *>(sb + L.ent);
  const double *massb = reinterpret_cast<const double *>(sb + L.mass);
  auto blk_addr = [&](unsigned a) { return pk_smem + (size_t)((a >> 8) & 7u) * Sh::ACC_D + (a & 255u) * 9; };

  // pairs held by several warps: reduce into the lowest warp's block, one extra warp per round (fixed order)
  if (n_merge > 0) {
    int lo = 0;
    for (int r = 0; r < 7 && lo < n_merge; r++) {
      const int hi = round_end[r];
      for (int t = threadIdx.x; t < (hi - lo) * 9; t += NT) {
        const int q = t / 9, i = t - 9 * q;
        const uint32_t op = ops[lo + q];
        blk_addr(op & 0xffffu)[i] += blk_addr(op >> 16)[i];
      }
      __syncthreads();
      lo = hi;
    }
  }

  double *Hb = p.H + s_desc.hbase;
  const int pos_first = 3 * n_first, pos_all = 3 * n_ent;
  for (int base = pos_first; base < pos_all || base == pos_first; base += NT * PK_NB) {
    // later touches: issue the loads of the old values ...
    double o[PK_NB][3];
#pragma unroll
    for (int b = 0; b < PK_NB; b++) {
      const int pos = base + b * NT + (int)threadIdx.x;
      if (pos < pos_all) {
        const int k = pos / 3, e = pos - 3 * k;
        const uint2 en = ent[k];
        const double *hp = Hb + en.x + e;
        const unsigned rl = en.y & 0xffffu;
        o[b][0] = hp[0];
        o[b][1] = hp[rl];
        o[b][2] = hp[2 * rl];
      }
    }
    // ... first touches run underneath (only once)
    if (base == pos_first) {
      for (int pos = threadIdx.x; pos < pos_first; pos += NT) {
        const int k = pos / 3, e = pos - 3 * k;
        const uint2 en = ent[k];
        double *hp = Hb + en.x + e;
        const unsigned rl = en.y & 0xffffu, src = en.y >> 16;
        const double *bsrc = blk_addr(src) + ((src & 0x800u) ? 3 * e : e);
        const int st = (src & 0x800u) ? 1 : 3;
        const double m = massb[k] * p.inv_h;
        hp[0] = (e == 0 ? m : 0.0) + bsrc[0];
        hp[rl] = (e == 1 ? m : 0.0) + bsrc[st];
        hp[2 * rl] = (e == 2 ? m : 0.0) + bsrc[2 * st];
      }
    }
#pragma unroll
    for (int b = 0; b < PK_NB; b++) {
      const int pos = base + b * NT + (int)threadIdx.x;
      if (pos < pos_all) {
        const int k = pos / 3, e = pos - 3 * k;
        const uint2 en = ent[k];
        double *hp = Hb + en.x + e;
        const unsigned rl = en.y & 0xffffu, src = en.y >> 16;
        const double *bsrc = blk_addr(src) + ((src & 0x800u) ? 3 * e : e);
        const int st = (src & 0x800u) ? 1 : 3;
        hp[0] = o[b][0] + bsrc[0];
        hp[rl] = o[b][1] + bsrc[st];
        hp[2 * rl] = o[b][2] + bsrc[2 * st];
      }
    }
  }
}

// mass section of every patch blob: mass[k] = M(row, column) of first-touch entry k.  The pattern block of an entry
// follows from its H position: 9*off[i] <= hbase + hoff < 9*off[i+1].
__global__ void pk_fill_mass_kernel(int n_patches, int n_coef, const PatchDesc *__restrict__ desc, unsigned char *blob,
                                    const int *__restrict__ off, const double *__restrict__ mass) {
  const int pch = blockIdx.x;
  if (pch >= n_patches) return;
  const PatchDesc d = desc[pch];
  const BlobLayout L(d.n_ent, d.n_first, d.n_merge);
  unsigned char *sb = blob + (size_t)d.blob_off * 16;
  const PEnt *ent = reinterpret_cast<const PEnt *>(sb + L.ent);
  double *massb = reinterpret_cast<double *>(sb + L.mass);
  for (int k = threadIdx.x; k < d.n_first; k += blockDim.x) {
    const int64_t habs = d.hbase + ent[k].hoff;
    const int64_t blk9 = habs / 9;
    int lo = 0, hi = n_coef;  // largest i with off[i] <= blk9
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (off[mid] <= blk9) lo = mid;
      else hi = mid;
    }
    const int slot = (int)((habs - 9 * (int64_t)off[lo]) / 3);
    massb[k] = mass[off[lo] + slot];
  }
}

// pattern blocks no element touches (constraint couplings, unused coefficients): H block = (M_ij / h) I
__global__ void pk_uncovered_kernel(int64_t n, const int64_t *__restrict__ blocks, const int *__restrict__ off,
                                    const int *__restrict__ row_of, const double *__restrict__ mass, double inv_h,
                                    double *H) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int64_t g = blocks[t];
  const int i = row_of[t];
  const int rs = off[i], deg = off[i + 1] - rs, slot = (int)(g - rs);
  const double m = mass[g] * inv_h;
  double *hp = H + 9 * (int64_t)rs + 3 * slot;
#pragma unroll
  for (int d = 0; d < 3; d++)
#pragma unroll
    for (int e = 0; e < 3; e++) hp[(int64_t)d * 3 * deg + e] = (d == e) ? m : 0.0;
}

}  // namespace tlfea
