// row2_kernel.cuh -- warp-per-row gather assembly with asynchronous operand staging (FEAT10-class
// elements with 3*S <= 32 lanes, SVK + optional Kelvin-Voigt).
//
// Same arithmetic and the same fixed summation order as row_kernel (kernels.cuh): a warp owns one
// block row of H at a time, lanes are (visit k, column b), three incident elements per pass, blocks
// added to the warp's shared-memory row in ascending element order, the finished row (+ mass/h)
// streamed to H exactly once.  What changes is how operands arrive:
//   * rows are taken in Morton-tile order (tiles.cu) in chunks of 32 per block, so the rows of a
//     block share their elements (L2/L1 reuse instead of DRAM re-reads),
//   * the three elements of the NEXT pass (gradN 1200 B + Gauss-point records 640 B + detJ each) are
//     copied global->shared with cp.async while the current pass computes, fully coalesced; lanes
//     then read their operands from shared memory (broadcast for the a side and the record),
//   * no block-level barrier anywhere: every warp runs its own two-stage pipeline.
// Replaces compute_hessian_assemble_csr (FEAT10DataFunc.cuh:512-791), assemble_sparse_hessian_mass
// (SyncedNewton.cu:213-259) and the force scatter (FEAT10DataFunc.cuh:397-458).
#pragma once
#include "kernels.cuh"
#include "tile_kernel.cuh"  // cp.async helpers

namespace tlfea {

constexpr int R2_WARPS = 4;    // warps per block
constexpr int R2_CHUNK = 32;   // rows (Morton order) per block iteration
constexpr int R2_GROUP = 4;    // rows a warp pipelines together (<= 32 passes)

template <int S, int Q>
struct Row2Shape {
  static constexpr int NG = Q * 3 * S;
  static constexpr int RAWK = ((NG + Q * 16 + Q) + 1) & ~1;  // one element: gradN | records | detJ
  static constexpr int AS = 32 / S;
  static constexpr int RAWP = AS * RAWK;                     // one pass
  static int mrow_doubles(int max_deg) { return (max_deg + 2) & ~1; }
  static int warp_doubles(int max_deg) { return 2 * RAWP + 2 * mrow_doubles(max_deg) + 4 + ((max_deg * 9 + 2) & ~1); }
};

template <int S, int Q>
__global__ void __launch_bounds__(R2_WARPS * 32) row2_kernel(RowArgs p, int warp_doubles, int mrow_doubles) {
  using Sh = Row2Shape<S, Q>;
  constexpr int NG = Sh::NG, RAWK = Sh::RAWK, AS = Sh::AS, RAWP = Sh::RAWP;
  extern __shared__ __align__(16) double sm2[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *wbase = sm2 + (size_t)warp * warp_doubles;
  double *buf = wbase;                               // [2][RAWP], 16-byte aligned
  double *mrow = buf + 2 * RAWP;                     // mass values of the current row
  double *frow = mrow + 2 * mrow_doubles;            // 3 force accumulators (+1 pad); mrow is double-buffered
  double *rowbuf = frow + 4;
  const int k = lane / S, b = lane - k * S;
  const bool lane_on = lane < AS * S;

  const int n_chunks = (p.n_coef + R2_CHUNK - 1) / R2_CHUNK;
  for (int ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    for (int half = 0; half < R2_CHUNK / (R2_WARPS * R2_GROUP); half++) {
      // ---- this warp's rows of the half chunk: lane t < R2_GROUP holds row t
      int row = -1, rs = 0, deg = 0, vb = 0, ve = 0;
      if (lane < R2_GROUP) {
        const int t8 = (half * R2_GROUP + lane) * R2_WARPS;
        // rotate the warp<->row assignment: in Morton order every 8th row is a hub (vertex) row
        const int ri = ch * R2_CHUNK + t8 + ((warp + R2_WARPS - ((t8 / R2_WARPS) % R2_WARPS)) % R2_WARPS);
        if (ri < p.n_coef && ri < (ch + 1) * R2_CHUNK) {
          row = p.perm ? p.perm[ri] : ri;
          rs = p.off[row];
          deg = p.off[row + 1] - rs;
          vb = p.inc_ptr[row];
          ve = p.inc_ptr[row + 1];
        }
      }
      const int npass_l = (row >= 0) ? (ve - vb + AS - 1) / AS : 0;
      int pre[R2_GROUP + 1];
      pre[0] = 0;
#pragma unroll
      for (int t = 0; t < R2_GROUP; t++) pre[t + 1] = pre[t] + __shfl_sync(0xffffffffu, npass_l, t);
      const int NP = pre[R2_GROUP];
      if (NP == 0) continue;
      // lane i describes pass i: row slot t, first visit, and the e*S+a of its visits
      int my_t = 0;
#pragma unroll
      for (int t = 1; t < R2_GROUP; t++) my_t += (lane >= pre[t]) ? 1 : 0;
      int my_pv = 0, my_ve = 0;
      {
        const int vb_t = __shfl_sync(0xffffffffu, vb, my_t), ve_t = __shfl_sync(0xffffffffu, ve, my_t);
        int first = 0;
#pragma unroll
        for (int t = 0; t < R2_GROUP; t++) first = (my_t == t) ? pre[t] : first;
        my_pv = vb_t + AS * (lane - first);
        my_ve = ve_t;
      }
      int my_ea[AS];
#pragma unroll
      for (int kk = 0; kk < AS; kk++) my_ea[kk] = (lane < NP && my_pv + kk < my_ve) ? p.inc_ea[my_pv + kk] : -1;

      // global -> shared copy of the AS elements of pass i (and the mass row when a row starts)
      auto prefetch = [&](int i) {
        double *dst = buf + (i & 1) * RAWP;
        constexpr int C16 = NG / 2 + Q * 8;
#pragma unroll
        for (int kk = 0; kk < AS; kk++) {
          const int ea = __shfl_sync(0xffffffffu, my_ea[kk], i);
          if (ea < 0) continue;  // warp-uniform
          const int e = ea / S;
          const double *g = p.gradN + (int64_t)e * NG;
          const double *r = p.rec + (int64_t)e * Q * 16;
          for (int c = lane; c < C16; c += 32)
            cp_async16(dst + kk * RAWK + 2 * c, (c < NG / 2) ? g + 2 * c : r + 2 * (c - NG / 2));
          if (lane < Q) cp_async8(dst + kk * RAWK + NG + Q * 16 + lane, p.detJ + (int64_t)e * Q + lane);
        }
        const int t = __shfl_sync(0xffffffffu, my_t, i);
        int first = 0;
#pragma unroll
        for (int tt = 0; tt < R2_GROUP; tt++) first = (t == tt) ? pre[tt] : first;
        if (i == first) {  // first pass of row t: its mass values
          const int rs_t = __shfl_sync(0xffffffffu, rs, t), deg_t = __shfl_sync(0xffffffffu, deg, t);
          for (int s2 = lane; s2 < deg_t; s2 += 32) cp_async8(mrow + (t & 1) * mrow_doubles + s2, p.mass + rs_t + s2);
        }
        cp_async_commit();
      };
      prefetch(0);

      for (int i = 0; i < NP; i++) {
        const int t = __shfl_sync(0xffffffffu, my_t, i);
        const int pv = __shfl_sync(0xffffffffu, my_pv, i);
        const int rs_t = __shfl_sync(0xffffffffu, rs, t), deg_t = __shfl_sync(0xffffffffu, deg, t);
        const int ve_t = __shfl_sync(0xffffffffu, ve, t), row_t = __shfl_sync(0xffffffffu, row, t);
        int first = 0, last = 0;
#pragma unroll
        for (int tt = 0; tt < R2_GROUP; tt++) {
          first = (t == tt) ? pre[tt] : first;
          last = (t == tt) ? pre[tt + 1] - 1 : last;
        }
        int ea_k = -1;
#pragma unroll
        for (int kk = 0; kk < AS; kk++) {
          const int v = __shfl_sync(0xffffffffu, my_ea[kk], i);
          ea_k = (k == kk) ? v : ea_k;
        }
        const bool valid = lane_on && ea_k >= 0;
        // per-visit global data needed only at commit time: issue now
        int slot = 0;
        double fadd = 0.0;
        if (valid) {
          slot = p.slot[(int64_t)(pv + k) * S + b];
          if (b < 3) fadd = p.fe[(int64_t)ea_k * 3 + b];
        }
        if (i == first) {  // new row: clear its buffer (the previous row has been written out)
          for (int idx = lane; idx < deg_t * 9; idx += 32) rowbuf[idx] = 0.0;
          if (lane < 3) frow[lane] = 0.0;
        }
        if (i + 1 < NP) {
          prefetch(i + 1);
          asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        } else {
          cp_async_wait_all();
        }
        __syncwarp();
        // ---- blocks of this pass (same arithmetic as row_kernel)
        double val[9];
        if (valid) {
          const int a = ea_k % S;
          const double *rw = buf + (i & 1) * RAWP + k * RAWK;
          double A[9], G[6], sacc = 0.0;
#pragma unroll
          for (int j = 0; j < 9; j++) A[j] = 0.0;
#pragma unroll
          for (int j = 0; j < 6; j++) G[j] = 0.0;
#pragma unroll
          for (int q = 0; q < Q; q++) {
            const double *g = rw + q * 3 * S;
            const double hb0 = g[b], hb1 = g[S + b], hb2 = g[2 * S + b];
            const double ha0 = g[a], ha1 = g[S + a], ha2 = g[2 * S + a];
            const double hab = ha0 * hb0 + ha1 * hb1 + ha2 * hb2;
            const double2 *r2 = reinterpret_cast<const double2 *>(rw + NG + q * 16);
            const double2 r0 = r2[0], r1 = r2[1], r2_ = r2[2], r3 = r2[3], r4 = r2[4], r5 = r2[5], r6 = r2[6],
                          r7 = r2[7];
            const double dV = rw[NG + Q * 16 + q] * p.qw.w[q];
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
          }
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
        }
        // ---- ordered accumulation: visit k = 0 first, then 1, ... (ascending element)
#pragma unroll
        for (int kk = 0; kk < AS; kk++) {
          if (valid && k == kk) {
            double *dst = rowbuf + slot * 9;
#pragma unroll
            for (int j = 0; j < 9; j++) dst[j] += val[j];
            if (b < 3) frow[b] += fadd;
          }
          __syncwarp();
        }
        // ---- last pass of the row: stream it out as three scalar CSR rows, adding M_ij/h
        if (i == last) {
          const int64_t base = 9 * (int64_t)rs_t;
          const int rowlen = 3 * deg_t;
          const double *mr = mrow + (t & 1) * mrow_doubles;
#pragma unroll
          for (int d = 0; d < 3; d++) {
            for (int c = lane; c < rowlen; c += 32) {
              const int s2 = c / 3, e2 = c - 3 * s2;
              double v = rowbuf[s2 * 9 + d * 3 + e2];
              if (e2 == d) v += mr[s2] * p.inv_h;
              p.H[base + d * rowlen + c] = v;
            }
          }
          if (lane < 3) p.f_int[(int64_t)row_t * 3 + lane] = frow[lane];
          __syncwarp();
        }
        (void)ve_t;
      }
    }
  }
}

}  // namespace tlfea
