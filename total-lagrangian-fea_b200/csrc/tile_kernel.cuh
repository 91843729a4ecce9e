// tile_kernel.cuh -- tangent + H assembly by spatial row tiles (SVK, optional Kelvin-Voigt).
//
// Replaces compute_hessian_assemble_csr (FEAT10DataFunc.cuh:512-791) + assemble_sparse_hessian_mass
// (SyncedNewton.cu:213-259) + the force scatter (FEAT10DataFunc.cuh:397-458).
//
// One CTA owns a TILE of H block rows (a Morton-compact set of coefficients, tiles.cu) and keeps
// them in shared memory; every row of the tile belongs to exactly ONE warp (local row % W), so no
// two warps ever add to the same row and no ordering protocol is needed.  The tile's incident
// elements are processed in BATCHES:
//   0. cp.async brings the next batch's gradN / Gauss-point records into shared memory,
//   1. all threads stage the batch: per Gauss point and local node c  u_c = F h_c and h_c, and the
//      symmetric 3x3  M_q = c3 dV F F^T + c4 dV (lambda trE - mu) I,
//   2. each warp walks ITS visits of the batch -- (element, a) pairs whose row it owns, a fixed
//      precomputed list -- three at a time, lane = (visit slot, b): 15 accumulators
//      A = sum_q (dV u_a) u_b^T,  Sg = sum_q (h_a.h_b) M_q,
//      K_ab = c1 A + c2 A^T + Sg + c4 mu tr(A) I   (SVK.cuh:35-55; the Kelvin-Voigt block
//      FEAT10DataFunc.cuh:695-762 is folded into c1..c3), added to the row buffer in list order.
// Fixed order => bit-reproducible; no FP64 atomics (the reference: FEAT10DataFunc.cuh:662-687).
// Finished rows (+ mass/h on the block diagonals) stream to H once.  DRAM traffic ~ gradN + records + H.
#pragma once
#include <cub/block/block_scan.cuh>

#include "common.cuh"

namespace tlfea {

struct TileArgs {
  int n_tiles;
  const int *trow, *trs, *tdeg, *tile_ptr, *te_ptr, *te_elem, *tb_ptr, *vptr;
  const unsigned *visit;
  const int *vea;
  const uint16_t *vslot;
  const double *mass;
  const double *rec;  // [n_elem*Q][16]
  const double *gradN, *detJ, *fe;
  QuadW qw;
  double inv_h, c1, c2, c3, c4, mu;
  double *H, *f_int;
  int buf_doubles;
  unsigned long long *timing;  // optional [8] phase cycle counters (TK_TIMING builds)
};

#ifdef TK_TIMING
#define TK_T(i) do { if (lane == 0) { long long _t = clock64(); atomicAdd(p.timing + (i), (unsigned long long)(_t - t_last)); t_last = _t; } } while (0)
#else
#define TK_T(i) do { } while (0)
#endif
constexpr int TK_MAX_ROWS = 254;
constexpr int TK_MAX_ENT = 768;  // elements incident to one tile (list kept in shared memory)

template <int S, int Q, int W, int B>
struct TileShape {
  static constexpr int NODE = Q * S * 6;                           // {u0,u1,u2,h0,h1,h2} per (q, node)
  static constexpr int STG0 = NODE + Q * 6 + ((Q + 1) & ~1);       // + M_q + dV_q
  static constexpr int STG = STG0 + ((STG0 % 16 == 2) ? 0 : ((18 - STG0 % 16) % 16));  // = 2 mod 16 doubles: broadcast reads of different elements fall in different banks
  static constexpr int RAW = ((Q * 3 * S + Q * 16 + Q) + 1) & ~1;  // gradN | records | detJ
  static size_t smem_bytes(int buf_doubles) {
    // row buffer | mass rows (buf/9) | frow | stage | raw | roff, grow, rstart, element list
    return ((size_t)buf_doubles + (buf_doubles / 9 + 2) + 256 * 3 + (size_t)B * (STG + RAW)) * sizeof(double) +
           (257 + 256 + 256 + TK_MAX_ENT) * sizeof(int);
  }
};

#define TL_HAVE_CP_ASYNC
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)),
               "l"(gmem));
}
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)),
               "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }


template <int S, int Q, int W, int B, int CTAS>
__global__ void __launch_bounds__(W * 32, CTAS) tile_kernel(TileArgs p) {
  static_assert(W * 32 >= 256, "block scan covers up to 256 rows");
  static_assert((Q * 3 * S) % 2 == 0, "gradN block of one element must be 16-byte granular");
  using Sh = TileShape<S, Q, W, B>;
  constexpr int AS = 32 / S;  // visits handled per pass
  constexpr int T = W * 32;
  constexpr int NODE = Sh::NODE, STG = Sh::STG, RAW = Sh::RAW;
  constexpr int NG = Q * 3 * S;  // gradN doubles per element
  extern __shared__ __align__(16) double sm[];
  double *rowbuf = sm;
  double *mrow = rowbuf + p.buf_doubles;            // mass values of the tile rows, row after row
  double *frow = mrow + (p.buf_doubles / 9 + 2);
  double *stage = frow + 256 * 3;                   // [B][STG]
  double *raw = stage + B * STG;                    // [B][RAW]
  int *roff = reinterpret_cast<int *>(raw + B * RAW);
  int *grow = roff + 257;
  int *rstart = grow + 256;
  int *tel = rstart + 256;                          // element ids of the tile

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef TK_TIMING
  long long t_last = clock64();
#endif
  const int tile = blockIdx.x;
  const int r0 = p.tile_ptr[tile], nrows = p.tile_ptr[tile + 1] - r0;
  const int e0 = p.te_ptr[tile], ne = p.te_ptr[tile + 1] - e0;
  const int gb0 = p.tb_ptr[tile], nb = p.tb_ptr[tile + 1] - gb0;

  // ---- row table: global row, CSR start, buffer offset (exclusive scan of 9*deg); element list
  {
    int row = 0, rs = 0, sz = 0;
    if (tid < nrows) {
      row = p.trow[r0 + tid];
      rs = p.trs[r0 + tid];
      sz = 9 * p.tdeg[r0 + tid];
    }
    for (int i = tid; i < ne; i += T) tel[i] = p.te_elem[e0 + i];
    using Scan = cub::BlockScan<int, T>;
    __shared__ typename Scan::TempStorage tmp;
    int excl;
    Scan(tmp).ExclusiveSum(sz, excl);
    if (tid < 256) {
      if (tid < nrows) {
        grow[tid] = row;
        rstart[tid] = rs;
        roff[tid] = excl;
        frow[3 * tid] = frow[3 * tid + 1] = frow[3 * tid + 2] = 0.0;
      }
      if (tid == nrows) roff[tid] = excl;
    }
  }
  __syncthreads();

  // raw data of batch bi -> shared memory, asynchronously (gradN, records: 16-B chunks; detJ: 8-B)
  auto prefetch = [&](int bi) {
    const int n_el = min(B, ne - bi * B);
    constexpr int C16 = NG / 2 + Q * 8;  // 16-byte chunks per element
    for (int i = tid; i < n_el * C16; i += T) {
      const int el = i / C16, c = i - el * C16;
      const int e = tel[bi * B + el];
      const double *src = (c < NG / 2) ? p.gradN + (int64_t)e * NG + 2 * c
                                       : p.rec + (int64_t)e * Q * 16 + 2 * (c - NG / 2);
      cp_async16(raw + el * RAW + 2 * c, src);
    }
    for (int i = tid; i < n_el * Q; i += T) {
      const int el = i / Q, q = i - el * Q;
      cp_async8(raw + el * RAW + NG + Q * 16 + q, p.detJ + (int64_t)tel[bi * B + el] * Q + q);
    }
    cp_async_commit();
  };
  // mass rows -> shared memory (needed only when the rows stream out)
  for (int r = warp; r < nrows; r += W) {
    const int deg = (roff[r + 1] - roff[r]) / 9;
    const double *src = p.mass + rstart[r];
    double *dst = mrow + roff[r] / 9;
    for (int s2 = lane; s2 < deg; s2 += 32) cp_async8(dst + s2, src + s2);
  }
  if (nb > 0) prefetch(0);
  else cp_async_commit();
  // ---- zero the row buffer (the mass term M_ij/h is added while streaming the rows out)
  {
    const int total = roff[nrows];
    double2 *z2 = reinterpret_cast<double2 *>(rowbuf);
    for (int i = tid; i < (total + 1) / 2; i += T) z2[i] = make_double2(0.0, 0.0);
  }

  TK_T(0);
  const int k = lane / S, b = lane - k * S;
  const bool lane_on = lane < AS * S;
  int v0n = 0, v1n = 0;
  if (nb > 0) {
    v0n = p.vptr[(int64_t)gb0 * W + warp];
    v1n = p.vptr[(int64_t)gb0 * W + warp + 1];
  }

  for (int bi = 0; bi < nb; bi++) {
    const int n_el = min(B, ne - bi * B);
    const int v0 = v0n, v1 = v1n;
    if (bi + 1 < nb) {  // next batch's visit range: in flight during this batch
      v0n = p.vptr[(int64_t)(gb0 + bi + 1) * W + warp];
      v1n = p.vptr[(int64_t)(gb0 + bi + 1) * W + warp + 1];
    }
    // per-visit data of this warp's batch share -> registers, issued BEFORE the barrier wait:
    // lane i holds the descriptor of visit v0+i; lanes 3j..3j+2 hold the element force of visit j (+10, +20)
    const int nvis = v1 - v0;
    unsigned d_reg = 0u;
    double fe_reg[3] = {0.0, 0.0, 0.0};
    int slot_nx = 0;
    if (lane < nvis) d_reg = p.visit[v0 + lane];
    {
      const int vj = lane / 3, comp = lane - 3 * vj;
#pragma unroll
      for (int g = 0; g < 3; g++) {
        const int j = g * 10 + vj;
        if (lane < 30 && j < nvis) fe_reg[g] = p.fe[(int64_t)p.vea[v0 + j] * 3 + comp];
      }
    }
    if (lane_on && k < nvis) slot_nx = p.vslot[(int64_t)(v0 + k) * S + b];
    TK_T(1);
    cp_async_wait_all();
    __syncthreads();  // raw(bi) complete for all threads; every warp is done with the previous stage
    TK_T(2);
    // ---- stage the batch
    for (int it = tid; it < n_el * Q * S; it += T) {
      const int el = it / (Q * S), qc = it - el * (Q * S);
      const int q = qc / S, c = qc - q * S;
      const double *rw = raw + el * RAW;
      const double *g = rw + q * 3 * S;
      const double h0 = g[c], h1 = g[S + c], h2 = g[2 * S + c];
      const double2 *r2 = reinterpret_cast<const double2 *>(rw + NG + q * 16);
      const double2 f01 = r2[0], f23 = r2[1], f45 = r2[2], f67 = r2[3];
      const double f8 = rw[NG + q * 16 + 8];
      double2 *N2 = reinterpret_cast<double2 *>(stage + el * STG + qc * 6);
      N2[0] = make_double2(f01.x * h0 + f01.y * h1 + f23.x * h2, f23.y * h0 + f45.x * h1 + f45.y * h2);
      N2[1] = make_double2(f67.x * h0 + f67.y * h1 + f8 * h2, h0);
      N2[2] = make_double2(h1, h2);
    }
    for (int it = tid; it < n_el * Q; it += T) {
      const int el = it / Q, q = it - el * Q;
      const double *r = raw + el * RAW + NG + q * 16;
      const double dg = p.c4 * r[15];
      double *st = stage + el * STG + NODE;
      double2 *M2 = reinterpret_cast<double2 *>(st + q * 6);
      M2[0] = make_double2(p.c3 * r[9] + dg, p.c3 * r[10]);
      M2[1] = make_double2(p.c3 * r[11], p.c3 * r[12] + dg);
      M2[2] = make_double2(p.c3 * r[13], p.c3 * r[14] + dg);
      st[Q * 6 + q] = raw[el * RAW + NG + Q * 16 + q] * p.qw.w[q];
    }
    TK_T(3);
    __syncthreads();  // stage visible; raw free
    TK_T(4);
    if (bi + 1 < nb) prefetch(bi + 1);
    TK_T(5);

    // ---- this warp's visits of the batch, AS at a time
    for (int vb = v0; vb < v1; vb += AS) {
      const int j = vb - v0 + k;  // visit index inside this warp's batch share
      const bool valid = lane_on && (j < nvis);
      unsigned d = __shfl_sync(0xffffffffu, d_reg, j & 31);
      const int slot = slot_nx;
      // element force component b of visit j sits in lane 3*(j%10)+b, register j/10
      const int jg = j / 10;
      const int fsrc = (3 * (j - 10 * jg) + b) & 31;
      const double f0 = __shfl_sync(0xffffffffu, fe_reg[0], fsrc);
      const double f1 = __shfl_sync(0xffffffffu, fe_reg[1], fsrc);
      const double f2 = __shfl_sync(0xffffffffu, fe_reg[2], fsrc);
      double fadd = (jg == 0) ? f0 : ((jg == 1) ? f1 : f2);
      if (valid && j >= 30) {  // beyond the register prefetch window
        if (b < 3) fadd = p.fe[(int64_t)p.vea[vb + k] * 3 + b];
      }
      if (valid && j >= 32) d = p.visit[vb + k];
      if (lane_on && j + AS < nvis) slot_nx = p.vslot[(int64_t)(vb + AS + k) * S + b];  // next pass
      if (!valid) d = 0u;
      const int el = d & 0xff, a = (d >> 8) & 0xff, lr = valid ? (int)(d >> 16) : -1;
      double val[9];
      if (valid) {
        const double *node = stage + el * STG;
        const double *Mq = node + NODE, *dVq = Mq + Q * 6;
        double A[9], Sg[6];
#pragma unroll
        for (int i = 0; i < 9; i++) A[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; i++) Sg[i] = 0.0;
#pragma unroll
        for (int q = 0; q < Q; q++) {
          const double2 *a2 = reinterpret_cast<const double2 *>(node + (q * S + a) * 6);
          const double2 *b2 = reinterpret_cast<const double2 *>(node + (q * S + b) * 6);
          const double2 *m2 = reinterpret_cast<const double2 *>(Mq + q * 6);
          const double2 a01 = a2[0], a23 = a2[1], a45 = a2[2];
          const double2 b01 = b2[0], b23 = b2[1], b45 = b2[2];
          const double2 m01 = m2[0], m23 = m2[1], m45 = m2[2];
          const double dV = dVq[q];
          const double ua0 = dV * a01.x, ua1 = dV * a01.y, ua2 = dV * a23.x;
          const double hab = a23.y * b23.y + a45.x * b45.x + a45.y * b45.y;
          A[0] += ua0 * b01.x; A[1] += ua0 * b01.y; A[2] += ua0 * b23.x;
          A[3] += ua1 * b01.x; A[4] += ua1 * b01.y; A[5] += ua1 * b23.x;
          A[6] += ua2 * b01.x; A[7] += ua2 * b01.y; A[8] += ua2 * b23.x;
          Sg[0] += hab * m01.x; Sg[1] += hab * m01.y; Sg[2] += hab * m23.x;
          Sg[3] += hab * m23.y; Sg[4] += hab * m45.x; Sg[5] += hab * m45.y;
        }
        const double dg = p.c4 * p.mu * (A[0] + A[4] + A[8]);
        val[0] = p.c1 * A[0] + p.c2 * A[0] + Sg[0] + dg;
        val[1] = p.c1 * A[1] + p.c2 * A[3] + Sg[1];
        val[2] = p.c1 * A[2] + p.c2 * A[6] + Sg[2];
        val[3] = p.c1 * A[3] + p.c2 * A[1] + Sg[1];
        val[4] = p.c1 * A[4] + p.c2 * A[4] + Sg[3] + dg;
        val[5] = p.c1 * A[5] + p.c2 * A[7] + Sg[4];
        val[6] = p.c1 * A[6] + p.c2 * A[2] + Sg[2];
        val[7] = p.c1 * A[7] + p.c2 * A[5] + Sg[4];
        val[8] = p.c1 * A[8] + p.c2 * A[8] + Sg[5] + dg;
      }
      // ---- add to the warp's own rows in list order.  Visit slots of one pass may hit the same
      //      row (two elements of the batch): then slot k waits for the slots before it.
      bool clash = false;
#pragma unroll
      for (int kk = 1; kk < AS; kk++) {
        const int lr_cur = __shfl_sync(0xffffffffu, lr, kk * S);
#pragma unroll
        for (int jj = 0; jj < kk; jj++) {
          const int lr_prev = __shfl_sync(0xffffffffu, lr, jj * S);
          clash |= (lr_cur >= 0) && (lr_cur == lr_prev);
        }
      }
      if (!clash) {
        if (valid) {
          double *dst = rowbuf + roff[lr] + slot * 9;
#pragma unroll
          for (int i = 0; i < 9; i++) dst[i] += val[i];
          if (b < 3) frow[3 * lr + b] += fadd;
        }
      } else {
#pragma unroll
        for (int kk = 0; kk < AS; kk++) {
          if (valid && k == kk) {
            double *dst = rowbuf + roff[lr] + slot * 9;
#pragma unroll
            for (int i = 0; i < 9; i++) dst[i] += val[i];
            if (b < 3) frow[3 * lr + b] += fadd;
          }
          __syncwarp();
        }
      }
      __syncwarp();
    }
  }
  TK_T(1);
  cp_async_wait_all();
  __syncthreads();
  TK_T(6);
  // ---- stream the finished block rows out as three scalar CSR rows of length 3*deg each,
  //      adding the mass term H(3i+d,3j+d) += M_ij/h (SyncedNewton.cu:213-259)
  for (int r = warp; r < nrows; r += W) {
    const int64_t base = 9 * (int64_t)rstart[r];
    const int rowlen = (roff[r + 1] - roff[r]) / 3;
    const double *src = rowbuf + roff[r];
    const double *mr = mrow + roff[r] / 9;
#pragma unroll
    for (int d = 0; d < 3; d++) {
#pragma unroll 2
      for (int c = lane; c < rowlen; c += 32) {
        const int s = c / 3, e2 = c - 3 * s;
        double v = src[s * 9 + d * 3 + e2];
        if (e2 == d) v += mr[s] * p.inv_h;
        p.H[base + d * rowlen + c] = v;
      }
    }
    if (lane < 3) p.f_int[(int64_t)grow[r] * 3 + lane] = frow[3 * r + lane];
  }
  TK_T(7);
}

}  // namespace tlfea
