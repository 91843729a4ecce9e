// tile_kernel.cuh -- tangent + H assembly by spatial row tiles (SVK, optional Kelvin-Voigt).
//
// Replaces compute_hessian_assemble_csr (FEAT10DataFunc.cuh:512-791) + assemble_sparse_hessian_mass
// (SyncedNewton.cu:213-259) + the force scatter (FEAT10DataFunc.cuh:397-458).
//
// One CTA owns a TILE of H block rows (a Morton-compact set of coefficients, tiles.cu) and keeps
// them in shared memory.  Its warps walk the tile's incident elements; for one element a warp
//   0. prefetches the NEXT element's gradN / Gauss-point records with cp.async while it works,
//   1. stages, per Gauss point and local node c:  u_c = F h_c  and  h_c,  and the symmetric
//      3x3  M_q = c3 dV F F^T + c4 dV (lambda trE - mu) I   (F, dV FF^T from the Gauss-point record),
//   2. computes the 3x3 blocks K_ab for the rows a that live in this tile: lane = (row slot, b),
//      15 accumulators  A = sum_q (dV u_a) u_b^T,  Sg = sum_q (h_a.h_b) M_q,
//      K_ab = c1 A + c2 A^T + Sg + c4 mu tr(A) I     (SVK.cuh:35-55 and the viscous block
//      FEAT10DataFunc.cuh:695-762 folded into c1..c3),
//   3. commits the blocks to the row buffer IN RANK ORDER: every contribution to a row carries
//      its position in the row's commit schedule (tiles.cu: conflict-free rounds); a warp waits
//      until the row's turn counter reaches its rank.  Fixed order => bit-reproducible, no FP64
//      atomics.
// Finished rows (+ mass/h on the block diagonals) stream to H once.  DRAM traffic ~ gradN + records + H.
#pragma once
#include <cub/block/block_scan.cuh>

#include "common.cuh"

namespace tlfea {

struct TileArgs {
  int n_tiles;
  const int *trow, *tile_ptr, *te_ptr, *te_elem;
  const uint8_t *te_lrow, *te_rank;
  const uint16_t *slot_e;
  const int *off;
  const double *mass;
  const double *rec;  // [n_elem*Q][16]
  const double *gradN, *detJ, *fe;
  QuadW qw;
  double inv_h, c1, c2, c3, c4, mu;
  double *H, *f_int;
  int *err;
  int buf_doubles;
};

constexpr int TK_MAX_ROWS = 254;
constexpr int TK_SPIN_LIMIT = 1 << 20;

template <int S, int Q, int W>
struct TileShape {
  static constexpr int NODE = Q * S * 6;                           // {u0,u1,u2,h0,h1,h2} per (q, node)
  static constexpr int STG = NODE + Q * 6 + ((Q + 1) & ~1);        // + M_q + dV_q
  static constexpr int RAW = ((Q * 3 * S + Q * 16 + Q) + 1) & ~1;  // gradN | records | detJ
  static size_t smem_bytes(int buf_doubles) {
    return ((size_t)buf_doubles + 256 * 3 + (size_t)W * (STG + RAW)) * sizeof(double) +
           (257 + 256 + 256 + 256) * sizeof(int);
  }
};

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

template <int S, int Q, int W>
__global__ void __launch_bounds__(W * 32, 1) tile_kernel(TileArgs p) {
  static_assert(W * 32 >= 256, "block scan covers up to 256 rows");
  static_assert((Q * 3 * S) % 2 == 0, "gradN block of one element must be 16-byte granular");
  using Sh = TileShape<S, Q, W>;
  constexpr int AS = 32 / S;  // rows of one element handled per pass
  constexpr int NODE = Sh::NODE, STG = Sh::STG, RAW = Sh::RAW;
  constexpr int NG = Q * 3 * S;  // gradN doubles per element
  extern __shared__ __align__(16) double sm[];
  double *rowbuf = sm;
  double *frow = rowbuf + p.buf_doubles;
  double *stage_all = frow + 256 * 3;
  double *raw_all = stage_all + W * STG;
  int *roff = reinterpret_cast<int *>(raw_all + W * RAW);
  int *grow = roff + 257;
  int *rstart = grow + 256;
  volatile int *turn = rstart + 256;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x;
  const int r0 = p.tile_ptr[tile], nrows = p.tile_ptr[tile + 1] - r0;
  const int e0 = p.te_ptr[tile], ne = p.te_ptr[tile + 1] - e0;

  double *stg = stage_all + warp * STG;
  double *node = stg, *Mq = stg + NODE, *dVq = Mq + Q * 6;
  double *raw = raw_all + warp * RAW;

  // raw element data -> shared memory, asynchronously (gradN and records in 16-B chunks, detJ 8-B)
  auto prefetch = [&](int e) {
    const double *g = p.gradN + (int64_t)e * NG;
    const double *r = p.rec + (int64_t)e * Q * 16;
    const double *dj = p.detJ + (int64_t)e * Q;
    for (int i = lane; i < NG / 2; i += 32) cp_async16(raw + 2 * i, g + 2 * i);
    for (int i = lane; i < Q * 8; i += 32) cp_async16(raw + NG + 2 * i, r + 2 * i);
    if (lane < Q) cp_async8(raw + NG + Q * 16 + lane, dj + lane);
    cp_async_commit();
  };
  // first entry of this warp: start its loads before the prologue
  int e_next = -1, lrow_next = 255, rank_next = 0;
  if (warp < ne) {
    e_next = p.te_elem[e0 + warp];
    if (lane < S) {
      lrow_next = p.te_lrow[(int64_t)(e0 + warp) * S + lane];
      rank_next = p.te_rank[(int64_t)(e0 + warp) * S + lane];
    }
    prefetch(e_next);
  }

  // ---- row table: global row, CSR start, buffer offset (exclusive scan of 9*deg)
  {
    int row = 0, rs = 0, sz = 0;
    if (tid < nrows) {
      row = p.trow[r0 + tid];
      rs = p.off[row];
      sz = 9 * (p.off[row + 1] - rs);
    }
    using Scan = cub::BlockScan<int, W * 32>;
    __shared__ typename Scan::TempStorage tmp;
    int excl;
    Scan(tmp).ExclusiveSum(sz, excl);
    if (tid < 256) {
      if (tid < nrows) {
        grow[tid] = row;
        rstart[tid] = rs;
        roff[tid] = excl;
        turn[tid] = 0;
        frow[3 * tid] = frow[3 * tid + 1] = frow[3 * tid + 2] = 0.0;
      }
      if (tid == nrows) roff[tid] = excl;
    }
  }
  __syncthreads();
  // ---- zero the row buffer (the mass term M_ij/h is added while streaming the rows out)
  {
    const int total = roff[nrows];
    double2 *z2 = reinterpret_cast<double2 *>(rowbuf);
    for (int i = tid; i < (total + 1) / 2; i += W * 32) z2[i] = make_double2(0.0, 0.0);
  }
  __syncthreads();

  const int k = lane / S, b = lane - k * S;
  const bool lane_on = lane < AS * S;

  for (int j = warp; j < ne; j += W) {
    const int e = e_next;
    const int my_lrow = lrow_next, my_rank = rank_next;
    const unsigned mask = __ballot_sync(0xffffffffu, my_lrow != 255);
    const int n_a = __popc(mask);
    // ---- stage the element from the prefetched raw data
    cp_async_wait_all();
    __syncwarp();
    for (int it = lane; it < Q * S; it += 32) {
      const int q = it / S, c = it - q * S;
      const double *g = raw + q * 3 * S;
      const double h0 = g[c], h1 = g[S + c], h2 = g[2 * S + c];
      const double2 *r2 = reinterpret_cast<const double2 *>(raw + NG + q * 16);
      const double2 f01 = r2[0], f23 = r2[1], f45 = r2[2], f67 = r2[3];
      const double f8 = raw[NG + q * 16 + 8];
      double2 *N2 = reinterpret_cast<double2 *>(node + (q * S + c) * 6);
      N2[0] = make_double2(f01.x * h0 + f01.y * h1 + f23.x * h2, f23.y * h0 + f45.x * h1 + f45.y * h2);
      N2[1] = make_double2(f67.x * h0 + f67.y * h1 + f8 * h2, h0);
      N2[2] = make_double2(h1, h2);
    }
    if (lane < Q) {
      const double *r = raw + NG + lane * 16;
      const double dg = p.c4 * r[15];
      double2 *M2 = reinterpret_cast<double2 *>(Mq + lane * 6);
      M2[0] = make_double2(p.c3 * r[9] + dg, p.c3 * r[10]);
      M2[1] = make_double2(p.c3 * r[11], p.c3 * r[12] + dg);
      M2[2] = make_double2(p.c3 * r[13], p.c3 * r[14] + dg);
      dVq[lane] = raw[NG + Q * 16 + lane] * p.qw.w[lane];
    }
    __syncwarp();
    // ---- next entry: metadata to registers, raw data to shared memory (in flight during the passes)
    if (j + W < ne) {
      e_next = p.te_elem[e0 + j + W];
      if (lane < S) {
        lrow_next = p.te_lrow[(int64_t)(e0 + j + W) * S + lane];
        rank_next = p.te_rank[(int64_t)(e0 + j + W) * S + lane];
      }
      prefetch(e_next);
    }

    // ---- phase 1: all blocks of this element that belong to the tile, kept in registers
    constexpr int MAXP = (S + AS - 1) / AS;
    double val[MAXP][9];
    double fadd[MAXP];
    int slot[MAXP], lrp[MAXP], rkp[MAXP];
#pragma unroll
    for (int pass = 0; pass < MAXP; pass++) {
      if (pass * AS >= n_a) break;  // warp-uniform
      const int idx = pass * AS + k;
      const bool valid = lane_on && idx < n_a;
      const int a = valid ? (int)__fns(mask, 0, idx + 1) : 0;
      const int lr_s = __shfl_sync(0xffffffffu, my_lrow, a);
      rkp[pass] = __shfl_sync(0xffffffffu, my_rank, a);
      lrp[pass] = valid ? lr_s : -1;
      slot[pass] = 0;
      fadd[pass] = 0.0;
      if (valid) {
        slot[pass] = p.slot_e[((int64_t)e * S + a) * S + b];
        if (b < 3) fadd[pass] = p.fe[((int64_t)e * S + a) * 3 + b];
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
        val[pass][0] = p.c1 * A[0] + p.c2 * A[0] + Sg[0] + dg;
        val[pass][1] = p.c1 * A[1] + p.c2 * A[3] + Sg[1];
        val[pass][2] = p.c1 * A[2] + p.c2 * A[6] + Sg[2];
        val[pass][3] = p.c1 * A[3] + p.c2 * A[1] + Sg[1];
        val[pass][4] = p.c1 * A[4] + p.c2 * A[4] + Sg[3] + dg;
        val[pass][5] = p.c1 * A[5] + p.c2 * A[7] + Sg[4];
        val[pass][6] = p.c1 * A[6] + p.c2 * A[2] + Sg[2];
        val[pass][7] = p.c1 * A[7] + p.c2 * A[5] + Sg[4];
        val[pass][8] = p.c1 * A[8] + p.c2 * A[8] + Sg[5] + dg;
      }
    }
    // ---- phase 2: ordered commits (wait for each row's turn; rows of one element are distinct)
#pragma unroll
    for (int pass = 0; pass < MAXP; pass++) {
      if (pass * AS >= n_a) break;
      const int lr = lrp[pass], rk = rkp[pass];
      const bool valid = lr >= 0;
      if (valid && b == 0) {
        int spins = 0;
        while (turn[lr] != rk) {
          __nanosleep(20);
          if (((++spins) & 255) == 0 && (spins > TK_SPIN_LIMIT || *((volatile int *)p.err) != 0)) {
            atomicExch(p.err, 1);
            break;
          }
        }
        __threadfence_block();
      }
      __syncwarp();
      if (valid) {
        double *dst = rowbuf + roff[lr] + slot[pass] * 9;
#pragma unroll
        for (int i = 0; i < 9; i++) dst[i] += val[pass][i];
        if (b < 3) frow[3 * lr + b] += fadd[pass];
      }
      __syncwarp();
      if (valid && b == 0) {
        __threadfence_block();
        turn[lr] = rk + 1;
      }
    }
  }
  __syncthreads();
  // ---- stream the finished block rows out as three scalar CSR rows of length 3*deg each,
  //      adding the mass term H(3i+d,3j+d) += M_ij/h (SyncedNewton.cu:213-259)
  for (int r = warp; r < nrows; r += W) {
    const int rs = rstart[r];
    const int64_t base = 9 * (int64_t)rs;
    const int rowlen = (roff[r + 1] - roff[r]) / 3;
    const double *src = rowbuf + roff[r];
#pragma unroll
    for (int d = 0; d < 3; d++) {
#pragma unroll 2
      for (int c = lane; c < rowlen; c += 32) {
        const int s = c / 3, e2 = c - 3 * s;
        double v = src[s * 9 + d * 3 + e2];
        if (e2 == d) v += p.mass[rs + s] * p.inv_h;
        p.H[base + d * rowlen + c] = v;
      }
    }
    if (lane < 3) p.f_int[(int64_t)grow[r] * 3 + lane] = frow[3 * r + lane];
  }
}

}  // namespace tlfea
