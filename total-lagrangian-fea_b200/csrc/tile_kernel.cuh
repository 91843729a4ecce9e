// tile_kernel.cuh -- tangent + H assembly by spatial row tiles (SVK, optional Kelvin-Voigt).
//
// Replaces compute_hessian_assemble_csr (FEAT10DataFunc.cuh:512-791) + assemble_sparse_hessian_mass
// (SyncedNewton.cu:213-259) + the force scatter (FEAT10DataFunc.cuh:397-458).
//
// One CTA owns a TILE of H block rows (a Morton-compact set of coefficients, tiles.cu) and keeps
// them in shared memory.  Its warps walk the tile's incident elements; for one element a warp
//   1. stages, per Gauss point and local node c:  u_c = F h_c,  dV u_c,  h_c  and the symmetric
//      3x3  M_q = c3 dV F F^T + c4 dV (lambda trE - mu) I   (F, dV FF^T from the Gauss-point record),
//   2. computes the 3x3 blocks K_ab for the rows a that live in this tile: lane = (row slot, b),
//      15 accumulators  A = sum_q (dV u_a) u_b^T,  Sg = sum_q (h_a.h_b) M_q   (18 DFMA / point),
//      K_ab = c1 A + c2 A^T + Sg + c4 mu tr(A) I     (SVK.cuh:35-55 and the viscous block
//      FEAT10DataFunc.cuh:695-762 folded into c1..c3),
//   3. commits the blocks to the row buffer IN RANK ORDER: every contribution to a row carries
//      its position in the row's (ascending element) visit list; a warp waits until the row's
//      turn counter reaches its rank.  Fixed order => bit-reproducible, no FP64 atomics.
// Finished rows (mass/h included) stream to H once.  DRAM traffic ~ gradN + records + H.
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
constexpr int TK_SPIN_LIMIT = 1 << 22;

template <int S, int Q, int W>
struct TileShape {
  static constexpr int NA = Q * S * 6;           // doubles of one node view
  static constexpr int STG = 2 * NA + Q * 6;     // a-side view | b-side view | M
  static size_t smem_bytes(int buf_doubles) {
    return ((size_t)buf_doubles + 256 * 3 + (size_t)W * STG) * sizeof(double) + (257 + 256 + 256) * sizeof(int);
  }
};

template <int S, int Q, int W>
__global__ void __launch_bounds__(W * 32, 1) tile_kernel(TileArgs p) {
  static_assert(W * 32 >= 256, "block scan covers up to 256 rows");
  constexpr int AS = 32 / S;  // rows of one element handled per pass
  constexpr int NA = TileShape<S, Q, W>::NA, STG = TileShape<S, Q, W>::STG;
  extern __shared__ __align__(16) double sm[];
  double *rowbuf = sm;
  double *frow = rowbuf + p.buf_doubles;
  double *stage_all = frow + 256 * 3;
  int *roff = reinterpret_cast<int *>(stage_all + W * STG);
  int *grow = roff + 257;
  volatile int *turn = grow + 256;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x;
  const int r0 = p.tile_ptr[tile], nrows = p.tile_ptr[tile + 1] - r0;

  // ---- row table: global row, buffer offset (exclusive scan of 9*deg)
  {
    int row = 0, sz = 0;
    if (tid < nrows) {
      row = p.trow[r0 + tid];
      sz = 9 * (p.off[row + 1] - p.off[row]);
    }
    using Scan = cub::BlockScan<int, W * 32>;
    __shared__ typename Scan::TempStorage tmp;
    int excl;
    Scan(tmp).ExclusiveSum(sz, excl);
    if (tid < 256) {
      if (tid < nrows) {
        grow[tid] = row;
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

  double *stg = stage_all + warp * STG;
  double *nodeA = stg, *nodeB = stg + NA, *Mq = stg + 2 * NA;
  const int k = lane / S, b = lane - k * S;
  const bool lane_on = lane < AS * S;
  const int e0 = p.te_ptr[tile], ne = p.te_ptr[tile + 1] - e0;

  for (int j = warp; j < ne; j += W) {
    const int e = p.te_elem[e0 + j];
    int my_lrow = 255, my_rank = 0;
    if (lane < S) {
      my_lrow = p.te_lrow[(int64_t)(e0 + j) * S + lane];
      my_rank = p.te_rank[(int64_t)(e0 + j) * S + lane];
    }
    const unsigned mask = __ballot_sync(0xffffffffu, my_lrow != 255);
    const int n_a = __popc(mask);
    // ---- stage the element
    __syncwarp();
    for (int it = lane; it < Q * S; it += 32) {
      const int q = it / S, c = it - q * S;
      const int64_t t = (int64_t)e * Q + q;
      const double *g = p.gradN + t * 3 * S;
      const double h0 = g[c], h1 = g[S + c], h2 = g[2 * S + c];
      const double2 *r2 = reinterpret_cast<const double2 *>(p.rec + t * 16);
      const double2 f01 = r2[0], f23 = r2[1], f45 = r2[2], f67 = r2[3];
      const double f8 = p.rec[t * 16 + 8];
      const double dV = p.detJ[t] * p.qw.w[q];
      const double u0 = f01.x * h0 + f01.y * h1 + f23.x * h2;
      const double u1 = f23.y * h0 + f45.x * h1 + f45.y * h2;
      const double u2 = f67.x * h0 + f67.y * h1 + f8 * h2;
      double2 *A2 = reinterpret_cast<double2 *>(nodeA + (q * S + c) * 6);
      double2 *B2 = reinterpret_cast<double2 *>(nodeB + (q * S + c) * 6);
      A2[0] = make_double2(dV * u0, dV * u1);
      A2[1] = make_double2(dV * u2, h0);
      A2[2] = make_double2(h1, h2);
      B2[0] = make_double2(u0, u1);
      B2[1] = make_double2(u2, h0);
      B2[2] = make_double2(h1, h2);
    }
    if (lane < Q) {
      const double *r = p.rec + ((int64_t)e * Q + lane) * 16;
      const double dg = p.c4 * r[15];
      double2 *M2 = reinterpret_cast<double2 *>(Mq + lane * 6);
      M2[0] = make_double2(p.c3 * r[9] + dg, p.c3 * r[10]);
      M2[1] = make_double2(p.c3 * r[11], p.c3 * r[12] + dg);
      M2[2] = make_double2(p.c3 * r[13], p.c3 * r[14] + dg);
    }
    __syncwarp();

    for (int pass = 0; pass * AS < n_a; pass++) {
      const int idx = pass * AS + k;
      const bool valid = lane_on && idx < n_a;
      const int a = valid ? (int)__fns(mask, 0, idx + 1) : 0;
      const int lr = __shfl_sync(0xffffffffu, my_lrow, a);
      const int rk = __shfl_sync(0xffffffffu, my_rank, a);
      double val[9];
      int slot = 0;
      double fadd = 0.0;
      if (valid) {
        slot = p.slot_e[((int64_t)e * S + a) * S + b];
        if (b < 3) fadd = p.fe[((int64_t)e * S + a) * 3 + b];
        double A[9], Sg[6];
#pragma unroll
        for (int i = 0; i < 9; i++) A[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; i++) Sg[i] = 0.0;
#pragma unroll
        for (int q = 0; q < Q; q++) {
          const double2 *a2 = reinterpret_cast<const double2 *>(nodeA + (q * S + a) * 6);
          const double2 *b2 = reinterpret_cast<const double2 *>(nodeB + (q * S + b) * 6);
          const double2 *m2 = reinterpret_cast<const double2 *>(Mq + q * 6);
          const double2 a01 = a2[0], a23 = a2[1], a45 = a2[2];
          const double2 b01 = b2[0], b23 = b2[1], b45 = b2[2];
          const double2 m01 = m2[0], m23 = m2[1], m45 = m2[2];
          const double hab = a23.y * b23.y + a45.x * b45.x + a45.y * b45.y;
          A[0] += a01.x * b01.x; A[1] += a01.x * b01.y; A[2] += a01.x * b23.x;
          A[3] += a01.y * b01.x; A[4] += a01.y * b01.y; A[5] += a01.y * b23.x;
          A[6] += a23.x * b01.x; A[7] += a23.x * b01.y; A[8] += a23.x * b23.x;
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
      // ---- ordered commit: wait for this row's turn
      if (valid && b == 0) {
        int spins = 0;
        while (turn[lr] != rk) {
          if (((++spins) & 1023) == 0 && (spins > TK_SPIN_LIMIT || *((volatile int *)p.err) != 0)) {
            atomicExch(p.err, 1);
            break;
          }
        }
        __threadfence_block();
      }
      __syncwarp();
      if (valid) {
        double *dst = rowbuf + roff[lr] + slot * 9;
#pragma unroll
        for (int i = 0; i < 9; i++) dst[i] += val[i];
        if (b < 3) frow[3 * lr + b] += fadd;
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
    const int row = grow[r];
    const int rs = p.off[row];
    const int64_t base = 9 * (int64_t)rs;
    const int rowlen = (roff[r + 1] - roff[r]) / 3;
    const double *src = rowbuf + roff[r];
#pragma unroll
    for (int d = 0; d < 3; d++) {
      for (int c = lane; c < rowlen; c += 32) {
        const int s = c / 3, e2 = c - 3 * s;
        double v = src[s * 9 + d * 3 + e2];
        if (e2 == d) v += p.mass[rs + s] * p.inv_h;
        p.H[base + d * rowlen + c] = v;
      }
    }
    if (lane < 3) p.f_int[(int64_t)row * 3 + lane] = frow[3 * r + lane];
  }
}

}  // namespace tlfea
