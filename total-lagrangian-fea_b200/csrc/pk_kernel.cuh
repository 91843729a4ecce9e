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
// Write-out, in place into the CSR values of H (lists: patch_plan.cuh; they are read straight from global memory,
// pulled into L2 by a bulk prefetch one patch ahead):
//   (a) as soon as a warp has finished its own elements it writes the H entries of the blocks it holds alone (83 %
//       of the entries on a Kuhn block) -- no CTA barrier, the other warps keep computing;
//   (b) after a CTA barrier the node pairs held by several warps are reduced in fixed warp order (merge rounds);
//   (c) all threads write the entries of the merged blocks.
// An entry either starts from M_ij/h (first touch of the H block in the colour sequence: plain stores) or is added to
// the value already in H (later touch: red.global.add.f64, performed by the L2 -- one reduction per address per
// launch, launches are stream-ordered, so the order per H entry is the colour order).  Patches of one launch share
// no coefficient (graph colouring): no two CTAs touch the same address, order per H entry = (colour, warp, element).
#pragma once
#include "ek_kernel.cuh"
#include "patch_plan.cuh"

namespace tlfea {

struct PKArgs {
  const PatchDesc *desc;  // first patch of this launch
  int n_patches;          // patches of this launch; CTA b takes b, b + gridDim.x, ...
  int debug;              // tuning aid (TLFEA_PK_DEBUG): 1 no write-out, 2 no accumulation, 4 no block step, 8 no Gauss-point step
  const PElem *elems;
  const unsigned char *blob;
  const double *x, *y, *z;
  const double *v;  // interleaved velocity (Kelvin-Voigt) or nullptr
  const double *gradN, *detJ;
  const double *aff;  // [n_elem][10] J^-T | detJ of affine elements (plan record bit 0 of pad0), or nullptr
  double dn[150];     // dN_a/dxi_k at the quadrature points, [q][a][k]: gradN of an affine element = J^-T dn
  QuadW qw;
  double lambda, mu, eta, lambda_d;
  double c1, c2, c3, c4;  // c1 = h lambda + lambda_d, c2 = c3 = h mu + eta, c4 = h
  double mu10, mu01, kappa;
  double inv_h;
  double *fe;  // [n_elem][10][3]
  double *H;
};

// Gauss-point results of the SVK path: row stride of the per-(element, point) force-partial and u rows, both laid out
// [a][i].  The Gauss-point step stores from lane = (element, point, row i): address = idx * 35 + 3 a + i, i.e. bank
// 3 idx + i + const -- the 16 lanes of a half warp hit 16 different banks (the stride 33 of ek_kernel.cuh is for its
// 10-lane step and gave a 9-way conflict here).
constexpr int PK_GQ = 35;
#ifndef PK_FU_OWN
#define PK_FU_OWN 4  // entries per thread per trip of the write-out: a warp's own segment / the merged blocks
#endif
#ifndef PK_FU_DEP
#define PK_FU_DEP 2
#endif

// Shared memory of a CTA: [accumulators x W][prefetch buffers x W][scratch x W].  The scratch areas (Gauss-point
// results of the pair in flight) are contiguous over the warps: the write-out lists of the patch land there.
template <bool DAMP, int MODEL>
struct PKShape {
  static constexpr bool MR = (MODEL == TL_MODEL_MR);
  static constexpr int ACC_D = PK_CAPU * 9;
  static constexpr int H_D = 300;   // gradN of an element pair
  static constexpr int DJ_D = 10;
  static constexpr int X_D = 60;
  static constexpr int PE_D = 32;   // two plan records
  static constexpr int U_D = MR ? 2 * 5 * EK_TS : 2 * 5 * PK_GQ;
  static constexpr int REC_D = 0;  // M_q | dV of the SVK path live in the spare doubles [30, 35) of the u and force-partial rows
  static constexpr int FP_D = MR ? 2 * 5 * EK_GQM : 2 * 5 * PK_GQ;
  static constexpr int OFF_H = 0;
  static constexpr int OFF_DJ = OFF_H + 2 * H_D;
  static constexpr int OFF_X = OFF_DJ + 2 * DJ_D;
  static constexpr int OFF_V = OFF_X + 2 * X_D;
  static constexpr int OFF_PE = OFF_V + (DAMP ? 2 * X_D : 0);
  static constexpr int PREF_D = OFF_PE + 2 * PE_D;  // prefetch doubles per warp (even)
  static constexpr int OFF_U = 0;
  static constexpr int OFF_REC = OFF_U + (U_D + 1) / 2 * 2;
  static constexpr int OFF_FP = OFF_REC + (REC_D + 1) / 2 * 2;
  static constexpr int SCR_D = OFF_FP + (FP_D + 1) / 2 * 2;  // scratch doubles per warp (even)
  static constexpr size_t smem_bytes(int warps) { return (size_t)warps * (ACC_D + PREF_D + SCR_D) * sizeof(double); }
  static constexpr int scratch_bytes(int warps) { return warps * SCR_D * (int)sizeof(double); }
};

// H[addr] += v performed by the L2 (no return value, nothing to wait for)
__device__ __forceinline__ void red_add_f64(double *addr, double v) {
  asm volatile("red.global.add.f64 [%0], %1;\n" ::"l"(addr), "d"(v) : "memory");
}

__device__ __forceinline__ void bulk_prefetch_l2(const void *gmem, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(gmem), "r"(bytes) : "memory");
}


// Write-out of one entry segment (patch_plan.cuh): thread = (entry slot, column e) of a 3x3 block of H, three rows d
// per thread: consecutive lanes store consecutive doubles of a scalar CSR row (an entry-per-thread mapping was
// measured: 2.3x the store sectors, 10 % slower).  The entries are read from shared memory (SMEM: the segment was
// staged by a TMA bulk copy) or straight from global memory (lists that do not fit the staging area; the blob was
// pulled into L2 when the patch started), FU per thread per trip with the next trip's loads in flight.
//   first touches of an H block in the colour sequence: H block = (M_ij / h) I + sum, plain stores;
//   later touches: H += sum as L2 reductions.  One reduction per address per launch (patches of a launch share no
//   coefficient) and launches are stream-ordered, so the additions per H entry happen in colour order: the same
//   IEEE additions a load-add-store would do, but with no load to wait for.
template <int FU, bool SMEM>
__device__ __forceinline__ void pk_write_segment(const unsigned char *seg, int nf, int nl, int slot, int nslots, int we,
                                                 const double *acc0, double *Hb, double inv_h,
                                                 unsigned hmask = 0xffffffffu) {
  // everything that depends on the column e is per-thread constant and branch-free (lanes of a warp differ in e)
  const unsigned acc_s = (unsigned)__cvta_generic_to_shared(acc0) + 8u * we;  // not transposed: + e, row stride 3
  const unsigned we16 = 16u * we;                                              // transposed: + 3 e, row stride 1
  const double m0 = (we == 0) ? inv_h : 0.0, m1 = (we == 1) ? inv_h : 0.0, m2 = (we == 2) ? inv_h : 0.0;
  double *Hw = Hb + we;
  auto block_col = [&](unsigned meta, double v[3]) {  // column e of the entry's block, rows 0..2
    const unsigned tr = meta >> 31;
    const unsigned a = acc_s + ((meta >> 13) & 0x1fff8u) + tr * we16, st = 24u - 16u * tr;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v[0]) : "r"(a));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v[1]) : "r"(a + st));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v[2]) : "r"(a + 2u * st));
  };
  // staged lists are read with explicit shared-memory loads: through the generic pointer they would be generic loads
  // (longer latency, long scoreboard)
  auto ld4 = [](const uint4 *q) {
    if (!SMEM) return __ldg(q);
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"((unsigned)__cvta_generic_to_shared(q)));
    return v;
  };
  auto ld2 = [](const uint2 *q) {
    if (!SMEM) return __ldg(q);
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"((unsigned)__cvta_generic_to_shared(q)));
    return v;
  };
  auto put_first = [&](const uint4 e, const double v[3]) {  // H block = (M_ij / h) I + sum
    const double m = __hiloint2double((int)e.w, (int)e.z);
    double *hp = Hw + (e.x & hmask);
    const unsigned rl = e.y & 0xffffu;
    hp[0] = fma(m, m0, v[0]);
    hp[rl] = fma(m, m1, v[1]);
    hp[2 * rl] = fma(m, m2, v[2]);
  };
  auto put_later = [&](const uint2 e, const double v[3]) {  // H block += sum
    double *hp = Hw + (e.x & hmask);
    const unsigned rl = e.y & 0xffffu;
    red_add_f64(hp, v[0]);
    red_add_f64(hp + rl, v[1]);
    red_add_f64(hp + 2 * rl, v[2]);
  };
  {
    const uint4 *ef = reinterpret_cast<const uint4 *>(seg);
    int k = slot;
    for (; k + (FU - 1) * nslots < nf; k += FU * nslots) {  // full trips: FU independent entries, no guards
      uint4 e[FU];
      double v[FU][3];
#pragma unroll
      for (int u = 0; u < FU; u++) e[u] = ld4(ef + k + u * nslots);
#pragma unroll
      for (int u = 0; u < FU; u++) block_col(e[u].y, v[u]);
#pragma unroll
      for (int u = 0; u < FU; u++) put_first(e[u], v[u]);
    }
    for (; k < nf; k += nslots) {
      const uint4 e = ld4(ef + k);
      double v[3];
      block_col(e.y, v);
      put_first(e, v);
    }
  }
  {
    const uint2 *el = reinterpret_cast<const uint2 *>(seg + (size_t)nf * 16);
    int k = slot;
    for (; k + (FU - 1) * nslots < nl; k += FU * nslots) {
      uint2 e[FU];
      double v[FU][3];
#pragma unroll
      for (int u = 0; u < FU; u++) e[u] = ld2(el + k + u * nslots);
#pragma unroll
      for (int u = 0; u < FU; u++) block_col(e[u].y, v[u]);
#pragma unroll
      for (int u = 0; u < FU; u++) put_later(e[u], v[u]);
    }
    for (; k < nl; k += nslots) {
      const uint2 e = ld2(el + k);
      double v[3];
      block_col(e.y, v);
      put_later(e, v);
    }
  }
}

template <int W, bool DAMP, bool TANGENT, int MODEL>
__global__ void __launch_bounds__(W * 32, (W <= 4) ? 2 : 1) pk_kernel(PKArgs p) {
  using Sh = PKShape<DAMP, MODEL>;
  constexpr bool MR = (MODEL == TL_MODEL_MR);
  constexpr int NT = W * 32;
  extern __shared__ __align__(16) double pk_smem[];
  __shared__ uint64_t s_bar[W][2];
  __shared__ uint64_t s_bar_wo[W];  // arrival of warp w's write-out chunk in its scratch area
  __shared__ PatchDesc s_desc[2];  // this CTA's current and next patch
  __shared__ double s_dn[150];
  // programmatic dependent launch (patch_asm.cu): let the next colour's CTAs be scheduled as ours exit
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  bool dep_waited = false;  // this thread has waited for the previous launch (needed before its first access to H)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double *acc = pk_smem + (size_t)warp * Sh::ACC_D;
  double *wb = pk_smem + (size_t)W * Sh::ACC_D + (size_t)warp * Sh::PREF_D;  // prefetch buffers of this warp
  double *scr0 = pk_smem + (size_t)W * (Sh::ACC_D + Sh::PREF_D);
  double *ws = scr0 + (size_t)warp * Sh::SCR_D;
  double *s_u = ws + Sh::OFF_U, *s_fp = ws + Sh::OFF_FP;

  auto load_desc = [&](int pk, int slot) {  // threads 0..15
    if (pk < p.n_patches)
      reinterpret_cast<int *>(&s_desc[slot])[threadIdx.x] = reinterpret_cast<const int *>(p.desc + pk)[threadIdx.x];
  };
  if (threadIdx.x < 16) {
    load_desc(blockIdx.x, 0);
    load_desc(blockIdx.x + gridDim.x, 1);
  }
  if (lane == 0) {
    mbar_init(&s_bar[warp][0], 1);
    mbar_init(&s_bar[warp][1], 1);
    mbar_init(&s_bar_wo[warp], 1);
  }
  for (int i = threadIdx.x; i < 150; i += NT) s_dn[i] = p.dn[i];
  __syncthreads();
  if (TANGENT && threadIdx.x == 0)  // the write-out lists are needed last: pull them into L2 now
    bulk_prefetch_l2(p.blob + (size_t)s_desc[0].blob_off * 16,
                     (unsigned)s_desc[0].blob_bytes);

  // ---- prefetch helpers (per warp): plan records / data of the element pair [kk, kk + ne)
  auto issue_pe = [&](int kk, int ne, int buf) {
    double *dst = wb + Sh::OFF_PE + buf * Sh::PE_D;
    const char *src = reinterpret_cast<const char *>(p.elems + kk);
    if (lane < ne * 8) cp_async16(reinterpret_cast<char *>(dst) + lane * 16, src + lane * 16);
  };
  auto issue_data = [&](int ne, int buf) {
    const PElem *pe = reinterpret_cast<const PElem *>(wb + Sh::OFF_PE + buf * Sh::PE_D);
    double *dh = wb + Sh::OFF_H + buf * Sh::H_D;
    // affine elements (plan record flag): 80 bytes of J^-T | detJ instead of 1200 bytes of gradients + 40 of detJ
    const bool af0 = p.aff && (pe[0].pad0 & 1), af1 = p.aff && ne > 1 && (pe[1].pad0 & 1);
    if (lane == 0) {
      const unsigned b = (unsigned)__cvta_generic_to_shared(&s_bar[warp][buf]);
      const unsigned n_bulk = (af0 ? 0u : 1u) + ((ne > 1 && !af1) ? 1u : 0u);
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // the buffer was last read by ordinary loads
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(n_bulk * 1200u) : "memory");
      for (int el = 0; el < ne; el++) {
        if (el == 0 ? af0 : af1) continue;
        const unsigned s = (unsigned)__cvta_generic_to_shared(dh + el * 150);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(s),
                     "l"(p.gradN + (int64_t)pe[el].elem * 150), "r"(1200u), "r"(b)
                     : "memory");
      }
    }
    if (lane < ne * 5) {
      const int el = lane / 5, q = lane - 5 * el;
      if (el == 0 ? af0 : af1)
        cp_async16(dh + el * 150 + 2 * q, p.aff + (int64_t)pe[el].elem * 10 + 2 * q);
      else
        cp_async8(wb + Sh::OFF_DJ + buf * Sh::DJ_D + lane, p.detJ + (int64_t)pe[el].elem * 5 + q);
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

  // The warp's element pairs form ONE stream over all patches of this CTA: while pair j computes, the data of
  // pair j+1 and the plan records of pair j+2 are in flight, across patch boundaries (the next patch's descriptor is
  // already in shared memory).  gi counts the pairs: buffer gi & 1, mbarrier phase (gi >> 1) & 1.
  int gi = 0;
  int pe_kk0 = -1, pe_kk1 = -1;  // first plan record held (or in flight) in each record buffer
  int dt_kk0 = -1, dt_kk1 = -1;  // ... whose data has been requested into each data buffer
  auto pe_get = [&](int b) { return b ? pe_kk1 : pe_kk0; };
  auto dt_get = [&](int b) { return b ? dt_kk1 : dt_kk0; };
  auto pe_set = [&](int b, int v) { if (b) pe_kk1 = v; else pe_kk0 = v; };
  auto dt_set = [&](int b, int v) { if (b) dt_kk1 = v; else dt_kk0 = v; };
  int n_done = 0;           // patches this CTA has finished (blob barrier phase)
  for (int pk = blockIdx.x; pk < p.n_patches; pk += gridDim.x, n_done++) {
    const PatchDesc &dc = s_desc[n_done & 1], &dn = s_desc[(n_done + 1) & 1];
    const bool has_next = pk + (int)gridDim.x < p.n_patches;
    const int k0 = dc.sub_ptr[warp], k1 = dc.sub_ptr[warp + 1];
    const int n_it = (k1 - k0 + 1) >> 1;
    const int k0n = has_next ? dn.sub_ptr[warp] : 0, k1n = has_next ? dn.sub_ptr[warp + 1] : 0;
    const int n_itn = (k1n - k0n + 1) >> 1;
    // write-out header of this patch (round ends, entry segments): lane l holds ints l and 32 + l; needed only after
    // the warp's last element, so the loads cost nothing here
    const unsigned char *blobp = p.blob + (size_t)dc.blob_off * 16;
    int hv0 = 0, hv1 = 0;
    if (TANGENT) {
      hv0 = __ldg(reinterpret_cast<const int *>(blobp) + lane);
      if (W > 4) hv1 = __ldg(reinterpret_cast<const int *>(blobp) + 32 + lane);
    }
    // descriptor of the patch after the next one: requested now, stored when the current descriptor is dead
    int desc_next = 0;
    if (threadIdx.x < 16 && pk + 2 * (int)gridDim.x < p.n_patches)
      desc_next = __ldg(reinterpret_cast<const int *>(p.desc + pk + 2 * gridDim.x) + threadIdx.x);
    // The warp's write-out chunk (own entries | merged entries | merge ops, patch_plan.cuh) is staged in its scratch
    // area by ONE TMA bulk copy as soon as the Gauss-point results of its last pair have been consumed: the copy
    // flies under the accumulation.  A chunk that does not fit is read from global memory instead.
    constexpr int SCR_BYTES = Sh::SCR_D * (int)sizeof(double);
    int ck_off = 0, ck_nf = 0, ck_nl = 0, ck_df = 0, ck_dl = 0, ck_ops = 0;
    bool ck_issued = false, ck_smem = false;
    auto stage_chunk = [&]() {  // whole warp
      const int hv = (warp < 4) ? hv0 : hv1, l0 = 8 * (warp & 3);
      ck_off = __shfl_sync(0xffffffffu, hv, l0);
      ck_nf = __shfl_sync(0xffffffffu, hv, l0 + 1);
      ck_nl = __shfl_sync(0xffffffffu, hv, l0 + 2);
      ck_df = __shfl_sync(0xffffffffu, hv, l0 + 3);
      ck_dl = __shfl_sync(0xffffffffu, hv, l0 + 4);
      ck_ops = __shfl_sync(0xffffffffu, hv, l0 + 5);
      const int bytes = 16 * (ck_nf + ck_df) + pk_al16(8 * ck_nl) + pk_al16(8 * ck_dl) + pk_al16(4 * ck_ops);
      ck_smem = bytes <= SCR_BYTES;
      ck_issued = true;
      __syncwarp();
      if (lane == 0) {
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // the scratch was last accessed by ordinary loads / stores
        if (ck_smem && bytes > 0) {
          bulk_load(ws, blobp + (size_t)ck_off * 16, (unsigned)bytes, &s_bar_wo[warp]);
        } else {
          const unsigned b = (unsigned)__cvta_generic_to_shared(&s_bar_wo[warp]);
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(b) : "memory");
        }
      }
    };

    // the j-th pair after pair `it` of this patch (possibly in the next patch); false = none known yet
    auto nth = [&](int it, int j, int &kk, int &ne) {
      if (it + j < n_it) {
        kk = k0 + 2 * (it + j);
        ne = min(2, k1 - kk);
        return true;
      }
      const int jj = it + j - n_it;
      if (jj < n_itn) {
        kk = k0n + 2 * jj;
        ne = min(2, k1n - kk);
        return true;
      }
      return false;
    };

  for (int it = 0; it < n_it; it++, gi++) {
    const int buf = gi & 1;
    const int kk = k0 + 2 * it, ne = min(2, k1 - kk);
    if (dt_get(buf) != kk) {  // cold start: nothing was requested ahead for this pair
      if (pe_get(buf) != kk) {
        issue_pe(kk, ne, buf);
        pe_set(buf, kk);
      }
      cp_async_commit();
      cp_async_wait_all();
      __syncwarp();
      issue_data(ne, buf);
      dt_set(buf, kk);
      cp_async_commit();
    }
    cp_async_wait_all();
    mbar_wait(&s_bar[warp][buf], (gi >> 1) & 1);
    __syncwarp();
    // what this lane needs of the plan records, before the buffer is recycled
    unsigned blkw = 0, trb = 0, fsb = 0;
    int eid[2], affb = 0;
    {
      const PElem *pe = reinterpret_cast<const PElem *>(wb + Sh::OFF_PE + buf * Sh::PE_D);
      const int el = min(el2, ne - 1);
      blkw = pe[el].blk4[tile];
      trb = (pe[el].tr[(tile * 4) >> 5] >> ((tile * 4) & 31)) & 15u;
      fsb = (pe[el].first[(tile * 4) >> 5] >> ((tile * 4) & 31)) & 15u;
      eid[0] = pe[0].elem;
      eid[1] = pe[ne - 1].elem;
      affb = p.aff ? ((pe[0].pad0 & 1) | ((ne > 1 && (pe[1].pad0 & 1)) ? 2 : 0)) : 0;
    }
    __syncwarp();
    {
      int kq, nq;
      if (nth(it, 1, kq, nq)) {  // data of the next pair: its plan records landed with the wait above
        if (pe_get(buf ^ 1) != kq) {  // (not requested ahead: the pair before was the cold start)
          issue_pe(kq, nq, buf ^ 1);
          pe_set(buf ^ 1, kq);
          cp_async_commit();
          cp_async_wait_all();
          __syncwarp();
        }
        issue_data(nq, buf ^ 1);
        dt_set(buf ^ 1, kq);
      }
      if (nth(it, 2, kq, nq)) {
        issue_pe(kq, nq, buf);
        pe_set(buf, kq);
      }
      cp_async_commit();
    }

    const double *s_h = wb + Sh::OFF_H + buf * Sh::H_D;
    const double *s_dj = wb + Sh::OFF_DJ + buf * Sh::DJ_D;
    const double *s_x = wb + Sh::OFF_X + buf * Sh::X_D;

    // affine elements: expand gradN[q][j][a] = sum_k Ginv[j][k] dN_a/dxi_k(q) and detJ[q] = detJ in place
    if (affb) {
#pragma unroll 1
      for (int el = 0; el < ne; el++) {
        if (!((affb >> el) & 1)) continue;
        double *hh = wb + Sh::OFF_H + buf * Sh::H_D + el * 150;
        double G[9];
#pragma unroll
        for (int i = 0; i < 9; i++) G[i] = hh[i];
        const double dj0 = hh[9];
        __syncwarp();
        for (int idx = lane; idx < 150; idx += 32) {
          const int q = idx / 30, r = idx - 30 * q, j = r / 10, a = r - 10 * j;
          const double *dn = s_dn + (q * 10 + a) * 3;
          const double g0 = j == 0 ? G[0] : (j == 1 ? G[3] : G[6]);
          const double g1 = j == 0 ? G[1] : (j == 1 ? G[4] : G[7]);
          const double g2 = j == 0 ? G[2] : (j == 1 ? G[5] : G[8]);
          hh[idx] = g0 * dn[0] + g1 * dn[1] + g2 * dn[2];
        }
        if (lane < 5) (wb + Sh::OFF_DJ + buf * Sh::DJ_D)[el * 5 + lane] = dj0;
      }
      __syncwarp();
    }

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
    // lane = (element, point, row i): the three lanes of a point build the same F (broadcast loads), then each
    // finishes row i of P dV, of the force partials P dV h_a and of u_a = F h_a; row 0 also leaves
    // M_q = dV [ (h mu + eta) F F^T + h (lambda trE - mu) I ] and dV for the block step.
    if (!MR && lane < ne * 15 && !(p.debug & 8)) {
      const int idx = lane / 3, part = lane - 3 * idx;
      const int el = idx / 5, q = idx - el * 5;
      const double *hq = s_h + el * 150 + q * 30;
      const double *xe = s_x + el * 30;
      double F[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
      double Fd[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
      double hr[10][3];  // the point's shape-function gradients stay in registers for the output loop below
#pragma unroll
      for (int aa = 0; aa < 10; aa++) {
        const double xa[3] = {xe[aa * 3], xe[aa * 3 + 1], xe[aa * 3 + 2]};
        hr[aa][0] = hq[aa];
        hr[aa][1] = hq[10 + aa];
        hr[aa][2] = hq[20 + aa];
        const double *ha = hr[aa];
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
      double Fi[3], FFTi[3], Pi[3];
#pragma unroll
      for (int j = 0; j < 3; j++) {
        Fi[j] = part == 0 ? F[0][j] : (part == 1 ? F[1][j] : F[2][j]);
        FFTi[j] = part == 0 ? FFT[0][j] : (part == 1 ? FFT[1][j] : FFT[2][j]);
      }
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const double FFtF = FFTi[0] * F[0][j] + FFTi[1] * F[1][j] + FFTi[2] * F[2][j];
        Pi[j] = lf * Fi[j] + p.mu * (FFtF - Fi[j]);
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
        for (int j = 0; j < 3; j++) {
          double sv = 0.0;
#pragma unroll
          for (int k = 0; k < 3; k++) sv += Fi[k] * (2.0 * p.eta * Ed[k][j] + ((k == j) ? p.lambda_d * trEd : 0.0));
          Pi[j] += sv;
        }
      }
#pragma unroll
      for (int j = 0; j < 3; j++) Pi[j] *= dV;
      double *fpq = s_fp + idx * PK_GQ;
      double *uq = s_u + idx * PK_GQ;
#pragma unroll
      for (int aa = 0; aa < 10; aa++) {
        const double *ha = hr[aa];
        fpq[aa * 3 + part] = Pi[0] * ha[0] + Pi[1] * ha[1] + Pi[2] * ha[2];
        if (TANGENT) uq[aa * 3 + part] = Fi[0] * ha[0] + Fi[1] * ha[1] + Fi[2] * ha[2];
      }
      if (TANGENT && part == 0) {
        // M_q (6) and dV go into the five spare doubles of the u row and the first two of the force-partial row
        // (the force partials are summed before the block step, but nothing overwrites the row before the next pair)
        const double sI = p.c4 * dV * (p.lambda * trE - p.mu);
        const double cf = p.c3 * dV;
        uq[30] = cf * FFT[0][0] + sI;
        uq[31] = cf * FFT[0][1];
        uq[32] = cf * FFT[0][2];
        uq[33] = cf * FFT[1][1] + sI;
        uq[34] = cf * FFT[1][2];
        fpq[30] = cf * FFT[2][2] + sI;
        fpq[31] = dV;
      }
    }
    __syncwarp();
    // element forces: sum of the 5 point partials in fixed order, one coalesced 240-byte row per element
    for (int el = 0; el < ne; el++) {
      if (lane < 30) {
        constexpr int FS = MR ? EK_GQM : PK_GQ;
        const double *fq = s_fp + el * 5 * FS + lane;
        p.fe[(int64_t)eid[el] * 30 + lane] = (((fq[0] + fq[FS]) + fq[2 * FS]) + fq[3 * FS]) + fq[4 * FS];
      }
    }

    // ------------------------------------------------------------ block step
    if (TANGENT && !(p.debug & 4)) {
      const int el = min(el2, ne - 1);  // odd tail: the second half shadows the first and never adds
      const bool on = lane_on && (el2 < ne);
      const double *he = s_h + el * 150;
      const double *ue = s_u + el * 5 * PK_GQ;
      const double *fpe = s_fp + el * 5 * PK_GQ;
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
        const double *hq = he + q * 30, *uq = ue + q * PK_GQ, *fq = fpe + q * PK_GQ;
        const double dV = fq[31];
        double ha[2][3], hb[2][3], ua[2][3], ub[2][3], M[6];
#pragma unroll
        for (int j = 0; j < 3; j++) {
          ha[0][j] = hq[j * 10 + a1];
          ha[1][j] = hq[j * 10 + a1 + 1];
          hb[0][j] = hq[j * 10 + b1];
          hb[1][j] = hq[j * 10 + b1 + 1];
          ua[0][j] = uq[a1 * 3 + j] * dV;
          ua[1][j] = uq[a1 * 3 + 3 + j] * dV;
          ub[0][j] = uq[b1 * 3 + j];
          ub[1][j] = uq[b1 * 3 + 3 + j];
        }
#pragma unroll
        for (int i = 0; i < 5; i++) M[i] = uq[30 + i];
        M[5] = fq[30];
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
          // a block that the accumulator holds transposed (global ids in the other order) is finished transposed:
          // (c1 A + c2 A^T + sym)^T = c2 A + c1 A^T + sym -- two selects here instead of nine in the accumulation
          const bool trt = (trb >> t) & 1u;
          const double ca = trt ? p.c2 : p.c1, cb = trt ? p.c1 : p.c2;
          double k[9];
          k[0] = ca * A[t][0] + cb * A[t][0] + Gs[t][0] + dg;
          k[1] = ca * A[t][1] + cb * A[t][3] + Gs[t][1];
          k[2] = ca * A[t][2] + cb * A[t][6] + Gs[t][2];
          k[3] = ca * A[t][3] + cb * A[t][1] + Gs[t][1];
          k[4] = ca * A[t][4] + cb * A[t][4] + Gs[t][3] + dg;
          k[5] = ca * A[t][5] + cb * A[t][7] + Gs[t][4];
          k[6] = ca * A[t][6] + cb * A[t][2] + Gs[t][2];
          k[7] = ca * A[t][7] + cb * A[t][5] + Gs[t][4];
          k[8] = ca * A[t][8] + cb * A[t][8] + Gs[t][5] + dg;
#pragma unroll
          for (int i = 0; i < 9; i++) A[t][i] = k[i];
        }
      }
      if (it == n_it - 1) stage_chunk();  // u, M_q, the force partials of this pair are dead: K sits in registers
      // ordered accumulation into the warp's private accumulator: first element of the pair, then the second.
      // Every lane walks the 9 MEMORY positions of its block in the same order (a transposed block picks the other
      // register), so the 15 lanes of a half warp hit 15 different banks when their slots differ modulo 16 -- which
      // the plan arranges (assign_slots, patch_plan.cu).
#pragma unroll 1
      for (int e = 0; e < 2; e++) {
        if (on && el2 == e && !(p.debug & 2)) {
          // the blocks of one element are distinct: load two blocks' old values, then add and store (the compiler
          // cannot know the four pointers differ and would serialise load -> store -> load otherwise)
#pragma unroll
          for (int t0 = 0; t0 < 4; t0 += 2) {
            double *d0 = acc + ((blkw >> (8 * t0)) & 255u) * 9;
            double *d1 = acc + ((blkw >> (8 * (t0 + 1))) & 255u) * 9;
            // Mooney-Rivlin blocks are transposed here; SVK blocks were finished in the stored orientation
            const bool tr0 = MR && ((trb >> t0) & 1u), tr1 = MR && ((trb >> (t0 + 1)) & 1u);
            // t0 = 2: (a1+1, a1) is the transpose of (a1, a1+1) on diagonal tiles -- the upper block carries it
            const bool skip0 = (t0 == 2) && (tI == tJ);
            // the first element of the sub-patch to touch a block stores it (no zero-fill of the accumulator)
            const bool f0 = (fsb >> t0) & 1u, f1 = (fsb >> (t0 + 1)) & 1u;
            double o0[9], o1[9];
#pragma unroll
            for (int m = 0; m < 9; m++) o0[m] = (skip0 || f0) ? 0.0 : d0[m];
#pragma unroll
            for (int m = 0; m < 9; m++) o1[m] = f1 ? 0.0 : d1[m];
            if (!skip0) {
#pragma unroll
              for (int m = 0; m < 9; m++) d0[m] = o0[m] + (tr0 ? A[t0][(m % 3) * 3 + m / 3] : A[t0][m]);
            }
#pragma unroll
            for (int m = 0; m < 9; m++) d1[m] = o1[m] + (tr1 ? A[t0 + 1][(m % 3) * 3 + m / 3] : A[t0 + 1][m]);
          }
        }
        __syncwarp();
      }
    }
    __syncwarp();
  }
  if (!TANGENT) {  // force only: no shared output, just keep the descriptor ring turning
    __syncthreads();
    if (threadIdx.x < 16 && pk + 2 * (int)gridDim.x < p.n_patches)
      reinterpret_cast<int *>(&s_desc[n_done & 1])[threadIdx.x] = desc_next;
    __syncthreads();
    continue;
  }

  // ------------------------------------------------------------ write-out
  // tuning aid (TLFEA_PK_DEBUG & 16): all H traffic of a CTA into a 128 KB window -- the assembly without its DRAM traffic
  double *Hb = p.H + ((p.debug & 16) ? (int64_t)(blockIdx.x & 255) * 65536 : dc.hbase);
  const unsigned hmask = (p.debug & 16) ? 0x3fffu : 0xffffffffu;
  const bool wo_on = !(p.debug & 1);
  if (!dep_waited) {  // H belongs to the previous colour until its launch has completed
    asm volatile("griddepcontrol.wait;\n" ::: "memory");
    dep_waited = true;
  }
  // (a) the blocks this warp holds alone are final: their entries leave now, while the other warps still compute
  if (!ck_issued) stage_chunk();  // a warp without elements in this patch, or the block step switched off
  mbar_wait(&s_bar_wo[warp], n_done & 1);
  const unsigned char *ck = ck_smem ? reinterpret_cast<const unsigned char *>(ws) : blobp + (size_t)ck_off * 16;
  const unsigned char *ck_dep = ck + 16 * ck_nf + pk_al16(8 * ck_nl);
  const uint32_t *ops = reinterpret_cast<const uint32_t *>(ck_dep + 16 * ck_df + pk_al16(8 * ck_dl));
  if (wo_on && lane < 30) {
    if (ck_smem)
      pk_write_segment<PK_FU_OWN, true>(ck, ck_nf, ck_nl, lane / 3, 10, lane % 3, pk_smem, Hb, p.inv_h, hmask);
    else
      pk_write_segment<PK_FU_OWN, false>(ck, ck_nf, ck_nl, lane / 3, 10, lane % 3, pk_smem, Hb, p.inv_h, hmask);
  }
  __syncthreads();  // every accumulator is final
  // (b) pairs held by several warps: dst += src in ascending warp order, one thread per dst (the ops of a dst are
  // consecutive; all but the first carry the continuation bit).  Warp w does the ops of chunk w ...
  auto ld_op = [&](const uint32_t *q) {  // staged: an explicit shared-memory load, else read-only global
    if (!ck_smem) return __ldg(q);
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(q)));
    return v;
  };
  if (wo_on) {
    for (int q = lane; q < ck_ops; q += 32) {
      uint32_t op = ld_op(ops + q);
      if (op & 0x8000u) continue;
      double *da = pk_smem + (op & 0x3fffu);
      double a[9];
#pragma unroll
      for (int i = 0; i < 9; i++) a[i] = da[i];
      int j = q;
      do {
        const double *sa = pk_smem + (op >> 16);
#pragma unroll
        for (int i = 0; i < 9; i++) a[i] += sa[i];
        j++;
        op = (j < ck_ops) ? ld_op(ops + j) : 0u;
      } while (op & 0x8000u);
#pragma unroll
      for (int i = 0; i < 9; i++) da[i] = a[i];
    }
    __syncwarp();
    // (c) ... and writes the H entries of exactly those dsts
    if (lane < 30) {
      if (ck_smem)
        pk_write_segment<PK_FU_DEP, true>(ck_dep, ck_df, ck_dl, lane / 3, 10, lane % 3, pk_smem, Hb, p.inv_h, hmask);
      else
        pk_write_segment<PK_FU_DEP, false>(ck_dep, ck_df, ck_dl, lane / 3, 10, lane % 3, pk_smem, Hb, p.inv_h, hmask);
    }
  }
  __syncthreads();  // nobody reads the accumulators, the lists or the current descriptor any more
  if (threadIdx.x < 16 && pk + 2 * (int)gridDim.x < p.n_patches)
    reinterpret_cast<int *>(&s_desc[n_done & 1])[threadIdx.x] = desc_next;
  if (has_next && threadIdx.x == 32 % NT)
    bulk_prefetch_l2(p.blob + (size_t)dn.blob_off * 16, (unsigned)dn.blob_bytes);
  __syncthreads();
  }  // patches
  cp_async_wait_all();
}

// mass field of every first-touch entry: M(row, column) of the pattern block the entry writes.  The block follows from
// the entry's H position: 9*off[i] <= hbase + hoff < 9*off[i+1].
__global__ void pk_fill_mass_kernel(int n_patches, int n_coef, const PatchDesc *__restrict__ desc, unsigned char *blob,
                                    const int *__restrict__ off, const double *__restrict__ mass) {
  const int pch = blockIdx.x;
  if (pch >= n_patches) return;
  const PatchDesc d = desc[pch];
  unsigned char *sb = blob + (size_t)d.blob_off * 16;
  const int *hdr = reinterpret_cast<const int *>(sb);
  for (int part = 0; part < 2 * PK_MAXW; part++) {  // own and merged first-touch entries of every chunk
    const int *hw = hdr + 8 * (part >> 1);
    unsigned char *at = sb + (size_t)hw[0] * 16;
    if (part & 1) at += 16 * hw[1] + pk_al16(8 * hw[2]);
    PEntF *ent = reinterpret_cast<PEntF *>(at);
    const int nf = (part & 1) ? hw[3] : hw[1];
    for (int k = threadIdx.x; k < nf; k += blockDim.x) {
      const int64_t habs = d.hbase + ent[k].hoff;
      const int64_t blk9 = habs / 9;
      int lo = 0, hi = n_coef;  // largest i with off[i] <= blk9
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= blk9) lo = mid;
        else hi = mid;
      }
      const int slot = (int)((habs - 9 * (int64_t)off[lo]) / 3);
      ent[k].mass = mass[off[lo] + slot];
    }
  }
}

// plan record flag: bit 0 of pad0 = the element is affine (elem.cu: t10_affine_kernel)
__global__ void pk_set_affine_kernel(int64_t n, PElem *elems, const unsigned char *__restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) elems[i].pad0 = flag[elems[i].elem] ? 1 : 0;
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
