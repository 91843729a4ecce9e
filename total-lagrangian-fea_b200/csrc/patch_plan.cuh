// patch_plan.cuh -- data structures of the single-pass FEAT10 assembly (pk_kernel.cuh).
//
// The mesh is cut into PATCHES of at most W SUB-PATCHES; a sub-patch is a short run of spatially adjacent
// elements whose distinct upper-triangle node pairs fit one warp-private accumulator (PK_CAPU 3x3 blocks in
// shared memory).  One CTA of W warps assembles one patch: warp w adds the blocks of its sub-patch's elements
// into its own accumulator (program order => fixed summation order, no atomics, no inter-warp hand-off), then
// the CTA sums the W accumulators per node pair in warp order and adds the patch's partial rows into the CSR
// values of H in place.  Patches that share a node never run in the same launch (graph colouring of the patch
// adjacency), so the in-place update needs no atomics either and the order per H entry is (colour, warp,
// element) -- fixed, hence bit-reproducible.  Replaces, for FEAT10, compute_hessian_assemble_csr
// (FEAT10DataFunc.cuh:512-791: 900 binary searches + 900 FP64 atomics per Gauss point) and
// assemble_sparse_hessian_mass (SyncedNewton.cu:213-259).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

namespace tlfea {

constexpr int PK_CAPU = 210;    // 3x3 blocks per warp accumulator (a Kuhn cell of 6 tets needs exactly 210)
constexpr int PK_MAXW = 8;      // sub-patches (= warps) per patch
constexpr int PK_MAX_ROWLEN = 65535;

struct PElem {          // 128 bytes, one per element, in processing order
  int elem;             // element id in the caller's numbering (rows of gradN, detJ, fe)
  int pad0;
  int node[10];         // global coefficient ids
  uint32_t blk4[15];    // per 2x2 tile of shape-function pairs (ek_kernel.cuh order): accumulator block of
                        // (a1+ta, b1+tb) in byte ta*2+tb; 255 = the redundant lower block of a diagonal tile
  uint32_t tr[2];       // bit tile*4+ta*2+tb: node[a] > node[b], the block is added transposed
  uint32_t first[2];    // same bit order: this element is the first of its sub-patch to touch the block (store, no add)
  int pad1;
};
static_assert(sizeof(PElem) == 128, "PElem is fetched as eight 16-byte pieces");

// Everything the write-out of a patch reads sits in one contiguous, 16-byte aligned BLOB in global memory (pulled
// into L2 by a bulk prefetch one patch ahead):
//   header int[PK_HDR_INTS] | chunk 0 | ... | chunk PK_MAXW-1
//   header: per chunk w the 8 ints {offset from the blob start in 16-byte units, n_first own, n_later own,
//           n_first merged, n_later merged, n_ops, 0, 0}
//   chunk w (what warp w needs, contiguous so that ONE TMA bulk copy stages it in the warp's scratch area):
//           PEntF[n_first own] PEnt[n_later own] pad16 | PEntF[n_first merged] PEnt[n_later merged] pad16 |
//           uint32 ops[n_ops] pad16
//   * own entries: the H entries of the accumulator blocks warp w holds ALONE.  Warp w writes them as soon as its
//     own elements are done, without waiting for the other warps.
//   * merge ops (dst | cont << 15 | src << 16, addresses in doubles from the start of the accumulators): a node pair
//     that several warps hold is reduced into the block of the lowest warp: dst += src, sources in ascending warp
//     order; the ops of one dst are consecutive, all but the first carry the continuation bit, and one thread does
//     them in sequence.  The merged pairs of a patch are dealt round-robin to the chunks; after the CTA barrier warp
//     w does the ops of chunk w and then writes the merged entries of chunk w (the entries of exactly those dsts).
//   * entry: one 3x3 block of H (both orientations of a pair are separate entries).  First touches (of the H block
//     in the launch sequence; they start from M_ij/h, carried in the entry and refreshed on the device when the mass
//     changes) before later touches (added to the value already in H), then (row, column).
//   entry src (16 bits): bits 0-13 offset of the accumulator block in doubles (warp * 9 PK_CAPU + 9 slot), bit 15 =
//   read transposed.
struct PEnt {            // 8 bytes: later touch
  uint32_t hoff;         // offset (doubles) of H(3i, 3j) from the patch's hbase: 9*off[i] + 3*slot - hbase
  uint16_t rowlen;       // 3*deg(i): distance between the three scalar rows of the block row
  uint16_t src;
};
static_assert(sizeof(PEnt) == 8, "PEnt layout");
struct PEntF {           // 16 bytes: first touch
  uint32_t hoff;
  uint16_t rowlen;
  uint16_t src;
  double mass;           // M_ij (scalar per node pair); the H block starts from (M_ij / h) I
};
static_assert(sizeof(PEntF) == 16, "PEntF layout");
constexpr int PK_HDR_INTS = 8 * PK_MAXW;  // 64 ints: lane l of a warp holds ints l and 32 + l
static_assert(9 * PK_CAPU * PK_MAXW < (1 << 14), "entry src: 14 bits of accumulator offset");

struct PatchDesc {       // 64 bytes
  int sub_ptr[PK_MAXW + 1];  // PElem range of sub-patch w: [sub_ptr[w], sub_ptr[w+1])
  int n_sub;
  int blob_bytes;            // whole blob of the patch (L2 prefetch)
  int n_first, n_merge;      // first-touch entries over all segments (statistics); merge ops
  uint32_t blob_off;         // in 16-byte units
  int64_t hbase;
};
static_assert(sizeof(PatchDesc) == 64, "PatchDesc layout");

__host__ __device__ inline int pk_al16(int b) { return (b + 15) & ~15; }

// Id-range chunks of the launch order (host-buffer pipelining): chunk r covers the coefficient ids
// [pk_chunk_bound(r), pk_chunk_bound(r + 1)).  The first and the last range are a quarter of the others: the upload
// of the first range and the download of the last one are the parts of a host-buffer step no kernel can hide.
inline int64_t pk_chunk_bound(int64_t n_coef, int K, int r) {
  if (K < 3) return n_coef * r / K;
  const int64_t tot = 4 * (int64_t)(K - 2) + 2;
  const int64_t w = (r <= 0) ? 0 : (r >= K ? tot : 1 + 4 * (int64_t)(r - 1));
  return n_coef * w / tot;
}
// The nodal forces travel back in finer pieces: every chunk range is cut into PK_GSUB sub-ranges; sub-range q is final
// after the last chunk that touches one of its ids (PatchPlanHost::range_ready[q]) -- only the sliver of a range next to
// the following one has to wait for the following chunk.
constexpr int PK_GSUB = 8;
inline int64_t pk_sub_bound(int64_t n_coef, int K, int q) {
  const int r = q / PK_GSUB, j = q - r * PK_GSUB;
  const int64_t b0 = pk_chunk_bound(n_coef, K, r), b1 = pk_chunk_bound(n_coef, K, r + 1);
  return b0 + (b1 - b0) * j / PK_GSUB;
}

struct PatchPlanHost {
  int W = 0;
  int n_patches = 0, n_colors = 0, n_chunks = 1;
  std::vector<int> color_ptr;    // launch l = chunk * n_colors + colour runs patches [color_ptr[l], color_ptr[l+1])
  std::vector<int> range_ready;  // [n_chunks * PK_GSUB] last chunk that touches a coefficient of the q-th id sub-range (-1: none)
  std::vector<PatchDesc> desc;
  std::vector<PElem> elems;
  std::vector<unsigned char> blob;  // all patches, each 16-byte aligned
  std::vector<int> row_node;        // host only: coefficient ids of every patch, patch by patch (tests)
  std::vector<int64_t> row_ptr;     // host only: patch q's ids are row_node[row_ptr[q] .. row_ptr[q+1])
  int max_blob_bytes = 0;
  std::vector<int64_t> uncovered;  // pattern blocks that no element touches (constraint couplings, unused nodes)
  int64_t sum_local_blocks = 0;    // sum over patches of their distinct node pairs (in-place update traffic)
};

// Pure host code (no CUDA call): ecoef [n_elem][10]; x, y, z reference coordinates per coefficient; off / cols the
// coefficient-level CSR pattern (sorted columns); n_chunks = id-range chunks of the launch order (1 = colours only;
// negative: at most |n_chunks|, fewer when the mesh has too few patches to fill the launches);
// Returns 0 or a negative tlfea_status with *err set.
int build_patch_plan(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y, const double *z,
                     const int *off, const int *cols, int W, int n_chunks, PatchPlanHost *out,
                     std::string *err);

}  // namespace tlfea
