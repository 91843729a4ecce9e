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

// Everything the write-out of a patch reads sits in one contiguous, 16-byte aligned BLOB so that it arrives by a
// single TMA bulk copy into scratch shared memory:
//   round_end int[8] | merge ops uint32[n_merge] | entries PEnt[n_ent] | mass double[n_first]
// (sections padded to 16 bytes).
//   * merge op (dst | src << 16): accumulator block src is added to accumulator block dst.  A node pair that
//     several warps of the patch hold is first reduced into the block of the lowest warp, one extra warp per
//     ROUND (ops of round r: [round_end[r-1], round_end[r]); within a round every dst is distinct).
//   * entry: one 3x3 block of H (both orientations of a pair are separate entries), ordered (first touch before
//     later touch, then row, then column).  Entry k < n_first starts from M_ij/h (mass[k], refreshed on the
//     device when the mass changes), the others from the value already in H.
//   accumulator block address (16 bits): bits 0-7 block, 8-10 warp; entry src adds bit 11 = read transposed.
struct PEnt {            // 8 bytes
  uint32_t hoff;         // offset (doubles) of H(3i, 3j) from the patch's hbase: 9*off[i] + 3*slot - hbase
  uint16_t rowlen;       // 3*deg(i): distance between the three scalar rows of the block row
  uint16_t src;
};
static_assert(sizeof(PEnt) == 8, "PEnt layout");

struct PatchDesc {       // 64 bytes
  int sub_ptr[PK_MAXW + 1];  // PElem range of sub-patch w: [sub_ptr[w], sub_ptr[w+1])
  int n_sub;
  int n_ent, n_first, n_merge;
  uint32_t blob_off;         // in 16-byte units
  int64_t hbase;
};
static_assert(sizeof(PatchDesc) == 64, "PatchDesc layout");

__host__ __device__ inline int pk_al16(int b) { return (b + 15) & ~15; }
struct BlobLayout {
  int ops, ent, mass, total;  // byte offsets; round_end sits at 0
  __host__ __device__ BlobLayout(int n_ent, int n_first, int n_merge) {
    ops = 32;
    ent = ops + pk_al16(n_merge * 4);
    mass = ent + pk_al16(n_ent * 8);
    total = mass + pk_al16(n_first * 8);
  }
};

struct PatchPlanHost {
  int W = 0;
  int n_patches = 0, n_colors = 0, n_chunks = 1;
  std::vector<int> color_ptr;    // launch l = chunk * n_colors + colour runs patches [color_ptr[l], color_ptr[l+1])
  std::vector<int> range_ready;  // [n_chunks] last chunk that touches a coefficient of the r-th id range (-1: none)
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
// coefficient-level CSR pattern (sorted columns); n_chunks = id-range chunks of the launch order (1 = colours only);
// blob_limit = largest write-out blob (bytes) a patch may have.
// Returns 0 or a negative tlfea_status with *err set.
int build_patch_plan(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y, const double *z,
                     const int *off, const int *cols, int W, int blob_limit, int n_chunks, PatchPlanHost *out,
                     std::string *err);

}  // namespace tlfea
