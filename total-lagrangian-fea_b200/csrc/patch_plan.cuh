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

constexpr int PK_CAPU = 216;    // 3x3 blocks per warp accumulator (a Kuhn cell of 6 tets needs 210)
constexpr int PK_MAXW = 8;      // sub-patches (= warps) per patch
constexpr int PK_MAX_SLOT = 8191;

struct PElem {          // 128 bytes, one per element, in processing order
  int elem;             // element id in the caller's numbering (rows of gradN, detJ, fe)
  int pad0;
  int node[10];         // global coefficient ids
  uint32_t blk4[15];    // per 2x2 tile of shape-function pairs (ek_kernel.cuh order): accumulator block of
                        // (a1+ta, b1+tb) in byte ta*2+tb; 255 = the redundant lower block of a diagonal tile
  uint32_t tr[2];       // bit tile*4+ta*2+tb: node[a] > node[b], the block is added transposed
  int pad1[3];
};
static_assert(sizeof(PElem) == 128, "PElem is fetched as eight 16-byte pieces");

struct PRow {           // 32 bytes, one per patch-local coefficient row
  int64_t hbase;        // 9*off[i]: start of the row's three scalar CSR rows in H
  int rowlen;           // 3*deg(i)
  int mass_base;        // off[i]
  int ent_begin;        // first entry of the row in ent[]
  int n_ent;
  int node;
  int pad;
};
static_assert(sizeof(PRow) == 32, "PRow layout");

struct PatchDesc {      // 64 bytes
  int sub_ptr[PK_MAXW + 1];  // PElem range of sub-patch w: [sub_ptr[w], sub_ptr[w+1])
  int row_begin, n_rows;     // PRow range
  int src_base;              // the patch's sources start here in src[]
  int n_sub;
  int pad[3];
};
static_assert(sizeof(PatchDesc) == 64, "PatchDesc layout");

// entry word: bits 0-12 slot of column j in row i, 13 first touch of this H block (start from M_ij/h instead of
// the value in memory), 14-16 number of sources - 1, 17-31 offset of the first source relative to src_base
// source word: bits 0-7 accumulator block, 8-10 warp, 11 read transposed (row id > column id)
constexpr uint32_t PK_ENT_FIRST = 1u << 13;

struct PatchPlanHost {
  int W = 0;
  int n_patches = 0, n_colors = 0;
  std::vector<int> color_ptr;  // launch c runs patches [color_ptr[c], color_ptr[c+1])
  std::vector<PatchDesc> desc;
  std::vector<PElem> elems;
  std::vector<PRow> rows;
  std::vector<uint32_t> ent;
  std::vector<uint16_t> src;
  std::vector<int64_t> uncovered;  // pattern blocks that no element touches (constraint couplings, unused nodes)
  int64_t sum_local_blocks = 0;    // sum over patches of their distinct node pairs (in-place update traffic)
};

// Pure host code (no CUDA call): ecoef [n_elem][10]; x, y, z reference coordinates per coefficient; off / cols the
// coefficient-level CSR pattern (sorted columns).  Returns 0 or a negative tlfea_status with *err set.
int build_patch_plan(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y, const double *z,
                     const int *off, const int *cols, int W, PatchPlanHost *out, std::string *err);

}  // namespace tlfea
