// patch_plan.cu -- host-side analysis for the single-pass FEAT10 assembly (see patch_plan.cuh, pk_kernel.cuh).
// Pure host code (OpenMP over octants / patches); runs once per sparsity analysis.
#include "patch_plan.cuh"

#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/tlfea_b200.h"

namespace tlfea {

namespace {

inline uint64_t spread21h(uint64_t v) {
  v &= 0x1fffffULL;
  v = (v | v << 32) & 0x1f00000000ffffULL;
  v = (v | v << 16) & 0x1f0000ff0000ffULL;
  v = (v | v << 8) & 0x100f00f00f00f00fULL;
  v = (v | v << 4) & 0x10c30c30c30c30c3ULL;
  v = (v | v << 2) & 0x1249249249249249ULL;
  return v;
}

// tile t of the symmetric 5x5 grid of 2x2 tiles, in the element kernel's lane order
inline void tile_ij(int t, int *I, int *J) {
  if (t < 5) *I = 0, *J = t;
  else if (t < 9) *I = 1, *J = t - 4;
  else if (t < 12) *I = 2, *J = t - 7;
  else if (t < 14) *I = 3, *J = t - 9;
  else *I = 4, *J = 4;
}

inline uint64_t upper_key(int i, int j) {
  const uint32_t a = (uint32_t)std::min(i, j), b = (uint32_t)std::max(i, j);
  return ((uint64_t)a << 32) | b;
}

struct SubPatch {
  int begin = 0, end = 0;  // range in the sorted element order
};

struct EntL {
  int64_t habs;   // 9*off[i] + 3*slot
  int64_t g;      // global pattern block
  int rowlen;
  uint16_t src;   // accumulator offset in doubles | transposed << 15
  bool first;
  int seg;        // chunk: the warp that owns the source block alone, or PK_MAXW + g: merged block of chunk g
};
struct PatchLocal {
  std::vector<PElem> elems;
  int sub_ptr[PK_MAXW + 1] = {0};
  int n_sub = 0;
  std::vector<int> nodes;  // distinct coefficient ids (ascending)
  std::vector<EntL> ent;
  std::vector<uint32_t> ops[PK_MAXW];  // merge ops of chunk g (patch_plan.cuh)
  int n_ops = 0;
  int color = -1;
  bool bad = false;
};

// Accumulator slots of a sub-patch's blocks.  The 15 lanes of a half warp add their t-th block at the same time:
// the adds are bank-conflict free when those 15 slots differ modulo 16 (a block is 9 doubles = odd stride).
// Greedy: blocks in order of first use take the residue class that collides least with the sets they belong to.
void assign_slots(int nb, const std::vector<std::vector<int>> &sets, std::vector<int> &slot) {
  // block -> the sets it belongs to, as one flat CSR (no per-block allocation: this runs once per sub-patch)
  const int ns = (int)sets.size();
  std::vector<int> mptr((size_t)nb + 1, 0), mset;
  for (auto &st : sets)
    for (int b : st) mptr[(size_t)b + 1]++;
  for (int b = 0; b < nb; b++) mptr[(size_t)b + 1] += mptr[(size_t)b];
  mset.resize((size_t)mptr[(size_t)nb]);
  {
    std::vector<int> fillp(mptr.begin(), mptr.end() - 1);
    for (int q = 0; q < ns; q++)
      for (int b : sets[(size_t)q]) mset[(size_t)fillp[(size_t)b]++] = q;
  }
  std::vector<uint16_t> used((size_t)ns, 0);  // residues taken per set
  int cap[16], fill[16] = {0};
  for (int r = 0; r < 16; r++) cap[r] = (PK_CAPU - r + 15) / 16;
  slot.assign((size_t)nb, -1);
  std::vector<int> order;
  order.reserve((size_t)nb);
  std::vector<char> seen((size_t)nb, 0);
  for (auto &st : sets)
    for (int b : st)
      if (!seen[(size_t)b]) seen[(size_t)b] = 1, order.push_back(b);
  for (int b = 0; b < nb; b++)
    if (!seen[(size_t)b]) order.push_back(b);
  for (int b : order) {
    // residues already taken in any set of this block, counted per residue
    int cnt[16] = {0};
    for (int k = mptr[(size_t)b]; k < mptr[(size_t)b + 1]; k++) {
      uint16_t u = used[(size_t)mset[(size_t)k]];
      while (u) {
        const int r = __builtin_ctz(u);
        cnt[r]++;
        u &= (uint16_t)(u - 1);
      }
    }
    int best = -1, best_cost = 1 << 30;
    for (int r = 0; r < 16; r++) {
      if (fill[r] >= cap[r]) continue;
      const int cost = cnt[r] * 64 + fill[r];  // ties: the emptiest class
      if (cost < best_cost) best_cost = cost, best = r;
    }
    slot[(size_t)b] = best + 16 * fill[best];
    fill[best]++;
    for (int k = mptr[(size_t)b]; k < mptr[(size_t)b + 1]; k++) used[(size_t)mset[(size_t)k]] |= (uint16_t)(1u << best);
  }
}

}  // namespace

int build_patch_plan(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y, const double *z,
                     const int *off, const int *cols, int W, int n_chunks, PatchPlanHost *out,
                     std::string *err) {
  auto failm = [&](int code, const char *m) {
    if (err) *err = m;
    return code;
  };
  if (W < 1 || W > PK_MAXW) return failm(TLFEA_ERR_INVALID, "patch plan: W out of range");
  if (n_elem <= 0) return failm(TLFEA_ERR_INVALID, "patch plan: no elements");
  const bool timing = getenv("TLFEA_PLAN_TIMING") != nullptr;
  auto t_last = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (!timing) return;
    const auto t = std::chrono::steady_clock::now();
    std::fprintf(stderr, "patch plan: %-28s %.3f s\n", what, std::chrono::duration<double>(t - t_last).count());
    t_last = t;
  };
  // ---- 1. spatial keys: Morton code of the cell that holds the centroid of the four vertices
  double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
  std::vector<double> cen((size_t)n_elem * 3);
  bool degenerate = false;
#pragma omp parallel for schedule(static) reduction(min : mn[:3]) reduction(max : mx[:3]) reduction(|| : degenerate)
  for (int e = 0; e < n_elem; e++) {
    const int *ec = ecoef + (size_t)e * 10;
    int s[10];
    for (int a = 0; a < 10; a++) s[a] = ec[a];
    std::sort(s, s + 10);
    for (int a = 0; a < 9; a++) degenerate = degenerate || (s[a] == s[a + 1]);
    degenerate = degenerate || s[0] < 0 || s[9] >= n_coef;
    double c[3] = {0, 0, 0};
    if (!(s[0] < 0 || s[9] >= n_coef))
      for (int a = 0; a < 4; a++) c[0] += 0.25 * x[ec[a]], c[1] += 0.25 * y[ec[a]], c[2] += 0.25 * z[ec[a]];
    for (int k = 0; k < 3; k++) {
      cen[(size_t)e * 3 + k] = c[k];
      mn[k] = std::min(mn[k], c[k]);
      mx[k] = std::max(mx[k], c[k]);
    }
  }
  if (degenerate)
    return failm(TLFEA_ERR_INVALID,
                 "an element repeats a node id or has one out of range: the ordered assembly needs pairwise "
                 "distinct ids per element");
  // cell edge: a cell should hold ~6 elements (one warp accumulator) on a uniform mesh; snapped so that the longest
  // axis of the node bounding box is a whole number of cells (exact for structured blocks)
  double nmn[3] = {1e300, 1e300, 1e300}, nmx[3] = {-1e300, -1e300, -1e300};
#pragma omp parallel for schedule(static) reduction(min : nmn[:3]) reduction(max : nmx[:3])
  for (int i = 0; i < n_coef; i++) {
    const double c[3] = {x[i], y[i], z[i]};
    for (int k = 0; k < 3; k++) nmn[k] = std::min(nmn[k], c[k]), nmx[k] = std::max(nmx[k], c[k]);
  }
  double ext[3], vol = 1.0, emax = 0.0;
  for (int k = 0; k < 3; k++) ext[k] = nmx[k] - nmn[k], emax = std::max(emax, ext[k]);
  int dims = 0;
  for (int k = 0; k < 3; k++)
    if (ext[k] > 1e-9 * emax) vol *= ext[k], dims++;
  double cell = 1.0;
  if (emax > 0.0 && dims > 0) {
    cell = std::pow(vol / std::max(1.0, n_elem / 6.0), 1.0 / dims);
    const double m = std::max(1.0, std::round(emax / cell));
    cell = emax / m;
  }
  for (int k = 0; k < 3; k++) mn[k] = nmn[k];
  const double inv_cell = (cell > 0.0) ? 1.0 / cell : 0.0;
  std::vector<std::pair<uint64_t, int>> order((size_t)n_elem);
#pragma omp parallel for schedule(static)
  for (int e = 0; e < n_elem; e++) {
    uint64_t key = 0;
    for (int k = 0; k < 3; k++) {
      double q = (cen[(size_t)e * 3 + k] - mn[k]) * inv_cell;
      q = std::min(std::max(q, 0.0), 2097151.0);
      key |= spread21h((uint64_t)q) << (2 - k);
    }
    order[e] = {key, e};
  }
  std::sort(order.begin(), order.end());
  std::vector<double>().swap(cen);

  lap("1 keys + sort");
  // ---- 2. octants (2x2x2 cells): the unit a patch never leaves
  std::vector<int> oct_ptr;
  for (int i = 0; i < n_elem; i++)
    if (i == 0 || (order[i].first >> 3) != (order[i - 1].first >> 3)) oct_ptr.push_back(i);
  oct_ptr.push_back(n_elem);
  const int n_oct = (int)oct_ptr.size() - 1;

  lap("2 octants");
  // ---- 3. sub-patches and patches per octant
  std::vector<std::vector<PatchLocal>> per_oct((size_t)n_oct);
  bool any_bad = false;
  double tacc[6] = {0, 0, 0, 0, 0, 0};
#pragma omp parallel for schedule(dynamic, 16) reduction(|| : any_bad) reduction(+ : tacc[:6])
  for (int o = 0; o < n_oct; o++) {
    auto t_prev = std::chrono::steady_clock::now();
    auto TT = [&](int k) { auto t = std::chrono::steady_clock::now(); tacc[k] += std::chrono::duration<double>(t - t_prev).count(); t_prev = t; };
    // greedy sub-patches: add elements while the distinct upper pairs fit one accumulator
    std::vector<SubPatch> subs;
    std::vector<uint64_t> cur, add, merged;
    SubPatch sp;
    sp.begin = oct_ptr[o];
    for (int i = oct_ptr[o]; i < oct_ptr[o + 1]; i++) {
      const int *ec = ecoef + (size_t)order[i].second * 10;
      add.clear();
      for (int a = 0; a < 10; a++)
        for (int b = a; b < 10; b++) add.push_back(upper_key(ec[a], ec[b]));
      std::sort(add.begin(), add.end());
      merged.resize(cur.size() + add.size());
      merged.resize(std::set_union(cur.begin(), cur.end(), add.begin(), add.end(), merged.begin()) - merged.begin());
      if ((int)merged.size() > PK_CAPU || i - sp.begin >= 250) {
        sp.end = i;
        subs.push_back(sp);
        sp.begin = i;
        cur = add;
      } else {
        cur.swap(merged);
      }
    }
    sp.end = oct_ptr[o + 1];
    subs.push_back(sp);

    TT(5);
    auto &plist = per_oct[o];
    for (size_t s0 = 0; s0 < subs.size();) {
      int ns = (int)std::min<size_t>(W, subs.size() - s0);
      PatchLocal P;
      P.n_sub = ns;
      struct Tup {
        uint64_t key;
        uint16_t s;  // accumulator block address in doubles: warp * 9 PK_CAPU + 9 slot
      };
      std::vector<Tup> tups;
      for (int w = 0; w < ns; w++) {
        const SubPatch &S = subs[s0 + w];
        P.sub_ptr[w] = (int)P.elems.size();
        std::vector<uint64_t> blocks;
        for (int i = S.begin; i < S.end; i++) {
          const int *ec = ecoef + (size_t)order[i].second * 10;
          for (int a = 0; a < 10; a++)
            for (int b = a; b < 10; b++) blocks.push_back(upper_key(ec[a], ec[b]));
        }
        std::sort(blocks.begin(), blocks.end());
        blocks.erase(std::unique(blocks.begin(), blocks.end()), blocks.end());
        // key -> index in `blocks`: a small open-addressing table (the 360 look-ups per sub-patch were binary searches)
        constexpr int HT = 1024;
        static_assert(HT >= 4 * PK_CAPU, "hash table of the sub-patch blocks");
        int ht_idx[HT];
        uint64_t ht_key[HT];
        for (int q = 0; q < HT; q++) ht_idx[q] = -1, ht_key[q] = 0;
        auto ht_slot = [&](uint64_t key) {
          uint64_t h = key * 0x9e3779b97f4a7c15ull;
          int q = (int)(h >> 54);
          while (ht_idx[q] >= 0 && ht_key[q] != key) q = (q + 1) & (HT - 1);
          return q;
        };
        const bool use_ht = blocks.size() <= (size_t)HT / 2;
        if (use_ht)
          for (size_t u = 0; u < blocks.size(); u++) {
            const int q = ht_slot(blocks[u]);
            ht_key[q] = blocks[u];
            ht_idx[q] = (int)u;
          }
        auto block_index = [&](uint64_t key) {
          if (use_ht) return ht_idx[ht_slot(key)];
          return (int)(std::lower_bound(blocks.begin(), blocks.end(), key) - blocks.begin());
        };
        // the block sets that are added simultaneously: (element, t = ta*2+tb) over the 15 tiles
        std::vector<std::vector<int>> sets;
        sets.reserve((size_t)(S.end - S.begin) * 4);
        std::vector<int> bidx((size_t)(S.end - S.begin) * 60, -1);
        for (int i = S.begin; i < S.end; i++) {
          const int *ec = ecoef + (size_t)order[i].second * 10;
          for (int tt = 0; tt < 4; tt++) {
            std::vector<int> st;
            st.reserve(15);
            for (int t = 0; t < 15; t++) {
              int I, J;
              tile_ij(t, &I, &J);
              if (I == J && tt == 2) continue;
              const int a = 2 * I + (tt >> 1), b = 2 * J + (tt & 1);
              const int bi = block_index(upper_key(ec[a], ec[b]));
              bidx[(size_t)(i - S.begin) * 60 + t * 4 + tt] = bi;
              st.push_back(bi);
            }
            sets.push_back(std::move(st));
          }
        }
        TT(0);
        std::vector<int> slot;
        assign_slots((int)blocks.size(), sets, slot);
        TT(1);
        std::vector<char> touched(blocks.size(), 0);
        for (int i = S.begin; i < S.end; i++) {
          const int e = order[i].second;
          const int *ec = ecoef + (size_t)e * 10;
          PElem pe;
          std::memset(&pe, 0, sizeof(pe));
          pe.elem = e;
          for (int a = 0; a < 10; a++) pe.node[a] = ec[a];
          for (int t = 0; t < 15; t++) {
            int I, J;
            tile_ij(t, &I, &J);
            uint32_t word = 0;
            for (int tt = 0; tt < 4; tt++) {
              const int a = 2 * I + (tt >> 1), b = 2 * J + (tt & 1), bit = t * 4 + tt;
              uint32_t u = 255;
              if (!(I == J && tt == 2)) {
                const int bi = bidx[(size_t)(i - S.begin) * 60 + t * 4 + tt];
                u = (uint32_t)slot[(size_t)bi];
                if (ec[a] > ec[b]) pe.tr[bit >> 5] |= 1u << (bit & 31);
                if (!touched[(size_t)bi]) touched[(size_t)bi] = 1, pe.first[bit >> 5] |= 1u << (bit & 31);
              }
              word |= u << (8 * tt);
            }
            pe.blk4[t] = word;
          }
          P.elems.push_back(pe);
        }
        for (size_t u = 0; u < blocks.size(); u++)
          tups.push_back({blocks[u], (uint16_t)(w * 9 * PK_CAPU + 9 * slot[u])});
      }
      TT(2);
      P.sub_ptr[ns] = (int)P.elems.size();
      for (int w = ns; w < PK_MAXW; w++) P.sub_ptr[w + 1] = P.sub_ptr[ns];
      // pairs held by several warps: reduced into the lowest warp's block, one extra warp per round
      std::sort(tups.begin(), tups.end(), [](const Tup &a, const Tup &b) {
        return a.key != b.key ? a.key < b.key : a.s < b.s;
      });
      int n_mergedst = 0, n_merged_total = 0;
      P.ent.reserve(2 * tups.size());
      P.nodes.reserve(2 * tups.size());
      for (size_t k = 0; k < tups.size();) {  // merged pairs of the patch (dealt to the chunks in contiguous runs)
        size_t k2 = k;
        while (k2 < tups.size() && tups[k2].key == tups[k].key) k2++;
        n_merged_total += (k2 - k > 1);
        k = k2;
      }
      for (size_t k = 0; k < tups.size();) {
        size_t k2 = k;
        while (k2 < tups.size() && tups[k2].key == tups[k].key) k2++;
        if (k2 - k > 8) P.bad = true;
        int seg = (int)(tups[k].s / (9 * PK_CAPU));
        if (k2 - k > 1) {  // merged pair: dealt round-robin to the chunks
          const int g = (int)((int64_t)(n_mergedst++) * W / n_merged_total);  // key order = row order: a chunk gets whole rows
          for (size_t q = k + 1; q < k2 && q - k <= 7; q++)
            P.ops[g].push_back((uint32_t)tups[k].s | (q > k + 1 ? 0x8000u : 0u) | ((uint32_t)tups[q].s << 16));
          seg = PK_MAXW + g;
        }
        const int i = (int)(tups[k].key >> 32), j = (int)(tups[k].key & 0xffffffffu);
        for (int side = 0; side < (i == j ? 1 : 2); side++) {
          const int r = side ? j : i, c = side ? i : j;
          const int deg = off[r + 1] - off[r];
          const int *cb = cols + off[r], *ce = cb + deg;
          const int *it = std::lower_bound(cb, ce, c);
          const int slot = (int)(it - cb);
          if (it == ce || *it != c || 3 * deg > PK_MAX_ROWLEN) P.bad = true;
          P.ent.push_back({9 * (int64_t)off[r] + 3 * slot, (int64_t)off[r] + slot, 3 * deg,
                           (uint16_t)(tups[k].s | (side ? 0x8000u : 0u)), false, seg});
          P.nodes.push_back(r);
        }
        k = k2;
      }
      TT(3);
      for (int g = 0; g < PK_MAXW; g++) P.n_ops += (int)P.ops[g].size();
      std::sort(P.ent.begin(), P.ent.end(), [](const EntL &a, const EntL &b) { return a.habs < b.habs; });
      std::sort(P.nodes.begin(), P.nodes.end());
      P.nodes.erase(std::unique(P.nodes.begin(), P.nodes.end()), P.nodes.end());
      any_bad = any_bad || P.bad;
      TT(4);
      plist.push_back(std::move(P));
      s0 += ns;
    }
  }
  if (any_bad)
    return failm(TLFEA_ERR_UNSUPPORTED, "patch plan: a node pair is missing from the pattern or a row is too long");
  std::vector<std::pair<uint64_t, int>>().swap(order);

  if (timing) std::fprintf(stderr, "  phase 3 cpu-seconds: blocks+sets %.2f slots %.2f pelem %.2f tups+entries %.2f sort %.2f greedy %.2f\n", tacc[0], tacc[1], tacc[2], tacc[3], tacc[4], tacc[5]);
  lap("3 sub-patches, slots, entries");
  // ---- 4. colour the patch adjacency (patches sharing a coefficient), first fit in Morton order
  std::vector<PatchLocal *> pl;
  for (auto &v : per_oct)
    for (auto &p : v) pl.push_back(&p);
  const int np = (int)pl.size();
  std::vector<int> nptr((size_t)n_coef + 1, 0);
  for (int p = 0; p < np; p++)
    for (int nd : pl[p]->nodes) nptr[(size_t)nd + 1]++;
  for (int i = 0; i < n_coef; i++) nptr[i + 1] += nptr[i];
  std::vector<int> npat((size_t)nptr[n_coef]);
  {
    std::vector<int> fill(nptr.begin(), nptr.end() - 1);
    for (int p = 0; p < np; p++)
      for (int nd : pl[p]->nodes) npat[fill[nd]++] = p;
  }
  int n_colors = 0;
  for (int p = 0; p < np; p++) {
    uint64_t used = 0;
    for (int nd : pl[p]->nodes)
      for (int k = nptr[nd]; k < nptr[nd + 1]; k++) {
        const int c = pl[npat[k]]->color;
        if (c >= 0) used |= 1ULL << c;
      }
    int c = 0;
    while (c < 64 && (used >> c & 1)) c++;
    if (c >= 64) return failm(TLFEA_ERR_UNSUPPORTED, "patch plan: more than 64 colours");
    pl[p]->color = c;
    n_colors = std::max(n_colors, c + 1);
  }
  std::vector<int>().swap(npat);
  std::vector<int>().swap(nptr);

  lap("4 colouring");
  // ---- 5. launch order, first-touch flags, concatenation.
  // Launch l = chunk * n_colors + colour.  Chunk c holds the patches whose LARGEST coefficient id falls into the c-th
  // of n_chunks id ranges (pk_chunk_bound: short first and last range): its kernels read no coordinate beyond that range (a host-buffer step starts them as
  // soon as that part of x, y, z has been uploaded), and once a chunk is done every coefficient whose patches all
  // lie in chunks <= c is final (its force can travel back to the host while later chunks compute).
  // n_chunks < 0: |n_chunks| is an upper bound -- every (chunk, colour) launch should give each of the ~300 resident CTAs a run of
  // patches long enough to amortise its start (small meshes and the slabs of a many-GPU partition get fewer chunks)
  const int K = (n_chunks < 0) ? std::max(1, std::min({-n_chunks, 16, np / 40000})) : std::max(1, std::min(n_chunks, 16));
  std::vector<int> chunk_of((size_t)np, 0), node_final((size_t)n_coef, -1);
  for (int p = 0; p < np; p++) {
    const int mxid = pl[p]->nodes.empty() ? 0 : pl[p]->nodes.back();
    int r = 0;
    while (r + 1 < K && mxid >= pk_chunk_bound(n_coef, K, r + 1)) r++;
    chunk_of[p] = r;
  }
  for (int p = 0; p < np; p++)
    for (int nd : pl[p]->nodes) node_final[nd] = std::max(node_final[nd], chunk_of[p]);
  std::vector<int> perm((size_t)np);
  for (int p = 0; p < np; p++) perm[p] = p;
  std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) {
    return chunk_of[a] != chunk_of[b] ? chunk_of[a] < chunk_of[b] : pl[a]->color < pl[b]->color;
  });
  PatchPlanHost &R = *out;
  R = PatchPlanHost();
  R.W = W;
  R.n_patches = np;
  R.n_colors = n_colors;
  R.n_chunks = K;
  const int n_launch = K * n_colors;
  R.color_ptr.assign((size_t)n_launch + 1, 0);
  for (int p = 0; p < np; p++) R.color_ptr[(size_t)chunk_of[p] * n_colors + pl[p]->color + 1]++;
  for (int c = 0; c < n_launch; c++) R.color_ptr[c + 1] += R.color_ptr[c];
  R.range_ready.assign((size_t)K * PK_GSUB, -1);
  for (int q = 0; q < K * PK_GSUB; q++) {
    const int64_t i0 = pk_sub_bound(n_coef, K, q), i1 = pk_sub_bound(n_coef, K, q + 1);
    int m = -1;
    for (int64_t i = i0; i < i1; i++) m = std::max(m, node_final[(size_t)i]);
    R.range_ready[q] = m;
  }

  const int64_t nnz = off[n_coef];
  std::vector<uint8_t> covered((size_t)nnz, 0);
  for (int c = 0; c < n_launch; c++) {
    // patches of one launch share no coefficient, hence no pattern block: no two threads touch the same byte
#pragma omp parallel for schedule(dynamic, 64)
    for (int q = R.color_ptr[c]; q < R.color_ptr[c + 1]; q++) {
      PatchLocal &P = *pl[perm[q]];
      for (auto &en : P.ent) {
        uint8_t &cv = covered[(size_t)en.g];
        if (!cv) {
          cv = 1;
          en.first = true;
        }
      }
      // by segment, first touches first, then (row, column)
      std::stable_sort(P.ent.begin(), P.ent.end(), [](const EntL &a, const EntL &b) {
        return a.seg != b.seg ? a.seg < b.seg : a.first > b.first;
      });
    }
  }
  for (int64_t g = 0; g < nnz; g++)
    if (!covered[(size_t)g]) R.uncovered.push_back(g);
  std::vector<uint8_t>().swap(covered);

  std::vector<int64_t> e_off((size_t)np + 1, 0), b_off((size_t)np + 1, 0), r_off((size_t)np + 1, 0);
  int64_t sum_local = 0;
  int max_blob = 0;
  bool span_bad = false;
  // blob of a patch: header | chunks (patch_plan.cuh)
  auto blob_layout = [](const PatchLocal &P, int *hdr) {  // returns the size; hdr (PK_HDR_INTS ints) may be null
    int nf[2 * PK_MAXW] = {0}, nl[2 * PK_MAXW] = {0};
    for (auto &en : P.ent) (en.first ? nf : nl)[en.seg]++;
    int pos = PK_HDR_INTS * 4;
    if (hdr) std::memset(hdr, 0, PK_HDR_INTS * sizeof(int));
    for (int w = 0; w < PK_MAXW; w++) {
      if (hdr) {
        int *hw = hdr + 8 * w;
        hw[0] = pos / 16, hw[1] = nf[w], hw[2] = nl[w], hw[3] = nf[PK_MAXW + w], hw[4] = nl[PK_MAXW + w];
        hw[5] = (int)P.ops[w].size();
      }
      pos += 16 * nf[w] + pk_al16(8 * nl[w]) + 16 * nf[PK_MAXW + w] + pk_al16(8 * nl[PK_MAXW + w]) +
             pk_al16(4 * (int)P.ops[w].size());
    }
    return pos;
  };
  for (int q = 0; q < np; q++) {
    const PatchLocal &P = *pl[perm[q]];
    const int total = blob_layout(P, nullptr);
    e_off[q + 1] = e_off[q] + (int64_t)P.elems.size();
    b_off[q + 1] = b_off[q] + total;
    r_off[q + 1] = r_off[q] + (int64_t)P.nodes.size();
    sum_local += (int64_t)P.ent.size();
    max_blob = std::max(max_blob, total);
  }
  if (b_off[np] / 16 > 4294967295LL) return failm(TLFEA_ERR_UNSUPPORTED, "patch plan: blob exceeds 64 GB");
  R.desc.resize((size_t)np);
  R.elems.resize((size_t)e_off[np]);
  R.blob.assign((size_t)b_off[np], 0);
  R.row_node.resize((size_t)r_off[np]);
  R.row_ptr.assign(r_off.begin(), r_off.end());
  R.sum_local_blocks = sum_local;
  R.max_blob_bytes = max_blob;
#pragma omp parallel for schedule(dynamic, 64) reduction(|| : span_bad)
  for (int q = 0; q < np; q++) {
    const PatchLocal &P = *pl[perm[q]];
    int nf = 0;
    int64_t hmin = INT64_MAX, hmax = 0;
    for (auto &en : P.ent) nf += en.first, hmin = std::min(hmin, en.habs), hmax = std::max(hmax, en.habs);
    if (P.ent.empty()) hmin = 0;
    span_bad = span_bad || (hmax - hmin > 4294967295LL);
    int hdr[PK_HDR_INTS];
    const int total = blob_layout(P, hdr);
    PatchDesc d;
    std::memset(&d, 0, sizeof(d));
    for (int w = 0; w <= PK_MAXW; w++) d.sub_ptr[w] = (int)e_off[q] + P.sub_ptr[w];
    d.n_sub = P.n_sub;
    d.blob_bytes = total;
    d.n_first = nf;
    d.n_merge = P.n_ops;
    d.blob_off = (uint32_t)(b_off[q] / 16);
    d.hbase = hmin;
    R.desc[q] = d;
    std::copy(P.elems.begin(), P.elems.end(), R.elems.begin() + e_off[q]);
    std::copy(P.nodes.begin(), P.nodes.end(), R.row_node.begin() + r_off[q]);
    unsigned char *bl = R.blob.data() + b_off[q];
    std::memcpy(bl, hdr, sizeof(hdr));
    // entries are sorted by (segment, first touch before later touch, H position); segment = w (own) or PK_MAXW + w
    size_t own_at[PK_MAXW + 1], dep_at[PK_MAXW + 1];
    {
      size_t k = 0;
      for (int w = 0; w < PK_MAXW; w++) own_at[w] = k, k += (size_t)(hdr[8 * w + 1] + hdr[8 * w + 2]);
      for (int w = 0; w < PK_MAXW; w++) dep_at[w] = k, k += (size_t)(hdr[8 * w + 3] + hdr[8 * w + 4]);
    }
    auto emit = [&](unsigned char *at, size_t k, int nf_, int nl_) {
      PEntF *ef = reinterpret_cast<PEntF *>(at);
      PEnt *el = reinterpret_cast<PEnt *>(ef + nf_);
      for (int i = 0; i < nf_; i++, k++) {
        const EntL &en = P.ent[k];
        ef[i].hoff = (uint32_t)(en.habs - hmin);
        ef[i].rowlen = (uint16_t)en.rowlen;
        ef[i].src = en.src;
        ef[i].mass = 0.0;  // filled on the device from the assembled mass values (pk_fill_mass_kernel)
      }
      for (int i = 0; i < nl_; i++, k++) {
        const EntL &en = P.ent[k];
        el[i].hoff = (uint32_t)(en.habs - hmin);
        el[i].rowlen = (uint16_t)en.rowlen;
        el[i].src = en.src;
      }
      return at + 16 * nf_ + pk_al16(8 * nl_);
    };
    for (int w = 0; w < PK_MAXW; w++) {
      const int *hw = hdr + 8 * w;
      unsigned char *at = bl + (size_t)hw[0] * 16;
      at = emit(at, own_at[w], hw[1], hw[2]);
      at = emit(at, dep_at[w], hw[3], hw[4]);
      if (hw[5]) std::memcpy(at, P.ops[w].data(), (size_t)hw[5] * sizeof(uint32_t));
    }
  }
  lap("5 order, flags, blobs");
  if (span_bad) return failm(TLFEA_ERR_UNSUPPORTED, "patch plan: a patch spans more than 2^32 entries of H");
  return TLFEA_OK;
}

}  // namespace tlfea

// ---------------------------------------------------------------------------------------------
// Test hooks: the plan is pure host data, so the CPU test-suite builds it without a GPU and replays the
// kernel's accumulation order in numpy (tests/test_patch_plan.py).
// ---------------------------------------------------------------------------------------------
#include "common.cuh"
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

int tlfea_debug_patch_plan_build(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y,
                                 const double *z, const int *off, const int *cols, int W, int n_chunks, void **plan_out) {
  TL_CHECK(ecoef && x && y && z && off && cols && plan_out, "null input");
  auto *p = new tlfea::PatchPlanHost();
  std::string err;
  const int rc = tlfea::build_patch_plan(n_elem, n_coef, ecoef, x, y, z, off, cols, W, n_chunks, p, &err);
  if (rc != TLFEA_OK) {
    delete p;
    return tlfea::fail(rc, "%s", err.c_str());
  }
  *plan_out = p;
  return TLFEA_OK;
}

int tlfea_debug_patch_plan_sizes(void *plan, int64_t *sizes) {
  TL_CHECK(plan && sizes, "null input");
  auto *p = static_cast<tlfea::PatchPlanHost *>(plan);
  sizes[0] = p->n_patches;
  sizes[1] = (int64_t)p->color_ptr.size() - 1;  // launches = chunks * colours
  sizes[2] = (int64_t)p->elems.size();
  sizes[3] = (int64_t)p->row_node.size();
  sizes[4] = (int64_t)p->blob.size();
  sizes[5] = p->max_blob_bytes;
  sizes[6] = (int64_t)p->uncovered.size();
  sizes[7] = p->sum_local_blocks;
  return TLFEA_OK;
}

int tlfea_debug_patch_plan_copy(void *plan, int *color_ptr, void *desc, void *elems, int *row_node, int64_t *row_ptr,
                                void *blob, int64_t *uncovered) {
  TL_CHECK(plan, "null input");
  auto *p = static_cast<tlfea::PatchPlanHost *>(plan);
  if (color_ptr) std::memcpy(color_ptr, p->color_ptr.data(), p->color_ptr.size() * sizeof(int));
  if (desc) std::memcpy(desc, p->desc.data(), p->desc.size() * sizeof(tlfea::PatchDesc));
  if (elems) std::memcpy(elems, p->elems.data(), p->elems.size() * sizeof(tlfea::PElem));
  if (row_node) std::memcpy(row_node, p->row_node.data(), p->row_node.size() * sizeof(int));
  if (row_ptr) std::memcpy(row_ptr, p->row_ptr.data(), p->row_ptr.size() * sizeof(int64_t));
  if (blob) std::memcpy(blob, p->blob.data(), p->blob.size());
  if (uncovered) std::memcpy(uncovered, p->uncovered.data(), p->uncovered.size() * sizeof(int64_t));
  return TLFEA_OK;
}

int tlfea_set_host_threads(int n) {
  TL_CHECK(n >= 0, "thread count must be >= 0 (0 = the OpenMP default)");
#ifdef _OPENMP
  omp_set_num_threads(n > 0 ? n : omp_get_num_procs());
#endif
  return TLFEA_OK;
}

int tlfea_debug_patch_plan_free(void *plan) {
  delete static_cast<tlfea::PatchPlanHost *>(plan);
  return TLFEA_OK;
}

}  // extern "C"
