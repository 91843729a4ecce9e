// patch_plan.cu -- host-side analysis for the single-pass FEAT10 assembly (see patch_plan.cuh, pk_kernel.cuh).
// Pure host code (OpenMP over octants / patches); runs once per sparsity analysis.
#include "patch_plan.cuh"

#include <algorithm>
#include <cmath>
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

struct PatchLocal {
  std::vector<PElem> elems;
  int sub_ptr[PK_MAXW + 1] = {0};
  int n_sub = 0;
  std::vector<PRow> rows;
  std::vector<uint32_t> ent;
  std::vector<int64_t> ent_g;  // global pattern block of each entry
  std::vector<uint16_t> src;
  std::vector<int> nodes;      // sorted distinct coefficient ids
  int color = -1;
  bool bad = false;
};

}  // namespace

int build_patch_plan(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y, const double *z,
                     const int *off, const int *cols, int W, PatchPlanHost *out, std::string *err) {
  auto failm = [&](int code, const char *m) {
    if (err) *err = m;
    return code;
  };
  if (W < 1 || W > PK_MAXW) return failm(TLFEA_ERR_INVALID, "patch plan: W out of range");
  if (n_elem <= 0) return failm(TLFEA_ERR_INVALID, "patch plan: no elements");
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

  // ---- 2. octants (2x2x2 cells): the unit a patch never leaves
  std::vector<int> oct_ptr;
  for (int i = 0; i < n_elem; i++)
    if (i == 0 || (order[i].first >> 3) != (order[i - 1].first >> 3)) oct_ptr.push_back(i);
  oct_ptr.push_back(n_elem);
  const int n_oct = (int)oct_ptr.size() - 1;

  // ---- 3. sub-patches and patches per octant
  std::vector<std::vector<PatchLocal>> per_oct((size_t)n_oct);
  bool any_bad = false;
#pragma omp parallel for schedule(dynamic, 16) reduction(|| : any_bad)
  for (int o = 0; o < n_oct; o++) {
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

    auto &plist = per_oct[o];
    for (size_t s0 = 0; s0 < subs.size(); s0 += W) {
      const int ns = (int)std::min<size_t>(W, subs.size() - s0);
      PatchLocal P;
      P.n_sub = ns;
      struct Tup {
        int i, j;
        uint16_t s;
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
            for (int ta = 0; ta < 2; ta++)
              for (int tb = 0; tb < 2; tb++) {
                const int a = 2 * I + ta, b = 2 * J + tb, bit = t * 4 + ta * 2 + tb;
                uint32_t u = 255;
                if (!(I == J && ta == 1 && tb == 0)) {
                  const uint64_t k = upper_key(ec[a], ec[b]);
                  u = (uint32_t)(std::lower_bound(blocks.begin(), blocks.end(), k) - blocks.begin());
                  if (ec[a] > ec[b]) pe.tr[bit >> 5] |= 1u << (bit & 31);
                }
                word |= u << (8 * (ta * 2 + tb));
              }
            pe.blk4[t] = word;
          }
          P.elems.push_back(pe);
        }
        for (size_t u = 0; u < blocks.size(); u++) {
          const int i = (int)(blocks[u] >> 32), j = (int)(blocks[u] & 0xffffffffu);
          tups.push_back({i, j, (uint16_t)(u | (w << 8))});
          if (i != j) tups.push_back({j, i, (uint16_t)(u | (w << 8) | (1u << 11))});
        }
      }
      P.sub_ptr[ns] = (int)P.elems.size();
      for (int w = ns; w < PK_MAXW; w++) P.sub_ptr[w + 1] = P.sub_ptr[ns];
      std::sort(tups.begin(), tups.end(), [](const Tup &a, const Tup &b) {
        if (a.i != b.i) return a.i < b.i;
        if (a.j != b.j) return a.j < b.j;
        return (a.s & 0x7ff) < (b.s & 0x7ff);  // warp, then block: the combine order
      });
      for (size_t k = 0; k < tups.size();) {
        const int i = tups[k].i;
        PRow r;
        std::memset(&r, 0, sizeof(r));
        r.node = i;
        r.hbase = 9 * (int64_t)off[i];
        const int deg = off[i + 1] - off[i];
        r.rowlen = 3 * deg;
        r.mass_base = off[i];
        r.ent_begin = (int)P.ent.size();
        P.nodes.push_back(i);
        while (k < tups.size() && tups[k].i == i) {
          const int j = tups[k].j;
          size_t k2 = k;
          while (k2 < tups.size() && tups[k2].i == i && tups[k2].j == j) k2++;
          const int nsrc = (int)(k2 - k);
          const int *cb = cols + off[i], *ce = cb + deg;
          const int *it = std::lower_bound(cb, ce, j);
          const int slot = (int)(it - cb);
          if (it == ce || *it != j || slot > PK_MAX_SLOT || nsrc > 8 || P.src.size() > 32767) P.bad = true;
          P.ent.push_back((uint32_t)slot | ((uint32_t)(nsrc - 1) << 14) | ((uint32_t)P.src.size() << 17));
          P.ent_g.push_back((int64_t)off[i] + slot);
          for (size_t q = k; q < k2; q++) P.src.push_back(tups[q].s);
          k = k2;
        }
        r.n_ent = (int)P.ent.size() - r.ent_begin;
        P.rows.push_back(r);
      }
      any_bad = any_bad || P.bad;
      plist.push_back(std::move(P));
    }
  }
  if (any_bad)
    return failm(TLFEA_ERR_UNSUPPORTED, "patch plan: a node pair is missing from the pattern or a row is too long");
  std::vector<std::pair<uint64_t, int>>().swap(order);

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

  // ---- 5. launch order (colour, Morton), first-touch flags, concatenation
  std::vector<int> perm((size_t)np);
  for (int p = 0; p < np; p++) perm[p] = p;
  std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return pl[a]->color < pl[b]->color; });
  PatchPlanHost &R = *out;
  R = PatchPlanHost();
  R.W = W;
  R.n_patches = np;
  R.n_colors = n_colors;
  R.color_ptr.assign((size_t)n_colors + 1, 0);
  for (int p = 0; p < np; p++) R.color_ptr[(size_t)pl[p]->color + 1]++;
  for (int c = 0; c < n_colors; c++) R.color_ptr[c + 1] += R.color_ptr[c];

  const int64_t nnz = off[n_coef];
  std::vector<uint8_t> covered((size_t)nnz, 0);
  for (int c = 0; c < n_colors; c++) {
    // patches of one colour share no coefficient, hence no pattern block: no two threads touch the same byte
#pragma omp parallel for schedule(dynamic, 64)
    for (int q = R.color_ptr[c]; q < R.color_ptr[c + 1]; q++) {
      PatchLocal &P = *pl[perm[q]];
      for (size_t k = 0; k < P.ent.size(); k++) {
        uint8_t &cv = covered[(size_t)P.ent_g[k]];
        if (!cv) {
          cv = 1;
          P.ent[k] |= PK_ENT_FIRST;
        }
      }
    }
  }
  for (int64_t g = 0; g < nnz; g++)
    if (!covered[(size_t)g]) R.uncovered.push_back(g);
  std::vector<uint8_t>().swap(covered);

  std::vector<int64_t> e_off((size_t)np + 1, 0), r_off((size_t)np + 1, 0), n_off((size_t)np + 1, 0),
      s_off((size_t)np + 1, 0);
  for (int q = 0; q < np; q++) {
    const PatchLocal &P = *pl[perm[q]];
    e_off[q + 1] = e_off[q] + (int64_t)P.elems.size();
    r_off[q + 1] = r_off[q] + (int64_t)P.rows.size();
    n_off[q + 1] = n_off[q] + (int64_t)P.ent.size();
    s_off[q + 1] = s_off[q] + (int64_t)P.src.size();
  }
  if (n_off[np] > 2147483647LL || s_off[np] > 2147483647LL)
    return failm(TLFEA_ERR_UNSUPPORTED, "patch plan: entry list exceeds int32");
  R.desc.resize((size_t)np);
  R.elems.resize((size_t)e_off[np]);
  R.rows.resize((size_t)r_off[np]);
  R.ent.resize((size_t)n_off[np]);
  R.src.resize((size_t)s_off[np]);
  R.sum_local_blocks = n_off[np];
#pragma omp parallel for schedule(dynamic, 64)
  for (int q = 0; q < np; q++) {
    const PatchLocal &P = *pl[perm[q]];
    PatchDesc d;
    std::memset(&d, 0, sizeof(d));
    for (int w = 0; w <= PK_MAXW; w++) d.sub_ptr[w] = (int)e_off[q] + P.sub_ptr[w];
    d.row_begin = (int)r_off[q];
    d.n_rows = (int)P.rows.size();
    d.src_base = (int)s_off[q];
    d.n_sub = P.n_sub;
    R.desc[q] = d;
    std::copy(P.elems.begin(), P.elems.end(), R.elems.begin() + e_off[q]);
    for (size_t k = 0; k < P.rows.size(); k++) {
      PRow r = P.rows[k];
      r.ent_begin += (int)n_off[q];
      R.rows[(size_t)r_off[q] + k] = r;
    }
    std::copy(P.ent.begin(), P.ent.end(), R.ent.begin() + n_off[q]);
    std::copy(P.src.begin(), P.src.end(), R.src.begin() + s_off[q]);
  }
  return TLFEA_OK;
}

}  // namespace tlfea

// ---------------------------------------------------------------------------------------------
// Test hooks: the plan is pure host data, so the CPU test-suite builds it without a GPU and replays the
// kernel's accumulation order in numpy (tests/test_patch_plan.py).
// ---------------------------------------------------------------------------------------------
#include "common.cuh"

extern "C" {

int tlfea_debug_patch_plan_build(int n_elem, int n_coef, const int *ecoef, const double *x, const double *y,
                                 const double *z, const int *off, const int *cols, int W, void **plan_out) {
  TL_CHECK(ecoef && x && y && z && off && cols && plan_out, "null input");
  auto *p = new tlfea::PatchPlanHost();
  std::string err;
  const int rc = tlfea::build_patch_plan(n_elem, n_coef, ecoef, x, y, z, off, cols, W, p, &err);
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
  sizes[1] = p->n_colors;
  sizes[2] = (int64_t)p->elems.size();
  sizes[3] = (int64_t)p->rows.size();
  sizes[4] = (int64_t)p->ent.size();
  sizes[5] = (int64_t)p->src.size();
  sizes[6] = (int64_t)p->uncovered.size();
  sizes[7] = p->sum_local_blocks;
  return TLFEA_OK;
}

int tlfea_debug_patch_plan_copy(void *plan, int *color_ptr, void *desc, void *elems, void *rows, uint32_t *ent,
                                uint16_t *src, int64_t *uncovered) {
  TL_CHECK(plan, "null input");
  auto *p = static_cast<tlfea::PatchPlanHost *>(plan);
  if (color_ptr) std::memcpy(color_ptr, p->color_ptr.data(), p->color_ptr.size() * sizeof(int));
  if (desc) std::memcpy(desc, p->desc.data(), p->desc.size() * sizeof(tlfea::PatchDesc));
  if (elems) std::memcpy(elems, p->elems.data(), p->elems.size() * sizeof(tlfea::PElem));
  if (rows) std::memcpy(rows, p->rows.data(), p->rows.size() * sizeof(tlfea::PRow));
  if (ent) std::memcpy(ent, p->ent.data(), p->ent.size() * sizeof(uint32_t));
  if (src) std::memcpy(src, p->src.data(), p->src.size() * sizeof(uint16_t));
  if (uncovered) std::memcpy(uncovered, p->uncovered.data(), p->uncovered.size() * sizeof(int64_t));
  return TLFEA_OK;
}

int tlfea_debug_patch_plan_free(void *plan) {
  delete static_cast<tlfea::PatchPlanHost *>(plan);
  return TLFEA_OK;
}

}  // extern "C"
