// tiles.cu -- one-time analysis for the tile assembly kernel (tile_kernel.cuh).
//
// Rows (coefficients) are grouped into spatially compact TILES by sorting them along a Morton
// curve of their reference position; a tile is a run of that order whose H block rows fit the
// shared-memory row buffer.  For every tile we list the elements incident to its rows
// ("entries") with, per local shape function a, the local row index inside the tile (or 255)
// and the commit rank of that contribution inside its row.  The commit rank fixes the order in
// which contributions are added to a row => bit-reproducible assembly without FP64 atomics
// (the reference uses atomicAdd per entry: FEAT10DataFunc.cuh:662-687).
#include <cub/cub.cuh>

#include <vector>

#include "common.cuh"

namespace tlfea {

__global__ void k_minmax3(int n, int cpn, const double *__restrict__ x, const double *__restrict__ y,
                          const double *__restrict__ z, unsigned long long *out /*[6] ordered-uint encodings*/) {
  // order-preserving map double -> uint64
  auto enc = [](double v) {
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u & 0x8000000000000000ULL) ? ~u : (u | 0x8000000000000000ULL);
  };
  unsigned long long mn[3] = {~0ULL, ~0ULL, ~0ULL}, mx[3] = {0, 0, 0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int c = i - i % cpn;
    const unsigned long long e[3] = {enc(x[c]), enc(y[c]), enc(z[c])};
#pragma unroll
    for (int k = 0; k < 3; k++) {
      mn[k] = min(mn[k], e[k]);
      mx[k] = max(mx[k], e[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[k] = min(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = max(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(out + k, mn[k]);
      atomicMax(out + 3 + k, mx[k]);
    }
  }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {
  v &= 0x1fffffULL;
  v = (v | v << 32) & 0x1f00000000ffffULL;
  v = (v | v << 16) & 0x1f0000ff0000ffULL;
  v = (v | v << 8) & 0x100f00f00f00f00fULL;
  v = (v | v << 4) & 0x10c30c30c30c30c3ULL;
  v = (v | v << 2) & 0x1249249249249249ULL;
  return v;
}

__global__ void k_morton(int n, int cpn, const double *__restrict__ x, const double *__restrict__ y,
                         const double *__restrict__ z, const unsigned long long *__restrict__ mm,
                         unsigned long long *keys, int *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  auto dec = [](unsigned long long u) {
    u = (u & 0x8000000000000000ULL) ? (u & 0x7fffffffffffffffULL) : ~u;
    return __longlong_as_double((long long)u);
  };
  const int c = i - i % cpn;
  const double p[3] = {x[c], y[c], z[c]};
  // one common cell size for the three axes keeps tiles cubic on anisotropic boxes
  double ext = 0.0;
#pragma unroll
  for (int k = 0; k < 3; k++) ext = fmax(ext, dec(mm[3 + k]) - dec(mm[k]));
  const double scale = (ext > 0.0) ? 2097151.0 / ext : 0.0;
  unsigned long long key = 0;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    double q = (p[k] - dec(mm[k])) * scale;
    q = fmin(fmax(q, 0.0), 2097151.0);
    key |= spread21((unsigned long long)q) << (2 - k);
  }
  keys[i] = key;
  vals[i] = i;
}

__global__ void k_perm_deg(int n, const int *__restrict__ trow, const int *__restrict__ off, int *deg) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) deg[i] = off[trow[i] + 1] - off[trow[i]];
}

__global__ void k_row_tiles(int n, int n_tiles, const int *__restrict__ trow, const int *__restrict__ tile_ptr,
                            int *tile_of_row, uint8_t *lrow_of_row) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int lo = 0, hi = n_tiles;  // last t with tile_ptr[t] <= i
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (tile_ptr[mid] <= i)
      lo = mid;
    else
      hi = mid;
  }
  tile_of_row[trow[i]] = lo;
  lrow_of_row[trow[i]] = (uint8_t)(i - tile_ptr[lo]);
}

__global__ void k_entry_keys(int64_t n_visits, int S, const int *__restrict__ inc_ea, const int *__restrict__ ecoef,
                             const int *__restrict__ tile_of_row, unsigned long long *keys, int *vidx) {
  int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_visits) return;
  const int ea = inc_ea[v];
  vidx[ea] = (int)v;
  keys[v] = ((unsigned long long)(unsigned)tile_of_row[ecoef[ea]] << 32) | (unsigned)(ea / S);
}

__global__ void k_split_words(const unsigned long long *__restrict__ keys, int64_t n, int *hi, int *lo) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  hi[i] = (int)(keys[i] >> 32);
  lo[i] = (int)(keys[i] & 0xffffffffULL);
}

__global__ void k_tile_lower_bound(const int *__restrict__ te_tile, int n, int n_tiles, int *te_ptr) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > n_tiles) return;
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (te_tile[mid] < t)
      lo = mid + 1;
    else
      hi = mid;
  }
  te_ptr[t] = lo;
}

// One descriptor per visit (row r of the tile, element e, local shape function a), grouped by
// (batch of the tile's element list, owner warp of the row):  elem-in-batch | a<<8 | local row<<16.
__global__ void k_visit_desc(int64_t n_visits, int S, int B, int W, const int *__restrict__ inc_ea,
                             const int *__restrict__ ecoef, const int *__restrict__ tile_of_row,
                             const uint8_t *__restrict__ lrow_of_row, const int *__restrict__ te_ptr,
                             const int *__restrict__ te_elem, const int *__restrict__ tb_ptr,
                             unsigned long long *keys, unsigned *desc, int *vals) {
  int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_visits) return;
  const int ea = inc_ea[v];
  const int e = ea / S, a = ea - e * S;
  const int r = ecoef[ea];
  const int t = tile_of_row[r], lr = lrow_of_row[r];
  int lo = te_ptr[t], hi = te_ptr[t + 1];
  const int base = lo;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (te_elem[mid] < e)
      lo = mid + 1;
    else
      hi = mid;
  }
  const int j = lo - base;
  const unsigned long long gbw = (unsigned long long)(tb_ptr[t] + j / B) * W + ((lr + lr / W) % W)  /* rotate: Morton-local indices that are multiples of 8 are the hub (vertex) rows */;
  const int ebi = j % B;
  keys[v] = (gbw << 16) | (unsigned)(ebi << 8) | (unsigned)a;
  desc[v] = (unsigned)ebi | ((unsigned)a << 8) | ((unsigned)lr << 16);
  vals[v] = (int)v;
}

// per-visit records in (batch, warp) order: descriptor, e*S+a, and the S column slots of the visit
__global__ void k_visit_fill(int64_t n_visits, int S, const int *__restrict__ order, const unsigned *__restrict__ desc0,
                             const int *__restrict__ inc_ea, const uint16_t *__restrict__ slot, unsigned *desc,
                             int *vea, uint16_t *vslot) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= n_visits * S) return;
  const int64_t i = tid / S;
  const int b = (int)(tid - i * S);
  const int v = order[i];
  vslot[tid] = slot[(int64_t)v * S + b];
  if (b == 0) {
    desc[i] = desc0[v];
    vea[i] = inc_ea[v];
  }
}

__global__ void k_tile_rows(int n, const int *__restrict__ trow, const int *__restrict__ off, int *trs, int *tdeg) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int r = trow[i];
  trs[i] = off[r];
  tdeg[i] = off[r + 1] - off[r];
}

__global__ void k_gbw_lower_bound(const unsigned long long *__restrict__ keys, int64_t n, int64_t n_groups, int *ptr) {
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g > n_groups) return;
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if ((int64_t)(keys[mid] >> 16) < g)
      lo = mid + 1;
    else
      hi = mid;
  }
  ptr[g] = (int)lo;
}

template <typename K>
static int sort_keys_t(K *d_in, K *d_out, int64_t n, int end_bit, cudaStream_t st) {
  size_t tmp_bytes = 0;
  TL_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, d_in, d_out, n, 0, end_bit, st));
  void *tmp = nullptr;
  TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
  cudaError_t e = cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, d_in, d_out, n, 0, end_bit, st);
  g_launches.fetch_add(1);
  cudaStreamSynchronize(st);
  cudaFree(tmp);
  TL_CUDA(e);
  return TLFEA_OK;
}

void Tiles::release() {
  dfree(d_trow);
  dfree(d_tile_ptr);
  dfree(d_te_ptr);
  dfree(d_te_elem);
  dfree(d_tb_ptr);
  dfree(d_vptr);
  dfree(d_visit);
  dfree(d_vea);
  dfree(d_vslot);
  dfree(d_trs);
  dfree(d_tdeg);
  dfree(d_err);
  n_tiles = 0;
}

// budget_doubles: capacity of the shared-memory row buffer; max_rows <= 254
int build_tiles(const tlfea_elem *h, int budget_doubles, int max_rows, int batch, int warps, Tiles *out) {
  const Pattern &pat = h->pat;
  const int n = pat.n_coef, S = pat.S;
  cudaStream_t st = h->stream;
  Tiles t;
  TL_CHECK(pat.max_inc <= 255, "tile assembly needs <= 255 elements per coefficient (have %d)", pat.max_inc);
  TL_CHECK(pat.max_deg * 9 <= budget_doubles, "coefficient row too long for the tile row buffer (max_deg=%d)",
           pat.max_deg);
  const int cpn = (h->type == TLFEA_TYPE_T10) ? 1 : 4;

  // ---- 1. Morton order of the rows
  unsigned long long *d_mm = nullptr, *k0 = nullptr, *k1 = nullptr;
  int *v0 = nullptr;
  TL_TRY(dmalloc(&d_mm, 6));
  const unsigned long long init[6] = {~0ULL, ~0ULL, ~0ULL, 0, 0, 0};
  TL_CUDA(cudaMemcpyAsync(d_mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
  TL_LAUNCH(k_minmax3, 592, 256, 0, st, n, cpn, h->d_xr, h->d_yr, h->d_zr, d_mm);
  TL_TRY(dmalloc(&k0, (size_t)n));
  TL_TRY(dmalloc(&k1, (size_t)n));
  TL_TRY(dmalloc(&v0, (size_t)n));
  TL_TRY(dmalloc(&t.d_trow, (size_t)n));
  TL_LAUNCH(k_morton, blocks_for(n, 256), 256, 0, st, n, cpn, h->d_xr, h->d_yr, h->d_zr, d_mm, k0, v0);
  {
    size_t tmp_bytes = 0;
    TL_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, t.d_trow, n, 0, 63, st));
    void *tmp = nullptr;
    TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, t.d_trow, n, 0, 63, st);
    g_launches.fetch_add(1);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    TL_CUDA(e);
  }
  cudaFree(k0);
  cudaFree(k1);
  cudaFree(d_mm);

  // ---- 2. greedy cut of the Morton order into tiles (host; one pass over n ints)
  TL_LAUNCH(k_perm_deg, blocks_for(n, 256), 256, 0, st, n, t.d_trow, pat.d_off, v0);
  std::vector<int> deg(n), tile_ptr;
  TL_CUDA(cudaMemcpyAsync(deg.data(), v0, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
  TL_CUDA(cudaStreamSynchronize(st));
  cudaFree(v0);
  tile_ptr.push_back(0);
  {
    int rows = 0, used = 0, mx_rows = 0, mx_used = 0;
    for (int i = 0; i < n; i++) {
      const int need = deg[i] * 9;
      if (rows > 0 && (rows + 1 > max_rows || used + need > budget_doubles)) {
        tile_ptr.push_back(i);
        rows = 0;
        used = 0;
      }
      rows++;
      used += need;
      mx_rows = std::max(mx_rows, rows);
      mx_used = std::max(mx_used, used);
    }
    tile_ptr.push_back(n);
    t.max_rows = mx_rows;
    t.max_buf = mx_used;
  }
  t.n_tiles = (int)tile_ptr.size() - 1;
  TL_TRY(dmalloc(&t.d_tile_ptr, tile_ptr.size()));
  TL_CUDA(cudaMemcpyAsync(t.d_tile_ptr, tile_ptr.data(), tile_ptr.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  int *tile_of_row = nullptr;
  uint8_t *lrow_of_row = nullptr;
  TL_TRY(dmalloc(&tile_of_row, (size_t)n));
  TL_TRY(dmalloc(&lrow_of_row, (size_t)n));
  TL_LAUNCH(k_row_tiles, blocks_for(n, 256), 256, 0, st, n, t.n_tiles, t.d_trow, t.d_tile_ptr, tile_of_row,
            lrow_of_row);

  // ---- 3. entries = unique (tile, element) over all visits
  const int64_t nv = pat.n_visits;
  int *vidx = nullptr;
  TL_TRY(dmalloc(&vidx, (size_t)nv));
  TL_TRY(dmalloc(&k0, (size_t)nv));
  TL_TRY(dmalloc(&k1, (size_t)nv));
  TL_LAUNCH(k_entry_keys, blocks_for(nv, 256), 256, 0, st, nv, S, pat.d_inc_ea, h->d_ecoef, tile_of_row, k0, vidx);
  int tile_bits = 1;
  while ((1LL << tile_bits) < t.n_tiles) tile_bits++;
  TL_TRY(sort_keys_t(k0, k1, nv, 32 + tile_bits, st));
  int64_t *d_num = nullptr;
  TL_TRY(dmalloc(&d_num, 1));
  {
    size_t tmp_bytes = 0;
    TL_CUDA(cub::DeviceSelect::Unique(nullptr, tmp_bytes, k1, k0, d_num, nv, st));
    void *tmp = nullptr;
    TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
    TL_CUDA(cub::DeviceSelect::Unique(tmp, tmp_bytes, k1, k0, d_num, nv, st));
    g_launches.fetch_add(1);
    TL_CUDA(cudaMemcpyAsync(&t.n_entries, d_num, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    TL_CUDA(cudaStreamSynchronize(st));
    cudaFree(tmp);
  }
  cudaFree(d_num);
  cudaFree(k1);
  TL_CHECK(t.n_entries < 2147483647LL, "too many tile entries");
  int *te_tile = nullptr;
  TL_TRY(dmalloc(&te_tile, (size_t)t.n_entries));
  TL_TRY(dmalloc(&t.d_te_elem, (size_t)t.n_entries));
  TL_TRY(dmalloc(&t.d_te_ptr, (size_t)t.n_tiles + 1));
  TL_LAUNCH(k_split_words, blocks_for(t.n_entries, 256), 256, 0, st, k0, t.n_entries, te_tile, t.d_te_elem);
  TL_LAUNCH(k_tile_lower_bound, blocks_for(t.n_tiles + 1, 256), 256, 0, st, te_tile, (int)t.n_entries, t.n_tiles,
            t.d_te_ptr);
  // ---- 4. batches of the tile element lists and the per-(batch, warp) visit lists
  t.batch = batch;
  t.warps = warps;
  {
    std::vector<int> te_ptr_h(t.n_tiles + 1), tb(t.n_tiles + 1);
    TL_CUDA(cudaMemcpyAsync(te_ptr_h.data(), t.d_te_ptr, te_ptr_h.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
    TL_CUDA(cudaStreamSynchronize(st));
    tb[0] = 0;
    for (int i = 0; i < t.n_tiles; i++) {
      tb[i + 1] = tb[i] + (te_ptr_h[i + 1] - te_ptr_h[i] + batch - 1) / batch;
      t.max_entries = std::max(t.max_entries, te_ptr_h[i + 1] - te_ptr_h[i]);
    }
    t.n_batches = tb[t.n_tiles];
    TL_TRY(dmalloc(&t.d_tb_ptr, tb.size()));
    TL_CUDA(cudaMemcpyAsync(t.d_tb_ptr, tb.data(), tb.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    TL_CUDA(cudaStreamSynchronize(st));
  }
  {
    unsigned long long *vk0 = nullptr, *vk1 = nullptr;
    unsigned *d0 = nullptr;
    int *o0 = nullptr, *o1 = nullptr;
    TL_TRY(dmalloc(&vk0, (size_t)nv));
    TL_TRY(dmalloc(&vk1, (size_t)nv));
    TL_TRY(dmalloc(&d0, (size_t)nv));
    TL_TRY(dmalloc(&o0, (size_t)nv));
    TL_TRY(dmalloc(&o1, (size_t)nv));
    TL_LAUNCH(k_visit_desc, blocks_for(nv, 256), 256, 0, st, nv, S, batch, warps, pat.d_inc_ea, h->d_ecoef, tile_of_row,
              lrow_of_row, t.d_te_ptr, t.d_te_elem, t.d_tb_ptr, vk0, d0, o0);
    const int64_t n_groups = (int64_t)t.n_batches * warps;
    int gbits = 1;
    while ((1LL << gbits) < n_groups + 1) gbits++;
    size_t tmp_bytes = 0;
    TL_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, vk0, vk1, o0, o1, nv, 0, 16 + gbits, st));
    void *tmp = nullptr;
    TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
    cudaError_t er = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, vk0, vk1, o0, o1, nv, 0, 16 + gbits, st);
    g_launches.fetch_add(1);
    TL_TRY(dmalloc(&t.d_vptr, (size_t)n_groups + 1));
    TL_LAUNCH(k_gbw_lower_bound, blocks_for(n_groups + 1, 256), 256, 0, st, vk1, nv, n_groups, t.d_vptr);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    cudaFree(vk0);
    cudaFree(vk1);
    cudaFree(o0);
    TL_CUDA(er);
    TL_TRY(dmalloc(&t.d_visit, (size_t)nv));
    TL_TRY(dmalloc(&t.d_vea, (size_t)nv));
    TL_TRY(dmalloc(&t.d_vslot, (size_t)nv * S));
    TL_LAUNCH(k_visit_fill, blocks_for(nv * S, 256), 256, 0, st, nv, S, o1, d0, pat.d_inc_ea, pat.d_slot, t.d_visit,
              t.d_vea, t.d_vslot);
    cudaStreamSynchronize(st);
    cudaFree(o1);
    cudaFree(d0);
  }
  // ---- 5. CSR start and degree of the rows in tile order (no dependent loads in the kernel prologue)
  TL_TRY(dmalloc(&t.d_trs, (size_t)n));
  TL_TRY(dmalloc(&t.d_tdeg, (size_t)n));
  TL_LAUNCH(k_tile_rows, blocks_for(n, 256), 256, 0, st, n, t.d_trow, pat.d_off, t.d_trs, t.d_tdeg);
  TL_TRY(dmalloc(&t.d_err, 4 + 16));  // 4 flags + 8 x 64-bit phase counters
  TL_CUDA(cudaMemsetAsync(t.d_err, 0, 20 * sizeof(int), st));
  TL_CUDA(cudaGetLastError());
  TL_CUDA(cudaStreamSynchronize(st));
  cudaFree(k0);
  cudaFree(vidx);
  cudaFree(te_tile);
  cudaFree(tile_of_row);
  cudaFree(lrow_of_row);
  *out = t;
  return TLFEA_OK;
}

// Morton order of the coefficient rows only (what the row-reduce kernel of ek_kernel.cuh walks).
int build_row_order(const tlfea_elem *h, int **d_order) {
  const int n = h->pat.n_coef;
  cudaStream_t st = h->stream;
  const int cpn = (h->type == TLFEA_TYPE_T10) ? 1 : 4;
  unsigned long long *d_mm = nullptr, *k0 = nullptr, *k1 = nullptr;
  int *v0 = nullptr;
  TL_TRY(dmalloc(&d_mm, 6));
  const unsigned long long init[6] = {~0ULL, ~0ULL, ~0ULL, 0, 0, 0};
  TL_CUDA(cudaMemcpyAsync(d_mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
  TL_LAUNCH(k_minmax3, 592, 256, 0, st, n, cpn, h->d_xr, h->d_yr, h->d_zr, d_mm);
  TL_TRY(dmalloc(&k0, (size_t)n));
  TL_TRY(dmalloc(&k1, (size_t)n));
  TL_TRY(dmalloc(&v0, (size_t)n));
  TL_TRY(dmalloc(d_order, (size_t)n));
  TL_LAUNCH(k_morton, blocks_for(n, 256), 256, 0, st, n, cpn, h->d_xr, h->d_yr, h->d_zr, d_mm, k0, v0);
  size_t tmp_bytes = 0;
  TL_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, *d_order, n, 0, 63, st));
  void *tmp = nullptr;
  TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
  cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, *d_order, n, 0, 63, st);
  g_launches.fetch_add(1);
  cudaStreamSynchronize(st);
  cudaFree(tmp);
  cudaFree(k0);
  cudaFree(k1);
  cudaFree(v0);
  cudaFree(d_mm);
  TL_CUDA(e);
  return TLFEA_OK;
}

// Rows in COMPLETION order for the chunked two-phase assembly: primary key = id of the row's last
// (largest) incident element, ties in Morton order.  chunk_ptr[k] = first position whose key >= k*chunk,
// so the rows of [chunk_ptr[k], chunk_ptr[k+1]) become complete when elements [k*chunk, (k+1)*chunk)
// have been processed.  Rows without elements land in chunk 0.
__global__ void k_completion_keys(int n, int S, const int *__restrict__ order, const int *__restrict__ inc_ptr,
                                  const int *__restrict__ inc_ea, unsigned long long *keys, int *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int row = order[i];
  const int vb = inc_ptr[row], ve = inc_ptr[row + 1];
  const unsigned long long last = (ve > vb) ? (unsigned long long)(inc_ea[ve - 1] / S) : 0ULL;
  keys[i] = (last << 32) | (unsigned)i;
  vals[i] = row;
}
__global__ void k_chunk_lower_bound(const unsigned long long *__restrict__ keys, int n, int chunk, int n_chunks,
                                    int *ptr) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_chunks) return;
  const unsigned long long target = (unsigned long long)k * (unsigned long long)chunk;
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if ((keys[mid] >> 32) < target)
      lo = mid + 1;
    else
      hi = mid;
  }
  ptr[k] = (k == n_chunks) ? n : lo;
}

int build_completion_order(const tlfea_elem *h, int chunk, int **d_order, std::vector<int> *chunk_ptr) {
  const int n = h->pat.n_coef, S = h->pat.S;
  cudaStream_t st = h->stream;
  int *d_morton = nullptr;
  TL_TRY(build_row_order(h, &d_morton));
  unsigned long long *k0 = nullptr, *k1 = nullptr;
  int *v0 = nullptr, *d_ptr = nullptr;
  TL_TRY(dmalloc(&k0, (size_t)n));
  TL_TRY(dmalloc(&k1, (size_t)n));
  TL_TRY(dmalloc(&v0, (size_t)n));
  TL_TRY(dmalloc(d_order, (size_t)n));
  TL_LAUNCH(k_completion_keys, blocks_for(n, 256), 256, 0, st, n, S, d_morton, h->pat.d_inc_ptr, h->pat.d_inc_ea, k0, v0);
  size_t tmp_bytes = 0;
  TL_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, *d_order, n, 0, 64, st));
  void *tmp = nullptr;
  TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
  cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, *d_order, n, 0, 64, st);
  g_launches.fetch_add(1);
  const int n_chunks = (h->n_elem + chunk - 1) / chunk;
  TL_TRY(dmalloc(&d_ptr, (size_t)n_chunks + 1));
  TL_LAUNCH(k_chunk_lower_bound, blocks_for(n_chunks + 1, 256), 256, 0, st, k1, n, chunk, n_chunks, d_ptr);
  chunk_ptr->resize(n_chunks + 1);
  cudaMemcpyAsync(chunk_ptr->data(), d_ptr, (size_t)(n_chunks + 1) * sizeof(int), cudaMemcpyDeviceToHost, st);
  cudaStreamSynchronize(st);
  cudaFree(tmp);
  cudaFree(k0);
  cudaFree(k1);
  cudaFree(v0);
  cudaFree(d_ptr);
  cudaFree(d_morton);
  TL_CUDA(e);
  return TLFEA_OK;
}

}  // namespace tlfea
