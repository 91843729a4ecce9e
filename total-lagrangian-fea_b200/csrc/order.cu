// order.cu -- one-time row orderings for the two-phase assembly (ek_kernel.cuh): Morton order of the
// coefficient rows by reference position, and the completion order (rows sorted by their last incident
// element, Morton ties) that lets a chunked schedule reduce rows right after their elements were computed.
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
