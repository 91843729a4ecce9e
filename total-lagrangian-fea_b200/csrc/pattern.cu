// pattern.cu -- one-time integer set-up (bit-exact vs the reference):
//   * coefficient-level adjacency CSR = BuildMassCSRPattern (FEAT10Data.cu:372-440,
//     ANCF3243Data.cu:401-469): 64-bit keys (i<<32 | j), sort, unique.
//   * coefficient -> (element, local index) incidence in ascending element order,
//     and the per-visit slot map.  Together they replace the reference's
//     900 binary searches + 900 FP64 atomics per Gauss point
//     (FEAT10DataFunc.cuh:662-687) with a fixed, deterministic gather order.
#include <cub/cub.cuh>

#include "common.cuh"

namespace tlfea {

__global__ void k_pair_keys(const int *__restrict__ ecoef, int64_t total, int S, unsigned long long *keys) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  int64_t e = tid / (S * S);
  int r = (int)(tid % (S * S));
  unsigned i = (unsigned)ecoef[e * S + r / S];
  unsigned j = (unsigned)ecoef[e * S + r % S];
  keys[tid] = ((unsigned long long)i << 32) | (unsigned long long)j;
}

__global__ void k_visit_keys(const int *__restrict__ ecoef, int64_t total, unsigned long long *keys) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  keys[tid] = ((unsigned long long)(unsigned)ecoef[tid] << 32) | (unsigned long long)(unsigned)tid;  // tid = e*S+a
}

// offsets[r] = first index whose key's high word is >= r (keys sorted): no atomics, no scan.
__global__ void k_row_lower_bound(const unsigned long long *__restrict__ keys, int64_t n, int n_rows, int *offsets) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n_rows) return;
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if ((int)(keys[mid] >> 32) < r)
      lo = mid + 1;
    else
      hi = mid;
  }
  offsets[r] = (int)lo;
}

__global__ void k_low_words(const unsigned long long *__restrict__ keys, int64_t n, int *out) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid < n) out[tid] = (int)(keys[tid] & 0xffffffffULL);
}

__global__ void k_slot_map(const int *__restrict__ ecoef, const int *__restrict__ inc_ea, const int *__restrict__ off,
                           const int *__restrict__ cols, int64_t n_visits, int S, uint16_t *slot, int *err) {
  int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= n_visits * S) return;
  int64_t v = tid / S;
  int b = (int)(tid % S);
  int ea = inc_ea[v];
  int e = ea / S;
  int row = ecoef[ea];
  int target = ecoef[(int64_t)e * S + b];
  int lo = off[row], hi = off[row + 1] - 1, base = lo, pos = -1;
  while (lo <= hi) {
    int mid = lo + ((hi - lo) >> 1);
    int c = cols[mid];
    if (c == target) {
      pos = mid - base;
      break;
    }
    if (c < target)
      lo = mid + 1;
    else
      hi = mid - 1;
  }
  if (pos < 0 || pos > 65535) {
    atomicExch(err, 1);
    pos = 0;
  }
  slot[tid] = (uint16_t)pos;
}

__global__ void k_max_diff(const int *__restrict__ off, int n, int *out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int v = (i < n) ? off[i + 1] - off[i] : 0;
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, v);
}

// The ordered (atomic-free) assembly adds the S x S blocks of an element to S distinct slots per row: a collapsed
// element that repeats a coefficient id would lose updates, so it is rejected at set-up.
__global__ void k_check_distinct(const int *__restrict__ ecoef, int n_elem, int S, int n_coef, int *err) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_elem) return;
  const int *ec = ecoef + (int64_t)e * S;
  bool bad = false;
  for (int a = 0; a < S; a++) {
    bad = bad || ec[a] < 0 || ec[a] >= n_coef;
    for (int b = a + 1; b < S; b++) bad = bad || (ec[a] == ec[b]);
  }
  if (bad) atomicExch(err, 1);
}

static int sort_keys(unsigned long long *d_in, unsigned long long *d_out, int64_t n, int end_bit, cudaStream_t st) {
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

int build_pattern(const int *d_ecoef, int n_elem, int S, int n_coef, cudaStream_t st, Pattern *out,
                  const unsigned long long *h_extra, int64_t n_extra) {
  Pattern p;
  p.n_coef = n_coef;
  p.S = S;
  p.n_visits = (int64_t)n_elem * S;
  int coef_bits = 1;
  while ((1LL << coef_bits) < n_coef) coef_bits++;
  {
    int *d_bad = nullptr, h_bad = 0;
    TL_TRY(dmalloc(&d_bad, 1));
    TL_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
    TL_LAUNCH(k_check_distinct, blocks_for(n_elem, 256), 256, 0, st, d_ecoef, n_elem, S, n_coef, d_bad);
    TL_CUDA(cudaMemcpyAsync(&h_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    TL_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_bad);
    if (h_bad)
      return fail(TLFEA_ERR_INVALID,
                  "an element repeats a coefficient id or has one out of range (the ordered assembly needs pairwise "
                  "distinct ids per element)");
  }

  // ---- adjacency: sorted unique (i,j)
  {
    const int64_t n_pairs = (int64_t)n_elem * S * S;
    const int64_t total = n_pairs + n_extra;  // extra (i, j) couplings: the J^T J fill of linear constraints
    unsigned long long *k0 = nullptr, *k1 = nullptr;
    TL_TRY(dmalloc(&k0, (size_t)total));
    TL_TRY(dmalloc(&k1, (size_t)total));
    TL_LAUNCH(k_pair_keys, blocks_for(n_pairs, 256), 256, 0, st, d_ecoef, n_pairs, S, k0);
    if (n_extra > 0)
      TL_CUDA(cudaMemcpyAsync(k0 + n_pairs, h_extra, (size_t)n_extra * sizeof(unsigned long long),
                              cudaMemcpyHostToDevice, st));
    TL_TRY(sort_keys(k0, k1, total, 32 + coef_bits, st));
    int64_t *d_num = nullptr;
    TL_TRY(dmalloc(&d_num, 1));
    size_t tmp_bytes = 0;
    TL_CUDA(cub::DeviceSelect::Unique(nullptr, tmp_bytes, k1, k0, d_num, total, st));
    void *tmp = nullptr;
    TL_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
    TL_CUDA(cub::DeviceSelect::Unique(tmp, tmp_bytes, k1, k0, d_num, total, st));
    g_launches.fetch_add(1);
    int64_t nnz = 0;
    TL_CUDA(cudaMemcpyAsync(&nnz, d_num, sizeof(nnz), cudaMemcpyDeviceToHost, st));
    TL_CUDA(cudaStreamSynchronize(st));
    cudaFree(tmp);
    cudaFree(d_num);
    cudaFree(k1);
    if (nnz > 2147483647LL) {
      cudaFree(k0);
      return fail(TLFEA_ERR_UNSUPPORTED, "coefficient-level nnz %lld exceeds int32", (long long)nnz);
    }
    p.nnz = nnz;
    TL_TRY(dmalloc(&p.d_off, (size_t)n_coef + 1));
    TL_TRY(dmalloc(&p.d_cols, (size_t)nnz));
    TL_LAUNCH(k_low_words, blocks_for(nnz, 256), 256, 0, st, k0, nnz, p.d_cols);
    TL_LAUNCH(k_row_lower_bound, blocks_for(n_coef + 1, 256), 256, 0, st, k0, nnz, n_coef, p.d_off);
    TL_CUDA(cudaStreamSynchronize(st));
    cudaFree(k0);
  }
  // ---- incidence: visits sorted by (coefficient, e*S+a)
  {
    int64_t total = p.n_visits;
    if (total > 2147483647LL) return fail(TLFEA_ERR_UNSUPPORTED, "n_elem*S exceeds int32");
    unsigned long long *k0 = nullptr, *k1 = nullptr;
    TL_TRY(dmalloc(&k0, (size_t)total));
    TL_TRY(dmalloc(&k1, (size_t)total));
    TL_LAUNCH(k_visit_keys, blocks_for(total, 256), 256, 0, st, d_ecoef, total, k0);
    TL_TRY(sort_keys(k0, k1, total, 32 + coef_bits, st));
    TL_TRY(dmalloc(&p.d_inc_ptr, (size_t)n_coef + 1));
    TL_TRY(dmalloc(&p.d_inc_ea, (size_t)total));
    TL_LAUNCH(k_low_words, blocks_for(total, 256), 256, 0, st, k1, total, p.d_inc_ea);
    TL_LAUNCH(k_row_lower_bound, blocks_for(n_coef + 1, 256), 256, 0, st, k1, total, n_coef, p.d_inc_ptr);
    TL_CUDA(cudaStreamSynchronize(st));
    cudaFree(k0);
    cudaFree(k1);
  }
  // ---- slot map + row statistics
  {
    int *d_err = nullptr;
    TL_TRY(dmalloc(&d_err, 3));
    TL_CUDA(cudaMemsetAsync(d_err, 0, 3 * sizeof(int), st));
    TL_TRY(dmalloc(&p.d_slot, (size_t)p.n_visits * S));
    TL_LAUNCH(k_slot_map, blocks_for(p.n_visits * S, 256), 256, 0, st, d_ecoef, p.d_inc_ea, p.d_off, p.d_cols,
              p.n_visits, S, p.d_slot, d_err);
    TL_LAUNCH(k_max_diff, blocks_for(n_coef, 256), 256, 0, st, p.d_off, n_coef, d_err + 1);
    TL_LAUNCH(k_max_diff, blocks_for(n_coef, 256), 256, 0, st, p.d_inc_ptr, n_coef, d_err + 2);
    int h[3];
    TL_CUDA(cudaMemcpyAsync(h, d_err, sizeof(h), cudaMemcpyDeviceToHost, st));
    TL_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_err);
    if (h[0]) return fail(TLFEA_ERR_INVALID, "slot map: column not found or row longer than 65535");
    p.max_deg = h[1];
    p.max_inc = h[2];
  }
  *out = p;
  return TLFEA_OK;
}

}  // namespace tlfea
