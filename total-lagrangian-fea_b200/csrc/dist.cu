// dist.cu -- multi-GPU data plane of the partitioned assembly / Newton iteration INSIDE the library (new design: the
// reference is single-GPU).  One rank per GPU; every rank assembles its own elements (sub-assembled H, no matrix
// exchange).  What moves between ranks:
//   * halo sum: the coefficients a rank shares with a neighbour form contiguous id ranges (the partitioner numbers
//     interface planes first / last); each range is exchanged with exactly that neighbour -- ncclSend / ncclRecv inside
//     one group call on the handle's stream -- and added in neighbour order (deterministic: a + b on one side, b + a on
//     the other);
//   * all-reduce (sum) of a few device-resident doubles: ||g||^2, r.z, p.Ap.
// NCCL is loaded at run time (dlopen of libnccl.so.2: inside a torch process this is the copy torch already loaded), so
// libtlfea_b200.so keeps linking against cudart only and single-GPU users need no NCCL at all.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace tlfea {

struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi *nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.lib ? &api : nullptr;
  tried = true;
  for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
    api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) return nullptr;
#define TL_SYM(field, sym)                                                  \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, sym)); \
  if (!api.field) {                                                         \
    api.lib = nullptr;                                                      \
    return nullptr;                                                         \
  }
  TL_SYM(GetUniqueId, "ncclGetUniqueId")
  TL_SYM(CommInitRank, "ncclCommInitRank")
  TL_SYM(CommDestroy, "ncclCommDestroy")
  TL_SYM(GroupStart, "ncclGroupStart")
  TL_SYM(GroupEnd, "ncclGroupEnd")
  TL_SYM(Send, "ncclSend")
  TL_SYM(Recv, "ncclRecv")
  TL_SYM(AllReduce, "ncclAllReduce")
  TL_SYM(GetErrorString, "ncclGetErrorString")
#undef TL_SYM
  return &api;
}

#define TL_NCCL(api, call)                                                                                        \
  do {                                                                                                            \
    ncclResult_t _r = (call);                                                                                     \
    if (_r != ncclSuccess) return fail(TLFEA_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, (api)->GetErrorString(_r)); \
  } while (0)

struct DistCtx {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  struct Nbr {
    int rank;
    int64_t c_lo, c_hi;  // shared coefficient ids [c_lo, c_hi)
    double *d_recv = nullptr;
    // peer-memory path (tlfea_dist_p2p_*): the mailbox the neighbour writes into (ours, exported by IPC handle) and the
    // neighbour's mailbox for messages from us (opened from its handle).  Layout: two 128-byte flag lines, then two
    // slots of (c_hi - c_lo) * HALO_MAX_WIDTH doubles; message q uses slot q & 1 and sets that slot's flag to q.
    unsigned char *d_mail = nullptr;   // local
    unsigned char *d_peer = nullptr;   // in the neighbour's memory
    unsigned int *d_done = nullptr;    // CTAs of the current put that have finished their part
  };
  std::vector<Nbr> nbr;
  int64_t halo_bytes = 0;  // bytes sent per halo sum of width 3 (reporting)
  unsigned long long seq = 0;  // halo sums done over the peer-memory path
  bool p2p = false;            // every neighbour connected
};

static constexpr int HALO_MAX_WIDTH = 9;

__global__ void k_halo_add(int64_t n, const double *__restrict__ buf, double *vec) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) vec[i] += buf[i];
}

// ---- halo sum over peer memory (NVLink loads / stores between the GPUs of one node, no NCCL kernel in the step).
// put: copy our shared range into the neighbour's mailbox slot; the last CTA to finish publishes the message number
// in the slot's flag (system-scope release).  take: wait for the neighbour's message number in OUR mailbox, add the
// slot to the vector.  A put never waits, so the exchange cannot deadlock; two slots suffice because the neighbour's
// put q + 1, which our take q + 1 waits for, comes after its take q in its stream.
constexpr int MAIL_FLAG_BYTES = 256;
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__global__ void __launch_bounds__(256) k_halo_put(int64_t n, const double *__restrict__ vec, unsigned char *peer_mail,
                                                  int64_t slot_doubles, unsigned long long q, unsigned int *done) {
  double *dst = reinterpret_cast<double *>(peer_mail + MAIL_FLAG_BYTES) + (q & 1ull) * slot_doubles;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = vec[i];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {
      *done = 0;
      __threadfence_system();
      st_release_sys(reinterpret_cast<unsigned long long *>(peer_mail + 128 * (q & 1ull)), q);
    }
  }
}
__global__ void __launch_bounds__(256) k_halo_take(int64_t n, double *vec, const unsigned char *mail, int64_t slot_doubles,
                                                   unsigned long long q) {
  if (threadIdx.x == 0) {
    const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(mail + 128 * (q & 1ull));
    while (ld_acquire_sys(flag) < q) __nanosleep(100);
  }
  __syncthreads();
  const double *src = reinterpret_cast<const double *>(mail + MAIL_FLAG_BYTES) + (q & 1ull) * slot_doubles;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    vec[i] += __ldcg(src + i);
}

void dist_halo_ranges(const tlfea_newton *s, std::vector<std::pair<int64_t, int64_t>> &out) {
  out.clear();
  const auto *dc = static_cast<const DistCtx *>(s->dist);
  if (!dc) return;
  for (auto &nb : dc->nbr)
    if (nb.c_hi > nb.c_lo) out.push_back({nb.c_lo, nb.c_hi});
}

void dist_release(tlfea_newton *s) {
  auto *dc = static_cast<DistCtx *>(s->dist);
  if (!dc) return;
  for (auto &nb : dc->nbr) {
    dfree(nb.d_recv);
    if (nb.d_peer) cudaIpcCloseMemHandle(nb.d_peer);
    dfree(nb.d_mail);
    dfree(nb.d_done);
  }
  NcclApi *api = nccl_api();
  if (api && dc->comm) api->CommDestroy(dc->comm);
  delete dc;
  s->dist = nullptr;
}

}  // namespace tlfea

using namespace tlfea;

extern "C" {

int tlfea_dist_unique_id(void *out128) {
  TL_CHECK(out128, "null out");
  NcclApi *api = nccl_api();
  if (!api) return fail(TLFEA_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
  ncclUniqueId id;
  TL_NCCL(api, api->GetUniqueId(&id));
  static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
  memcpy(out128, &id, sizeof(id));
  return TLFEA_OK;
}

int tlfea_dist_init(tlfea_newton *s, const void *id128, int rank, int world) {
  TL_CHECK(s && id128 && world >= 1 && rank >= 0 && rank < world, "bad input");
  TL_CHECK(!s->dist, "the solver already has a communicator");
  NcclApi *api = nccl_api();
  if (!api) return fail(TLFEA_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
  TL_CUDA(cudaSetDevice(s->e->device));
  auto *dc = new DistCtx();
  dc->rank = rank;
  dc->world = world;
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclResult_t r = api->CommInitRank(&dc->comm, world, id, rank);
  if (r != ncclSuccess) {
    delete dc;
    return fail(TLFEA_ERR_CUDA, "ncclCommInitRank -> %s", api->GetErrorString(r));
  }
  s->dist = dc;
  return TLFEA_OK;
}

int tlfea_dist_set_halo(tlfea_newton *s, int n_nbr, const int *nbr_rank, const int64_t *coef_lo, const int64_t *coef_hi) {
  TL_CHECK(s && s->dist, "tlfea_dist_init must run first");
  TL_CHECK(n_nbr >= 0 && (n_nbr == 0 || (nbr_rank && coef_lo && coef_hi)), "bad input");
  auto *dc = static_cast<DistCtx *>(s->dist);
  TL_CUDA(cudaSetDevice(s->e->device));
  for (auto &nb : dc->nbr) {
    dfree(nb.d_recv);
    if (nb.d_peer) cudaIpcCloseMemHandle(nb.d_peer);
    dfree(nb.d_mail);
    dfree(nb.d_done);
  }
  dc->nbr.clear();
  dc->halo_bytes = 0;
  dc->p2p = false;
  dc->seq = 0;
  for (int k = 0; k < n_nbr; k++) {
    TL_CHECK(nbr_rank[k] >= 0 && nbr_rank[k] < dc->world && nbr_rank[k] != dc->rank, "neighbour rank out of range");
    TL_CHECK(coef_lo[k] >= 0 && coef_lo[k] <= coef_hi[k] && coef_hi[k] <= s->e->n_coef, "shared range out of bounds");
    DistCtx::Nbr nb{nbr_rank[k], coef_lo[k], coef_hi[k], nullptr, nullptr, nullptr, nullptr};
    TL_TRY(dmalloc(&nb.d_recv, (size_t)(nb.c_hi - nb.c_lo) * HALO_MAX_WIDTH));
    dc->halo_bytes += (nb.c_hi - nb.c_lo) * 3 * (int64_t)sizeof(double);
    dc->nbr.push_back(nb);
  }
  return TLFEA_OK;
}

// d_vec holds `width` doubles per coefficient (3: nodal vectors, 9: diagonal blocks); shared entries become the sum
// over the two ranks that hold them
int tlfea_dist_halo_sum(tlfea_newton *s, double *d_vec, int width) {
  TL_CHECK(s && s->dist && d_vec, "tlfea_dist_init must run first");
  TL_CHECK(width >= 1 && width <= HALO_MAX_WIDTH, "width must be 1..9");
  auto *dc = static_cast<DistCtx *>(s->dist);
  if (dc->nbr.empty()) return TLFEA_OK;
  NcclApi *api = nccl_api();
  TL_CUDA(cudaSetDevice(s->e->device));
  cudaStream_t st = s->e->stream;
  if (dc->p2p) {  // peer-memory path: puts first (they never wait), then the takes in neighbour order
    const unsigned long long q = ++dc->seq;
    for (auto &nb : dc->nbr) {
      const int64_t n = (nb.c_hi - nb.c_lo) * width, slot = (nb.c_hi - nb.c_lo) * HALO_MAX_WIDTH;
      const unsigned grid = (unsigned)std::min<int64_t>(std::max<int64_t>((n + 1023) / 1024, 1), 64);
      TL_LAUNCH(k_halo_put, grid, 256, 0, st, n, d_vec + nb.c_lo * width, nb.d_peer, slot, q, nb.d_done);
    }
    for (auto &nb : dc->nbr) {
      const int64_t n = (nb.c_hi - nb.c_lo) * width, slot = (nb.c_hi - nb.c_lo) * HALO_MAX_WIDTH;
      const unsigned grid = (unsigned)std::min<int64_t>(std::max<int64_t>((n + 1023) / 1024, 1), 64);
      TL_LAUNCH(k_halo_take, grid, 256, 0, st, n, d_vec + nb.c_lo * width, nb.d_mail, slot, q);
    }
    TL_CUDA(cudaGetLastError());
    return TLFEA_OK;
  }
  TL_NCCL(api, api->GroupStart());
  for (auto &nb : dc->nbr) {
    const size_t n = (size_t)(nb.c_hi - nb.c_lo) * width;
    if (n == 0) continue;
    TL_NCCL(api, api->Send(d_vec + nb.c_lo * width, n, ncclDouble, nb.rank, dc->comm, st));
    TL_NCCL(api, api->Recv(nb.d_recv, n, ncclDouble, nb.rank, dc->comm, st));
  }
  TL_NCCL(api, api->GroupEnd());
  for (auto &nb : dc->nbr) {  // fixed neighbour order
    const int64_t n = (nb.c_hi - nb.c_lo) * width;
    if (n > 0) TL_LAUNCH(k_halo_add, blocks_for(n, 256), 256, 0, st, n, nb.d_recv, d_vec + nb.c_lo * width);
  }
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

int tlfea_dist_allreduce_sum(tlfea_newton *s, double *d_buf, int n) {
  TL_CHECK(s && s->dist && d_buf && n > 0, "tlfea_dist_init must run first");
  auto *dc = static_cast<DistCtx *>(s->dist);
  if (dc->world == 1) return TLFEA_OK;
  NcclApi *api = nccl_api();
  TL_CUDA(cudaSetDevice(s->e->device));
  TL_NCCL(api, api->AllReduce(d_buf, d_buf, (size_t)n, ncclDouble, ncclSum, dc->comm, s->e->stream));
  return TLFEA_OK;
}

int tlfea_dist_p2p_export(tlfea_newton *s, int k, void *out64) {
  TL_CHECK(s && s->dist && out64, "tlfea_dist_set_halo must run first");
  auto *dc = static_cast<DistCtx *>(s->dist);
  TL_CHECK(k >= 0 && k < (int)dc->nbr.size(), "neighbour index out of range");
  TL_CUDA(cudaSetDevice(s->e->device));
  auto &nb = dc->nbr[(size_t)k];
  if (!nb.d_mail) {
    const size_t bytes = MAIL_FLAG_BYTES + 2 * (size_t)(nb.c_hi - nb.c_lo) * HALO_MAX_WIDTH * sizeof(double);
    TL_TRY(dmalloc(&nb.d_mail, bytes));
    TL_TRY(dmalloc(&nb.d_done, 1));
    TL_CUDA(cudaMemset(nb.d_mail, 0, bytes));
    TL_CUDA(cudaMemset(nb.d_done, 0, sizeof(unsigned int)));
    TL_CUDA(cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t h;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  TL_CUDA(cudaIpcGetMemHandle(&h, nb.d_mail));
  memcpy(out64, &h, sizeof(h));
  return TLFEA_OK;
}

int tlfea_dist_p2p_connect(tlfea_newton *s, int k, const void *handle64) {
  TL_CHECK(s && s->dist && handle64, "tlfea_dist_set_halo must run first");
  auto *dc = static_cast<DistCtx *>(s->dist);
  TL_CHECK(k >= 0 && k < (int)dc->nbr.size(), "neighbour index out of range");
  TL_CUDA(cudaSetDevice(s->e->device));
  auto &nb = dc->nbr[(size_t)k];
  TL_CHECK(nb.d_mail, "tlfea_dist_p2p_export must run first for this neighbour");
  if (nb.d_peer) {
    cudaIpcCloseMemHandle(nb.d_peer);
    nb.d_peer = nullptr;
  }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void *ptr = nullptr;
  TL_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
  nb.d_peer = static_cast<unsigned char *>(ptr);
  bool all = true;
  for (auto &x : dc->nbr) all = all && x.d_peer && x.d_mail;
  dc->p2p = all;
  return TLFEA_OK;
}

int64_t tlfea_dist_halo_bytes(tlfea_newton *s) {
  return (s && s->dist) ? static_cast<DistCtx *>(s->dist)->halo_bytes : 0;
}

int tlfea_dist_finalize(tlfea_newton *s) {
  TL_CHECK(s, "null handle");
  if (s->dist) {
    cudaSetDevice(s->e->device);
    cudaStreamSynchronize(s->e->stream);
    dist_release(s);
  }
  return TLFEA_OK;
}

}  // extern "C"
