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
  };
  std::vector<Nbr> nbr;
  int64_t halo_bytes = 0;  // bytes sent per halo sum of width 3 (reporting)
};

static constexpr int HALO_MAX_WIDTH = 9;

__global__ void k_halo_add(int64_t n, const double *__restrict__ buf, double *vec) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) vec[i] += buf[i];
}

void dist_release(tlfea_newton *s) {
  auto *dc = static_cast<DistCtx *>(s->dist);
  if (!dc) return;
  for (auto &nb : dc->nbr) dfree(nb.d_recv);
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
  for (auto &nb : dc->nbr) dfree(nb.d_recv);
  dc->nbr.clear();
  dc->halo_bytes = 0;
  for (int k = 0; k < n_nbr; k++) {
    TL_CHECK(nbr_rank[k] >= 0 && nbr_rank[k] < dc->world && nbr_rank[k] != dc->rank, "neighbour rank out of range");
    TL_CHECK(coef_lo[k] >= 0 && coef_lo[k] <= coef_hi[k] && coef_hi[k] <= s->e->n_coef, "shared range out of bounds");
    DistCtx::Nbr nb{nbr_rank[k], coef_lo[k], coef_hi[k], nullptr};
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
