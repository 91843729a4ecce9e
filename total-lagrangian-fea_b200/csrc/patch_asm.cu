// patch_asm.cu -- host side of the single-pass FEAT10 assembly (pk_kernel.cuh): builds the patch plan once per
// sparsity pattern, keeps it on the device, and launches one kernel per colour.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "pk_kernel.cuh"

namespace tlfea {

struct PatchPlanDev {
  int W = 0;
  int n_patches = 0, n_colors = 0, n_chunks = 1;
  std::vector<int> color_ptr, range_ready;
  PatchDesc *d_desc = nullptr;
  PElem *d_elems = nullptr;
  unsigned char *d_blob = nullptr;
  int64_t n_unc = 0;
  int64_t *d_unc = nullptr;
  int *d_unc_row = nullptr;
  int max_blob = 0;
  int mass_version_seen = -1;
  int aff_version_seen = -1;
  int64_t n_elems = 0;
  int64_t sum_local_blocks = 0, blob_bytes = 0;
};

void patch_release(tlfea_newton *s) {
  auto *pp = static_cast<PatchPlanDev *>(s->patch);
  if (!pp) return;
  dfree(pp->d_desc);
  dfree(pp->d_elems);
  dfree(pp->d_blob);
  dfree(pp->d_unc);
  dfree(pp->d_unc_row);
  delete pp;
  s->patch = nullptr;
}

static int patch_chunk_count() {
  static int k = [] {
    const char *e = getenv("TLFEA_PK_CHUNKS");  // id-range chunks of the launch order (host-buffer pipelining)
    if (e) return std::min(std::max(atoi(e), 1), 16);  // exact
    return -4;                                         // at most 4, fewer for small meshes (patch_plan.cu)
  }();
  return k;
}

int patch_warps() {
  static int w = [] {
    const char *e = getenv("TLFEA_PK_W");  // warps (= sub-patches) per patch: 4 (two CTAs per SM) or 8
    const int v = e ? atoi(e) : 4;
    return (v == 8) ? 8 : 4;
  }();
  return w;
}

// Builds the plan on the host from the element's connectivity, reference coordinates and pattern.
int patch_prepare(tlfea_newton *s) {
  tlfea_elem *e = s->e;
  if (s->patch) return TLFEA_OK;
  TL_CHECK(e->type == TLFEA_TYPE_T10, "the patch assembly is the FEAT10 path");
  const int W = (e->model == TL_MODEL_MR) ? 4 : (s->pk_w ? s->pk_w : patch_warps());
  TL_CUDA(cudaStreamSynchronize(e->stream));
  std::vector<int> ecoef((size_t)e->n_elem * 10), off((size_t)e->n_coef + 1), cols((size_t)e->pat.nnz);
  std::vector<double> x((size_t)e->n_coef), y((size_t)e->n_coef), z((size_t)e->n_coef);
  TL_CUDA(cudaMemcpy(ecoef.data(), e->d_ecoef, ecoef.size() * sizeof(int), cudaMemcpyDeviceToHost));
  TL_CUDA(cudaMemcpy(off.data(), e->pat.d_off, off.size() * sizeof(int), cudaMemcpyDeviceToHost));
  TL_CUDA(cudaMemcpy(cols.data(), e->pat.d_cols, cols.size() * sizeof(int), cudaMemcpyDeviceToHost));
  TL_CUDA(cudaMemcpy(x.data(), e->d_xr, x.size() * sizeof(double), cudaMemcpyDeviceToHost));
  TL_CUDA(cudaMemcpy(y.data(), e->d_yr, y.size() * sizeof(double), cudaMemcpyDeviceToHost));
  TL_CUDA(cudaMemcpy(z.data(), e->d_zr, z.size() * sizeof(double), cudaMemcpyDeviceToHost));
  PatchPlanHost hp;
  std::string err;
  const int rc = build_patch_plan(e->n_elem, e->n_coef, ecoef.data(), x.data(), y.data(), z.data(), off.data(),
                                  cols.data(), W, s->pk_chunks ? s->pk_chunks : patch_chunk_count(), &hp, &err);
  if (rc != TLFEA_OK) return fail(rc, "%s", err.c_str());
  auto *pp = new PatchPlanDev();
  s->patch = pp;
  pp->W = W;
  pp->n_patches = hp.n_patches;
  pp->n_colors = hp.n_colors;
  pp->color_ptr = hp.color_ptr;
  pp->n_chunks = hp.n_chunks;
  pp->range_ready = hp.range_ready;
  pp->max_blob = hp.max_blob_bytes;
  pp->sum_local_blocks = hp.sum_local_blocks;
  pp->blob_bytes = (int64_t)hp.blob.size();
  pp->n_elems = (int64_t)hp.elems.size();
  TL_TRY(dmalloc(&pp->d_desc, hp.desc.size()));
  TL_TRY(dmalloc(&pp->d_elems, hp.elems.size()));
  TL_TRY(dmalloc(&pp->d_blob, hp.blob.size()));
  // The uploads go on the handle's stream: it is a non-blocking stream, so a plain cudaMemcpy (legacy stream, and
  // from pageable memory it may return while the DMA is still in flight) is NOT ordered against the kernels that
  // follow -- pk_fill_mass_kernel could write the mass fields first and the tail of the copy overwrite them.
  TL_CUDA(cudaMemcpyAsync(pp->d_desc, hp.desc.data(), hp.desc.size() * sizeof(PatchDesc), cudaMemcpyHostToDevice, e->stream));
  TL_CUDA(cudaMemcpyAsync(pp->d_elems, hp.elems.data(), hp.elems.size() * sizeof(PElem), cudaMemcpyHostToDevice, e->stream));
  TL_CUDA(cudaMemcpyAsync(pp->d_blob, hp.blob.data(), hp.blob.size(), cudaMemcpyHostToDevice, e->stream));
  pp->n_unc = (int64_t)hp.uncovered.size();
  if (pp->n_unc > 0) {
    std::vector<int> row(hp.uncovered.size());
    for (size_t k = 0; k < row.size(); k++)
      row[k] = (int)(std::upper_bound(off.begin(), off.end(), (int)hp.uncovered[k]) - off.begin()) - 1;
    TL_TRY(dmalloc(&pp->d_unc, hp.uncovered.size()));
    TL_TRY(dmalloc(&pp->d_unc_row, row.size()));
    TL_CUDA(cudaMemcpyAsync(pp->d_unc, hp.uncovered.data(), hp.uncovered.size() * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
    TL_CUDA(cudaMemcpyAsync(pp->d_unc_row, row.data(), row.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    TL_CUDA(cudaStreamSynchronize(e->stream));  // row is a local
  }
  TL_CUDA(cudaStreamSynchronize(e->stream));  // the host plan goes out of scope
  return TLFEA_OK;
}

// the plan for the configuration that will run (patch width follows the material model)
static int ensure_plan(tlfea_newton *s) {
  TL_TRY(patch_prepare(s));
  auto *pp = static_cast<PatchPlanDev *>(s->patch);
  const int want_w = (s->e->model == TL_MODEL_MR) ? 4 : (s->pk_w ? s->pk_w : patch_warps());
  if (want_w != pp->W) {  // the material model changed under a live plan
    patch_release(s);
    TL_TRY(patch_prepare(s));
  }
  return TLFEA_OK;
}

template <int W, bool DAMP, bool TANGENT, int MODEL>
static int launch_pk_t(tlfea_elem *e, const PatchPlanDev *pp, PKArgs a, int p0, int np, bool pdl) {
  using Sh = PKShape<DAMP, MODEL>;
  constexpr size_t smem = Sh::smem_bytes(W);
  static_assert(smem <= 227 * 1024, "patch kernel configuration exceeds the shared memory of one SM");
  static thread_local int attr_dev = -1;
  if (attr_dev != e->device) {
    TL_CUDA(cudaFuncSetAttribute(pk_kernel<W, DAMP, TANGENT, MODEL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    attr_dev = e->device;
  }
  if (np <= 0) return TLFEA_OK;
  a.desc = pp->d_desc + p0;
  a.n_patches = np;
  // persistent CTAs, as many as fit the SMs at once: each walks its patches with the prefetch stream running
  const int per_sm = std::max(1, std::min((int)((228 * 1024) / (smem + 1024)), 2048 / (W * 32)));
  const int grid = std::min(np, per_sm * sm_count(e->device));
  // Programmatic dependent launch: the CTAs of this launch take the SMs as those of the previous colour drain and
  // compute their first patch (which reads nothing an earlier colour writes); they wait for the previous launch
  // (griddepcontrol.wait in the kernel) only before their first access to H.  Hides the tail of every launch.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(W * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = e->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  TL_CUDA(cudaLaunchKernelEx(&cfg, pk_kernel<W, DAMP, TANGENT, MODEL>, a));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return TLFEA_OK;
}

template <bool DAMP, bool TANGENT>
static int launch_pk(tlfea_elem *e, const PatchPlanDev *pp, const PKArgs &a, int p0, int np, bool pdl) {
  if (e->model == TL_MODEL_MR) return launch_pk_t<4, DAMP, TANGENT, TL_MODEL_MR>(e, pp, a, p0, np, pdl);
  if (pp->W == 8) return launch_pk_t<8, DAMP, TANGENT, TL_MODEL_SVK>(e, pp, a, p0, np, pdl);
  return launch_pk_t<4, DAMP, TANGENT, TL_MODEL_SVK>(e, pp, a, p0, np, pdl);
}

// Element forces fe (always) and, with tangent, H = M/h + h K_t (+ C_visc) in place, colour by colour.
int patch_assemble(tlfea_newton *s, bool tangent, bool damp, const PatchHooks *hooks) {
  tlfea_elem *e = s->e;
  TL_TRY(ensure_plan(s));
  auto *pp = static_cast<PatchPlanDev *>(s->patch);
  const double h = s->prm.time_step;
  if (tangent && pp->mass_version_seen != e->mass_version) {
    TL_LAUNCH(pk_fill_mass_kernel, pp->n_patches, 128, 0, e->stream, pp->n_patches, e->n_coef, pp->d_desc, pp->d_blob,
              e->pat.d_off, e->d_mass);
    pp->mass_version_seen = e->mass_version;
  }
  // Affine-element compression is opt-in: it removes 1.16 KB of the 1.28 KB an element reads from HBM (10.1 M tets:
  // 105 -> 94 GB per assemble) but the in-kernel expansion of the gradients costs more than the TMA copy it replaces
  // while the kernel is issue / latency bound (43.7 ms against 38.5 ms), see DESIGN.md.
  static const bool env_aff = [] {
    const char *v = getenv("TLFEA_PK_AFFINE");
    return v && atoi(v) != 0;
  }();
  const bool use_aff = s->pk_affine < 0 ? env_aff : (s->pk_affine != 0);
  if (use_aff && e->d_aff && pp->aff_version_seen != e->aff_version) {  // copy the affine flags into the plan records
    TL_LAUNCH(pk_set_affine_kernel, blocks_for(pp->n_elems, 256), 256, 0, e->stream, pp->n_elems, pp->d_elems,
              e->d_aff_flag);
    pp->aff_version_seen = e->aff_version;
  }
  PKArgs a;
  a.aff = (use_aff && e->d_aff) ? e->d_aff : nullptr;
  for (int i = 0; i < 150; i++) a.dn[i] = e->t10_dn[i];
  a.desc = pp->d_desc;
  a.elems = pp->d_elems;
  a.blob = pp->d_blob;
  a.x = e->d_x;
  a.y = e->d_y;
  a.z = e->d_z;
  a.v = damp ? s->d_v : nullptr;
  a.gradN = e->d_gradN;
  a.detJ = e->d_detJ;
  a.qw = e->qw;
  a.lambda = e->lambda;
  a.mu = e->mu;
  a.eta = e->eta;
  a.lambda_d = e->lambda_d;
  a.c1 = h * e->lambda + e->lambda_d;
  a.c2 = h * e->mu + e->eta;
  a.c3 = h * e->mu + e->eta;
  a.c4 = h;
  a.mu10 = e->mu10;
  a.mu01 = e->mu01;
  a.kappa = e->kappa;
  a.inv_h = tangent ? 1.0 / h : 0.0;
  {
    static const int dbg = [] { const char *v = getenv("TLFEA_PK_DEBUG"); return v ? atoi(v) : 0; }();
    a.debug = dbg;
  }
  a.fe = e->d_fe;
  a.H = s->d_H;
  if (!tangent) {  // no shared output: every patch in one launch
    TL_TRY((damp ? launch_pk<true, false>(e, pp, a, 0, pp->n_patches, false) : launch_pk<false, false>(e, pp, a, 0, pp->n_patches, false)));
    TL_CUDA(cudaGetLastError());
    return TLFEA_OK;
  }
  static const bool use_pdl = [] { const char *v = getenv("TLFEA_PK_PDL"); return !v || atoi(v) != 0; }();
  if (pp->n_unc > 0)
    TL_LAUNCH(pk_uncovered_kernel, blocks_for(pp->n_unc, 256), 256, 0, e->stream, pp->n_unc, pp->d_unc, e->pat.d_off,
              pp->d_unc_row, e->d_mass, 1.0 / h, s->d_H);
  for (int ch = 0; ch < pp->n_chunks; ch++) {
    if (hooks && hooks->before_chunk) TL_TRY(hooks->before_chunk(hooks->ctx, ch));
    for (int c = 0; c < pp->n_colors; c++) {
      const int l = ch * pp->n_colors + c;
      const int p0 = pp->color_ptr[l], np = pp->color_ptr[l + 1] - p0;
      // the first launch of a chunk follows other work in the stream (uploads, event waits): no early start for it
      const bool pdl = use_pdl && (c > 0 || (ch > 0 && !hooks));
      TL_TRY((damp ? launch_pk<true, true>(e, pp, a, p0, np, pdl) : launch_pk<false, true>(e, pp, a, p0, np, pdl)));
    }
    if (hooks && hooks->after_chunk) TL_TRY(hooks->after_chunk(hooks->ctx, ch));
  }
  TL_CUDA(cudaGetLastError());
  return TLFEA_OK;
}

int patch_chunks(tlfea_newton *s, int *n_chunks, const int **range_ready) {
  TL_TRY(ensure_plan(s));
  auto *pp = static_cast<PatchPlanDev *>(s->patch);
  *n_chunks = pp->n_chunks;
  *range_ready = pp->range_ready.data();
  return TLFEA_OK;
}

int patch_stats(tlfea_newton *s, int64_t *out) {
  auto *pp = static_cast<PatchPlanDev *>(s->patch);
  if (!pp) return fail(TLFEA_ERR_INVALID, "no patch plan (FEAT10 single-pass assembly has not run)");
  out[0] = pp->n_patches;
  out[1] = pp->n_colors;
  out[2] = pp->W;
  out[3] = pp->sum_local_blocks;
  out[4] = pp->blob_bytes;
  out[5] = pp->n_unc;
  out[6] = pp->n_chunks;
  return TLFEA_OK;
}

}  // namespace tlfea
