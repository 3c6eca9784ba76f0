// abi.cu -- extern "C" entry points declared in include/gsr_b200.h (orchestration only).
//
// Replaces the host side of CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (DGR/cuda_rasterizer/rasterizer_impl.cu:141-153,198-433) and the pybind glue
// (DGR/rasterize_points.cu:35-217).  No allocation, no torch: the caller owns every buffer.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <nvtx3/nvToolsExt.h>
#include "common.cuh"

namespace gsr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- NVTX ranges named after the stages (for ncu --nvtx / Nsight Systems), enabled with GSR_NVTX=1 ----
static int nvtx_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("GSR_NVTX"); v = (e && e[0] == '1') ? 1 : 0; }
  return v;
}
void nvtx_push(int stage) {
  static const char* names[ST_COUNT] = {"gsr:preprocess_fwd", "gsr:duplicate_keys", "gsr:sort", "gsr:tile_ranges",
                                        "gsr:blend_fwd", "gsr:blend_bwd", "gsr:preprocess_bwd"};
  if (nvtx_enabled() && stage >= 0 && stage < ST_COUNT) nvtxRangePushA(names[stage]);
}
void nvtx_pop() { if (nvtx_enabled()) nvtxRangePop(); }

// ---- profiling state: process-wide, serialised by g_prof_mu (calls may come from several host threads) ----------
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static long long g_launches = 0;
struct ProfRec { int stage; cudaEvent_t a, b; };
static ProfRec g_prof[8192];
static int g_prof_n = 0, g_prof_made = 0;

void count_launches(int n) { __atomic_fetch_add(&g_launches, (long long)n, __ATOMIC_RELAXED); }
bool profile_begin(int stage, cudaStream_t st) {
  if (!g_prof_on) return false;
  g_prof_mu.lock();  // held until profile_end: one stage is recorded at a time
  if (g_prof_n >= 8192) return true;
  ProfRec& r = g_prof[g_prof_n];
  if (g_prof_n >= g_prof_made) { cudaEventCreate(&r.a); cudaEventCreate(&r.b); g_prof_made++; }
  r.stage = stage;
  cudaEventRecord(r.a, st);
  return true;
}
void profile_end(bool recording, cudaStream_t st) {
  if (!recording) return;
  if (g_prof_n < 8192) {
    cudaEventRecord(g_prof[g_prof_n].b, st);
    g_prof_n++;
  }
  g_prof_mu.unlock();
}

static inline bool misaligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) != 0; }

static int check_inputs(const gsr_inputs* in) {
  if (!in) { set_error("inputs is NULL"); return GSR_ERR_INVALID; }
  if (in->P < 0 || in->W <= 0 || in->H <= 0) { set_error("bad sizes P=%d W=%d H=%d", in->P, in->W, in->H); return GSR_ERR_INVALID; }
  if (in->W > 65535 * TILE_X || in->H > 65535 * TILE_Y) { set_error("image too large"); return GSR_ERR_INVALID; }
  if (!in->means3D || !in->opacities || !in->bg || !in->viewmatrix || !in->projmatrix || !in->campos) {
    set_error("a required input pointer is NULL"); return GSR_ERR_INVALID;
  }
  if ((in->shs == nullptr) == (in->colors_precomp == nullptr)) {
    set_error("Please provide excatly one of either SHs or precomputed colors!"); return GSR_ERR_INVALID;
  }
  const bool sr = in->scales != nullptr && in->rotations != nullptr;
  if (sr == (in->cov3D_precomp != nullptr) || ((in->scales != nullptr) != (in->rotations != nullptr))) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return GSR_ERR_INVALID;
  }
  if (in->shs && (in->M < 1 || in->D < 0 || in->D > 3 || (in->D + 1) * (in->D + 1) > in->M)) {
    set_error("SH degree %d needs %d coefficients, M=%d", in->D, (in->D + 1) * (in->D + 1), in->M); return GSR_ERR_INVALID;
  }
  if (in->shs_rest && (!in->shs || in->M < 2)) { set_error("shs_rest needs shs (the DC block) and M >= 2"); return GSR_ERR_INVALID; }
  if (in->raw_params != 0 && in->raw_params != 1) { set_error("raw_params must be 0 or 1"); return GSR_ERR_INVALID; }
  if (in->K != 0 && (in->K != MAX_OBJ || !in->obj_feats)) { set_error("object channels: K must be 0 or 16 with obj_feats"); return GSR_ERR_INVALID; }
  // arrays the kernels move with 16-byte accesses (include/gsr_b200.h: "device pointers are 16-byte aligned")
  if (misaligned16(in->shs) || misaligned16(in->rotations) || misaligned16(in->obj_feats)) {
    set_error("shs, rotations and obj_feats must be 16-byte aligned"); return GSR_ERR_INVALID;
  }
  return GSR_OK;
}

static int debug_sync(const gsr_inputs* in, cudaStream_t st, const char* stage) {
  if (!in->debug) return GSR_OK;
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { set_error("[CUDA ERROR] after %s: %s", stage, cudaGetErrorString(e)); return GSR_ERR_CUDA; }
  return GSR_OK;
}
#define GSR_TRY(x) do { int _r = (x); if (_r != GSR_OK) return _r; } while (0)

}  // namespace gsr

using namespace gsr;

extern "C" {

int gsr_abi_version(void) { return GSR_ABI_VERSION; }
const char* gsr_last_error(void) { return g_err; }
uint32_t gsr_higher_msb(uint32_t n) { return higher_msb(n); }

int gsr_profile_enable(int on) { g_prof_on = on != 0; g_prof_n = 0; return GSR_OK; }
int64_t gsr_launch_count(int reset) { long long v = g_launches; if (reset) g_launches = 0; return v; }
int gsr_profile_read(float* ms_host, int32_t* calls_host) {
  for (int i = 0; i < ST_COUNT; i++) { ms_host[i] = 0.f; if (calls_host) calls_host[i] = 0; }
  for (int i = 0; i < g_prof_n; i++) {
    float ms = 0.f;
    GSR_CUDA_CHECK(cudaEventSynchronize(g_prof[i].b));
    GSR_CUDA_CHECK(cudaEventElapsedTime(&ms, g_prof[i].a, g_prof[i].b));
    ms_host[g_prof[i].stage] += ms;
    if (calls_host) calls_host[g_prof[i].stage]++;
  }
  g_prof_n = 0;
  return GSR_OK;
}

int gsr_debug_counters(uint64_t* out_host, int reset) {
  if (!out_host) { set_error("debug_counters: NULL"); return GSR_ERR_INVALID; }
  GSR_TRY(blend_fwd_stats(out_host, reset));
  return blend_bwd_stats(out_host + 8, reset);
}

int gsr_debug_sort_profile(uint64_t* out_host, int reset) {
  if (!out_host) { set_error("debug_sort_profile: NULL"); return GSR_ERR_INVALID; }
  return sort_profile_read(out_host, reset);
}

size_t gsr_geom_bytes(int32_t P) { return P <= 0 ? 0 : Geom(nullptr, P).bytes; }
size_t gsr_image_bytes(int32_t W, int32_t H) { return Image(nullptr, W, H).bytes; }
size_t gsr_scratch_bytes(int32_t P, int64_t L_cap, int32_t, int32_t) { return Scratch(nullptr, P, L_cap).bytes; }
size_t gsr_sort_temp_bytes(int64_t n) { return sort_temp_bytes(n, 64, 8) + align_up((size_t)n * 8) + align_up((size_t)n * 4); }

int gsr_forward_geometry(const gsr_inputs* in, void* geom, int32_t* radii, void* scratch, size_t scratch_bytes,
                         uint32_t* num_rendered_dev, uint32_t* num_rendered_host, void* stream) {
  GSR_TRY(check_inputs(in));
  cudaStream_t st = (cudaStream_t)stream;
  if (!geom || !radii || !scratch || !num_rendered_dev) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (scratch_bytes < Scratch(nullptr, in->P, 0).bytes) { set_error("scratch too small"); return GSR_ERR_CAPACITY; }
  if (in->P == 0) {
    GSR_CUDA_CHECK(cudaMemsetAsync(num_rendered_dev, 0, 4, st));
    if (num_rendered_host) *num_rendered_host = 0;
    return GSR_OK;
  }
  Geom g(geom, in->P);
  Scratch s(scratch, in->P, 0);
  { StageScope ps(ST_PREPROCESS_FWD, st); GSR_TRY(launch_preprocess_fwd(*in, g, radii, s, num_rendered_dev, st)); }
  GSR_TRY(debug_sync(in, st, "preprocess"));
  // order the Gaussians by (depth bits, index): 4 passes, the result lands back in (dkeys_a, geom.perm)
  { StageScope ps(ST_SORT, st);
    GSR_TRY(launch_sort_pairs_pingpong_u32(s.dkeys_a, g.perm, s.dkeys_b, s.dvals_b, in->P, nullptr,
                                           make_sort_plan(in->P, 32, 4), s.sort_temp, false, true, st)); }
  GSR_TRY(debug_sync(in, st, "depth_sort"));
  { StageScope ps(ST_DUPLICATE, st); GSR_TRY(launch_sorted_scan(in->P, g, s, num_rendered_dev, num_rendered_host, st)); }
  GSR_TRY(debug_sync(in, st, "sorted_scan"));
  if (in->prefiltered) {
    uint32_t flag = 0;
    GSR_CUDA_CHECK(cudaMemcpyAsync(&flag, s.flags, 4, cudaMemcpyDeviceToHost, st));
    GSR_CUDA_CHECK(cudaStreamSynchronize(st));
    if (flag) { set_error("Point is filtered although prefiltered is set. This shouldn't happen!"); return GSR_ERR_PREFILTERED; }
  }
  return GSR_OK;
}

int gsr_forward_preprocess_views(const gsr_inputs* in, int32_t n_views, const gsr_view_fwd* views, void* stream) {
  if (!in) { set_error("inputs is NULL"); return GSR_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  if (in->P < 0) { set_error("bad size P=%d", in->P); return GSR_ERR_INVALID; }
  if (in->P == 0 || n_views == 0) return GSR_OK;
  if (n_views < 0 || n_views > GSR_MAX_FUSED_VIEWS || !views) {
    set_error("forward_preprocess_views: 1..%d views per call, got %d", GSR_MAX_FUSED_VIEWS, n_views); return GSR_ERR_INVALID;
  }
  if (in->raw_params || in->shs_rest || in->K > 0 || in->prefiltered) {
    set_error("forward_preprocess_views: raw_params / shs_rest / object channels / prefiltered are not supported");
    return GSR_ERR_INVALID;
  }
  if (!in->means3D || !in->opacities || (in->shs == nullptr) == (in->colors_precomp == nullptr)) {
    set_error("Please provide excatly one of either SHs or precomputed colors!"); return GSR_ERR_INVALID;
  }
  const bool sr = in->scales != nullptr && in->rotations != nullptr;
  if (sr == (in->cov3D_precomp != nullptr) || ((in->scales != nullptr) != (in->rotations != nullptr))) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return GSR_ERR_INVALID;
  }
  if (in->shs && (in->M < 1 || in->D < 0 || in->D > 3 || (in->D + 1) * (in->D + 1) > in->M)) {
    set_error("SH degree %d needs %d coefficients, M=%d", in->D, (in->D + 1) * (in->D + 1), in->M); return GSR_ERR_INVALID;
  }
  if (misaligned16(in->shs) || misaligned16(in->rotations)) { set_error("shs and rotations must be 16-byte aligned"); return GSR_ERR_INVALID; }
  for (int v = 0; v < n_views; v++) {
    const gsr_view_fwd& r = views[v];
    if (!r.viewmatrix || !r.projmatrix || !r.campos || !r.geom || !r.radii || r.W <= 0 || r.H <= 0 ||
        r.W > 65535 * TILE_X || r.H > 65535 * TILE_Y || misaligned16(r.geom)) {
      set_error("forward_preprocess_views: view %d has a NULL / misaligned pointer or a bad size", v); return GSR_ERR_INVALID;
    }
  }
  { StageScope ps(ST_PREPROCESS_FWD, st); GSR_TRY(launch_preprocess_fwd_views(*in, n_views, views, st)); }
  return debug_sync(in, st, "preprocess_views");
}

int gsr_forward_order(const gsr_inputs* in, void* geom, void* scratch, size_t scratch_bytes,
                      uint32_t* num_rendered_dev, uint32_t* num_rendered_host, void* stream) {
  if (!in) { set_error("inputs is NULL"); return GSR_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  if (!geom || !scratch || !num_rendered_dev) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (in->P < 0) { set_error("bad size P=%d", in->P); return GSR_ERR_INVALID; }
  if (scratch_bytes < Scratch(nullptr, in->P, 0).bytes) { set_error("scratch too small"); return GSR_ERR_CAPACITY; }
  if (in->P == 0) {
    GSR_CUDA_CHECK(cudaMemsetAsync(num_rendered_dev, 0, 4, st));
    if (num_rendered_host) *num_rendered_host = 0;
    return GSR_OK;
  }
  Geom g(geom, in->P);
  Scratch s(scratch, in->P, 0);
  GSR_CUDA_CHECK(cudaMemsetAsync(s.scan_status, 0, s.scan_zero_bytes, st));
  // the depth keys were left in the view's offsets_sorted array (gsr_forward_preprocess_views); 4 passes bring the
  // sorted keys back there and the order into geom.perm; the scan then overwrites offsets_sorted with its result
  { StageScope ps(ST_SORT, st);
    GSR_TRY(launch_sort_pairs_pingpong_u32(g.offsets_sorted, g.perm, s.dkeys_b, s.dvals_b, in->P, nullptr,
                                           make_sort_plan(in->P, 32, 4), s.sort_temp, false, true, st)); }
  GSR_TRY(debug_sync(in, st, "depth_sort"));
  { StageScope ps(ST_DUPLICATE, st); GSR_TRY(launch_sorted_scan(in->P, g, s, num_rendered_dev, num_rendered_host, st)); }
  return debug_sync(in, st, "sorted_scan");
}

// L = capacity of point_list / the tile-sort buffers in pairs.  n_dev == nullptr: L is also the exact pair count
// (the host read it back).  n_dev != nullptr: the count is read on the device (speculative launch, the host has not
// seen it yet); if it exceeds L the result is incomplete -- never out of bounds -- and the caller redoes the call.
static int forward_binning_impl(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                                const uint32_t* n_dev, uint32_t* point_list, void* image, void* scratch,
                                size_t scratch_bytes, cudaStream_t st) {
  if (!image || L < 0) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (in->P == 0) return GSR_OK;
  if (!geom || !radii || !scratch || (L > 0 && !point_list)) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (scratch_bytes < Scratch(nullptr, in->P, L).bytes) {
    set_error("scratch too small: need %zu bytes for L=%lld", Scratch(nullptr, in->P, L).bytes, (long long)L);
    return GSR_ERR_CAPACITY;
  }
  Geom g(const_cast<void*>(geom), in->P);
  Image im(image, in->W, in->H);
  Scratch s(scratch, in->P, L);
  const int gx = (in->W + TILE_X - 1) / TILE_X, gy = (in->H + TILE_Y - 1) / TILE_Y;
  const int end_bit = (int)higher_msb((uint32_t)(gx * gy));
  const SortPlan plan = make_sort_plan(L, end_bit, 4);
  const int passes = plan.passes;
  // pass p reads pair (p&1) and writes the other: start in the pair that makes the result land in point_list
  uint32_t* k0 = (passes & 1) ? s.tkeys_b : s.tkeys_a;
  uint32_t* k1 = (passes & 1) ? s.tkeys_a : s.tkeys_b;
  uint32_t* v0 = (passes & 1) ? s.vals_a : point_list;
  uint32_t* v1 = (passes & 1) ? point_list : s.vals_a;
  if (L > 0) {
    // up to two digits (65536 tiles): duplicate_tiles builds the sort's digit histograms
    const bool fused_hist = passes <= 2;
    uint32_t* hist = nullptr;
    { StageScope ps(ST_DUPLICATE, st);
      if (fused_hist) GSR_TRY(sort_prepare(s.sort_temp, L, plan, &hist, st));
      GSR_TRY(launch_duplicate_tiles(in->P, g, gx, k0, v0, hist, plan, L, st)); }
    GSR_TRY(debug_sync(in, st, "duplicate_tiles"));
    { StageScope ps(ST_SORT, st);
      GSR_TRY(launch_sort_pairs_pingpong_u32(k0, v0, k1, v1, L, n_dev, plan, s.sort_temp, fused_hist, false, st)); }
    GSR_TRY(debug_sync(in, st, "sort"));
  }
  { StageScope ps(ST_RANGES, st); GSR_TRY(launch_tile_ranges(L, n_dev, s.tkeys_a, im.ranges, gx * gy, st)); }
  return debug_sync(in, st, "tile_ranges");
}

static int forward_blend_impl(const gsr_inputs* in, const void* geom, const uint32_t* point_list, void* image,
                              float* out_color, float* out_objects, cudaStream_t st) {
  if (!image || !out_color || (in->K > 0 && !out_objects)) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  const size_t HW = (size_t)in->W * in->H;
  if (in->P == 0) {  // rasterize_points.cu:68,81: zero-filled image, nothing launched
    GSR_CUDA_CHECK(cudaMemsetAsync(out_color, 0, 3 * HW * sizeof(float), st));
    if (in->K > 0) GSR_CUDA_CHECK(cudaMemsetAsync(out_objects, 0, in->K * HW * sizeof(float), st));
    return GSR_OK;
  }
  if (!geom) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  Geom g(const_cast<void*>(geom), in->P);
  Image im(image, in->W, in->H);
  { StageScope ps(ST_BLEND_FWD, st); GSR_TRY(launch_blend_fwd(*in, g, im, point_list, out_color, out_objects, st)); }
  return debug_sync(in, st, "blend_fwd");
}

int gsr_forward_binning(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L, uint32_t* point_list,
                        void* image, void* scratch, size_t scratch_bytes, void* stream) {
  GSR_TRY(check_inputs(in));
  return forward_binning_impl(in, geom, radii, L, nullptr, point_list, image, scratch, scratch_bytes, (cudaStream_t)stream);
}

int gsr_forward_blend(const gsr_inputs* in, const void* geom, const uint32_t* point_list, void* image,
                      float* out_color, float* out_objects, void* stream) {
  GSR_TRY(check_inputs(in));
  return forward_blend_impl(in, geom, point_list, image, out_color, out_objects, (cudaStream_t)stream);
}

int gsr_forward_render(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L, uint32_t* point_list,
                       void* image, void* scratch, size_t scratch_bytes, float* out_color, float* out_objects,
                       void* stream) {
  GSR_TRY(check_inputs(in));
  if (!out_color || (in->K > 0 && !out_objects)) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  GSR_TRY(forward_binning_impl(in, geom, radii, L, nullptr, point_list, image, scratch, scratch_bytes, (cudaStream_t)stream));
  return forward_blend_impl(in, geom, point_list, image, out_color, out_objects, (cudaStream_t)stream);
}

int gsr_forward_render_speculative(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L_cap,
                                   const uint32_t* num_rendered_dev, uint32_t* point_list, void* image, void* scratch,
                                   size_t scratch_bytes, float* out_color, float* out_objects, void* stream) {
  GSR_TRY(check_inputs(in));
  if (!out_color || (in->K > 0 && !out_objects) || !num_rendered_dev) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (L_cap < 1) { set_error("forward_render_speculative: capacity must be >= 1"); return GSR_ERR_INVALID; }
  GSR_TRY(forward_binning_impl(in, geom, radii, L_cap, num_rendered_dev, point_list, image, scratch, scratch_bytes,
                               (cudaStream_t)stream));
  return forward_blend_impl(in, geom, point_list, image, out_color, out_objects, (cudaStream_t)stream);
}

static int backward_impl(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                         const uint32_t* point_list, const void* image, void* scratch, size_t scratch_bytes,
                         const float* dL_dpix, const float* dL_dobjpix, float* dL_dmean2D, float* dL_dopacity,
                         float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                         float* dL_drot, float* dL_dobj, bool accumulate, int phase, void* stream) {
  GSR_TRY(check_inputs(in));
  cudaStream_t st = (cudaStream_t)stream;
  if (in->P == 0) return GSR_OK;
  if (!geom || !radii || !image || !scratch || !dL_dpix || (!dL_dmean2D && !accumulate) || !dL_dopacity || !dL_dmean3D ||
      (L > 0 && !point_list)) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (in->shs && !dL_dsh) { set_error("dL_dsh is NULL but shs given"); return GSR_ERR_INVALID; }
  if (in->shs_rest && !in->dL_dshs_rest) { set_error("dL_dshs_rest is NULL but shs_rest given"); return GSR_ERR_INVALID; }
  if (in->scales && (!dL_dscale || !dL_drot)) { set_error("dL_dscale/dL_drot NULL but scales given"); return GSR_ERR_INVALID; }
  if (in->cov3D_precomp && !dL_dcov3D) { set_error("dL_dcov3D NULL but cov3D_precomp given"); return GSR_ERR_INVALID; }
  if (in->colors_precomp && !dL_dcolor) { set_error("dL_dcolor NULL but colors_precomp given"); return GSR_ERR_INVALID; }
  if (in->K > 0 && (!dL_dobjpix || !dL_dobj)) { set_error("object gradients NULL but K>0"); return GSR_ERR_INVALID; }
  if (accumulate && in->K > 0) { set_error("backward_accumulate: object channels are not supported"); return GSR_ERR_INVALID; }
  if (misaligned16(dL_dsh) || misaligned16(dL_drot) || misaligned16(dL_dobj) || misaligned16(scratch)) {
    set_error("dL_dsh, dL_drot, dL_dobj and scratch must be 16-byte aligned (vector stores / reductions)"); return GSR_ERR_INVALID;
  }
  if (scratch_bytes < Scratch(nullptr, in->P, 0).bytes) { set_error("scratch too small"); return GSR_ERR_CAPACITY; }
  Geom g(const_cast<void*>(geom), in->P);
  Image im(const_cast<void*>(image), in->W, in->H);
  Scratch s(scratch, in->P, 0);
  if (phase & 1) {
    StageScope ps(ST_BLEND_BWD, st);
    GSR_TRY(launch_blend_bwd(*in, g, im, point_list, dL_dpix, dL_dobjpix, s.accum, dL_dobj, st));
  }
  GSR_TRY(debug_sync(in, st, "blend_bwd"));
  if (phase & 2) {
    StageScope ps(ST_PREPROCESS_BWD, st);
    GSR_TRY(launch_preprocess_bwd(*in, g, radii, s.accum, dL_dmean2D, dL_dopacity, in->colors_precomp ? dL_dcolor : nullptr,
                                  dL_dmean3D, in->cov3D_precomp ? dL_dcov3D : nullptr, in->shs ? dL_dsh : nullptr,
                                  in->scales ? dL_dscale : nullptr, in->scales ? dL_drot : nullptr, accumulate, st));
  }
  GSR_TRY(debug_sync(in, st, "preprocess_bwd"));
  return GSR_OK;
}

int gsr_backward(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L, const uint32_t* point_list,
                 const void* image, void* scratch, size_t scratch_bytes, const float* dL_dpix, const float* dL_dobjpix,
                 float* dL_dmean2D, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dobj, void* stream) {
  return backward_impl(in, geom, radii, L, point_list, image, scratch, scratch_bytes, dL_dpix, dL_dobjpix, dL_dmean2D,
                       dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dobj, false, 3, stream);
}

int gsr_backward_accumulate(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                            const uint32_t* point_list, const void* image, void* scratch, size_t scratch_bytes,
                            const float* dL_dpix, float* dL_dmean2D, float* dL_dopacity, float* dL_dcolor,
                            float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                            void* stream) {
  return backward_impl(in, geom, radii, L, point_list, image, scratch, scratch_bytes, dL_dpix, nullptr, dL_dmean2D,
                       dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, nullptr, true, 3, stream);
}

int gsr_backward_accumulate_phase(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                                  const uint32_t* point_list, const void* image, void* scratch, size_t scratch_bytes,
                                  const float* dL_dpix, float* dL_dmean2D, float* dL_dopacity, float* dL_dcolor,
                                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                  int32_t phase, void* stream) {
  if (phase < 1 || phase > 3) { set_error("backward_accumulate_phase: phase must be 1, 2 or 3"); return GSR_ERR_INVALID; }
  return backward_impl(in, geom, radii, L, point_list, image, scratch, scratch_bytes, dL_dpix, nullptr, dL_dmean2D,
                       dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, nullptr, true, phase, stream);
}

size_t gsr_view_accum_bytes(int32_t P) { return align_up((size_t)(P > 0 ? P : 0) * 12 * sizeof(float)); }

int gsr_backward_blend_view(const gsr_inputs* in, const void* geom, int64_t L, const uint32_t* point_list,
                            const void* image, const float* dL_dpix, float* accum, void* stream) {
  GSR_TRY(check_inputs(in));
  cudaStream_t st = (cudaStream_t)stream;
  if (in->P == 0) return GSR_OK;
  if (!geom || !image || !dL_dpix || !accum || (L > 0 && !point_list)) { set_error("NULL buffer"); return GSR_ERR_INVALID; }
  if (in->K > 0) { set_error("backward_blend_view: object channels are not supported"); return GSR_ERR_INVALID; }
  if (misaligned16(accum)) { set_error("accum must be 16-byte aligned"); return GSR_ERR_INVALID; }
  Geom g(const_cast<void*>(geom), in->P);
  Image im(const_cast<void*>(image), in->W, in->H);
  {
    StageScope ps(ST_BLEND_BWD, st);
    GSR_TRY(launch_blend_bwd(*in, g, im, point_list, dL_dpix, nullptr, accum, nullptr, st));
  }
  return debug_sync(in, st, "blend_bwd");
}

int gsr_backward_preprocess_views(const gsr_inputs* in, int32_t n_views, const gsr_view_ref* views,
                                  int32_t accumulate, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                                  float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, void* stream) {
  if (!in) { set_error("inputs is NULL"); return GSR_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  if (in->P < 0) { set_error("bad size P=%d", in->P); return GSR_ERR_INVALID; }
  if (in->P == 0 || n_views == 0) return GSR_OK;
  if (n_views < 0 || n_views > GSR_MAX_FUSED_VIEWS || !views) {
    set_error("backward_preprocess_views: 1..%d views per call, got %d", GSR_MAX_FUSED_VIEWS, n_views); return GSR_ERR_INVALID;
  }
  if (in->raw_params || in->shs_rest || in->K > 0) {
    set_error("backward_preprocess_views: raw_params / shs_rest / object channels are not supported"); return GSR_ERR_INVALID;
  }
  if (!in->means3D || (in->shs == nullptr) == (in->colors_precomp == nullptr)) {
    set_error("Please provide excatly one of either SHs or precomputed colors!"); return GSR_ERR_INVALID;
  }
  const bool sr = in->scales != nullptr && in->rotations != nullptr;
  if (sr == (in->cov3D_precomp != nullptr) || ((in->scales != nullptr) != (in->rotations != nullptr))) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return GSR_ERR_INVALID;
  }
  if (in->shs && (in->M < 1 || in->M > 16 || in->D < 0 || in->D > 3 || (in->D + 1) * (in->D + 1) > in->M)) {
    set_error("SH degree %d / M=%d not supported by the multi-view backward (M <= 16)", in->D, in->M); return GSR_ERR_INVALID;
  }
  if (!dL_dopacity || !dL_dmean3D || (in->shs && !dL_dsh) || (in->scales && (!dL_dscale || !dL_drot)) ||
      (in->cov3D_precomp && !dL_dcov3D) || (in->colors_precomp && !dL_dcolor)) {
    set_error("a gradient buffer is NULL for an input that was given"); return GSR_ERR_INVALID;
  }
  if (misaligned16(in->shs) || misaligned16(in->rotations) || misaligned16(dL_dsh) || misaligned16(dL_drot)) {
    set_error("shs, rotations, dL_dsh and dL_drot must be 16-byte aligned"); return GSR_ERR_INVALID;
  }
  for (int v = 0; v < n_views; v++) {
    const gsr_view_ref& r = views[v];
    if (!r.viewmatrix || !r.projmatrix || !r.campos || !r.geom || !r.radii || !r.accum || r.W <= 0 || r.H <= 0) {
      set_error("backward_preprocess_views: view %d has a NULL pointer or a bad size", v); return GSR_ERR_INVALID;
    }
    if (misaligned16(r.accum) || misaligned16(r.geom)) { set_error("view %d: accum / geom must be 16-byte aligned", v); return GSR_ERR_INVALID; }
  }
  {
    StageScope ps(ST_PREPROCESS_BWD, st);
    GSR_TRY(launch_preprocess_bwd_views(*in, n_views, views, accumulate != 0, dL_dopacity,
                                        in->colors_precomp ? dL_dcolor : nullptr, dL_dmean3D,
                                        in->cov3D_precomp ? dL_dcov3D : nullptr, in->shs ? dL_dsh : nullptr,
                                        in->scales ? dL_dscale : nullptr, in->scales ? dL_drot : nullptr, st));
  }
  return debug_sync(in, st, "preprocess_bwd_views");
}

int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* /*projmatrix*/,
                     uint8_t* present, void* stream) {
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("mark_visible: bad arguments"); return GSR_ERR_INVALID; }
  if (P == 0) return GSR_OK;
  return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

int gsr_pgd_linf_update(float* param, const float* grad, const float* orig, int64_t n, float alpha, float epsilon,
                        void* stream) {
  if (n < 0 || (n > 0 && (!param || !grad || !orig))) { set_error("pgd_linf_update: bad arguments"); return GSR_ERR_INVALID; }
  return launch_pgd_linf(param, grad, orig, n, alpha, epsilon, (cudaStream_t)stream);
}

int gsr_pgd_l2_update(float* param, const float* grad, const float* orig, int64_t rows, int32_t width, float alpha,
                      float epsilon, double* scratch8, void* stream) {
  if (rows < 0 || width < 0 || (rows * width > 0 && (!param || !grad || !orig || !scratch8))) {
    set_error("pgd_l2_update: bad arguments"); return GSR_ERR_INVALID;
  }
  return launch_pgd_l2(param, grad, orig, rows, width, alpha, epsilon, scratch8, (cudaStream_t)stream);
}

int gsr_image_to_u8_hwc(const float* chw, uint8_t* hwc, int32_t W, int32_t H, void* stream) {
  if (W < 0 || H < 0 || ((int64_t)W * H > 0 && (!chw || !hwc))) { set_error("image_to_u8_hwc: bad arguments"); return GSR_ERR_INVALID; }
  return launch_image_u8_hwc(chw, hwc, W, H, (cudaStream_t)stream);
}

int gsr_detector_scale(const float* x, float* y, const float* dL_dy, float* dL_dx, int64_t n, void* stream) {
  if (n < 0 || (n > 0 && (!x || (!y && !dL_dx) || (dL_dx && !dL_dy)))) { set_error("detector_scale: bad arguments"); return GSR_ERR_INVALID; }
  return launch_detector_scale(x, y, dL_dy, dL_dx, n, (cudaStream_t)stream);
}

int gsr_sort_pairs_u64(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                       int64_t n, int32_t end_bit, void* temp, size_t temp_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (n < 0 || end_bit < 1 || end_bit > 64) { set_error("sort: bad n/end_bit"); return GSR_ERR_INVALID; }
  if (n == 0) return GSR_OK;
  if (temp_bytes < gsr_sort_temp_bytes(n)) { set_error("sort temp too small"); return GSR_ERR_CAPACITY; }
  Carver c(temp);
  const SortPlan plan = make_sort_plan(n, end_bit, 8);
  void* st_temp = c.take<char>(sort_temp_bytes(n, 64, 8));
  uint64_t* ktmp = c.take<uint64_t>((size_t)n);
  uint32_t* vtmp = c.take<uint32_t>((size_t)n);
  const int passes = plan.passes;
  uint64_t* k0 = (passes & 1) ? ktmp : keys_out;
  uint32_t* v0 = (passes & 1) ? vtmp : vals_out;
  uint64_t* k1 = (passes & 1) ? keys_out : ktmp;
  uint32_t* v1 = (passes & 1) ? vals_out : vtmp;
  GSR_CUDA_CHECK(cudaMemcpyAsync(k0, keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
  GSR_CUDA_CHECK(cudaMemcpyAsync(v0, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
  return launch_sort_pairs_pingpong(k0, v0, k1, v1, n, plan, st_temp, st);
}

size_t gsr_sort_u32_temp_bytes(int64_t n_cap) { return sort_temp_bytes(n_cap, 32, 4) + 2 * align_up((size_t)n_cap * 4); }

int gsr_sort_pairs_u32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                       int64_t n_cap, const uint32_t* n_dev, int32_t end_bit, void* temp, size_t temp_bytes,
                       void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (n_cap < 0 || end_bit < 1 || end_bit > 32) { set_error("sort: bad n/end_bit"); return GSR_ERR_INVALID; }
  if (n_cap == 0) return GSR_OK;
  if (temp_bytes < gsr_sort_u32_temp_bytes(n_cap)) { set_error("sort temp too small"); return GSR_ERR_CAPACITY; }
  const SortPlan plan = make_sort_plan(n_cap, end_bit, 4);
  Carver c(temp);
  void* st_temp = c.take<char>(sort_temp_bytes(n_cap, 32, 4));
  uint32_t* ktmp = c.take<uint32_t>((size_t)n_cap);
  uint32_t* vtmp = c.take<uint32_t>((size_t)n_cap);
  const int passes = plan.passes;
  uint32_t* k0 = (passes & 1) ? ktmp : keys_out;
  uint32_t* v0 = (passes & 1) ? vtmp : vals_out;
  uint32_t* k1 = (passes & 1) ? keys_out : ktmp;
  uint32_t* v1 = (passes & 1) ? vals_out : vtmp;
  if (keys_in != k0) GSR_CUDA_CHECK(cudaMemcpyAsync(k0, keys_in, (size_t)n_cap * 4, cudaMemcpyDeviceToDevice, st));
  if (vals_in != v0) GSR_CUDA_CHECK(cudaMemcpyAsync(v0, vals_in, (size_t)n_cap * 4, cudaMemcpyDeviceToDevice, st));
  return launch_sort_pairs_pingpong_u32(k0, v0, k1, v1, n_cap, n_dev, plan, st_temp, false, false, st);
}

int64_t gsr_debug_fetch(const char* name, int32_t P, int32_t W, int32_t H, int64_t L, const void* geom,
                        const void* image, const void* scratch, const uint32_t* point_list, void* dst, int64_t cap,
                        void* stream) {
  Geom g(const_cast<void*>(geom), P);
  Image im(const_cast<void*>(image), W, H);
  Scratch s(const_cast<void*>(scratch), P, L);
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
  const size_t n = (size_t)P, HW = (size_t)W * H;
  const void* src = nullptr;
  size_t bytes = 0;
#define F(nm, ptr, cnt) if (!strcmp(name, nm)) { src = ptr; bytes = (cnt); }
  F("depths", g.depths, n * 4) F("means2D", g.means2D, n * 8) F("cov3D", g.cov3D, n * 24)
  F("conic_opacity", g.conic_opacity, n * 16) F("rgb", g.rgb_thr, n * 16) F("clamped", g.clamped, n)
  F("tiles_touched", g.tiles_touched, n * 4) F("rects", g.rects, n * 8)
  F("perm", g.perm, n * 4) F("offsets_sorted", g.offsets_sorted, n * 4)
  F("final_T", im.final_T, HW * 4) F("n_contrib", im.n_contrib, HW * 4) F("ranges", im.ranges, (size_t)gx * gy * 8)
  F("tiles_sorted", s.tkeys_a, (size_t)L * 4)
#undef F
  if (!strcmp(name, "point_offsets")) {
    // the reference's InclusiveSum over tiles_touched in index order (rasterizer_impl.cu:277), computed on demand
    bytes = n * 4;
    if (dst) {
      if (!scratch || cap < (int64_t)bytes) { set_error("debug_fetch: point_offsets needs scratch and cap >= 4 P"); return GSR_ERR_INVALID; }
      GSR_TRY(launch_index_scan(P, g, s, (uint32_t*)dst, (cudaStream_t)stream));
    }
    return (int64_t)bytes;
  }
  if (!strcmp(name, "keys_sorted")) {
    // the reference's sorted 64-bit keys (tile << 32 | depth bits), rebuilt from the sorted tile ids + point list
    bytes = (size_t)L * 8;
    if (dst) {
      if (!scratch || !point_list || cap < (int64_t)bytes) { set_error("debug_fetch: keys_sorted needs scratch, point_list and cap >= 8 L"); return GSR_ERR_INVALID; }
      GSR_TRY(launch_rebuild_keys(L, s.tkeys_a, point_list, g.depths, (uint64_t*)dst, (cudaStream_t)stream));
    }
    return (int64_t)bytes;
  }
  if (!src && bytes == 0 && strcmp(name, "tiles_sorted")) { set_error("debug_fetch: unknown array '%s'", name); return GSR_ERR_INVALID; }
  if (dst && bytes) {
    const size_t ncopy = bytes < (size_t)cap ? bytes : (size_t)cap;
    if (cudaMemcpyAsync(dst, src, ncopy, cudaMemcpyDeviceToDevice, (cudaStream_t)stream) != cudaSuccess) {
      set_error("debug_fetch: copy failed"); return GSR_ERR_CUDA;
    }
  }
  return (int64_t)bytes;
}

}  // extern "C"
