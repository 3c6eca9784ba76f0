// common.cuh -- shared declarations of the sm_100a rasterizer kernels (internal; the public
// boundary is include/gsr_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/gsr_b200.h"

namespace gsr {

constexpr int TILE_X = 16;  // DGR/cuda_rasterizer/config.h:16-17 (BLOCK_X/BLOCK_Y); part of the result
constexpr int TILE_Y = 16;  // definition: tile lists and n_contrib depend on it
constexpr int TILE_PIX = TILE_X * TILE_Y;
constexpr int MAX_OBJ = 16;  // scene/gaussian_model.py:52 num_objects

// ---- thread-local error string ------------------------------------------------------------
void set_error(const char* fmt, ...);
void count_launches(int n);
#define GSR_CUDA_CHECK(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      gsr::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return GSR_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)
#define GSR_LAUNCH_CHECK(name)                                                                 \
  do {                                                                                         \
    cudaError_t _e = cudaGetLastError();                                                       \
    gsr::count_launches(1);                                                                    \
    if (_e != cudaSuccess) {                                                                   \
      gsr::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));                 \
      return GSR_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)

// ---- buffer carving (all sub-arrays 256-byte aligned) ---------------------------------------
inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

struct Carver {
  char* base;
  size_t off = 0;
  explicit Carver(void* b) : base((char*)b) {}
  template <typename T>
  T* take(size_t n) {
    T* p = (T*)(base ? base + off : nullptr);
    off = align_up(off + n * sizeof(T));
    return p;
  }
};

// A splat next to the camera can cover the whole image (32400 tiles at 4K): expanded by one warp, a few of them in the
// same 256-Gaussian chunk are the tail of the whole emission kernel (C4: 0.79 ms).  Rects above GIANT_TILES are
// listed by the scan (up to GIANT_CAP; beyond that everything is large -- the dense-overlap case -- and the work is
// balanced anyway) and dealt to the blocks one at a time, each expanded by a whole block.
#ifndef GSR_GIANT_TILES
#define GSR_GIANT_TILES 512
#endif
constexpr uint32_t GIANT_TILES = GSR_GIANT_TILES;
constexpr int GIANT_CAP = 16384;

// Per-view geometry state (what the reference keeps in GeometryState, rasterizer_impl.h:29-44,
// minus CUB temp, plus the tile rect and the packed blend record).
struct Geom {
  float* depths;           // [P]
  float2* means2D;         // [P]
  float4* conic_opacity;   // [P]
  float4* rgb_thr;         // [P]  (r, g, b, power threshold below which alpha < 1/255)
  float* cov3D;            // [6P]
  uint8_t* clamped;        // [P]  bit c set <=> channel c was clamped at 0 (forward.cu:66-69)
  uint32_t* tiles_touched; // [P]
  ushort4* rects;          // [P]  (x0, y0, x1, y1) tile rect (auxiliary.h:46-56)
  uint32_t* perm;          // [P]  Gaussian indices in (depth bits, index) order
  uint32_t* offsets_sorted;// [P]  inclusive prefix sum of tiles_touched[perm[.]]
  uint32_t* giant_count;   // [1]  rects above GIANT_TILES (listed by the depth-order scan for the emission kernel; they
  uint4* giants;           // [GIANT_CAP] (Gaussian, first output slot, tiles, x0 | y0 << 16)   live here, not in scratch,
  uint32_t* giants_w;      // [GIANT_CAP] width in tiles                                       because scratch may be
  size_t bytes;            //                                                    re-allocated between the two phases)
  Geom(void* buf, int P) {
    Carver c(buf);
    size_t n = (size_t)P;
    depths = c.take<float>(n);
    means2D = c.take<float2>(n);
    conic_opacity = c.take<float4>(n);
    rgb_thr = c.take<float4>(n);
    cov3D = c.take<float>(6 * n);
    clamped = c.take<uint8_t>(n);
    tiles_touched = c.take<uint32_t>(n);
    rects = c.take<ushort4>(n);
    perm = c.take<uint32_t>(n);
    offsets_sorted = c.take<uint32_t>(n);
    giant_count = c.take<uint32_t>(1);
    giants = c.take<uint4>(GIANT_CAP);
    giants_w = c.take<uint32_t>(GIANT_CAP);
    bytes = c.off;
  }
};

// Per-view image state (ImageState, rasterizer_impl.h:46-53).
struct Image {
  float* final_T;       // [W*H]
  uint32_t* n_contrib;  // [W*H]
  uint2* ranges;        // [tiles]
  size_t bytes;
  Image(void* buf, int W, int H) {
    Carver c(buf);
    size_t n = (size_t)W * H;
    size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    final_T = c.take<float>(n);
    n_contrib = c.take<uint32_t>(n);
    ranges = c.take<uint2>(tiles);
    bytes = c.off;
  }
};

// ---- radix sort geometry ---------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
#ifndef GSR_SORT_ITEMS
#define GSR_SORT_ITEMS 16
#endif
#ifndef GSR_SORT_ITEMS_SMALL
#define GSR_SORT_ITEMS_SMALL 8
#endif
constexpr int SORT_ITEMS = GSR_SORT_ITEMS;              // 4096 pairs per block
constexpr int SORT_ITEMS_SMALL = GSR_SORT_ITEMS_SMALL;  // 2048 pairs per block for sorts of <= SORT_SMALL_TILE_LIMIT pairs
#ifndef GSR_SORT_SMALL_LIMIT_LOG2
#define GSR_SORT_SMALL_LIMIT_LOG2 21
#endif
constexpr int64_t SORT_SMALL_TILE_LIMIT = (int64_t)1 << GSR_SORT_SMALL_LIMIT_LOG2;
constexpr int64_t SORT_SMALL_LIMIT = (int64_t)1 << 30;  // below: 32-bit look-back words, above: 64-bit
constexpr int SORT_MAX_PASSES = 8;
constexpr int SORT_LB_GROUP = 8;   // tiles per look-back group (radix_sort.cu)
inline int64_t sort_tiles(int64_t n, int items) { return (n + SORT_THREADS * items - 1) / (SORT_THREADS * items); }

// Digit layout and tile size of one LSD sort (radix_sort.cu: make_sort_plan).
struct SortPlan {
  int end_bit = 0, passes = 0, items = SORT_ITEMS;
  int shift[SORT_MAX_PASSES] = {0}, bits[SORT_MAX_PASSES] = {0};
};
SortPlan make_sort_plan(int64_t n_cap, int end_bit, int key_bytes);

struct SortTemp {
  uint32_t* hist;     // [SORT_MAX_PASSES][256] digit histograms (raw counts; every scatter block scans them)
  uint32_t* tickets;  // [SORT_MAX_PASSES] dynamic tile ids
  uint32_t* status;   // [passes]{[tiles][256] tile words, [groups][256] group words}: decoupled look-back
  size_t status_off, pass_stride, group_off;
  size_t bytes;
  SortTemp(void* buf, int64_t n_cap, int passes, int items) {
    Carver c(buf);
    hist = c.take<uint32_t>(SORT_MAX_PASSES * 256);
    tickets = c.take<uint32_t>(64);
    status_off = c.off;
    const size_t word = n_cap >= SORT_SMALL_LIMIT ? 8 : 4;
    const size_t tiles = (size_t)sort_tiles(n_cap, items);
    group_off = tiles * 256 * word;
    pass_stride = group_off + (tiles + SORT_LB_GROUP - 1) / SORT_LB_GROUP * 256 * word;
    status = (uint32_t*)c.take<char>((size_t)passes * pass_stride);
    bytes = c.off;
  }
  // prefix of the buffer that must be zeroed before a sort of `passes` passes
  size_t zero_bytes(int passes) const { return status_off + (size_t)passes * pass_stride; }
};
inline size_t sort_temp_bytes(int64_t n_cap, int end_bit, int key_bytes) {
  const SortPlan p = make_sort_plan(n_cap, end_bit, key_bytes);
  return SortTemp(nullptr, n_cap, p.passes, p.items).bytes;
}

// Reusable scratch (not needed after the call that used it returns).
//
// Binning works in two stable sorts instead of the reference's one sort of 64-bit (tile | depth) keys
// (rasterizer_impl.cu:289-308): Gaussians are first ordered by (depth bits, index) -- a 32-bit sort of P
// items -- then (tile, Gaussian) pairs are emitted in that order and stably sorted by tile id alone.
// Within a tile the result is ordered by (depth bits, index), i.e. exactly the reference's order.
constexpr int PRE_THREADS = 256;  // Gaussians per preprocess block
#ifndef GSR_SCAN_ITEMS
#define GSR_SCAN_ITEMS 2048
#endif
constexpr int SCAN_ITEMS = GSR_SCAN_ITEMS;  // Gaussians per block of the sorted-order scan (a multiple of 1024)
struct Scratch {
  unsigned long long* scan_status;   // [pre_blocks] look-back words of the index-order prefix sum (debug fetch only)
  unsigned long long* scan2_status;  // [scan_blocks] look-back words of the sorted-order prefix sum
  uint32_t* scan_ticket;             // [64] ([0] preprocess, [1] sorted-order scan)
  uint32_t* flags;                   // [64] flags[0] = prefiltered violation
  float* accum;                      // [P][12] backward accumulators
  uint32_t* dkeys_a;                 // [P] depth keys (ping)
  uint32_t* dkeys_b;                 // [P] depth keys (pong)
  uint32_t* dvals_b;                 // [P] pong of geom.perm
  uint32_t* tkeys_a;                 // [L] tile ids (ping)
  uint32_t* tkeys_b;                 // [L] tile ids (pong)
  uint32_t* vals_a;                  // [L] partner of point_list
  void* sort_temp;
  size_t scan_zero_bytes;
  size_t bytes;
  Scratch(void* buf, int P, int64_t L) {
    Carver c(buf);
    size_t pre_blocks = ((size_t)P + PRE_THREADS - 1) / PRE_THREADS;
    size_t scan_blocks = ((size_t)P + SCAN_ITEMS - 1) / SCAN_ITEMS;
    scan_status = c.take<unsigned long long>(pre_blocks + 1);
    scan2_status = c.take<unsigned long long>(scan_blocks + 1);
    scan_ticket = c.take<uint32_t>(64);
    flags = c.take<uint32_t>(64);
    scan_zero_bytes = c.off;
    accum = c.take<float>((size_t)P * 12);
    dkeys_a = c.take<uint32_t>((size_t)P);
    dkeys_b = c.take<uint32_t>((size_t)P);
    dvals_b = c.take<uint32_t>((size_t)P);
    tkeys_a = c.take<uint32_t>((size_t)L);
    tkeys_b = c.take<uint32_t>((size_t)L);
    vals_a = c.take<uint32_t>((size_t)L);
    // shared by the depth sort (P pairs, 32 key bits) and the tile sort (L pairs, <= 32 key bits)
    const size_t ta = sort_temp_bytes((int64_t)P, 32, 4), tb = sort_temp_bytes(L, 32, 4);
    sort_temp = c.take<char>(ta > tb ? ta : tb);
    bytes = c.off;
  }
};

// ---- activations of the raw model parameters (gsr_inputs::raw_params; scene/gaussian_model.py:97-121), written
// like the torch kernels they replace: sigmoid = 1 / (1 + exp(-x)), exp, F.normalize = x / max(||x||_2, 1e-12)
#ifdef __CUDACC__
__device__ __forceinline__ float act_sigmoid(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }
__device__ __forceinline__ float4 act_normalize(float4 q, float* inv_norm_out) {
  const float n = sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.x, q.x), __fmul_rn(q.y, q.y)), __fmul_rn(q.z, q.z)),
                                  __fmul_rn(q.w, q.w)));
  const float d = fmaxf(n, 1e-12f);
  if (inv_norm_out) *inv_norm_out = 1.0f / d;
  return make_float4(__fdiv_rn(q.x, d), __fdiv_rn(q.y, d), __fdiv_rn(q.z, d), __fdiv_rn(q.w, d));
}
// The (P, 15, 3) SH "rest" block has 180-byte rows: only 4-byte aligned, but row idx begins exactly (idx & 3) floats
// after a 16-byte boundary (45 = 1 mod 4), so the 48-float window that starts at that boundary covers the row and is
// moved with 12 aligned 16-byte accesses; the up to 3 floats either side belong to the neighbouring rows (loads ignore
// them, reductions add 0 to them).  Not for the last row (its window would end past the array) or an unaligned base.
__device__ __forceinline__ bool rest_window_ok(const void* base, size_t idx, int P) {
  return (reinterpret_cast<uintptr_t>(base) & 15) == 0 && idx + 1 < (size_t)P;
}
// shv[3 .. 47] <- row idx of the rest block (shv[0 .. 2], the DC coefficient, is the caller's)
__device__ __forceinline__ void load_rest_row(const float* __restrict__ rest, size_t idx, float* shv) {
  const int s = (int)(idx & 3);
  const float4* wp = reinterpret_cast<const float4*>(rest + 45 * idx - s);
  float w[48], t[47];
#pragma unroll
  for (int k = 0; k < 12; k++) {
    const float4 v = __ldg(wp + k);
    w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
  }
#pragma unroll
  for (int m = 0; m < 47; m++) t[m] = (s & 1) ? w[m + 1] : w[m];
#pragma unroll
  for (int m = 0; m < 45; m++) shv[3 + m] = (s & 2) ? t[m + 2] : t[m];
}
// coefficient i (0 .. 3M-1, coefficient-major, channel-minor) of Gaussian idx from either SH layout
__device__ __forceinline__ float sh_coeff(const gsr_inputs& in, size_t idx, int i) {
  if (in.shs_rest) return i < 3 ? __ldg(in.shs + 3 * idx + i) : __ldg(in.shs_rest + (size_t)(in.M - 1) * 3 * idx + (i - 3));
  return __ldg(in.shs + (size_t)in.M * 3 * idx + i);
}
#endif

// ---- stage launchers (each returns a gsr_status) ---------------------------------------------
int launch_preprocess_fwd(const gsr_inputs& in, Geom& g, int32_t* radii, Scratch& s, uint32_t* num_rendered_dev,
                          cudaStream_t st);
int launch_preprocess_fwd_views(const gsr_inputs& in, int n_views, const gsr_view_fwd* views, cudaStream_t st);
int launch_sorted_scan(int P, const Geom& g, Scratch& s, uint32_t* num_rendered_dev, uint32_t* num_rendered_host,
                       cudaStream_t st);
int launch_index_scan(int P, const Geom& g, Scratch& s, uint32_t* out, cudaStream_t st);
int launch_duplicate_tiles(int P, const Geom& g, int gx, uint32_t* tile_keys, uint32_t* vals, uint32_t* hist,
                           const SortPlan& plan, int64_t cap, cudaStream_t st);
int sort_prepare(void* temp, int64_t n_cap, const SortPlan& plan, uint32_t** hist_out, cudaStream_t st);
int launch_rebuild_keys(int64_t L, const uint32_t* tile_keys, const uint32_t* point_list, const float* depths,
                        uint64_t* keys_out, cudaStream_t st);
int launch_sort_pairs_pingpong(uint64_t* keys0, uint32_t* vals0, uint64_t* keys1, uint32_t* vals1, int64_t n,
                               const SortPlan& plan, void* temp, cudaStream_t st);
int launch_sort_pairs_pingpong_u32(uint32_t* keys0, uint32_t* vals0, uint32_t* keys1, uint32_t* vals1, int64_t n_cap,
                                   const uint32_t* n_dev, const SortPlan& plan, void* temp, bool hist_ready,
                                   bool iota_vals, cudaStream_t st);
int launch_tile_ranges(int64_t L_cap, const uint32_t* n_dev, const uint32_t* tile_keys_sorted, uint2* ranges, int tiles,
                       cudaStream_t st);
int launch_blend_fwd(const gsr_inputs& in, const Geom& g, const Image& im, const uint32_t* point_list,
                     float* out_color, float* out_objects, cudaStream_t st);
int launch_blend_bwd(const gsr_inputs& in, const Geom& g, const Image& im, const uint32_t* point_list,
                     const float* dL_dpix, const float* dL_dobjpix, float* accum, float* dL_dobj, cudaStream_t st);
int launch_preprocess_bwd(const gsr_inputs& in, const Geom& g, const int32_t* radii, const float* accum,
                          float* dL_dmean2D, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                          float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, bool accumulate,
                          cudaStream_t st);
int launch_preprocess_bwd_views(const gsr_inputs& in, int n_views, const gsr_view_ref* views, bool accumulate,
                                float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                                float* dL_dsh, float* dL_dscale, float* dL_drot, cudaStream_t st);
int launch_pgd_linf(float* p, const float* g, const float* p0, int64_t n, float alpha, float eps, cudaStream_t st);
int launch_pgd_l2(float* p, const float* g, const float* p0, int64_t rows, int width, float alpha, float eps,
                  double* sumsq_scratch, cudaStream_t st);
int launch_image_u8_hwc(const float* chw, uint8_t* hwc, int W, int H, cudaStream_t st);
int launch_detector_scale(const float* x, float* y, const float* dy, float* dx, int64_t n, cudaStream_t st);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);

uint32_t higher_msb(uint32_t n);
int sort_profile_read(uint64_t* out_host, int reset);
int blend_fwd_stats(uint64_t* out_host, int reset);
int blend_bwd_stats(uint64_t* out_host, int reset);

// ---- per-stage profiling (gsr_profile_*): CUDA events on the launching stream around each stage,
// and a count of kernel launches (gsr_launch_count) ------------------------------------------------
enum Stage { ST_PREPROCESS_FWD = 0, ST_DUPLICATE, ST_SORT, ST_RANGES, ST_BLEND_FWD, ST_BLEND_BWD, ST_PREPROCESS_BWD,
             ST_COUNT };
bool profile_begin(int stage, cudaStream_t st);  // true: a record is open (and the profiling lock held) until profile_end
void profile_end(bool recording, cudaStream_t st);
void count_launches(int n);
void nvtx_push(int stage);  // NVTX range around the host-side launches of a stage (GSR_NVTX=1; abi.cu)
void nvtx_pop();
struct StageScope {
  int stage; cudaStream_t st; bool rec;
  StageScope(int s, cudaStream_t t) : stage(s), st(t) { nvtx_push(stage); rec = profile_begin(stage, st); }
  ~StageScope() { profile_end(rec, st); nvtx_pop(); }
};

}  // namespace gsr
