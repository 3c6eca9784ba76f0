/* gsr_b200.h -- C ABI of the B200-native differentiable Gaussian-splat rasterizer.
 *
 * This is the drop-in boundary for the hot path of poloclub/3d-gaussian-splat-attack:
 * it replaces the private pybind module `diff_gaussian_rasterization._C`
 *   (DGR/ext.cpp:15-19, DGR/rasterize_points.h:18-67;
 *    DGR = submodules/gaussian-splatting/submodules/diff-gaussian-rasterization)
 * and the C++ class behind it, CudaRasterizer::Rasterizer (DGR/cuda_rasterizer/rasterizer.h:20-85).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no torch types.  Every pointer named in gsr_inputs and
 *    every buffer argument is a DEVICE pointer unless the name ends in _host.
 *  - NULL pointer = absent input (the reference passes empty tensors -> nullptr,
 *    DGR/diff_gaussian_rasterization/__init__.py:197-207, forward.cu:205,241).
 *  - The caller owns every buffer; the library never allocates device memory.  Buffer sizes come
 *    from the gsr_*_bytes() queries.  geom / image / scratch buffers must be 256-byte aligned.
 *    Input and gradient arrays keep the reference's layouts and may start anywhere a float may,
 *    EXCEPT the ones moved with 16-byte vector accesses: shs, rotations, obj_feats, dL_dsh,
 *    dL_drot, dL_dobj must be 16-byte aligned (GSR_ERR_INVALID otherwise; torch allocations are).
 *  - Every launch goes to the cudaStream_t passed in (as void*); the reference uses the legacy
 *    default stream.
 *  - Every entry point returns 0 on success or a negative gsr_status; gsr_last_error() returns
 *    a message for the last failure on the calling thread.
 */
#ifndef GSR_B200_H
#define GSR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_ABI_VERSION 4

typedef enum {
  GSR_OK = 0,
  GSR_ERR_INVALID = -1,   /* bad argument (shape / NULL / exclusivity)                         */
  GSR_ERR_CUDA = -2,      /* a CUDA runtime call or launch failed                              */
  GSR_ERR_CAPACITY = -3,  /* a caller buffer is too small (needed size reported by the call)   */
  GSR_ERR_PREFILTERED = -4 /* prefiltered=1 but a Gaussian was culled (auxiliary.h:156-160)    */
} gsr_status;

/* Inputs of one view; mirrors the argument list of Rasterizer::forward
 * (DGR/cuda_rasterizer/rasterizer.h:27-50) plus the optional object-feature channel that
 * gaussian_renderer/__init__.py:86-95 passes as sh_objs. */
typedef struct gsr_inputs {
  int32_t P;              /* number of Gaussians                                            */
  int32_t D;              /* active SH degree (0..3)                                        */
  int32_t M;              /* SH coefficients per Gaussian in `shs` ((max_degree+1)^2), 0 if none */
  int32_t W, H;           /* image size                                                     */
  int32_t K;              /* object-feature channels (0 = none, <= 16)                      */
  int32_t prefiltered;    /* reference flag: trap if a Gaussian is near-culled              */
  int32_t debug;          /* reference flag: synchronise + check after every stage          */
  float scale_modifier, tanfovx, tanfovy;
  int32_t raw_params;     /* 0: inputs are what render() passes (gaussian_renderer/__init__.py:53-81).  1: `opacities`,
                           * `scales`, `rotations` are the model's RAW parameters and the activations of
                           * scene/gaussian_model.py:97-121 (sigmoid, exp, normalize) are applied inside preprocess;
                           * their gradient outputs are then w.r.t. the raw parameters           */
  const float* bg;             /* (>=3)                                                     */
  const float* means3D;        /* (P,3)                                                     */
  const float* shs;            /* (P,M,3) or NULL                                           */
  const float* colors_precomp; /* (P,3) or NULL                                             */
  const float* opacities;      /* (P,1)                                                     */
  const float* scales;         /* (P,3) or NULL                                             */
  const float* rotations;      /* (P,4) or NULL                                             */
  const float* cov3D_precomp;  /* (P,6) or NULL                                             */
  const float* viewmatrix;     /* (4,4) column-major in memory (scene/cameras.py:54)        */
  const float* projmatrix;     /* (4,4) column-major in memory                              */
  const float* campos;         /* (3)                                                       */
  const float* obj_feats;      /* (P,K) or NULL                                             */
  const float* shs_rest;       /* NULL, or (P,M-1,3): then `shs` is only the DC block (P,1,3) -- the model's
                                * _features_dc / _features_rest without the per-call torch.cat of
                                * scene/gaussian_model.py:110-113                                              */
  float* dL_dshs_rest;         /* output of the backward calls when shs_rest is given: (P,M-1,3); dL_dsh is then
                                * (P,1,3)                                                                      */
} gsr_inputs;

int gsr_abi_version(void);
const char* gsr_last_error(void);

/* ---- buffer sizing (replaces required<GeometryState/ImageState/BinningState>,
 *      DGR/cuda_rasterizer/rasterizer_impl.h:67-72) ---- */
size_t gsr_geom_bytes(int32_t P);               /* per-view, kept until that view's backward   */
size_t gsr_image_bytes(int32_t W, int32_t H);   /* per-view, kept until backward               */
size_t gsr_scratch_bytes(int32_t P, int64_t L_cap, int32_t W, int32_t H);  /* reusable across views */

/* ---- forward, phase 1: preprocess + fused prefix sum (forward.cu:155-256 preprocessCUDA,
 *      rasterizer_impl.cu:277 InclusiveSum).  Writes radii (P) and the geometry buffer, and the
 *      number of (tile, Gaussian) duplicates L to *num_rendered_dev (device u32); when
 *      num_rendered_host is non-NULL it is also stored there by a kernel in stream order, so it must be
 *      page-locked, device-mapped host memory (cudaHostAlloc / torch pin_memory under UVA); the store
 *      does not use a copy engine and therefore never queues behind another stream's bulk transfer. ---- */
int gsr_forward_geometry(const gsr_inputs* in, void* geom, int32_t* radii, void* scratch, size_t scratch_bytes,
                         uint32_t* num_rendered_dev, uint32_t* num_rendered_host, void* stream);

/* ---- forward, phase 2: duplicateWithKeys + radix sort + identifyTileRanges + blending
 *      (rasterizer_impl.cu:289-334, forward.cu:261-374).  L must be the value produced by phase 1.
 *      point_list (L u32) is an output kept for backward.  out_color (3,H,W); out_objects (K,H,W)
 *      when in->K > 0. ---- */
int gsr_forward_render(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                       uint32_t* point_list, void* image, void* scratch, size_t scratch_bytes,
                       float* out_color, float* out_objects, void* stream);

/* ---- forward, phase 2 WITHOUT the host knowing L (replaces the blocking read-back of rasterizer_impl.cu:281 in
 *      steady state): point_list and scratch are sized for L_cap pairs (e.g. 1.25 x the previous view's L) and the
 *      kernels read the actual count from num_rendered_dev (the word phase 1 wrote).  Grids are sized from L_cap,
 *      surplus blocks leave at once.  If the count turns out larger than L_cap nothing is written out of bounds,
 *      but the tile lists are incomplete: the caller -- which reads num_rendered_host once phase 1's kernels have
 *      run, while phase 2 is already executing -- must then repeat phase 2 through gsr_forward_render with the
 *      exact L and buffers of that size. ---- */
int gsr_forward_render_speculative(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L_cap,
                                   const uint32_t* num_rendered_dev, uint32_t* point_list, void* image,
                                   void* scratch, size_t scratch_bytes, float* out_color, float* out_objects,
                                   void* stream);

/* ---- phase 2 in two calls, so that a driver can put them on different streams: binning (duplicateWithKeys + sort +
 *      identifyTileRanges; uses `scratch`, writes point_list and the tile ranges inside `image`) and blending
 *      (reads geom / point_list / image only).  gsr_forward_render == gsr_forward_binning then gsr_forward_blend.
 *      views.ViewPipeline runs geometry + binning of the next camera on a high-priority stream underneath the
 *      blending kernels of the current one. ---- */
int gsr_forward_binning(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L, uint32_t* point_list,
                        void* image, void* scratch, size_t scratch_bytes, void* stream);
int gsr_forward_blend(const gsr_inputs* in, const void* geom, const uint32_t* point_list, void* image,
                      float* out_color, float* out_objects, void* stream);

/* ---- backward (Rasterizer::backward, rasterizer_impl.cu:340-433; rasterize_points.cu:117-196).
 *      All gradient outputs are fully written (no pre-zeroing by the caller).  Pass NULL for
 *      dL_dsh / dL_dscale+dL_drot / dL_dcov3D / dL_dcolor when the matching input is absent.
 *      dL_dcov3D is written only when cov3D_precomp was given, dL_dcolor only when colors_precomp
 *      was given (otherwise they are internal). ---- */
int gsr_backward(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                 const uint32_t* point_list, const void* image, void* scratch, size_t scratch_bytes,
                 const float* dL_dpix, const float* dL_dobjpix,
                 float* dL_dmean2D /*(P,3)*/, float* dL_dopacity /*(P,1)*/, float* dL_dcolor /*(P,3)*/,
                 float* dL_dmean3D /*(P,3)*/, float* dL_dcov3D /*(P,6)*/, float* dL_dsh /*(P,M,3)*/,
                 float* dL_dscale /*(P,3)*/, float* dL_drot /*(P,4)*/, float* dL_dobj /*(P,K)*/, void* stream);

/* ---- backward of one view ADDED to the caller's gradient buffers (new; replaces the reference's pattern of
 *      nine zero-filled tensors per view, rasterize_points.cu:151-159, followed by autograd's AccumulateGrad
 *      add per attribute when a step sums several cameras, attack.py:481-494).  Rows of Gaussians the view did
 *      not render are not touched; rendered rows are read-modify-written.  dL_dmean2D may be NULL.  The
 *      buffers must have been initialised by the caller (zeroed at the start of a step).  K must be 0. ---- */
int gsr_backward_accumulate(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                            const uint32_t* point_list, const void* image, void* scratch, size_t scratch_bytes,
                            const float* dL_dpix, float* dL_dmean2D /*(P,3) or NULL*/, float* dL_dopacity,
                            float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                            float* dL_drot, void* stream);

/* Same, one half at a time so a driver can put the halves on different streams: phase 1 = the per-tile blending
 * gradients into the scratch accumulators (issue-bound), phase 2 = the per-Gaussian backward that adds into the
 * caller's buffers (HBM-bound), phase 3 = both.  Phase 2 of a view must follow its phase 1, and the next view's
 * phase 1 must follow this view's phase 2 (they share the accumulators in `scratch`). */
int gsr_backward_accumulate_phase(const gsr_inputs* in, const void* geom, const int32_t* radii, int64_t L,
                                  const uint32_t* point_list, const void* image, void* scratch, size_t scratch_bytes,
                                  const float* dL_dpix, float* dL_dmean2D, float* dL_dopacity, float* dL_dcolor,
                                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                  int32_t phase, void* stream);

/* ---- backward of SEVERAL views with the per-Gaussian half done once for all of them (new; no counterpart in the
 *      reference, whose backward.cu:144-396 kernels run once per camera and whose caller sums cameras through
 *      autograd, attack.py:481-494).  Step 1, per view: gsr_backward_blend_view = the per-tile blending gradients
 *      (renderCUDA backward, backward.cu:399-557) into that view's own accumulator `accum`
 *      (gsr_view_accum_bytes(P) bytes, 256-byte aligned, zeroed by the call).  Step 2, once per group of up to
 *      GSR_MAX_FUSED_VIEWS views: gsr_backward_preprocess_views reads each Gaussian's attributes once, loops over the
 *      views that rendered it and writes (accumulate = 0: every row, zeros for Gaussians no view rendered) or adds
 *      (accumulate = 1: rows of Gaussians at least one view rendered) the gradients summed over the views.
 *      `in` supplies the scene (P, D, M, scale_modifier, means3D, shs, colors_precomp, scales, rotations,
 *      cov3D_precomp); its camera fields are ignored, each view brings its own.  `views` is a HOST array.
 *      The fused-activation inputs (raw_params, shs_rest) and K > 0 are not supported here (GSR_ERR_INVALID).
 *      Results equal the sum of gsr_backward_accumulate over the views up to summation order. ---- */
#define GSR_MAX_FUSED_VIEWS 16
typedef struct gsr_view_ref {
  const float* viewmatrix;  /* the view's camera, as in gsr_inputs                                        */
  const float* projmatrix;
  const float* campos;
  float tanfovx, tanfovy;
  int32_t W, H;
  const void* geom;         /* that view's geometry state (gsr_forward_geometry)                           */
  const int32_t* radii;     /* that view's radii (P)                                                       */
  float* accum;             /* that view's accumulator: written by gsr_backward_blend_view                 */
  float* dL_dmean2D;        /* optional per-view (P,3) screen-space gradient (written / added), or NULL    */
} gsr_view_ref;
size_t gsr_view_accum_bytes(int32_t P);
int gsr_backward_blend_view(const gsr_inputs* in, const void* geom, int64_t L, const uint32_t* point_list,
                            const void* image, const float* dL_dpix, float* accum, void* stream);
int gsr_backward_preprocess_views(const gsr_inputs* in, int32_t n_views, const gsr_view_ref* views_host,
                                  int32_t accumulate, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                                  float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, void* stream);

/* ---- forward phase 1 of SEVERAL views of the same Gaussians (new; the reference's preprocessCUDA, forward.cu:155-256,
 *      runs once per camera and re-reads every Gaussian's attributes -- 236 bytes, 192 of them the SH row -- each time).
 *      gsr_forward_preprocess_views reads the attributes once per Gaussian and writes the geometry state and radii of
 *      up to GSR_MAX_FUSED_VIEWS views (`views_host` is a HOST array); per view the arithmetic and therefore every
 *      output is bit-identical to gsr_forward_geometry's.  gsr_forward_order then finishes phase 1 of ONE view (depth
 *      order, prefix sum, duplicate count -- the rest of gsr_forward_geometry) and is followed by phase 2 as usual.
 *      `in` supplies the scene (camera fields ignored by the first call; P by the second).  Standard inputs only:
 *      raw_params / shs_rest / prefiltered / K > 0 return GSR_ERR_INVALID. ---- */
typedef struct gsr_view_fwd {
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  float tanfovx, tanfovy;
  int32_t W, H;
  void* geom;               /* output: that view's geometry state (gsr_geom_bytes(P))                      */
  int32_t* radii;           /* output: that view's radii (P)                                               */
} gsr_view_fwd;
int gsr_forward_preprocess_views(const gsr_inputs* in, int32_t n_views, const gsr_view_fwd* views_host, void* stream);
int gsr_forward_order(const gsr_inputs* in, void* geom, void* scratch, size_t scratch_bytes,
                      uint32_t* num_rendered_dev, uint32_t* num_rendered_host, void* stream);

/* ---- markVisible (rasterizer_impl.cu:141-153, checkFrustum :54-66); present is (P) bytes ---- */
int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* ---- the elementwise steps either side of the rasterizer in an attack iteration ----
 * PGD updates of one attribute tensor, in place (attack.py:25-173).  `orig` is the un-attacked copy.
 *   L-inf (gaussian_*_linf_attack):  p <- clamp(p - alpha sign(g) - orig, -eps, eps) + orig, rounded like the torch
 *                                    op sequence it replaces.
 *   L2   (gaussian_*_l2_attack):     p <- p - alpha g/||g||_2 (no step if the norm is 0); each row (one Gaussian,
 *                                    `width` floats) of p - orig longer than eps is scaled by eps/(len + 1e-7)
 *                                    (Tensor.renorm(2, 0, eps)); p <- orig + that.  scratch8 = 8 device bytes. */
int gsr_pgd_linf_update(float* param, const float* grad, const float* orig, int64_t n, float alpha, float epsilon,
                        void* stream);
int gsr_pgd_l2_update(float* param, const float* grad, const float* orig, int64_t rows, int32_t width, float alpha,
                      float epsilon, double* scratch8, void* stream);
/* Image hand-off to the detector: uint8 HWC = byte(clamp(x,0,1)*255) of a float CHW render (attack.py:438-443,
 * :538-543); y = clip(x*255 + 0.5, 0, 255) and/or dL/dx = 255 dL/dy inside the clip range
 * (detectors/detectron2_detector.py:78).  In gsr_detector_scale pass y = NULL or dL_dx = NULL to skip that half. */
int gsr_image_to_u8_hwc(const float* chw, uint8_t* hwc, int32_t W, int32_t H, void* stream);
int gsr_detector_scale(const float* x, float* y, const float* dL_dy, float* dL_dx, int64_t n, void* stream);

/* ---- stage-level entry points (parity tests and per-stage roofline timing) ---- */

/* Stable LSD radix sort of (u64 key, u32 value) pairs on bits [0, end_bit); replaces
 * cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:303-308). temp >= gsr_sort_temp_bytes(n). */
size_t gsr_sort_temp_bytes(int64_t n);
int gsr_sort_pairs_u64(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                       int64_t n, int32_t end_bit, void* temp, size_t temp_bytes, void* stream);
/* The same sort on (u32 key, u32 value) pairs -- the form the forward pass uses (depth bits of P Gaussians, then tile
 * ids of L pairs).  The pair count may live on the device: n_dev (device u32, or NULL) holds it, n_cap bounds it and
 * sizes every buffer (keys_in / vals_in / keys_out / vals_out hold n_cap elements; the first min(*n_dev, n_cap) are
 * sorted).  keys_in == keys_out (and vals) is allowed when the pass count is even.  temp >= gsr_sort_u32_temp_bytes. */
size_t gsr_sort_u32_temp_bytes(int64_t n_cap);
int gsr_sort_pairs_u32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                       int64_t n_cap, const uint32_t* n_dev, int32_t end_bit, void* temp, size_t temp_bytes,
                       void* stream);

/* The same (u32 key, u32 value) sort for callers that cannot allocate device memory or hand over a temp buffer (code
 * under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The temp block lives in a file-scope __device__ array inside the library
 * (csrc/static_scratch.cu): GSR_STATIC_SORT_SCRATCH_BYTES on every GPU that loads the library, held until the process
 * exits.  n > GSR_STATIC_SORT_MAX_PAIRS returns GSR_ERR_CAPACITY.  Calls nothing that allocates.  All calls on a
 * device share the array (each call clears the part it needs on `stream` first): calls that overlap on the GPU --
 * different streams without an ordering between them -- race; order them or use gsr_sort_pairs_u32.  keys_in ==
 * keys_out (and vals) is allowed when ceil(end_bit / 8) is even.  Same result as gsr_sort_pairs_u32, bit for bit. */
#define GSR_STATIC_SORT_MAX_PAIRS (1 << 20)
#define GSR_STATIC_SORT_SCRATCH_BYTES (21 * 512 * 1024) /* 10.5 MiB >= gsr_sort_u32_temp_bytes(GSR_STATIC_SORT_MAX_PAIRS) */
int gsr_sort_pairs_u32_static(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                              int64_t n, int32_t end_bit, void* stream);

/* Copies internal arrays of a finished forward out for inspection. name is one of: "depths",
 * "means2D", "cov3D", "conic_opacity", "rgb", "clamped", "tiles_touched", "point_offsets", "rects",
 * "perm", "offsets_sorted", "final_T", "n_contrib", "ranges", "tiles_sorted", "keys_sorted".
 * "keys_sorted" is the reference's sorted 64-bit key list (tile << 32 | depth bits,
 * rasterizer_impl.cu:303-308), rebuilt from the sorted tile ids, point_list and depths (needs
 * scratch, point_list and cap >= 8 L).  Returns the byte count (copying min(count, cap) bytes
 * into dst, a device pointer), <0 on error. */
int64_t gsr_debug_fetch(const char* name, int32_t P, int32_t W, int32_t H, int64_t L, const void* geom,
                        const void* image, const void* scratch, const uint32_t* point_list, void* dst,
                        int64_t cap, void* stream);

/* ---- measurement hooks (bench.py): per-stage CUDA-event timing on the launching stream and a
 *      count of kernel launches.  Stage order: preprocess_fwd, duplicate_keys, sort, tile_ranges,
 *      blend_fwd, blend_bwd, preprocess_bwd (GSR_STAGE_COUNT entries).  gsr_profile_read
 *      synchronises on the recorded events, sums milliseconds / calls per stage since the last
 *      read and clears the records. ---- */
#define GSR_STAGE_COUNT 7
int gsr_profile_enable(int on);
int gsr_profile_read(float* ms_host, int32_t* calls_host);
int64_t gsr_launch_count(int reset);
/* 16 counters of the blend kernels (8 forward, 8 backward), non-zero only in a `make STATS=1` build:
 * cull rounds, entries tested, entries kept, warp evaluations with a hit, contributing pairs, active lanes. */
int gsr_debug_counters(uint64_t* out_host, int reset);
/* Cycle counts per phase of the sort's scatter kernel (zeros unless built with -DGSR_SORT_PROFILE):
 * out_host[8] = tiles, load, rank, count+scan, look-back, stage, write-out, -. */
int gsr_debug_sort_profile(uint64_t* out_host, int reset);

/* Host helper: number of key bits the tile id needs (getHigherMsb, rasterizer_impl.cu:35-50). */
uint32_t gsr_higher_msb(uint32_t n);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H */
