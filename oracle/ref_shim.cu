// oracle/ref_shim.cu -- TEST / BASELINE INFRASTRUCTURE ONLY.
//
// A thin extern "C" wrapper (our own code) around the UNMODIFIED reference rasterizer
//   CudaRasterizer::Rasterizer::{forward,backward,markVisible}
//   (DGR/cuda_rasterizer/rasterizer.h:20-85, rasterizer_impl.cu), DGR =
//   /root/reference/submodules/gaussian-splatting/submodules/diff-gaussian-rasterization,
// whose sources oracle/Makefile compiles for sm_100a where they lie (nothing is copied into this
// repo) into oracle/_ref/libdgr_ref.so.  It stands in for the reference's torch glue
// (DGR/rasterize_points.cu), which needs ~8 min of torch headers to compile and adds nothing to
// the arithmetic.  Used on the GPU box for (i) bit-exact parity of keys / ranges / n_contrib /
// images, (ii) generating tests/golden fixtures that pin oracle/cpu_ref.cpp, (iii) timing the
// reference CUDA on the same B200.  The product never loads it.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include "rasterizer.h"
#include "rasterizer_impl.h"

namespace {
struct Buf {
  char* p = nullptr;
  size_t cap = 0, used = 0;
  char* get(size_t n) {
    if (n > cap) {
      if (p) cudaFree(p);
      cap = n + n / 4 + 256;
      if (cudaMalloc(&p, cap) != cudaSuccess) { p = nullptr; cap = 0; }
    }
    used = n;
    return p;
  }
};
Buf g_geom, g_bin, g_img;
int g_P = 0, g_R = 0, g_W = 0, g_H = 0;
}  // namespace

extern "C" {

// Returns num_rendered (or -1 on a CUDA error).  All pointers are device pointers.
int ref_forward(int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const float* view, const float* proj,
                const float* campos, float tanfovx, float tanfovy, int prefiltered, float* out_color, int* radii,
                int debug) {
  std::function<char*(size_t)> fg = [](size_t n) { return g_geom.get(n); };
  std::function<char*(size_t)> fb = [](size_t n) { return g_bin.get(n); };
  std::function<char*(size_t)> fi = [](size_t n) { return g_img.get(n); };
  int R = 0;
  try {
    R = CudaRasterizer::Rasterizer::forward(fg, fb, fi, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities,
                                            scales, scale_modifier, rotations, cov3D_precomp, view, proj, campos,
                                            tanfovx, tanfovy, prefiltered != 0, out_color, radii, debug != 0);
  } catch (const std::exception& e) {
    fprintf(stderr, "ref_forward: %s\n", e.what());
    return -1;
  }
  g_P = P; g_R = R; g_W = W; g_H = H;
  return cudaGetLastError() == cudaSuccess ? R : -1;
}

int ref_backward(int P, int D, int M, int R, const float* bg, int W, int H, const float* means3D, const float* shs,
                 const float* colors_precomp, const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* view, const float* proj, const float* campos, float tanfovx,
                 float tanfovy, const int* radii, const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                 float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscale, float* dL_drot, int debug) {
  try {
    CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales, scale_modifier,
                                         rotations, cov3D_precomp, view, proj, campos, tanfovx, tanfovy, radii,
                                         g_geom.p, g_bin.p, g_img.p, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity,
                                         dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
  } catch (const std::exception& e) {
    fprintf(stderr, "ref_backward: %s\n", e.what());
    return -1;
  }
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int ref_mark_visible(int P, float* means3D, float* view, float* proj, bool* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// Copies an internal array of the last forward into dst (device pointer). Returns its byte size.
long long ref_fetch(const char* name, void* dst, long long cap) {
  using namespace CudaRasterizer;
  char* c = g_geom.p;
  GeometryState gs = GeometryState::fromChunk(c, g_P);
  c = g_bin.p;
  BinningState bs = BinningState::fromChunk(c, g_R);
  c = g_img.p;
  ImageState is = ImageState::fromChunk(c, (size_t)g_W * g_H);
  const size_t P = g_P, R = g_R, N = (size_t)g_W * g_H;
  const size_t tiles = (size_t)((g_W + 15) / 16) * ((g_H + 15) / 16);
  const void* src = nullptr;
  size_t bytes = 0;
#define F(nm, ptr, cnt) if (!strcmp(name, nm)) { src = ptr; bytes = (cnt); }
  F("depths", gs.depths, P * 4) F("clamped", gs.clamped, P * 3) F("means2D", gs.means2D, P * 8)
  F("cov3D", gs.cov3D, P * 24) F("conic_opacity", gs.conic_opacity, P * 16) F("rgb", gs.rgb, P * 12)
  F("point_offsets", gs.point_offsets, P * 4) F("tiles_touched", gs.tiles_touched, P * 4)
  F("keys_unsorted", bs.point_list_keys_unsorted, R * 8) F("keys_sorted", bs.point_list_keys, R * 8)
  F("vals_unsorted", bs.point_list_unsorted, R * 4) F("point_list", bs.point_list, R * 4)
  F("ranges", is.ranges, tiles * 8) F("n_contrib", is.n_contrib, N * 4) F("final_T", is.accum_alpha, N * 4)
#undef F
  if (!src) return -1;
  if (dst) cudaMemcpy(dst, src, bytes < (size_t)cap ? bytes : (size_t)cap, cudaMemcpyDeviceToDevice);
  return (long long)bytes;
}

}  // extern "C"
