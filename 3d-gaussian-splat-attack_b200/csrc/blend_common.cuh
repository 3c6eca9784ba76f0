// blend_common.cuh -- pieces shared by the forward and backward blending kernels.
#pragma once
#include "common.cuh"
#include "fp_exact.cuh"

namespace gsr {

constexpr int BLEND_BATCH = TILE_PIX;  // list entries staged per block iteration
// A 256-thread block owns one 16x16 tile; each warp owns an 8x4 pixel rectangle of it
// (warp w: columns (w&1)*8.., rows (w>>1)*4..; lane l: column l&7, row l>>3).
constexpr int WARP_W = 8, WARP_H = 4;

// Optional instrumentation (make STATS=1): [0] cull rounds, [1] entries tested, [2] entries kept,
// [3] (warp, entry) evaluations with >= 1 contributing lane, [4] contributing (pixel, entry) pairs,
// [5] lanes active at evaluation.  One copy per translation unit (forward / backward).
static __device__ unsigned long long g_blend_stats[8];
inline int read_blend_stats(uint64_t* out_host, int reset) {
  if (cudaMemcpyFromSymbol(out_host, g_blend_stats, 8 * sizeof(uint64_t)) != cudaSuccess) return GSR_ERR_CUDA;
  if (reset) {
    const unsigned long long z[8] = {0};
    if (cudaMemcpyToSymbol(g_blend_stats, z, sizeof(z)) != cudaSuccess) return GSR_ERR_CUDA;
  }
  return GSR_OK;
}
#ifdef GSR_STATS
#define GSR_STAT(i, n) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_blend_stats[i], (unsigned long long)(n)); } while (0)
#else
#define GSR_STAT(i, n) do { } while (0)
#endif

// power = -0.5 (A dx^2 + C dy^2) - B dx dy with the roundings of the reference build
// (forward.cu:335, backward.cu:490); see fp_exact.cuh.
__device__ __forceinline__ float gauss_power(float dx, float dy, float A, float B, float C) {
  using namespace fpx;
  const float q = fma_(dx, mul(dx, A), mul(dy, mul(dy, C)));
  return sub(mul(q, -0.5f), mul(dy, mul(dx, B)));
}

// power from a staged record with sm_100a's packed-pair FP32 instructions: 1 FADD2 + 2 FMUL2 + FMUL + 2 FFMA instead
// of 2 FADD + 5 FMUL + 2 FFMA, every intermediate rounded exactly as in gauss_power (same bits).  a = BlendStage::a
// (y, x, C, B), A = BlendStage::b.x, (npy, npx) = minus the pixel centre.
__device__ __forceinline__ float gauss_power_packed(const float4 a, float A, float npy, float npx) {
  using namespace fpx;
  float dy, dx, t2, t4, t3, u;
  add2(dy, dx, a.x, a.y, npy, npx);  // (y - py, x - px)
  mul2(t2, t4, dy, dx, a.z, a.w);    // (dy C, dx B)
  mul2_bcast(t3, u, t2, t4, dy);     // (dy (dy C), dy (dx B))
  const float q = fma_(dx, mul(dx, A), t3);
  return sub(mul(q, -0.5f), u);
}

// One staged list entry (shared memory, SoA of vector words).  The order inside a word puts the operands of the
// packed-pair instructions in even-aligned register pairs after a 16-byte shared load.
struct BlendStage {
  float4 a[BLEND_BATCH];  // y, x, conic C, conic B
  float4 b[BLEND_BATCH];  // conic A, cull threshold (see stage_entry), -B/C, -B/A   (a and b are all the cull reads)
  float4 c[BLEND_BATCH];  // r, g, b, opacity
};

// (tx0, ty0) = the tile's first pixel.  The rounding margin of the rectangle cull and its "not safely positive
// definite" escape depend on the entry and the tile only, so the staging thread folds them into the threshold once
// for all eight warps: thr_cull = thr - (2e-6 * S_tile + 1e-4), or -inf for a degenerate conic (never culled);
// S_tile bounds the term magnitudes over the whole tile, hence over every warp rectangle.
__device__ __forceinline__ void stage_record(BlendStage& S, int slot, const float2 xy, const float4 co, const float4 ct,
                                             float tx0, float ty0) {
  const float ux = tx0 - xy.x, vy = ty0 - xy.y;
  const float um = fmaxf(fabsf(ux), fabsf(ux + (float)(TILE_X - 1))), vm = fmaxf(fabsf(vy), fabsf(vy + (float)(TILE_Y - 1)));
  const float s_tile = 0.5f * (co.x * um * um + co.z * vm * vm) + fabsf(co.y) * um * vm;
  const bool degenerate = !(co.y * co.y < 0.99999f * (co.x * co.z));
  const float thr_cull = degenerate ? -INFINITY : ct.w - (2e-6f * s_tile + 1e-4f);
  S.a[slot] = make_float4(xy.y, xy.x, co.z, co.y);
  S.b[slot] = make_float4(co.x, thr_cull, __fdividef(-co.y, co.z), __fdividef(-co.y, co.x));
  S.c[slot] = make_float4(ct.x, ct.y, ct.z, co.w);
}
__device__ __forceinline__ void stage_entry(BlendStage& S, int slot, uint32_t g, const float2* __restrict__ means2D,
                                            const float4* __restrict__ conic_opacity,
                                            const float4* __restrict__ rgb_thr, float tx0, float ty0) {
  stage_record(S, slot, means2D[g], conic_opacity[g], rgb_thr[g], tx0, ty0);
}

// Conservative test of one list entry against a warp's pixel rectangle [x0, x0+WARP_W-1] x [y0, y0+WARP_H-1]:
// returns false only if NO pixel of the rectangle can pass the per-pixel test `power >= threshold`.
// f(u,v) = -0.5 (A u^2 + C v^2) - B u v is concave (the conic is positive definite), so its maximum over the
// rectangle is 0 when the centre is inside and otherwise lies on an edge visible from the centre; the two
// candidates below are the exact edge maxima (clamped 1-D maximisers).  The margin covers the rounding of
// both this evaluation and the per-pixel one (each <= ~5e-7 * S, S = sum of the magnitudes of the terms,
// bounded over the rectangle by Smax), so a pair that passes per pixel is never culled here.  NaNs compare
// false and are kept.
__device__ __forceinline__ bool rect_may_contribute(const float4 a, const float4 b, float x0, float y0) {
  const float A = b.x, B = a.w, C = a.z, thr = b.y;
  const float ux0 = x0 - a.y, ux1 = ux0 + (float)(WARP_W - 1);
  const float vy0 = y0 - a.x, vy1 = vy0 + (float)(WARP_H - 1);
  const float u1 = fminf(fmaxf(0.f, ux0), ux1);
  const float v1 = fminf(fmaxf(b.z * u1, vy0), vy1);
  const float v2 = fminf(fmaxf(0.f, vy0), vy1);
  const float u2 = fminf(fmaxf(b.w * v2, ux0), ux1);
  const float f1 = -0.5f * (A * u1 * u1 + C * v1 * v1) - B * u1 * v1;
  const float f2 = -0.5f * (A * u2 * u2 + C * v2 * v2) - B * u2 * v2;
  return !(fmaxf(f1, f2) < thr);  // thr carries the margin and the degenerate-conic escape (stage_entry)
}

}  // namespace gsr
