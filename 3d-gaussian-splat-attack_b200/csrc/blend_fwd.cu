// blend_fwd.cu -- per-tile front-to-back alpha blending.
//
// Replaces renderCUDA<3> (DGR/cuda_rasterizer/forward.cu:261-374).  One 256-thread block per
// 16x16 tile, one thread per pixel; the tile's depth-sorted list is consumed in batches of 256
// records staged in shared memory.  Differences from the reference that do not change results:
//   * the per-Gaussian record (xy, conic, opacity, RGB) is staged once per batch as three vector
//     words, so the inner loop reads only shared memory (the reference re-reads RGB from global
//     memory for every contributing pixel, forward.cu:355);
//   * a per-Gaussian power threshold (computed in preprocess) rejects pairs whose alpha cannot
//     reach 1/255 before the exponential is evaluated; pairs that pass go through exactly the
//     reference's sequence (same expf, same roundings), so T, n_contrib and the colour are
//     bit-identical.
// Bound: FP32 issue + MUFU.EX2 (nothing here is bandwidth- or tensor-shaped).
#include "common.cuh"
#include "fp_exact.cuh"

namespace gsr {
namespace {

constexpr int BATCH = TILE_PIX;

// power = -0.5 (A dx^2 + C dy^2) - B dx dy, rounded step by step as the reference build does (see fp_exact.cuh)
__device__ __forceinline__ float gauss_power(float dx, float dy, float A, float B, float C) {
  using namespace fpx;
  const float q = fma_(dx, mul(dx, A), mul(dy, mul(dy, C)));
  return sub(mul(q, -0.5f), mul(dy, mul(dx, B)));
}

template <int K>
__global__ void __launch_bounds__(TILE_PIX)
blend_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                 const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                 const float4* __restrict__ rgb_thr, const float* __restrict__ obj_feats,
                 const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                 float* __restrict__ out_color, float* __restrict__ out_objects) {
  __shared__ float4 s_a[BATCH];  // x, y, conic.x, conic.y
  __shared__ float4 s_b[BATCH];  // conic.z, opacity, threshold, r
  __shared__ float2 s_c[BATCH];  // g, b
  __shared__ uint32_t s_id[K > 0 ? BATCH : 1];

  const int tid = threadIdx.y * TILE_X + threadIdx.x;
  const uint32_t px = blockIdx.x * TILE_X + threadIdx.x, py = blockIdx.y * TILE_Y + threadIdx.y;
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float pxf = (float)px, pyf = (float)py;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  int todo = (int)(range.y - range.x);

  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
  float O[K > 0 ? K : 1];
#pragma unroll
  for (int k = 0; k < (K > 0 ? K : 1); k++) O[k] = 0.f;
  uint32_t last_contributor = 0;

  for (uint32_t base = range.x; base < range.y; base += BATCH, todo -= BATCH) {
    if (__syncthreads_count(done) == TILE_PIX) break;
    if (base + tid < range.y) {
      const uint32_t g = point_list[base + tid];
      const float2 xy = means2D[g];
      const float4 co = conic_opacity[g];
      const float4 ct = rgb_thr[g];
      s_a[tid] = make_float4(xy.x, xy.y, co.x, co.y);
      s_b[tid] = make_float4(co.z, co.w, ct.w, ct.x);
      s_c[tid] = make_float2(ct.y, ct.z);
      if (K > 0) s_id[tid] = g;
    }
    __syncthreads();
    const int nb = min(BATCH, todo);
    const uint32_t first = base - range.x;
    for (int j = 0; !done && j < nb; j++) {
      const float4 a = s_a[j];
      const float4 b = s_b[j];
      const float dx = fpx::sub(a.x, pxf), dy = fpx::sub(a.y, pyf);
      const float power = gauss_power(dx, dy, a.z, a.w, b.x);
      if (power > 0.0f || power < b.z) continue;
      const float alpha = fminf(0.99f, fpx::mul(b.y, expf(power)));
      if (alpha < 1.0f / 255.0f) continue;
      const float test_T = fpx::mul(T, fpx::sub(1.0f, alpha));
      if (test_T < 0.0001f) { done = true; continue; }
      const float2 c = s_c[j];
      C0 = fpx::fma_(T, fpx::mul(alpha, b.w), C0);
      C1 = fpx::fma_(T, fpx::mul(alpha, c.x), C1);
      C2 = fpx::fma_(T, fpx::mul(alpha, c.y), C2);
      if (K > 0) {
        const float* f = obj_feats + (size_t)s_id[j] * K;
#pragma unroll
        for (int k = 0; k < K; k++) O[k] = fpx::fma_(T, fpx::mul(alpha, __ldg(f + k)), O[k]);
      }
      T = test_T;
      last_contributor = first + j + 1;
    }
  }
  if (inside) {
    const size_t HW = (size_t)W * H, pid = (size_t)py * W + px;
    final_T[pid] = T;
    n_contrib[pid] = last_contributor;
    out_color[pid] = fpx::fma_(T, bg[0], C0);
    out_color[HW + pid] = fpx::fma_(T, bg[1], C1);
    out_color[2 * HW + pid] = fpx::fma_(T, bg[2], C2);
    if (K > 0) {
#pragma unroll
      for (int k = 0; k < K; k++) out_objects[k * HW + pid] = O[k];
    }
  }
}

}  // namespace

int launch_blend_fwd(const gsr_inputs& in, const Geom& g, const Image& im, const uint32_t* point_list,
                     float* out_color, float* out_objects, cudaStream_t st) {
  const dim3 grid((in.W + TILE_X - 1) / TILE_X, (in.H + TILE_Y - 1) / TILE_Y), block(TILE_X, TILE_Y);
#define GSR_BLEND_FWD(KK)                                                                                     \
  blend_fwd_kernel<KK><<<grid, block, 0, st>>>(im.ranges, point_list, in.W, in.H, g.means2D, g.conic_opacity, \
                                               g.rgb_thr, in.obj_feats, in.bg, im.final_T, im.n_contrib,      \
                                               out_color, out_objects)
  if (in.K == 0) GSR_BLEND_FWD(0);
  else if (in.K == 16) GSR_BLEND_FWD(16);
  else { set_error("object channels K=%d unsupported (0 or 16)", in.K); return GSR_ERR_INVALID; }
#undef GSR_BLEND_FWD
  GSR_LAUNCH_CHECK("blend_fwd_kernel");
  return GSR_OK;
}

}  // namespace gsr
