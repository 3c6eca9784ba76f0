// blend_bwd.cu -- per-tile back-to-front blending gradients.
//
// Replaces renderCUDA<3> backward (DGR/cuda_rasterizer/backward.cu:399-557).  Same per-pixel
// recurrence (recompute alpha, rebuild T by division, running accum_rec), but:
//   * the walk starts at the tile's largest n_contrib instead of the end of the tile list, and a
//     warp skips entries beyond its own largest n_contrib;
//   * the 9 per-Gaussian partials (3 colour, 2 mean2D, 3 conic, 1 opacity) are reduced across the
//     warp with shuffles before one RED per value per warp (the reference issues 9 atomics per
//     contributing pixel, backward.cu:523-554), and only when some lane contributes;
//   * partials land in one packed [P][12] accumulator, so a Gaussian's 9 words share two sectors.
// The contribution test uses exactly the forward arithmetic, so the set of contributing pairs is
// the forward's.  Sum order differs from the reference's atomics -> gradients are compared with
// a tolerance.  Bound: FP32 issue + shuffle + L2 reduction throughput.
#include "common.cuh"
#include "fp_exact.cuh"

namespace gsr {
namespace {

constexpr int BATCH = TILE_PIX;

__device__ __forceinline__ float gauss_power(float dx, float dy, float A, float B, float C) {
  using namespace fpx;
  const float q = fma_(dx, mul(dx, A), mul(dy, mul(dy, C)));
  return sub(mul(q, -0.5f), mul(dy, mul(dx, B)));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int K>
__global__ void __launch_bounds__(TILE_PIX)
blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                 const float* __restrict__ bg, const float2* __restrict__ means2D,
                 const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb_thr,
                 const float* __restrict__ obj_feats, const float* __restrict__ final_T,
                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                 const float* __restrict__ dL_dobjpix, float* __restrict__ accum, float* __restrict__ dL_dobj) {
  __shared__ float4 s_a[BATCH];  // x, y, conic.x, conic.y
  __shared__ float4 s_b[BATCH];  // conic.z, opacity, threshold, r
  __shared__ float2 s_c[BATCH];  // g, b
  __shared__ uint32_t s_id[BATCH];
  __shared__ uint32_t s_max;

  const int tid = threadIdx.y * TILE_X + threadIdx.x, lane = tid & 31;
  const uint32_t px = blockIdx.x * TILE_X + threadIdx.x, py = blockIdx.y * TILE_Y + threadIdx.y;
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float pxf = (float)px, pyf = (float)py;
  const size_t HW = (size_t)W * H, pid = (size_t)py * W + px;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];

  const float T_final = inside ? final_T[pid] : 0.f;
  const uint32_t my_last = inside ? n_contrib[pid] : 0u;
  float T = T_final;
  float dpix[3] = {0.f, 0.f, 0.f};
  if (inside) { dpix[0] = dL_dpix[pid]; dpix[1] = dL_dpix[HW + pid]; dpix[2] = dL_dpix[2 * HW + pid]; }
  float dpo[K > 0 ? K : 1], acc_o[K > 0 ? K : 1], last_o[K > 0 ? K : 1];
#pragma unroll
  for (int k = 0; k < (K > 0 ? K : 1); k++) {
    dpo[k] = (K > 0 && inside) ? dL_dobjpix[k * HW + pid] : 0.f;
    acc_o[k] = 0.f; last_o[k] = 0.f;
  }
  const float bg_dot = bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2];
  float acc_r[3] = {0.f, 0.f, 0.f}, last_c[3] = {0.f, 0.f, 0.f}, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, my_last);
  if (tid == 0) s_max = 0;
  __syncthreads();
  if (lane == 0) atomicMax(&s_max, warp_last);
  __syncthreads();
  const uint32_t n_live = min(s_max, range.y - range.x);  // entries [0, n_live) can matter to this tile

  // batches walk the list back to front: batch b covers list positions n_live-1-b*BATCH-j
  for (uint32_t done_cnt = 0; done_cnt < n_live; done_cnt += BATCH) {
    __syncthreads();
    if (done_cnt + tid < n_live) {
      const uint32_t pos = n_live - 1 - done_cnt - tid;
      const uint32_t g = point_list[range.x + pos];
      const float2 xy = means2D[g];
      const float4 co = conic_opacity[g];
      const float4 ct = rgb_thr[g];
      s_a[tid] = make_float4(xy.x, xy.y, co.x, co.y);
      s_b[tid] = make_float4(co.z, co.w, ct.w, ct.x);
      s_c[tid] = make_float2(ct.y, ct.z);
      s_id[tid] = g;
    }
    __syncthreads();
    const int nb = (int)min((uint32_t)BATCH, n_live - done_cnt);
    for (int j = 0; j < nb; j++) {
      const uint32_t pos = n_live - 1 - done_cnt - j;  // 0-based position in the tile list
      if (pos >= warp_last) continue;                   // warp-uniform
      const float4 a = s_a[j];
      const float4 b = s_b[j];
      const float dx = fpx::sub(a.x, pxf), dy = fpx::sub(a.y, pyf);
      const float power = gauss_power(dx, dy, a.z, a.w, b.x);
      bool hit = pos < my_last && !(power > 0.0f || power < b.z);
      float G = 0.f, alpha = 0.f;
      if (hit) {
        G = expf(power);
        alpha = fminf(0.99f, fpx::mul(b.y, G));
        hit = !(alpha < 1.0f / 255.0f);
      }
      if (!__any_sync(0xffffffffu, hit)) continue;
      float v[9 + (K > 0 ? K : 0)];
#pragma unroll
      for (int k = 0; k < 9 + (K > 0 ? K : 0); k++) v[k] = 0.f;
      if (hit) {
        const float2 c = s_c[j];
        const float col[3] = {b.w, c.x, c.y};
        const float inv = __frcp_rn(1.f - alpha);
        T = T * inv;
        const float w = alpha * T;
        float dL_dalpha = 0.f;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          acc_r[ch] = last_alpha * last_c[ch] + (1.f - last_alpha) * acc_r[ch];
          last_c[ch] = col[ch];
          dL_dalpha += (col[ch] - acc_r[ch]) * dpix[ch];
          v[ch] = w * dpix[ch];
        }
        if (K > 0) {
          const float* f = obj_feats + (size_t)s_id[j] * K;
#pragma unroll
          for (int k = 0; k < K; k++) {
            const float o = __ldg(f + k);
            acc_o[k] = last_alpha * last_o[k] + (1.f - last_alpha) * acc_o[k];
            last_o[k] = o;
            dL_dalpha += (o - acc_o[k]) * dpo[k];
            v[9 + k] = w * dpo[k];
          }
        }
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final * inv) * bg_dot;
        const float dL_dG = b.y * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        const float dG_ddelx = -gdx * a.z - gdy * a.w;
        const float dG_ddely = -gdy * b.x - gdx * a.w;
        v[3] = dL_dG * dG_ddelx * ddelx_dx;
        v[4] = dL_dG * dG_ddely * ddely_dy;
        v[5] = -0.5f * gdx * dx * dL_dG;
        v[6] = -0.5f * gdx * dy * dL_dG;
        v[7] = -0.5f * gdy * dy * dL_dG;
        v[8] = G * dL_dalpha;
      }
#pragma unroll
      for (int k = 0; k < 9 + (K > 0 ? K : 0); k++) v[k] = warp_sum(v[k]);
      const uint32_t g = s_id[j];
      if (lane < 9) {
        // lane k adds value k (select without dynamic register indexing)
        float out = v[0];
#pragma unroll
        for (int k = 1; k < 9; k++) out = lane == k ? v[k] : out;
        atomicAdd(accum + (size_t)g * 12 + lane, out);
      }
      if (K > 0 && lane >= 16) {
        float out = v[9];
#pragma unroll
        for (int k = 1; k < K; k++) out = (lane - 16) == k ? v[9 + k] : out;
        atomicAdd(dL_dobj + (size_t)g * K + (lane - 16), out);
      }
    }
  }
}

}  // namespace

int launch_blend_bwd(const gsr_inputs& in, const Geom& g, const Image& im, const uint32_t* point_list,
                     const float* dL_dpix, const float* dL_dobjpix, float* accum, float* dL_dobj, cudaStream_t st) {
  const dim3 grid((in.W + TILE_X - 1) / TILE_X, (in.H + TILE_Y - 1) / TILE_Y), block(TILE_X, TILE_Y);
  GSR_CUDA_CHECK(cudaMemsetAsync(accum, 0, (size_t)in.P * 12 * sizeof(float), st));
  if (in.K > 0) GSR_CUDA_CHECK(cudaMemsetAsync(dL_dobj, 0, (size_t)in.P * in.K * sizeof(float), st));
#define GSR_BLEND_BWD(KK)                                                                                      \
  blend_bwd_kernel<KK><<<grid, block, 0, st>>>(im.ranges, point_list, in.W, in.H, in.bg, g.means2D,            \
                                               g.conic_opacity, g.rgb_thr, in.obj_feats, im.final_T,          \
                                               im.n_contrib, dL_dpix, dL_dobjpix, accum, dL_dobj)
  if (in.K == 0) GSR_BLEND_BWD(0);
  else if (in.K == 16) GSR_BLEND_BWD(16);
  else { set_error("object channels K=%d unsupported (0 or 16)", in.K); return GSR_ERR_INVALID; }
#undef GSR_BLEND_BWD
  GSR_LAUNCH_CHECK("blend_bwd_kernel");
  return GSR_OK;
}

}  // namespace gsr
