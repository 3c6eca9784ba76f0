// blend_fwd.cu -- per-tile front-to-back alpha blending.
//
// Replaces renderCUDA<3> (DGR/cuda_rasterizer/forward.cu:261-374).  One 256-thread block per
// 16x16 tile, one thread per pixel; the tile's depth-sorted list is consumed in batches of 256
// records staged in shared memory.  Differences from the reference that do not change results:
//   * the per-Gaussian record (xy, conic, opacity, RGB) is staged once per batch as three vector
//     words, so the inner loop reads only shared memory (the reference re-reads RGB from global
//     memory for every contributing pixel, forward.cu:355);
//   * each warp owns an 8x4 pixel rectangle and first tests 32 staged entries at a time, one per
//     lane, against that rectangle (rect_may_contribute, conservative); only entries that can reach
//     alpha >= 1/255 somewhere in the rectangle are then evaluated per pixel, in list order;
//   * a per-Gaussian power threshold (computed in preprocess) rejects pairs whose alpha cannot
//     reach 1/255 before the exponential is evaluated; pairs that pass go through exactly the
//     reference's sequence (same expf, same roundings), so T, n_contrib and the colour are
//     bit-identical.
// Bound: FP32 issue + MUFU.EX2 (nothing here is bandwidth- or tensor-shaped).
#include "blend_common.cuh"

namespace gsr {
namespace {

#ifndef GSR_FWD_MINBLOCKS
#define GSR_FWD_MINBLOCKS 6
#endif
#ifndef GSR_FWD_OBJ_MINBLOCKS
#define GSR_FWD_OBJ_MINBLOCKS 4  // 64 registers; 3 blocks (79 regs) 0.589 ms, 5 blocks (48 regs, spills) 0.561 ms, 4: 0.524 ms
#endif
template <int K>
__global__ void __launch_bounds__(TILE_PIX, (K > 0 ? GSR_FWD_OBJ_MINBLOCKS : GSR_FWD_MINBLOCKS))
blend_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                 const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                 const float4* __restrict__ rgb_thr, const float* __restrict__ obj_feats,
                 const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                 float* __restrict__ out_color, float* __restrict__ out_objects) {
  __shared__ BlendStage S;
  // object-identity features of the staged entries, four 16-byte words per entry (K = 16): staged with the batch, so
  // the blending loop reads them as broadcast shared loads instead of 16 global loads per contributing pair
  __shared__ float4 s_obj[K > 0 ? K / 4 : 1][K > 0 ? BLEND_BATCH : 1];
  __shared__ uint8_t s_q[TILE_PIX / 32][32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t wx0 = blockIdx.x * TILE_X + (warp & 1) * WARP_W, wy0 = blockIdx.y * TILE_Y + (warp >> 1) * WARP_H;
  const uint32_t px = wx0 + (lane & 7), py = wy0 + (lane >> 3);
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float npxf = -(float)px, npyf = -(float)py, wx0f = (float)wx0, wy0f = (float)wy0;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  int todo = (int)(range.y - range.x);

  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
  // object channels as eight packed pairs: sm_100a's FFMA2 (fma.rn.f32x2) does two channels per issue slot
  float O[K > 0 ? K : 1];
#pragma unroll
  for (int k = 0; k < (K > 0 ? K : 1); k++) O[k] = 0.f;
  uint32_t last_contributor = 0;

  const float tx0f = (float)(blockIdx.x * TILE_X), ty0f = (float)(blockIdx.y * TILE_Y);
  for (uint32_t base = range.x; base < range.y; base += BLEND_BATCH, todo -= BLEND_BATCH) {
    if (__syncthreads_count(done) == TILE_PIX) break;
    if (base + tid < range.y) {
      const uint32_t g = point_list[base + tid];
      stage_entry(S, tid, g, means2D, conic_opacity, rgb_thr, tx0f, ty0f);
      if (K > 0) {
        const float4* f = reinterpret_cast<const float4*>(obj_feats + (size_t)g * K);
#pragma unroll
        for (int q = 0; q < K / 4; q++) s_obj[q][tid] = __ldg(f + q);
      }
    }
    __syncthreads();
    const int nb = min(BLEND_BATCH, todo);
    const uint32_t first = base - range.x;
    for (int r = 0; r < nb; r += 32) {
      if (__all_sync(0xffffffffu, done)) break;
      const int jl = r + lane;
      bool keep = false;
      if (jl < nb) keep = rect_may_contribute(S.a[jl], S.b[jl], wx0f, wy0f);
      const uint32_t mask = __ballot_sync(0xffffffffu, keep);
      GSR_STAT(0, 1); GSR_STAT(1, min(32, nb - r)); GSR_STAT(2, __popc(mask));
      // compact the kept slots into this warp's queue (list order is preserved)
      if (keep) s_q[warp][__popc(mask & lt_mask)] = (uint8_t)jl;
      __syncwarp();
      const int nk = __popc(mask);
      for (int i = 0; i < nk; i++) {
        const int j = s_q[warp][i];
#ifdef GSR_STATS
        { const uint32_t act = __ballot_sync(__activemask(), !done); GSR_STAT(5, __popc(act)); }
#endif
        const float4 a = S.a[j];
        const float power = gauss_power_packed(a, S.b[j].x, npyf, npxf);
        // Straight-line evaluation: every lane computes alpha, and the reference's three `continue`s
        // (forward.cu:336-349) become predicates on the accumulation -- same values, same decisions, no divergent
        // region in the loop (the per-Gaussian threshold pre-test is left to the rectangle cull).
        const float4 c = S.c[j];
        float wobj = 0.f;  // alpha * T of this pair when it contributes (object channels only)
        const bool p1 = !done && !(power > 0.0f);
        const float alpha = fminf(0.99f, fpx::mul(c.w, expf(power)));
        const bool p2 = p1 && !(alpha < 1.0f / 255.0f);
        const float test_T = fpx::mul(T, fpx::sub(1.0f, alpha));
        const bool stop = p2 && test_T < 0.0001f;
        done = done || stop;
        if (p2 && !stop) {
          float m0, m1;
          fpx::mul2_bcast(m0, m1, c.x, c.y, alpha);  // (alpha r, alpha g), then (C0, C1) += T * (.,.): FMUL2 + FFMA2
          fpx::fma2_bcast(C0, C1, m0, m1, T);
          C2 = fpx::fma_(T, fpx::mul(alpha, c.z), C2);
          if (K > 0) wobj = T * alpha;
          T = test_T;
          last_contributor = first + j + 1;
#ifdef GSR_STATS
          atomicAdd(&g_blend_stats[4], 1ull);
#endif
        }
        if (K > 0) {  // every lane, wobj = 0 where the pair does not contribute (skipping entries without any
                      // contributing lane measured slower: 0.555 vs 0.524 ms on C2, most kept entries contribute)
          // O_k += f_k * alpha * T (Gaussian-Grouping's extra channels: same weights as the colour, no background)
#pragma unroll
          for (int q = 0; q < K / 4; q++) {
            const float4 f = s_obj[q][j];
            fpx::fma2_bcast(O[4 * q], O[4 * q + 1], f.x, f.y, wobj);
            fpx::fma2_bcast(O[4 * q + 2], O[4 * q + 3], f.z, f.w, wobj);
          }
        }
      }
      __syncwarp();  // the queue is rewritten next round: votes and shuffles are not memory barriers
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

int blend_fwd_stats(uint64_t* out_host, int reset) { return read_blend_stats(out_host, reset); }

int launch_blend_fwd(const gsr_inputs& in, const Geom& g, const Image& im, const uint32_t* point_list,
                     float* out_color, float* out_objects, cudaStream_t st) {
  const dim3 grid((in.W + TILE_X - 1) / TILE_X, (in.H + TILE_Y - 1) / TILE_Y), block(TILE_PIX);
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
