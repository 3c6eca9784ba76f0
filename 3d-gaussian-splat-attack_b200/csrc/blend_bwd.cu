// blend_bwd.cu -- per-tile back-to-front blending gradients.
//
// Replaces renderCUDA<3> backward (DGR/cuda_rasterizer/backward.cu:399-557).  Same per-pixel
// recurrence (recompute alpha, rebuild T by division, running accum_rec), but:
//   * the walk starts at the tile's largest n_contrib instead of the end of the tile list, and a
//     warp (an 8x4 pixel rectangle) skips entries beyond its own largest n_contrib and entries
//     that cannot reach alpha >= 1/255 inside its rectangle (rect_may_contribute);
//   * what is accumulated per Gaussian is linear in two per-pixel weights, wc = alpha*T and
//     w = G dL/dalpha: 3 colour sums (wc dL/dpixel), sum w, and the first / second moments of w about
//     the centre (w dx, w dy, w dx^2, w dx dy, w dy^2); preprocess_bwd turns the moments into
//     dL/dmean2D, dL/dconic and dL/dopacity (same algebra as backward.cu:523-554);
//   * the sum over the pixels of a Gaussian is done WITHOUT a per-entry cross-lane reduction:
//     phase 1 (lane = pixel) runs the sequential per-pixel recurrence and writes (wc, w) of every
//     (pixel, entry) pair of a group of GRP = 8 kept entries to a per-warp shared-memory panel;
//     phase 2 (lane = entry x pixel-row) sweeps the panel, each lane accumulating the 9 sums of its
//     own entry over one 8-pixel row in registers, two shuffles per value join the four rows, and the
//     9 values go to the packed [P][12] accumulator as two 16-byte vector REDs + one scalar RED per
//     (warp, entry) (the reference issues 9 atomics per contributing pixel).
//   * the reference's three per-channel running colours (accum_rec, backward.cu:499-521) collapse into one
//     running scalar: dL/dalpha_j = T_j (c_j . dL/dpixel) - behind_j / (1 - alpha_j), where behind_j =
//     sum_{k behind j} (c_k . dL/dpixel) alpha_k T_k + T_final (bg . dL/dpixel)  (same algebra, 11 instead of
//     ~24 instructions per pair);
//   * exp() is MUFU.EX2 of the scaled argument; the decision "did this pair contribute" is still the
//     forward's, bit for bit (an alpha next to the 1/255 threshold is redone with the forward's expf).
// Sum order and roundings differ from the reference's atomics -> gradients are compared with a tolerance
// (measured on C2 against the reference CUDA: the same distance as two runs of the reference from each other,
// tools/grad_ratio.py).  Bound: warp-instruction issue (82 %) and the L1TEX data pipe (69 % of its wavefront rate).
#include "blend_common.cuh"

namespace gsr {
namespace {

#ifndef GSR_BWD_GRP
#define GSR_BWD_GRP 8
#endif
#ifndef GSR_BWD_MINBLOCKS
#define GSR_BWD_MINBLOCKS 5
#endif
constexpr int GRP = GSR_BWD_GRP;  // kept entries per phase-2 sweep (8 or 16); each lane then sweeps GRP pixels
constexpr int PANEL_STRIDE = 33; // float2 row stride of the panel (bank-conflict-free in both phases)
constexpr int NWARP = TILE_PIX / 32;

struct BwdWarp {
  float2 panel[GRP][PANEL_STRIDE];  // [group slot][pixel lane] = (alpha*T, G*dL/dalpha); 0 when the pair does not contribute
  float2 pix_rg[32];                // dL/dpixel (r, g) and (b, 1) of each pixel lane as two 8-byte words: operand pairs of
  float2 pix_b1[32];                // the sweep's packed FFMA2s (one 16-byte read with four distinct addresses costs more)
};
struct BwdSmem {
  BlendStage S;
  uint32_t id[BLEND_BATCH];
  BwdWarp w[NWARP];
  uint32_t smax;
};

__device__ __forceinline__ float xchg_add(float lo, float hi, int off, bool upper) {
  const float send = upper ? lo : hi, mine = upper ? hi : lo;
  return mine + __shfl_xor_sync(0xffffffffu, send, off);
}
// Sums 8 values across the warp with 9 shuffles: the lane whose low two bits are 0 ends up holding the
// warp total of value index (lane >> 2).  Used for the object-feature channels only.
__device__ __forceinline__ float warp_sum8(const float* v, int lane) {
  const bool u16 = lane & 16, u8 = lane & 8, u4 = lane & 4;
  const float w0 = xchg_add(v[0], v[4], 16, u16), w1 = xchg_add(v[1], v[5], 16, u16);
  const float w2 = xchg_add(v[2], v[6], 16, u16), w3 = xchg_add(v[3], v[7], 16, u16);
  const float x0 = xchg_add(w0, w2, 8, u8), x1 = xchg_add(w1, w3, 8, u8);
  float y = xchg_add(x0, x1, 4, u4);
  y += __shfl_xor_sync(0xffffffffu, y, 2);
  y += __shfl_xor_sync(0xffffffffu, y, 1);
  return y;
}
__device__ __forceinline__ void red_add4(float* p, float x, float y, float z, float w) {  // 16-byte aligned p
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ float fast_exp(float x) {  // ex2.approx(x log2 e): 2 ulp, enough for gradients
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));
  return r;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// Phase 2: lane (e = lane % GRP, h = lane / GRP) sums, over pixel lanes h*GRP .. h*GRP+GRP-1, the 9 per-Gaussian
// quantities of group entry e; the 32/GRP partial sums are joined with shuffles and lanes < ng issue the REDs.
// slot_j: lane e < ng holds the batch slot of group entry e (kept in a register, not in shared memory).
__device__ __forceinline__ void flush_group(BwdSmem& M, BwdWarp& Wp, int ng, int lane, int slot_j, float wx0f,
                                            float wy0f, float* __restrict__ accum) {
  __syncwarp();
  const int e = lane & (GRP - 1), h = lane / GRP;
  float cx = 0.f, cy = 0.f;
  uint32_t g = 0;
  const int j = __shfl_sync(0xffffffffu, slot_j, e);
  if (e < ng) {
    const float2 cxy = *reinterpret_cast<const float2*>(&M.S.a[j]);
    cx = cxy.y; cy = cxy.x; g = M.id[j];  // staged as (y, x)
  }
  const float2* row = &Wp.panel[e][h * GRP];
  const float2* pix_rg = &Wp.pix_rg[h * GRP];
  const float2* pix_b1 = &Wp.pix_b1[h * GRP];
  // raw moments of w in the lane's local pixel coordinates (lx = 0..7, r = row inside the lane's strip): the
  // coordinates are compile-time constants, so each sum is one FFMA/FADD per pixel (none where the factor is 0);
  // (s0, s1) += wc (dL/dr, dL/dg) and (s2, M0) += (wc, w) (dL/db, 1) are one FFMA2 each
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, M0 = 0.f, Mx = 0.f, My = 0.f, Mxx = 0.f, Mxy = 0.f;
#pragma unroll
  for (int i = 0; i < GRP; i++) {
    const float2 pr = row[i];
    const float2 rg = pix_rg[i];
    const float2 b1 = pix_b1[i];
    const int lx = i & 7, r = i >> 3;
    fpx::fma2_bcast(s0, s1, rg.x, rg.y, pr.x);
    fpx::fma2_pair(s2, M0, pr.x, pr.y, b1.x, b1.y);
    if (lx == 1) { Mx += pr.y; Mxx += pr.y; }
    if (lx > 1) { Mx = fmaf(pr.y, (float)lx, Mx); Mxx = fmaf(pr.y, (float)(lx * lx), Mxx); }
    if (r == 1) {
      My += pr.y;  // r is 0 or 1, so sum w r^2 = sum w r
      if (lx == 1) Mxy += pr.y;
      if (lx > 1) Mxy = fmaf(pr.y, (float)lx, Mxy);
    }
  }
  // moments about the Gaussian's centre: dx = Dx - lx, dy = Dy - r
  const float Dx = cx - wx0f, Dy = cy - (wy0f + (float)(h * (GRP / 8)));
  const float a1 = fmaf(Dx, M0, -Mx), a2 = fmaf(Dy, M0, -My);          // sum w dx, sum w dy
  float m0 = M0, m1 = a1, m2 = a2;
  float mxx = fmaf(Dx, a1, fmaf(-Dx, Mx, Mxx));                          // Dx^2 M0 - 2 Dx Mx + Mxx
  float mxy = fmaf(Dy, a1, fmaf(-Dx, My, Mxy));                          // Dx Dy M0 - Dy Mx - Dx My + Mxy
  float myy = fmaf(Dy, a2, fmaf(-Dy, My, My));                           // Dy^2 M0 - 2 Dy My + Myy (Myy = My)
#pragma unroll
  for (int off = GRP; off < 32; off <<= 1) {  // join the row strips: nine shuffles, five (packed) additions per level
    const float t0 = __shfl_xor_sync(0xffffffffu, s0, off), t1 = __shfl_xor_sync(0xffffffffu, s1, off);
    const float t2 = __shfl_xor_sync(0xffffffffu, s2, off), t3 = __shfl_xor_sync(0xffffffffu, m0, off);
    const float t4 = __shfl_xor_sync(0xffffffffu, m1, off), t5 = __shfl_xor_sync(0xffffffffu, m2, off);
    const float t6 = __shfl_xor_sync(0xffffffffu, mxx, off), t7 = __shfl_xor_sync(0xffffffffu, mxy, off);
    const float t8 = __shfl_xor_sync(0xffffffffu, myy, off);
    fpx::add2(s0, s1, s0, s1, t0, t1); fpx::add2(s2, m0, s2, m0, t2, t3);
    fpx::add2(m1, m2, m1, m2, t4, t5); fpx::add2(mxx, mxy, mxx, mxy, t6, t7);
    myy += t8;
  }
  if (lane < ng) {
    float* dst = accum + (size_t)g * 12;  // 48-byte rows: two 16-byte vector reductions + one scalar
    red_add4(dst, s0, s1, s2, m1);
    red_add4(dst + 4, m2, mxx, mxy, myy);
    atomicAdd(dst + 8, m0);
  }
  __syncwarp();
}

template <int K>
__global__ void __launch_bounds__(TILE_PIX, (K > 0 ? 2 : GSR_BWD_MINBLOCKS))
blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                 const float* __restrict__ bg, const float2* __restrict__ means2D,
                 const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb_thr,
                 const float* __restrict__ obj_feats, const float* __restrict__ final_T,
                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                 const float* __restrict__ dL_dobjpix, float* __restrict__ accum, float* __restrict__ dL_dobj) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BwdSmem& M = *reinterpret_cast<BwdSmem*>(smem_raw);
  BlendStage& S = M.S;
  // K = 16: the staged entries' object features, four 16-byte words per entry, behind BwdSmem
  float4 (*s_obj)[BLEND_BATCH] = reinterpret_cast<float4 (*)[BLEND_BATCH]>(smem_raw + sizeof(BwdSmem));

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  BwdWarp& Wp = M.w[warp];
  const uint32_t wx0 = blockIdx.x * TILE_X + (warp & 1) * WARP_W, wy0 = blockIdx.y * TILE_Y + (warp >> 1) * WARP_H;
  const uint32_t px = wx0 + (lane & 7), py = wy0 + (lane >> 3);
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float npxf = -(float)px, npyf = -(float)py, wx0f = (float)wx0, wy0f = (float)wy0;
  const size_t HW = (size_t)W * H, pid = (size_t)py * W + px;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];

  const float T_final = inside ? final_T[pid] : 0.f;
  const uint32_t my_last = inside ? n_contrib[pid] : 0u;
  float T = T_final;
  float dpix[3] = {0.f, 0.f, 0.f};
  if (inside) { dpix[0] = dL_dpix[pid]; dpix[1] = dL_dpix[HW + pid]; dpix[2] = dL_dpix[2 * HW + pid]; }
  Wp.pix_rg[lane] = make_float2(dpix[0], dpix[1]);
  Wp.pix_b1[lane] = make_float2(dpix[2], 1.0f);
  float dpo[K > 0 ? K : 1];
#pragma unroll
  for (int k = 0; k < (K > 0 ? K : 1); k++) dpo[k] = (K > 0 && inside) ? dL_dobjpix[k * HW + pid] : 0.f;
  // dL/dpixel . (everything blended behind the current entry): starts as the background's share
  float behind = T_final * (bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2]);

  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, my_last);
  if (tid == 0) M.smax = 0;
  __syncthreads();
  if (lane == 0) atomicMax(&M.smax, warp_last);
  __syncthreads();
  const uint32_t n_live = min(M.smax, range.y - range.x);  // entries [0, n_live) can matter to this tile

  int ng = 0;      // entries in this warp's current group (warp-uniform)
  int slot_j = 0;  // lane e: batch slot of group entry e
  // batches walk the list back to front: slot j of a batch holds list position n_live-1-done_cnt-j
  for (uint32_t done_cnt = 0; done_cnt < n_live; done_cnt += BLEND_BATCH) {
    __syncthreads();
    if (done_cnt + tid < n_live) {
      const uint32_t g = point_list[range.x + (n_live - 1 - done_cnt - tid)];
      stage_entry(S, tid, g, means2D, conic_opacity, rgb_thr, (float)(blockIdx.x * TILE_X), (float)(blockIdx.y * TILE_Y));
      M.id[tid] = g;
      if (K > 0) {
        const float4* f = reinterpret_cast<const float4*>(obj_feats + (size_t)g * K);
#pragma unroll
        for (int q = 0; q < K / 4; q++) s_obj[q][tid] = __ldg(f + q);
      }
    }
    __syncthreads();
    const int nb = (int)min((uint32_t)BLEND_BATCH, n_live - done_cnt);
    const uint32_t top = n_live - 1 - done_cnt;  // list position of slot 0
    for (int r = 0; r < nb; r += 32) {
      const int jl = r + 31 - lane;  // lane 31 tests the round's first slot: the ballot's leading bit is the next slot
      bool keep = false;
      if (jl < nb && top - jl < warp_last) keep = rect_may_contribute(S.a[jl], S.b[jl], wx0f, wy0f);
      const uint32_t mask = __ballot_sync(0xffffffffu, keep);
      GSR_STAT(0, 1); GSR_STAT(1, min(32, nb - r)); GSR_STAT(2, __popc(mask));
      // kept slots in ascending order straight from the ballot (no queue in shared memory)
      for (uint32_t rest = mask; rest != 0;) {
        const int lz = __clz(rest);
        rest &= ~(0x80000000u >> lz);
        const int j = r + lz;
        const float4 a = S.a[j];
        const uint32_t pos = top - j;
        const float power = gauss_power_packed(a, S.b[j].x, npyf, npxf);  // the forward's bits
        // (no `power < threshold` pre-test here: alpha is evaluated on every lane anyway and decides alone)
        bool hit = pos < my_last && !(power > 0.0f);
        const float4 c = S.c[j];
#ifdef GSR_BWD_EXACT_EXP
        float G = 0.f, alpha = 0.f;
        if (hit) {
          G = expf(power);
          alpha = fminf(0.99f, fpx::mul(c.w, G));
          hit = !(alpha < 1.0f / 255.0f);
        }
#else
        // MUFU.EX2 of the scaled argument (2 instructions instead of expf's 9), evaluated by every lane (no
        // divergent region).  Whether a pair contributed must be decided exactly as the forward decided it, so
        // when some lane's alpha lands next to the 1/255 threshold (rare) the warp redoes it with the forward's
        // arithmetic.
        float G = fast_exp(power);
        float alpha = fminf(0.99f, c.w * G);
        const bool near = hit && fabsf(alpha - 1.0f / 255.0f) < 4e-8f;
        if (__any_sync(0xffffffffu, near)) {
          const float Ge = expf(fminf(power, 0.f));
          const float ae = fminf(0.99f, fpx::mul(c.w, Ge));
          G = near ? Ge : G;
          alpha = near ? ae : alpha;
        }
        hit = hit && !(alpha < 1.0f / 255.0f);
#endif
        if (!__any_sync(0xffffffffu, hit)) continue;
#ifdef GSR_STATS
        { const uint32_t hm = __ballot_sync(0xffffffffu, hit); GSR_STAT(3, 1); GSR_STAT(4, __popc(hm)); }
#endif
        float wc = 0.f, wg = 0.f;  // alpha*T and G*dL/dalpha of this pair
        float vo[K > 0 ? K : 1];
#pragma unroll
        for (int k = 0; k < (K > 0 ? K : 1); k++) vo[k] = 0.f;
        if (K == 0) {
          // straight-line form of the block below (no divergent region): a pair that does not contribute runs the
          // same arithmetic with alpha = 0 and G = 0, which leaves T (x 1/(1-0)) and `behind` (+ cd*0) unchanged
          const float ae = hit ? alpha : 0.f, Ge = hit ? G : 0.f;
          const float cd = c.x * dpix[0] + c.y * dpix[1] + c.z * dpix[2];
          const float inv = fast_rcp(1.f - ae);
          T = T * inv;
          wc = ae * T;
          const float dL_dalpha = fmaf(T, cd, -behind * inv);
          behind = fmaf(cd, wc, behind);
          wg = Ge * dL_dalpha;
        } else if (hit) {
          // dC/dalpha_j = T_j c_j - (sum_{k behind j} c_k alpha_k T_k + T_final bg) / (1 - alpha_j); dotted with
          // dL/dpixel the sum is ONE running scalar `behind` (the reference carries the same quantity per
          // channel, normalised, as accum_rec: backward.cu:499-521)
          float cd = c.x * dpix[0] + c.y * dpix[1] + c.z * dpix[2];
          const float inv = fast_rcp(1.f - alpha);
          T = T * inv;
          wc = alpha * T;
          if (K > 0) {
#pragma unroll
            for (int q = 0; q < K / 4; q++) {
              const float4 f = s_obj[q][j];
              cd = fmaf(f.x, dpo[4 * q], cd); cd = fmaf(f.y, dpo[4 * q + 1], cd);
              cd = fmaf(f.z, dpo[4 * q + 2], cd); cd = fmaf(f.w, dpo[4 * q + 3], cd);
            }
#pragma unroll
            for (int k = 0; k < K; k++) vo[k] = wc * dpo[k];
          }
          const float dL_dalpha = fmaf(T, cd, -behind * inv);
          behind = fmaf(cd, wc, behind);
          wg = G * dL_dalpha;
        }
        Wp.panel[ng][lane] = make_float2(wc, wg);
        if (lane == ng) slot_j = j;
        if (K > 0) {
          const uint32_t g = M.id[j];
          const float o8a = warp_sum8(vo, lane), o8b = warp_sum8(vo + (K > 8 ? 8 : 0), lane);
          if ((lane & 3) == 0) {
            atomicAdd(dL_dobj + (size_t)g * K + (lane >> 2), o8a);
            if (K > 8) atomicAdd(dL_dobj + (size_t)g * K + 8 + (lane >> 2), o8b);
          }
        }
        if (++ng == GRP) { flush_group(M, Wp, ng, lane, slot_j, wx0f, wy0f, accum); ng = 0; }
      }
    }
    // the staged batch is about to be replaced: finish the partial group that refers to it
    if (ng > 0) { flush_group(M, Wp, ng, lane, slot_j, wx0f, wy0f, accum); ng = 0; }
  }
}

}  // namespace

int blend_bwd_stats(uint64_t* out_host, int reset) { return read_blend_stats(out_host, reset); }

int launch_blend_bwd(const gsr_inputs& in, const Geom& g, const Image& im, const uint32_t* point_list,
                     const float* dL_dpix, const float* dL_dobjpix, float* accum, float* dL_dobj, cudaStream_t st) {
  const dim3 grid((in.W + TILE_X - 1) / TILE_X, (in.H + TILE_Y - 1) / TILE_Y), block(TILE_PIX);
  GSR_CUDA_CHECK(cudaMemsetAsync(accum, 0, (size_t)in.P * 12 * sizeof(float), st));
  if (in.K > 0) GSR_CUDA_CHECK(cudaMemsetAsync(dL_dobj, 0, (size_t)in.P * in.K * sizeof(float), st));
#define GSR_BLEND_BWD(KK)                                                                                       \
  do {                                                                                                          \
    constexpr size_t smem = sizeof(BwdSmem) + (size_t)KK * BLEND_BATCH * sizeof(float);                         \
    static bool attr_set = false;                                                                               \
    if (!attr_set) {                                                                                            \
      GSR_CUDA_CHECK(cudaFuncSetAttribute(blend_bwd_kernel<KK>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                          (int)smem));                                                          \
      attr_set = true;                                                                                          \
    }                                                                                                           \
    blend_bwd_kernel<KK><<<grid, block, smem, st>>>(im.ranges, point_list, in.W, in.H, in.bg,                   \
                                                               g.means2D, g.conic_opacity, g.rgb_thr,           \
                                                               in.obj_feats, im.final_T, im.n_contrib, dL_dpix, \
                                                               dL_dobjpix, accum, dL_dobj);                     \
  } while (0)
  if (in.K == 0) GSR_BLEND_BWD(0);
  else if (in.K == 16) GSR_BLEND_BWD(16);
  else { set_error("object channels K=%d unsupported (0 or 16)", in.K); return GSR_ERR_INVALID; }
#undef GSR_BLEND_BWD
  GSR_LAUNCH_CHECK("blend_bwd_kernel");
  return GSR_OK;
}

}  // namespace gsr
