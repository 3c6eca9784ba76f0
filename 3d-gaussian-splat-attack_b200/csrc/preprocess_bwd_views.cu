// preprocess_bwd_views.cu -- per-Gaussian backward of SEVERAL views in one pass (new; no counterpart in the reference).
//
// A gradient step of the attack sums a loss over the cameras of a batch (attack.py:481-494).  The reference (and
// gsr_backward_accumulate) runs the per-Gaussian backward once per camera: computeCov2DCUDA + preprocessCUDA backward
// (DGR/cuda_rasterizer/backward.cu:144-396) read the Gaussian's mean / scale / rotation / 192-byte SH row and
// read-modify-write its 236-byte gradient row for EVERY view -- about 800 bytes per (Gaussian, view), and at 1M+
// Gaussians the gradient buffers do not fit L2.  Here the per-tile half of each view's backward (blend_bwd) leaves its
// twelve per-Gaussian sums in a per-view accumulator, and this kernel then visits each Gaussian ONCE: it reads the
// view-independent attributes once, loops over the views that rendered it (48-byte accumulator row + 16-byte conic
// per view), sums the gradients of all views in registers and writes (or adds) each gradient row once.
// Bytes per Gaussian for n views: ~232 + 69 n read + 236 written, instead of ~800 n.
// Same arithmetic per view as preprocess_bwd_kernel; the sum over views is taken in registers in view order instead
// of by reductions in memory.  cov3D is recomputed from scale / rotation (computeCov3D, forward.cu:118-150) rather
// than read from every view's geometry state.
#include "sh_bwd.cuh"

namespace gsr {
namespace {

constexpr int FV_MAX = GSR_MAX_FUSED_VIEWS;
constexpr int FV_THREADS = 128;  // (16 views x 3 x 128 floats of per-thread dL/dRGB = 24 KB of static shared memory)

struct FusedViewDev {
  const float4* conic_opacity;
  const uint8_t* clamped;
  const int32_t* radii;
  const float* accum;
  float* dL_dmean2D;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  float hx, hy, limx, limy, halfW, halfH;  // W / (2 tanfovx), H / (2 tanfovy), 1.3 tanfov, W/2, H/2
};
struct FusedViewsDev {
  int n;
  FusedViewDev v[FV_MAX];
};

// per-view pointers and scalars, staged in shared memory so the rolled view loop can index them
struct ViewShared {
  const float4* conic_opacity;
  const uint8_t* clamped;
  const int32_t* radii;
  const float* accum;
  float* dL_dmean2D;
  float hx, hy, limx, limy, halfW, halfH;
  float V[16], Pm[16], cam[3];
};

// The view loop is rolled and each iteration's loads depend on nothing but the Gaussian index, so what the next
// iteration (and the SH phase) will read is requested into L2 ahead of time: no registers held, and the demand
// load then pays an L2 hit instead of a DRAM round trip.
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_view_rows(const ViewShared& S, size_t idx) {
  const char* ap = reinterpret_cast<const char*>(S.accum) + idx * 48;
  prefetch_l2(ap);
  prefetch_l2(ap + 32);  // a 48-byte row can straddle a 32-byte sector boundary pair / a 128-byte line
  prefetch_l2(S.conic_opacity + idx);
  prefetch_l2(S.clamped + idx);
}

struct ViewRows { float4 a0, a1, co; float a8; uint8_t cb; };
__device__ __forceinline__ void load_view_rows(ViewRows& r, const ViewShared& S, size_t idx, bool sh) {
  const float4* ap = reinterpret_cast<const float4*>(S.accum) + idx * 3;
  r.a0 = ap[0]; r.a1 = ap[1]; r.a8 = ap[2].x;
  r.co = S.conic_opacity[idx];
  r.cb = sh ? S.clamped[idx] : (uint8_t)0;
}

template <bool ACC>
__global__ void __launch_bounds__(FV_THREADS, 4)  // 128 registers
preprocess_bwd_views_kernel(gsr_inputs in, FusedViewsDev fv, float* __restrict__ dL_dopacity,
                            float* __restrict__ dL_dcolor, float* __restrict__ dL_dmean3D,
                            float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh, float* __restrict__ dL_dscale,
                            float* __restrict__ dL_drot) {
  __shared__ ViewShared sv[FV_MAX];
  __shared__ float s_rgb[FV_MAX][3][FV_THREADS];  // dL/dRGB of this thread's Gaussian in view v (clamp mask applied)
  const int tid = threadIdx.x, nv = fv.n;
#pragma unroll
  for (int v = 0; v < FV_MAX; v++) {
    if (v < nv) {
      if (tid == 0) {
        sv[v].conic_opacity = fv.v[v].conic_opacity; sv[v].clamped = fv.v[v].clamped; sv[v].radii = fv.v[v].radii;
        sv[v].accum = fv.v[v].accum; sv[v].dL_dmean2D = fv.v[v].dL_dmean2D;
        sv[v].hx = fv.v[v].hx; sv[v].hy = fv.v[v].hy; sv[v].limx = fv.v[v].limx; sv[v].limy = fv.v[v].limy;
        sv[v].halfW = fv.v[v].halfW; sv[v].halfH = fv.v[v].halfH;
      }
      if (tid < 16) { sv[v].V[tid] = fv.v[v].viewmatrix[tid]; sv[v].Pm[tid] = fv.v[v].projmatrix[tid]; }
      if (tid < 3) sv[v].cam[tid] = fv.v[v].campos[tid];
    }
  }
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + tid;
  if (idx >= in.P) return;
  const int M = in.M;

  // every load that depends on the Gaussian index only is issued here, before anything waits: the radii of all views
  // (unrolled), the mean, and scale / rotation or the precomputed covariance
  int32_t rad[FV_MAX];
#pragma unroll
  for (int v = 0; v < FV_MAX; v++) rad[v] = v < nv ? __ldg(sv[v].radii + idx) : 0;
  const float mx = __ldg(in.means3D + 3 * (size_t)idx), my = __ldg(in.means3D + 3 * (size_t)idx + 1),
              mz = __ldg(in.means3D + 3 * (size_t)idx + 2);
  float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
  float s[3] = {0.f, 0.f, 0.f};   // scale_modifier * scale
  float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f;  // 3D covariance (upper triangle)
  if (in.scales) {
    q = __ldg(reinterpret_cast<const float4*>(in.rotations) + idx);
#pragma unroll
    for (int k = 0; k < 3; k++) s[k] = in.scale_modifier * __ldg(in.scales + 3 * (size_t)idx + k);
  } else {
    const float* c3 = in.cov3D_precomp + 6 * (size_t)idx;
    v0 = __ldg(c3); v1 = __ldg(c3 + 1); v2 = __ldg(c3 + 2); v3 = __ldg(c3 + 3); v4 = __ldg(c3 + 4); v5 = __ldg(c3 + 5);
  }
  uint32_t live = 0;
#pragma unroll
  for (int v = 0; v < FV_MAX; v++) live |= (uint32_t)(rad[v] > 0) << v;

  if (!ACC) {  // per-view screen-space gradients of views that did not render this Gaussian: zero rows
    for (int v = 0; v < nv; v++) {
      float* d2 = sv[v].dL_dmean2D;
      if (d2 && !((live >> v) & 1)) { d2[3 * (size_t)idx] = 0.f; d2[3 * (size_t)idx + 1] = 0.f; d2[3 * (size_t)idx + 2] = 0.f; }
    }
  }
  if (live == 0) {
    if (ACC) return;
    dL_dopacity[idx] = 0.f;
#pragma unroll
    for (int k = 0; k < 3; k++) dL_dmean3D[3 * (size_t)idx + k] = 0.f;
    if (dL_dcolor) { for (int k = 0; k < 3; k++) dL_dcolor[3 * (size_t)idx + k] = 0.f; }
    if (dL_dcov3D) { for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = 0.f; }
    if (dL_dscale) {
      for (int k = 0; k < 3; k++) dL_dscale[3 * (size_t)idx + k] = 0.f;
      reinterpret_cast<float4*>(dL_drot)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (dL_dsh && M > 0) {
      if (M == 16) {
        float4* dp = reinterpret_cast<float4*>(dL_dsh) + (size_t)idx * 12;
#pragma unroll
        for (int k = 0; k < 12; k++) dp[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        float* dp = dL_dsh + (size_t)idx * M * 3;
        for (int k = 0; k < 3 * M; k++) dp[k] = 0.f;
      }
    }
    return;
  }

#ifndef GSR_VIEWS_NO_PREFETCH
  if (in.shs) {  // the SH row is read after the view loop: request it now
    const char* sp = reinterpret_cast<const char*>(in.shs) + (size_t)idx * M * 12;
    prefetch_l2(sp);
    if (M * 12 > 128) prefetch_l2(sp + 128);
    if (M * 12 > 64) prefetch_l2(sp + M * 12 - 4);
  }
  {
    const uint32_t later = live >> __ffs(live);
    if (later) prefetch_view_rows(sv[__ffs(live) + __ffs(later) - 1], (size_t)idx);
  }
#endif
  // ---- view-independent attributes (loaded above) ----
  float R[3][3];  // R[c][k]: column c, row k
  if (in.scales) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y - r * z); R[0][2] = 2.f * (x * z + r * y);
    R[1][0] = 2.f * (x * y + r * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z - r * x);
    R[2][0] = 2.f * (x * z - r * y); R[2][1] = 2.f * (y * z + r * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
    // Sigma = M^T M with M[c][k] = s_k R[c][k]  (computeCov3D, forward.cu:118-150)
    const float s0 = s[0] * s[0], s1 = s[1] * s[1], s2 = s[2] * s[2];
    v0 = s0 * R[0][0] * R[0][0] + s1 * R[0][1] * R[0][1] + s2 * R[0][2] * R[0][2];
    v1 = s0 * R[0][0] * R[1][0] + s1 * R[0][1] * R[1][1] + s2 * R[0][2] * R[1][2];
    v2 = s0 * R[0][0] * R[2][0] + s1 * R[0][1] * R[2][1] + s2 * R[0][2] * R[2][2];
    v3 = s0 * R[1][0] * R[1][0] + s1 * R[1][1] * R[1][1] + s2 * R[1][2] * R[1][2];
    v4 = s0 * R[1][0] * R[2][0] + s1 * R[1][1] * R[2][1] + s2 * R[1][2] * R[2][2];
    v5 = s0 * R[2][0] * R[2][0] + s1 * R[2][1] * R[2][1] + s2 * R[2][2] * R[2][2];
  }

  // ---- phase A: screen-space gradients of every view -> 3D mean, 3D covariance, opacity, colour ----
  float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, dcol[3] = {0.f, 0.f, 0.f};
  float dop = 0.f;
  // the rows of the NEXT live view are loaded (into registers) before the current view's arithmetic starts, and the
  // rows of the one after that are requested into L2
  ViewRows nxt;
  int v = __ffs(live) - 1;
  load_view_rows(nxt, sv[v], (size_t)idx, in.shs != nullptr);
  while (v >= 0) {
    const ViewShared& S = sv[v];
    const float4 a0 = nxt.a0, a1 = nxt.a1, co = nxt.co;
    const float a8 = nxt.a8;
    const uint8_t cb = nxt.cb;
    const uint32_t later = live >> (v + 1);
    const int vn = later ? v + __ffs(later) : -1;
    if (vn >= 0) {
      load_view_rows(nxt, sv[vn], (size_t)idx, in.shs != nullptr);
#ifndef GSR_VIEWS_NO_PREFETCH
      const uint32_t later2 = live >> (vn + 1);
      if (later2) prefetch_view_rows(sv[vn + __ffs(later2)], (size_t)idx);
#endif
    }
    // a0 = dcolor.rgb, sum w dx ; a1 = sum w dy, sum w dx^2, sum w dx dy, sum w dy^2 ; a8 = sum w  (blend_bwd.cu)
    const float op = co.w;
    const float g2x = op * S.halfW * (-co.x * a0.w - co.y * a1.x);
    const float g2y = op * S.halfH * (-co.z * a1.x - co.y * a0.w);
    const float dcx = -0.5f * op * a1.y, dcy = -0.5f * op * a1.z, dcz = -0.5f * op * a1.w;
    dop += a8;
    if (S.dL_dmean2D) {
      float* d2 = S.dL_dmean2D + 3 * (size_t)idx;
      if (ACC) { atomicAdd(d2, g2x); atomicAdd(d2 + 1, g2y); }
      else { d2[0] = g2x; d2[1] = g2y; d2[2] = 0.f; }
    }
    const float* V = S.V;
    const float* Pm = S.Pm;
    // ---- 2D covariance / conic backward (backward.cu:144-274) ----
    float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
    float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
    const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(S.limx, fmaxf(-S.limx, txtz)) * tz;
    ty = fminf(S.limy, fmaxf(-S.limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -S.limx || txtz > S.limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -S.limy || tytz > S.limy) ? 0.f : 1.f;
    const float hx = S.hx, hy = S.hy;
    const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    const float J00 = hx * itz, J02 = -(hx * tx) * itz2, J11 = hy * itz, J12 = -(hy * ty) * itz2;
    const float T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
    const float T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
    const float TV00 = T00 * v0 + T01 * v1 + T02 * v2, TV01 = T00 * v1 + T01 * v3 + T02 * v4, TV02 = T00 * v2 + T01 * v4 + T02 * v5;
    const float TV10 = T10 * v0 + T11 * v1 + T12 * v2, TV11 = T10 * v1 + T11 * v3 + T12 * v4, TV12 = T10 * v2 + T11 * v4 + T12 * v5;
    const float a = TV00 * T00 + TV01 * T01 + TV02 * T02 + 0.3f;
    const float b = TV00 * T10 + TV01 * T11 + TV02 * T12;
    const float c = TV10 * T10 + TV11 * T11 + TV12 * T12 + 0.3f;
    const float denom = a * c - b * b;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    if (denom2inv != 0.f) {
      dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
      dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
      dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
      dcov[0] += T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
      dcov[3] += T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
      dcov[5] += T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
      dcov[1] += 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
      dcov[2] += 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
      dcov[4] += 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
    }
    const float dT00 = 2 * TV00 * dL_da + TV10 * dL_db, dT01 = 2 * TV01 * dL_da + TV11 * dL_db, dT02 = 2 * TV02 * dL_da + TV12 * dL_db;
    const float dT10 = 2 * TV10 * dL_dc + TV00 * dL_db, dT11 = 2 * TV11 * dL_dc + TV01 * dL_db, dT12 = 2 * TV12 * dL_dc + TV02 * dL_db;
    const float dJ00 = V[0] * dT00 + V[4] * dT01 + V[8] * dT02;
    const float dJ02 = V[2] * dT00 + V[6] * dT01 + V[10] * dT02;
    const float dJ11 = V[1] * dT10 + V[5] * dT11 + V[9] * dT12;
    const float dJ12 = V[2] * dT10 + V[6] * dT11 + V[10] * dT12;
    const float dtx = x_grad_mul * -hx * itz2 * dJ02;
    const float dty = y_grad_mul * -hy * itz2 * dJ12;
    const float dtz = -hx * itz2 * dJ00 - hy * itz2 * dJ11 + (2 * hx * tx) * itz3 * dJ02 + (2 * hy * ty) * itz3 * dJ12;
    dmean[0] += V[0] * dtx + V[1] * dty + V[2] * dtz;
    dmean[1] += V[4] * dtx + V[5] * dty + V[6] * dtz;
    dmean[2] += V[8] * dtx + V[9] * dty + V[10] * dtz;
    // ---- screen-space mean -> 3D mean through the projection (backward.cu:367-384) ----
    const float m_w = 1.0f / (Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15] + 0.0000001f);
    const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
    const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
    dmean[0] += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
    dmean[1] += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
    dmean[2] += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
    // ---- colour ----
    if (in.shs) {
      s_rgb[v][0][tid] = (cb & 1) ? 0.f : a0.x;
      s_rgb[v][1][tid] = (cb & 2) ? 0.f : a0.y;
      s_rgb[v][2][tid] = (cb & 4) ? 0.f : a0.z;
    } else {
      dcol[0] += a0.x; dcol[1] += a0.y; dcol[2] += a0.z;
    }
    v = vn;
  }

  // ---- geometry rows out: mean (without the SH direction term yet), opacity, colour, covariance / scale + rotation ----
  if (ACC) {
    atomicAdd(dL_dopacity + idx, dop);
    if (dL_dcolor) { for (int k = 0; k < 3; k++) atomicAdd(dL_dcolor + 3 * (size_t)idx + k, dcol[k]); }
    if (dL_dcov3D) { for (int k = 0; k < 6; k++) atomicAdd(dL_dcov3D + 6 * (size_t)idx + k, dcov[k]); }
  } else {
    dL_dopacity[idx] = dop;
    if (dL_dcolor) { for (int k = 0; k < 3; k++) dL_dcolor[3 * (size_t)idx + k] = dcol[k]; }
    if (dL_dcov3D) { for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = dcov[k]; }
  }
  if (in.scales) {  // 3D covariance -> scale / rotation (backward.cu:278-341), once for the sum over the views
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                            {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                            {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
    float dMt[3][3], dscale[3], drot[4];
#pragma unroll
    for (int c2 = 0; c2 < 3; c2++)
#pragma unroll
      for (int k = 0; k < 3; k++)
        dMt[k][c2] = 2.0f * s[k] * (R[0][k] * dS[c2][0] + R[1][k] * dS[c2][1] + R[2][k] * dS[c2][2]);
#pragma unroll
    for (int k = 0; k < 3; k++) dscale[k] = R[0][k] * dMt[k][0] + R[1][k] * dMt[k][1] + R[2][k] * dMt[k][2];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
      for (int c2 = 0; c2 < 3; c2++) dMt[k][c2] *= s[k];
    drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
    drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
    drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
    if (ACC) {
#pragma unroll
      for (int k = 0; k < 3; k++) atomicAdd(dL_dscale + 3 * (size_t)idx + k, dscale[k]);
      red_add4(reinterpret_cast<float4*>(dL_drot) + idx, make_float4(drot[0], drot[1], drot[2], drot[3]));
    } else {
#pragma unroll
      for (int k = 0; k < 3; k++) dL_dscale[3 * (size_t)idx + k] = dscale[k];
      reinterpret_cast<float4*>(dL_drot)[idx] = make_float4(drot[0], drot[1], drot[2], drot[3]);
    }
  }

  // ---- phase B: SH colour backward, summed over the views (backward.cu:20-139) ----
  if (in.shs) {
    float shv[48], outv[48];
    if (M == 16) {
      const float4* sp = reinterpret_cast<const float4*>(in.shs) + (size_t)idx * 12;
#pragma unroll
      for (int k = 0; k < 12; k++) {
        const float4 t = __ldg(sp + k);
        shv[4 * k] = t.x; shv[4 * k + 1] = t.y; shv[4 * k + 2] = t.z; shv[4 * k + 3] = t.w;
      }
    } else {
      const float* sp = in.shs + (size_t)idx * M * 3;
#pragma unroll
      for (int k = 0; k < 48; k++) shv[k] = k < 3 * M ? __ldg(sp + k) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 48; k++) outv[k] = 0.f;
    for (int v = 0; v < nv; v++) {
      if (!((live >> v) & 1)) continue;
      const V3 dRGB = {s_rgb[v][0][tid], s_rgb[v][1][tid], s_rgb[v][2][tid]};
      const float ox = mx - sv[v].cam[0], oy = my - sv[v].cam[1], oz = mz - sv[v].cam[2];
      const float sum2 = ox * ox + oy * oy + oz * oz;
      const float inv_len = 1.0f / sqrtf(sum2);
      const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
      const V3 ddir = sh_backward(in.D, x, y, z, dRGB,
                                  [&](int i) { return V3{shv[3 * i], shv[3 * i + 1], shv[3 * i + 2]}; },
                                  [&](int i, float w) {
                                    outv[3 * i] = fmaf(w, dRGB.x, outv[3 * i]);
                                    outv[3 * i + 1] = fmaf(w, dRGB.y, outv[3 * i + 1]);
                                    outv[3 * i + 2] = fmaf(w, dRGB.z, outv[3 * i + 2]);
                                  });
      // d(normalised direction)/d(mean)
      const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
      dmean[0] += ((sum2 - ox * ox) * ddir.x - oy * ox * ddir.y - oz * ox * ddir.z) * inv32;
      dmean[1] += (-ox * oy * ddir.x + (sum2 - oy * oy) * ddir.y - oz * oy * ddir.z) * inv32;
      dmean[2] += (-ox * oz * ddir.x - oy * oz * ddir.y + (sum2 - oz * oz) * ddir.z) * inv32;
    }
    if (M == 16) {
      float4* dp = reinterpret_cast<float4*>(dL_dsh) + (size_t)idx * 12;
      const int nq = ACC ? (3 * (in.D + 1) * (in.D + 1) + 3) / 4 : 12;  // ACC: only the active coefficients change
#pragma unroll
      for (int k = 0; k < 12; k++) {
        if (k < nq) {
          const float4 o = make_float4(outv[4 * k], outv[4 * k + 1], outv[4 * k + 2], outv[4 * k + 3]);
          if (ACC) red_add4(dp + k, o);
          else dp[k] = o;
        }
      }
    } else {
      float* dp = dL_dsh + (size_t)idx * M * 3;
      const int nc = ACC ? 3 * (in.D + 1) * (in.D + 1) : 3 * M;
#pragma unroll
      for (int k = 0; k < 48; k++) {
        if (k < nc) {
          if (ACC) atomicAdd(dp + k, outv[k]);
          else dp[k] = outv[k];
        }
      }
    }
  }
  if (ACC) {
#pragma unroll
    for (int k = 0; k < 3; k++) atomicAdd(dL_dmean3D + 3 * (size_t)idx + k, dmean[k]);
  } else {
#pragma unroll
    for (int k = 0; k < 3; k++) dL_dmean3D[3 * (size_t)idx + k] = dmean[k];
  }
}

}  // namespace

int launch_preprocess_bwd_views(const gsr_inputs& in, int n_views, const gsr_view_ref* views, bool accumulate,
                                float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                                float* dL_dsh, float* dL_dscale, float* dL_drot, cudaStream_t st) {
  FusedViewsDev fv;
  fv.n = n_views;
  for (int v = 0; v < n_views; v++) {
    const gsr_view_ref& r = views[v];
    Geom g(const_cast<void*>(r.geom), in.P);
    FusedViewDev& d = fv.v[v];
    d.conic_opacity = g.conic_opacity; d.clamped = g.clamped; d.radii = r.radii; d.accum = r.accum;
    d.dL_dmean2D = r.dL_dmean2D; d.viewmatrix = r.viewmatrix; d.projmatrix = r.projmatrix; d.campos = r.campos;
    d.hx = r.W / (2.0f * r.tanfovx); d.hy = r.H / (2.0f * r.tanfovy);
    d.limx = 1.3f * r.tanfovx; d.limy = 1.3f * r.tanfovy; d.halfW = 0.5f * r.W; d.halfH = 0.5f * r.H;
  }
  const int blocks = (in.P + FV_THREADS - 1) / FV_THREADS;
  if (accumulate)
    preprocess_bwd_views_kernel<true><<<blocks, FV_THREADS, 0, st>>>(in, fv, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D,
                                                                     dL_dsh, dL_dscale, dL_drot);
  else
    preprocess_bwd_views_kernel<false><<<blocks, FV_THREADS, 0, st>>>(in, fv, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D,
                                                                      dL_dsh, dL_dscale, dL_drot);
  GSR_LAUNCH_CHECK("preprocess_bwd_views_kernel");
  return GSR_OK;
}

}  // namespace gsr
