// preprocess_bwd.cu -- per-Gaussian backward of the preprocess stage, one fused kernel.
//
// Replaces computeCov2DCUDA (DGR/cuda_rasterizer/backward.cu:144-274), preprocessCUDA backward
// (:346-396), computeColorFromSH backward (:20-139), computeCov3D backward (:278-341) and the
// nine torch::zeros of rasterize_points.cu:151-159 (every output row is written here, zeros for
// Gaussians that were not rendered).  The reference splits cov2D from the rest "due to length"
// and round-trips dL_dcov3D / dL_dconic / dL_dcolor through memory; here they stay in registers
// unless the caller asked for them (cov3D_precomp / colors_precomp given).
// HBM-bound: reads ~150 B + 192 B SH, writes 192 B dL_dsh + ~60 B per rendered Gaussian.
#include "sh_bwd.cuh"

namespace gsr {
namespace {

// ACC = false: every output row is written (zeros for Gaussians that were not rendered).
// ACC = true : the gradients of this view are ADDED to the caller's buffers (a multi-view step accumulates
//              into one flat bucket without per-view temporaries); rows of unrendered Gaussians are untouched,
//              dL_dmean2D may be NULL.
// SPLIT / RAW: the N3 extensions (gsr_inputs::shs_rest, ::raw_params) are compiled out of the standard kernels
template <bool ACC, bool SPLIT, bool RAW>
__global__ void __launch_bounds__(256, 2)  // 128 registers: two resident blocks (some variants took 134-194 and ran one per SM)
preprocess_bwd_kernel(gsr_inputs in, Geom g, const int32_t* __restrict__ radii, const float* __restrict__ accum,
                      float* __restrict__ dL_dmean2D, float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolor,
                      float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
                      float* __restrict__ dL_dscale, float* __restrict__ dL_drot) {
  __shared__ float V[16], Pm[16], cam[3];
  if (threadIdx.x < 16) { V[threadIdx.x] = in.viewmatrix[threadIdx.x]; Pm[threadIdx.x] = in.projmatrix[threadIdx.x]; }
  if (threadIdx.x < 3) cam[threadIdx.x] = in.campos[threadIdx.x];
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= in.P) return;
  const bool live = radii[idx] > 0;
  if (ACC && !live) return;
  const int M = in.M;

#ifndef GSR_PREB_NO_HOIST
  // the SH row is the largest read of this kernel: start it first
  float shv[48];
  if (live && in.shs && M == 16) {
    if (SPLIT) {  // DC / rest blocks where the model keeps them (4-byte aligned rows)
      const int ncoef = 3 * (in.D + 1) * (in.D + 1);
#pragma unroll
      for (int k = 0; k < 3; k++) shv[k] = __ldg(in.shs + 3 * (size_t)idx + k);
      if (rest_window_ok(in.shs_rest, (size_t)idx, in.P)) {
        load_rest_row(in.shs_rest, (size_t)idx, shv);
      } else {
#pragma unroll
        for (int k = 3; k < 48; k++) shv[k] = k < ncoef ? sh_coeff(in, (size_t)idx, k) : 0.f;
      }
    } else {
      const float4* sp0 = reinterpret_cast<const float4*>(in.shs) + (size_t)idx * 12;
#pragma unroll
      for (int k = 0; k < 12; k++) {
        const float4 t = __ldg(sp0 + k);
        shv[4 * k] = t.x; shv[4 * k + 1] = t.y; shv[4 * k + 2] = t.z; shv[4 * k + 3] = t.w;
      }
    }
  }
#endif
  float4 a0 = make_float4(0, 0, 0, 0), a1 = a0;
  float a8 = 0.f;
  if (live) {
    const float4* ap = reinterpret_cast<const float4*>(accum) + (size_t)idx * 3;
    a0 = ap[0]; a1 = ap[1]; a8 = ap[2].x;
  }
  // accumulated by blend_bwd: a0 = dcolor.rgb, sum w dx ; a1 = sum w dy, sum w dx^2, sum w dx dy, sum w dy^2 ;
  // a8 = sum w  (w = G dL/dalpha).  Convert to dL/dmean2D, dL/dconic, dL/dopacity (backward.cu:523-554):
  if (live) {
    const float4 co = g.conic_opacity[idx];
    const float m1 = a0.w, m2 = a1.x, op = co.w;
    a0.w = op * (0.5f * in.W) * (-co.x * m1 - co.y * m2);
    a1.x = op * (0.5f * in.H) * (-co.z * m2 - co.y * m1);
    a1.y *= -0.5f * op; a1.z *= -0.5f * op; a1.w *= -0.5f * op;
    if (RAW) a8 *= op * (1.f - op);  // through the sigmoid: d op / d raw
  }
  if (ACC) {
    if (dL_dmean2D) { atomicAdd(dL_dmean2D + 3 * (size_t)idx, a0.w); atomicAdd(dL_dmean2D + 3 * (size_t)idx + 1, a1.x); }
    atomicAdd(dL_dopacity + idx, a8);
    if (dL_dcolor) { atomicAdd(dL_dcolor + 3 * (size_t)idx, a0.x); atomicAdd(dL_dcolor + 3 * (size_t)idx + 1, a0.y); atomicAdd(dL_dcolor + 3 * (size_t)idx + 2, a0.z); }
  } else {
    dL_dmean2D[3 * (size_t)idx] = a0.w; dL_dmean2D[3 * (size_t)idx + 1] = a1.x; dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
    dL_dopacity[idx] = a8;
    if (dL_dcolor) { dL_dcolor[3 * (size_t)idx] = a0.x; dL_dcolor[3 * (size_t)idx + 1] = a0.y; dL_dcolor[3 * (size_t)idx + 2] = a0.z; }
  }

  float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
  if (live) {
    const float mx = in.means3D[3 * (size_t)idx], my = in.means3D[3 * (size_t)idx + 1], mz = in.means3D[3 * (size_t)idx + 2];
    const float* c3 = in.cov3D_precomp ? in.cov3D_precomp + 6 * (size_t)idx : g.cov3D + 6 * (size_t)idx;
    const float v0 = c3[0], v1 = c3[1], v2 = c3[2], v3 = c3[3], v4 = c3[4], v5 = c3[5];
    const float hx = in.W / (2.0f * in.tanfovx), hy = in.H / (2.0f * in.tanfovy);
    // ---- 2D covariance / conic backward ----
    float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
    float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
    const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
    const float limx = 1.3f * in.tanfovx, limy = 1.3f * in.tanfovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    const float J00 = hx * itz, J02 = -(hx * tx) * itz2, J11 = hy * itz, J12 = -(hy * ty) * itz2;
    // rows of T = W J (2x3): T0k = W[k][0] J00 + W[k][2] J02 ; T1k = W[k][1] J11 + W[k][2] J12
    const float T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
    const float T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
    // TV_rk = sum_j T[r][j] Vrk[j][k]
    const float TV00 = T00 * v0 + T01 * v1 + T02 * v2, TV01 = T00 * v1 + T01 * v3 + T02 * v4, TV02 = T00 * v2 + T01 * v4 + T02 * v5;
    const float TV10 = T10 * v0 + T11 * v1 + T12 * v2, TV11 = T10 * v1 + T11 * v3 + T12 * v4, TV12 = T10 * v2 + T11 * v4 + T12 * v5;
    const float a = TV00 * T00 + TV01 * T01 + TV02 * T02 + 0.3f;
    const float b = TV00 * T10 + TV01 * T11 + TV02 * T12;
    const float c = TV10 * T10 + TV11 * T11 + TV12 * T12 + 0.3f;
    const float dcx = a1.y, dcy = a1.z, dcz = a1.w;
    const float denom = a * c - b * b;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    if (denom2inv != 0.f) {
      dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
      dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
      dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
      dcov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
      dcov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
      dcov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
      dcov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
      dcov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
      dcov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
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
    dmean[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;
    dmean[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
    dmean[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;
    // ---- screen-space mean -> 3D mean through the projection ----
    const float m_w = 1.0f / (Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15] + 0.0000001f);
    const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
    const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
    const float g2x = a0.w, g2y = a1.x;
    dmean[0] += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
    dmean[1] += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
    dmean[2] += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
    // ---- SH colour backward ----
    if (in.shs) {
      const uint8_t cb = g.clamped[idx];
      const V3 dRGB = {(cb & 1) ? 0.f : a0.x, (cb & 2) ? 0.f : a0.y, (cb & 4) ? 0.f : a0.z};
      const float ox = mx - cam[0], oy = my - cam[1], oz = mz - cam[2];
      const float sum2 = ox * ox + oy * oy + oz * oz;
      const float inv_len = 1.0f / sqrtf(sum2);
      const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
      V3 ddir;
      if (M == 16) {
        const float4* sp = reinterpret_cast<const float4*>(in.shs) + (size_t)idx * 12;
        float4* dp = reinterpret_cast<float4*>(dL_dsh) + (size_t)idx * 12;
        float outv[48];
#ifdef GSR_PREB_NO_HOIST
        float shv[48];
#pragma unroll
        for (int k = 0; k < 12; k++) {
          const float4 t = __ldg(sp + k);
          shv[4 * k] = t.x; shv[4 * k + 1] = t.y; shv[4 * k + 2] = t.z; shv[4 * k + 3] = t.w;
        }
#else
        (void)sp;
#endif
#pragma unroll
        for (int k = 0; k < 48; k++) outv[k] = 0.f;
        ddir = sh_backward(in.D, x, y, z, dRGB, [&](int i) { return V3{shv[3 * i], shv[3 * i + 1], shv[3 * i + 2]}; },
                           [&](int i, float w) { outv[3 * i] = w * dRGB.x; outv[3 * i + 1] = w * dRGB.y; outv[3 * i + 2] = w * dRGB.z; });
        if (SPLIT && ACC && in.D >= 3 && rest_window_ok(in.dL_dshs_rest, (size_t)idx, in.P)) {
          // DC: 3 scalar reductions; rest row: its aligned 48-float window, zeros outside the row (see common.cuh)
#pragma unroll
          for (int c2 = 0; c2 < 3; c2++) atomicAdd(dL_dsh + 3 * (size_t)idx + c2, outv[c2]);
          const int s = idx & 3;
          auto ext = [&](int j) { return (j >= 0 && j < 45) ? outv[3 + j] : 0.f; };  // j is a constant after unrolling
          float4* wp = reinterpret_cast<float4*>(in.dL_dshs_rest + (size_t)45 * idx - s);
#pragma unroll
          for (int k = 0; k < 12; k++) {
            float o[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const int j = 4 * k + e;
              const float t0 = (s & 1) ? ext(j - 1) : ext(j), t2 = (s & 1) ? ext(j - 3) : ext(j - 2);
              o[e] = (s & 2) ? t2 : t0;
            }
            red_add4(wp + k, make_float4(o[0], o[1], o[2], o[3]));
          }
        } else if (SPLIT) {
          const int nc = (in.D + 1) * (in.D + 1);
#pragma unroll
          for (int k = 0; k < 16; k++) {
            if (ACC && k >= nc) continue;  // ACC: only the active coefficients change
            float* q3 = k == 0 ? dL_dsh + 3 * (size_t)idx : in.dL_dshs_rest + ((size_t)15 * idx + (k - 1)) * 3;
#pragma unroll
            for (int c2 = 0; c2 < 3; c2++) {
              if (ACC) atomicAdd(q3 + c2, outv[3 * k + c2]);
              else q3[c2] = outv[3 * k + c2];
            }
          }
        } else {
        const int nq = ACC ? (3 * (in.D + 1) * (in.D + 1) + 3) / 4 : 12;  // ACC: only the active coefficients change
#pragma unroll
        for (int k = 0; k < 12; k++) {
          if (k < nq) {
            float4 o = make_float4(outv[4 * k], outv[4 * k + 1], outv[4 * k + 2], outv[4 * k + 3]);
            if (ACC) red_add4(dp + k, o);  // RED.ADD.F32x4: no read round trip through the SM
            else dp[k] = o;
          }
        }
        }
      } else {
        auto coef = [&](int i) -> float* {  // where the 3 channels of coefficient i go, in either SH layout
          if (SPLIT) return i == 0 ? dL_dsh + 3 * (size_t)idx : in.dL_dshs_rest + ((size_t)(M - 1) * idx + (i - 1)) * 3;
          return dL_dsh + ((size_t)M * idx + i) * 3;
        };
        if (!ACC) for (int i = 0; i < M; i++) { float* q3 = coef(i); q3[0] = 0.f; q3[1] = 0.f; q3[2] = 0.f; }
        ddir = sh_backward(in.D, x, y, z, dRGB,
                           [&](int i) {
                             if (SPLIT) return V3{sh_coeff(in, (size_t)idx, 3 * i), sh_coeff(in, (size_t)idx, 3 * i + 1), sh_coeff(in, (size_t)idx, 3 * i + 2)};
                             const float* sp = in.shs + (size_t)idx * M * 3;
                             return V3{sp[3 * i], sp[3 * i + 1], sp[3 * i + 2]};
                           },
                           [&](int i, float w) {
                             float* q3 = coef(i);
                             if (ACC) { atomicAdd(q3, w * dRGB.x); atomicAdd(q3 + 1, w * dRGB.y); atomicAdd(q3 + 2, w * dRGB.z); }
                             else { q3[0] = w * dRGB.x; q3[1] = w * dRGB.y; q3[2] = w * dRGB.z; }
                           });
      }
      // d(normalised direction)/d(mean)
      const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
      dmean[0] += ((sum2 - ox * ox) * ddir.x - oy * ox * ddir.y - oz * ox * ddir.z) * inv32;
      dmean[1] += (-ox * oy * ddir.x + (sum2 - oy * oy) * ddir.y - oz * oy * ddir.z) * inv32;
      dmean[2] += (-ox * oz * ddir.x - oy * oz * ddir.y + (sum2 - oz * oz) * ddir.z) * inv32;
    }
    // ---- 3D covariance -> scale / rotation ----
    if (in.scales) {
      float4 q = reinterpret_cast<const float4*>(in.rotations)[idx];
      float sa[3] = {in.scales[3 * (size_t)idx], in.scales[3 * (size_t)idx + 1], in.scales[3 * (size_t)idx + 2]};
      float inv_qn = 1.f;
      if (RAW) {  // same activations as the forward (scene/gaussian_model.py:97-105)
        sa[0] = expf(sa[0]); sa[1] = expf(sa[1]); sa[2] = expf(sa[2]);
        q = act_normalize(q, &inv_qn);
      }
      const float r = q.x, x = q.y, y = q.z, z = q.w;
      const float s[3] = {in.scale_modifier * sa[0], in.scale_modifier * sa[1], in.scale_modifier * sa[2]};
      // R[c][k]: column c, row k
      const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
      const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                              {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                              {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
      // dM[c][k] = 2 sum_j M[j][k] dS[c][j], with M[j][k] = s_k R[j][k];  dMt[c][k] = dM[k][c]
      float dMt[3][3];
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
      if (RAW) {
        // back through exp (d s / d raw = s) and through normalize (d qhat / d q = (I - qhat qhat^T) / ||q||)
#pragma unroll
        for (int k = 0; k < 3; k++) dscale[k] *= sa[k];
        const float qd = r * drot[0] + x * drot[1] + y * drot[2] + z * drot[3];
        drot[0] = (drot[0] - r * qd) * inv_qn; drot[1] = (drot[1] - x * qd) * inv_qn;
        drot[2] = (drot[2] - y * qd) * inv_qn; drot[3] = (drot[3] - z * qd) * inv_qn;
      }
    }
  } else if (dL_dsh && M > 0) {
    if (SPLIT) {
      float* d3 = dL_dsh + 3 * (size_t)idx;
      d3[0] = 0.f; d3[1] = 0.f; d3[2] = 0.f;
      float* dr = in.dL_dshs_rest + (size_t)(M - 1) * 3 * idx;
      for (int k = 0; k < 3 * (M - 1); k++) dr[k] = 0.f;
    } else if (M == 16) {
      float4* dp = reinterpret_cast<float4*>(dL_dsh) + (size_t)idx * 12;
#pragma unroll
      for (int k = 0; k < 12; k++) dp[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      float* dp = dL_dsh + (size_t)idx * M * 3;
      for (int k = 0; k < 3 * M; k++) dp[k] = 0.f;
    }
  }
  if (ACC) {
#pragma unroll
    for (int k = 0; k < 3; k++) atomicAdd(dL_dmean3D + 3 * (size_t)idx + k, dmean[k]);
    if (dL_dcov3D) {
#pragma unroll
      for (int k = 0; k < 6; k++) atomicAdd(dL_dcov3D + 6 * (size_t)idx + k, dcov[k]);
    }
    if (dL_dscale) {
#pragma unroll
      for (int k = 0; k < 3; k++) atomicAdd(dL_dscale + 3 * (size_t)idx + k, dscale[k]);
      red_add4(reinterpret_cast<float4*>(dL_drot) + idx, make_float4(drot[0], drot[1], drot[2], drot[3]));
    }
    return;
  }
#pragma unroll
  for (int k = 0; k < 3; k++) dL_dmean3D[3 * (size_t)idx + k] = dmean[k];
  if (dL_dcov3D) {
#pragma unroll
    for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
  }
  if (dL_dscale) {
#pragma unroll
    for (int k = 0; k < 3; k++) dL_dscale[3 * (size_t)idx + k] = dscale[k];
    reinterpret_cast<float4*>(dL_drot)[idx] = make_float4(drot[0], drot[1], drot[2], drot[3]);
  }
}

}  // namespace

int launch_preprocess_bwd(const gsr_inputs& in, const Geom& g, const int32_t* radii, const float* accum,
                          float* dL_dmean2D, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                          float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, bool accumulate,
                          cudaStream_t st) {
#define GSR_PRE_BWD(A, S, R)                                                                                      \
  preprocess_bwd_kernel<A, S, R><<<(in.P + 255) / 256, 256, 0, st>>>(in, g, radii, accum, dL_dmean2D, dL_dopacity,      \
                                                                      dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot)
  const bool ext = in.shs_rest != nullptr || in.raw_params != 0;  // N3 paths: one extended kernel handles any mix
  if (!ext) { if (accumulate) GSR_PRE_BWD(true, false, false); else GSR_PRE_BWD(false, false, false); }
  else if (in.shs_rest) {
    if (in.raw_params) { if (accumulate) GSR_PRE_BWD(true, true, true); else GSR_PRE_BWD(false, true, true); }
    else { if (accumulate) GSR_PRE_BWD(true, true, false); else GSR_PRE_BWD(false, true, false); }
  } else { if (accumulate) GSR_PRE_BWD(true, false, true); else GSR_PRE_BWD(false, false, true); }
#undef GSR_PRE_BWD
  GSR_LAUNCH_CHECK("preprocess_bwd_kernel");
  return GSR_OK;
}

}  // namespace gsr
