// preprocess_fwd.cu -- per-Gaussian forward preprocess.
//
// Replaces preprocessCUDA (DGR/cuda_rasterizer/forward.cu:155-256: in_frustum auxiliary.h:139-164,
// computeCov3D forward.cu:118-152, computeCov2D :74-113, ndc2Pix auxiliary.h:41-44, getRect :46-56,
// computeColorFromSH forward.cu:20-71) and checkFrustum (rasterizer_impl.cu:54-66).  The reference's
// cub::DeviceScan::InclusiveSum over tiles_touched in index order (rasterizer_impl.cu:277) is not needed by
// this design's binning (binning.cu scans in depth order, which also yields num_rendered); index_scan in
// binning.cu reproduces it on demand for the parity tests.
//
// Bit-exactness: depths, radii, pixel centres, conics and RGB feed the sort keys, the tile lists and
// the alpha thresholds.  Which mul/add pairs end up fused is decided twice in the reference build:
// by the front end (visible in its PTX as fma.rn vs mul/add) and again by ptxas, which may fuse the
// remaining un-rounded mul.f32/add.f32 pairs.  fp_exact.cuh therefore emits the arithmetic as inline
// PTX with the same rounding qualifiers and the same dataflow (which products are shared), so ptxas
// is given the same choice and makes it the same way.  Verified bit-for-bit against the reference
// CUDA build on the B200 (tests/test_gpu_parity.py).
#include "common.cuh"
#include "fp_exact.cuh"

namespace gsr {
namespace {

using namespace fpx;

#define SHC0 0.28209479177387814f
#define SHC1 0.4886025119029199f
#define SHC2_0 1.0925484305920792f
#define SHC2_1 -1.0925484305920792f
#define SHC2_2 0.31539156525252005f
#define SHC2_3 -1.0925484305920792f
#define SHC2_4 0.5462742152960396f
#define SHC3_0 -0.5900435899266435f
#define SHC3_1 2.890611442640554f
#define SHC3_2 -0.4570457994644658f
#define SHC3_3 0.3731763325901154f
#define SHC3_4 -0.4570457994644658f
#define SHC3_5 1.445305721320277f
#define SHC3_6 -0.5900435899266435f

// three-term inner product, middle product rounded first, outer two fused onto it
__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
  return fma_(a2, b2, fma_(a0, b0, mul(a1, b1)));
}
// row r of a column-major 4x4 applied to (x, y, z, 1)
__device__ __forceinline__ float affine_row(const float* m, int r, float x, float y, float z) {
  return add(m[12 + r], dot3(x, m[r], y, m[4 + r], z, m[8 + r]));
}

struct Cam {
  float view[16];
  float proj[16];
  float campos[3];
};

// Sigma = (S R)^T (S R) from post-activation scale and (un-normalised) quaternion (r, x, y, z).
__device__ __forceinline__ void covariance3d(float3 s, float mod, float4 q, float* c) {
  const float sx = mul(mod, s.x), sy = mul(mod, s.y), sz = mul(mod, s.z);
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  const float yy = mul(y, y), zz = mul(z, z);
  const float xy = mul(x, y), rz = mul(r, z), xz = mul(x, z), ry = mul(r, y), yz = mul(y, z), rx = mul(r, x);
  const float a_yyzz = add(yy, zz);
  const float a_xxzz = fma_(x, x, zz);
  const float a_xxyy = fma_(x, x, yy);
  const float d0 = sub(xy, rz), d1 = add(ry, xz), d2 = add(xy, rz);
  const float d3 = sub(yz, rx), d4 = sub(xz, ry), d5 = add(rx, yz);
  // rotation columns scaled row-wise: m[c][k] = s_k * R[c][k]
  const float m00 = mul(sx, sub(1.0f, add(a_yyzz, a_yyzz))), m01 = mul(sy, add(d0, d0)), m02 = mul(sz, add(d1, d1));
  const float m10 = mul(sx, add(d2, d2)), m11 = mul(sy, sub(1.0f, add(a_xxzz, a_xxzz))), m12 = mul(sz, add(d3, d3));
  const float m20 = mul(sx, add(d4, d4)), m21 = mul(sy, add(d5, d5)), m22 = mul(sz, sub(1.0f, add(a_xxyy, a_xxyy)));
  c[0] = dot3(m00, m00, m01, m01, m02, m02);
  c[1] = dot3(m10, m00, m11, m01, m12, m02);
  c[2] = dot3(m20, m00, m21, m01, m22, m02);
  c[3] = dot3(m10, m10, m11, m11, m12, m12);
  c[4] = dot3(m20, m10, m21, m11, m22, m12);
  c[5] = dot3(m20, m20, m21, m21, m22, m22);
}

// EWA projection of the 3D covariance: returns (a, b, c) of the dilated 2D covariance.
__device__ __forceinline__ float3 covariance2d(float x, float y, float z, float fx, float fy, float tanx, float tany,
                                               const float* c3, const float* V) {
  const float tx = affine_row(V, 0, x, y, z), ty = affine_row(V, 1, x, y, z), tz = affine_row(V, 2, x, y, z);
  const float limx = mul(tanx, 1.3f), limy = mul(tany, 1.3f);
  const float cx = fminf(limx, fmaxf(-limx, fdiv(tx, tz)));
  const float cy = fminf(limy, fmaxf(-limy, fdiv(ty, tz)));
  const float ntz = -tz, tz2 = mul(tz, tz);
  const float j00 = fdiv(fx, tz);
  const float j02 = fdiv(mul(fx, mul(cx, ntz)), tz2);
  const float j11 = fdiv(fy, tz);
  const float j12 = fdiv(mul(fy, mul(cy, ntz)), tz2);
  const float t00 = fma_(V[2], j02, mul(V[0], j00));
  const float t01 = fma_(V[6], j02, mul(V[4], j00));
  const float t02 = fma_(j02, V[10], mul(V[8], j00));
  const float t10 = fma_(V[2], j12, mul(j11, V[1]));
  const float t11 = fma_(V[6], j12, mul(j11, V[5]));
  const float t12 = fma_(j12, V[10], mul(j11, V[9]));
  const float a00 = dot3(t00, c3[0], t01, c3[1], t02, c3[2]);
  const float a01 = dot3(t10, c3[0], t11, c3[1], t12, c3[2]);
  const float a10 = dot3(t00, c3[1], t01, c3[3], t02, c3[4]);
  const float a11 = dot3(t10, c3[1], t11, c3[3], t12, c3[4]);
  const float a20 = dot3(t00, c3[2], t01, c3[4], t02, c3[5]);
  const float a21 = dot3(t10, c3[2], t11, c3[4], t12, c3[5]);
  float3 o;
  o.x = add(dot3(t00, a00, t01, a10, t02, a20), 0.3f);
  o.y = dot3(t00, a01, t01, a11, t02, a21);
  o.z = add(dot3(t10, a01, t11, a11, t12, a21), 0.3f);
  return o;
}

// SH (degree <= 3) -> RGB + 0.5, clamp at 0.  sh[] holds 3*M floats of this Gaussian, coefficient-major.
template <typename ShFetch>
__device__ __forceinline__ void sh_to_rgb(int deg, float px, float py, float pz, const float* campos, ShFetch sh,
                                          float* rgb, uint8_t& clamp_bits) {
  const float dx = sub(px, campos[0]), dy = sub(py, campos[1]), dz = sub(pz, campos[2]);
  const float len = fsqrt(dot3(dx, dx, dy, dy, dz, dz));
  const float x = fdiv(dx, len), y = fdiv(dy, len), z = fdiv(dz, len);
  float res[3];
#pragma unroll
  for (int c = 0; c < 3; c++) res[c] = mul(sh(c), SHC0);
  if (deg > 0) {
    const float ky = mul(y, SHC1), kz = mul(z, SHC1), kx = mul(x, SHC1);
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float t = sub(res[c], mul(ky, sh(3 + c)));
      t = fma_(kz, sh(6 + c), t);
      res[c] = sub(t, mul(kx, sh(9 + c)));
    }
    if (deg > 1) {
      const float xx = mul(x, x), yy = mul(y, y), zz = mul(z, z), xy = mul(x, y), yz = mul(y, z), xz = mul(x, z);
      const float zz2 = add(zz, zz), xx_yy = sub(xx, yy);
      const float k4 = mul(xy, SHC2_0), k5 = mul(yz, SHC2_1);
      const float k6 = mul(sub(sub(zz2, xx), yy), SHC2_2);
      const float k7 = mul(xz, SHC2_3), k8 = mul(xx_yy, SHC2_4);
#pragma unroll
      for (int c = 0; c < 3; c++) {
        float t = fma_(k4, sh(12 + c), res[c]);
        t = fma_(k5, sh(15 + c), t);
        t = fma_(k6, sh(18 + c), t);
        t = fma_(k7, sh(21 + c), t);
        res[c] = fma_(k8, sh(24 + c), t);
      }
      if (deg > 2) {
        const float xx3 = mul(xx, 3.0f), yy3 = mul(yy, 3.0f);
        const float q = sub(sub(mul(zz, 4.0f), xx), yy);
        const float k9 = mul(mul(y, SHC3_0), sub(xx3, yy));
        const float k10 = mul(z, mul(xy, SHC3_1));
        const float k11 = mul(mul(y, SHC3_2), q);
        const float k12 = mul(mul(z, SHC3_3), sub(sub(zz2, xx3), yy3));
        const float k13 = mul(mul(x, SHC3_4), q);
        const float k14 = mul(mul(z, SHC3_5), xx_yy);
        const float k15 = mul(mul(x, SHC3_6), sub(xx, yy3));
#pragma unroll
        for (int c = 0; c < 3; c++) {
          float t = fma_(k9, sh(27 + c), res[c]);
          t = fma_(k10, sh(30 + c), t);
          t = fma_(k11, sh(33 + c), t);
          t = fma_(k12, sh(36 + c), t);
          t = fma_(k13, sh(39 + c), t);
          t = fma_(k14, sh(42 + c), t);
          res[c] = fma_(k15, sh(45 + c), t);
        }
      }
    }
  }
  clamp_bits = 0;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float v = add(res[c], 0.5f);
    if (v < 0.0f) clamp_bits |= (uint8_t)(1u << c);
    rgb[c] = v < 0.0f ? 0.0f : v;
  }
}

// EXT: the N3 inputs (split SH blocks, raw parameters) -- compiled out of the standard kernel, as in preprocess_bwd
template <bool EXT>
__global__ void __launch_bounds__(PRE_THREADS)
preprocess_fwd_kernel(gsr_inputs in, Geom g, int32_t* __restrict__ radii, uint32_t* flags,
                      uint32_t* __restrict__ dkeys) {
  __shared__ Cam cam;
  const int tid = threadIdx.x;
  if (tid < 16) { cam.view[tid] = in.viewmatrix[tid]; cam.proj[tid] = in.projmatrix[tid]; }
  if (tid < 3) cam.campos[tid] = in.campos[tid];
  __syncthreads();
  const int idx = (int)(blockIdx.x * PRE_THREADS + tid);
  const int W = in.W, H = in.H;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;

  uint32_t touched = 0;
  if (idx < in.P) {
    int radius = 0;
    ushort4 rect = make_ushort4(0, 0, 0, 0);
    const float px = in.means3D[3 * idx], py = in.means3D[3 * idx + 1], pz = in.means3D[3 * idx + 2];
    // the small per-Gaussian reads are issued together with the mean (one DRAM latency instead of two)
    float3 s_in = make_float3(0.f, 0.f, 0.f);
    float4 q_in = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!in.cov3D_precomp) {
      s_in = make_float3(in.scales[3 * idx], in.scales[3 * idx + 1], in.scales[3 * idx + 2]);
      q_in = reinterpret_cast<const float4*>(in.rotations)[idx];
    }
    const float op_in = in.opacities[idx];
    const float depth = affine_row(cam.view, 2, px, py, pz);
    if (depth > 0.2f) {
      const float hx = affine_row(cam.proj, 0, px, py, pz);
      const float hy = affine_row(cam.proj, 1, px, py, pz);
      const float hw = affine_row(cam.proj, 3, px, py, pz);
      const float p_w = frcp(add(hw, 0.0000001f));
      const float ndc_x = mul(hx, p_w), ndc_y = mul(hy, p_w);
      float c3[6];
      if (in.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) c3[k] = in.cov3D_precomp[6 * (size_t)idx + k];
      } else {
        float3 s = s_in;
        float4 q = q_in;
        if (EXT && in.raw_params) {  // exp / normalize activations applied here (scene/gaussian_model.py:97-105)
          s = make_float3(expf(s.x), expf(s.y), expf(s.z));
          q = act_normalize(q, nullptr);
        }
        covariance3d(s, in.scale_modifier, q, c3);
#pragma unroll
        for (int k = 0; k < 6; k++) g.cov3D[6 * (size_t)idx + k] = c3[k];
      }
      const float fy = fdiv((float)H, mul(2.0f, in.tanfovy)), fx = fdiv((float)W, mul(2.0f, in.tanfovx));
      const float3 cov = covariance2d(px, py, pz, fx, fy, in.tanfovx, in.tanfovy, c3, cam.view);
      const float det = sub(mul(cov.x, cov.z), mul(cov.y, cov.y));
      if (det != 0.0f) {
        const float det_inv = frcp(det);
        const float mid = mul(add(cov.x, cov.z), 0.5f);
        const float disc = fsqrt(fmaxf(sub(mul(mid, mid), det), 0.1f));
        const float lam = fmaxf(add(mid, disc), sub(mid, disc));
        const float my_radius = ceilf(mul(fsqrt(lam), 3.0f));
        // pixel centre in FP64, as the reference's double literals force it (auxiliary.h:41-44)
        const float pix_x = (float)__dmul_rn(__fma_rn(__dadd_rn((double)ndc_x, 1.0), (double)W, -1.0), 0.5);
        const float pix_y = (float)__dmul_rn(__fma_rn(__dadd_rn((double)ndc_y, 1.0), (double)H, -1.0), 0.5);
        const int ir = (int)my_radius;
        const float fr = (float)ir;
        const int x0 = min(gx, max(0, (int)mul(sub(pix_x, fr), 0.0625f)));
        const int y0 = min(gy, max(0, (int)mul(sub(pix_y, fr), 0.0625f)));
        const int x1 = min(gx, max(0, (int)mul(add(add(add(pix_x, fr), 16.0f), -1.0f), 0.0625f)));
        const int y1 = min(gy, max(0, (int)mul(add(add(add(pix_y, fr), 16.0f), -1.0f), 0.0625f)));
        const uint32_t area = (uint32_t)(x1 - x0) * (uint32_t)(y1 - y0);
        if (area != 0) {
          float rgb[3];
          uint8_t cbits = 0;
          if (in.colors_precomp) {
            rgb[0] = in.colors_precomp[3 * (size_t)idx];
            rgb[1] = in.colors_precomp[3 * (size_t)idx + 1];
            rgb[2] = in.colors_precomp[3 * (size_t)idx + 2];
          } else if (EXT && in.shs_rest) {
            // DC and rest blocks read where the model keeps them (no torch.cat): rows are only 4-byte aligned
            float shv[48];
            const int ncoef = 3 * (in.D + 1) * (in.D + 1);
            if (in.M == 16) {
#pragma unroll
              for (int k = 0; k < 3; k++) shv[k] = __ldg(in.shs + 3 * (size_t)idx + k);
              if (rest_window_ok(in.shs_rest, (size_t)idx, in.P)) {
                load_rest_row(in.shs_rest, (size_t)idx, shv);
              } else {
#pragma unroll
                for (int k = 3; k < 48; k++) shv[k] = k < ncoef ? sh_coeff(in, (size_t)idx, k) : 0.f;
              }
              sh_to_rgb(in.D, px, py, pz, cam.campos, [&](int i) { return shv[i]; }, rgb, cbits);
            } else {
              sh_to_rgb(in.D, px, py, pz, cam.campos, [&](int i) { return sh_coeff(in, (size_t)idx, i); }, rgb, cbits);
            }
          } else if (in.M == 16) {
            // 48 floats = 12 aligned 16-byte loads
            float shv[48];
            const float4* sp = reinterpret_cast<const float4*>(in.shs) + (size_t)idx * 12;
            const int nload = in.D >= 3 ? 12 : (in.D == 2 ? 7 : (in.D == 1 ? 3 : 1));
#pragma unroll
            for (int k = 0; k < 12; k++) {
              float4 v = k < nload ? __ldg(sp + k) : make_float4(0.f, 0.f, 0.f, 0.f);
              shv[4 * k] = v.x; shv[4 * k + 1] = v.y; shv[4 * k + 2] = v.z; shv[4 * k + 3] = v.w;
            }
            sh_to_rgb(in.D, px, py, pz, cam.campos, [&](int i) { return shv[i]; }, rgb, cbits);
          } else {
            const float* sp = in.shs + (size_t)idx * in.M * 3;
            sh_to_rgb(in.D, px, py, pz, cam.campos, [&](int i) { return __ldg(sp + i); }, rgb, cbits);
          }
          const float op = (EXT && in.raw_params) ? act_sigmoid(op_in) : op_in;
          g.depths[idx] = depth;
          g.means2D[idx] = make_float2(pix_x, pix_y);
          g.conic_opacity[idx] = make_float4(mul(cov.z, det_inv), mul(det_inv, -cov.y), mul(cov.x, det_inv), op);
          // pairs with power below this can never reach alpha >= 1/255 (1e-4 covers every rounding error)
          g.rgb_thr[idx] = make_float4(rgb[0], rgb[1], rgb[2], -__logf(255.0f * op) - 1e-4f);
          g.clamped[idx] = cbits;
          radius = ir;
          rect = make_ushort4((unsigned short)x0, (unsigned short)y0, (unsigned short)x1, (unsigned short)y1);
          touched = area;
        }
      }
    } else if (in.prefiltered) {
      flags[0] = 1u;  // the reference __trap()s here (auxiliary.h:156-160); reported by the host instead
    }
    radii[idx] = radius;
    g.tiles_touched[idx] = touched;
    g.rects[idx] = rect;
    // input of the (depth bits, index) ordering: rendered Gaussians sort by their depth bits (positive floats
    // order like unsigned integers, as in the reference's keys), the others go last
    dkeys[idx] = touched ? __float_as_uint(depth) : 0xffffffffu;  // (the sort generates the indices itself)
  }

}

// ---- several views in one pass (new; no counterpart in the reference, whose preprocessCUDA runs once per camera) ----
// A multi-camera step renders the SAME Gaussians from several cameras.  Per camera the kernel above reads 236 bytes of
// attributes per Gaussian (192 of them the SH row) and writes 89; here a thread reads its Gaussian's attributes once,
// builds the 3D covariance once and loops over the cameras, writing each camera's geometry state.  The per-view
// arithmetic is the block above, statement for statement (same helpers, same dataflow), so every per-view output is
// bit-identical to the single-view kernel's (tests/test_gpu_views_fused.py compares the buffers byte for byte).
// Depth keys go to the view's own offsets_sorted array (free until the depth-order scan overwrites it), from where
// gsr_forward_order sorts them.  Standard inputs only (no raw_params / shs_rest / prefiltered).
constexpr int PV_MAX = GSR_MAX_FUSED_VIEWS;
struct FwdViewDev {
  float* depths; float2* means2D; float4* conic_opacity; float4* rgb_thr; float* cov3D; uint8_t* clamped;
  uint32_t* tiles_touched; ushort4* rects; uint32_t* dkeys; int32_t* radii;
  const float* viewmatrix; const float* projmatrix; const float* campos;
  float tanfovx, tanfovy;
  int W, H;
};
struct FwdViewsDev {
  int n;
  FwdViewDev v[PV_MAX];
};
struct FwdViewShared {
  FwdViewDev d;
  Cam cam;
};

// Measured on C2, 8 views per launch: SH row kept in registers across the views (121 registers, two resident blocks)
// 318 us; re-read per view from L1 / L2 at three resident blocks (80 registers) 420 us, at four 466 us.
#ifndef GSR_PV_NO_SH_CACHE
#define GSR_PV_SH_CACHE 1
#endif
#ifndef GSR_PV_MINBLOCKS
#define GSR_PV_MINBLOCKS 2
#endif
__global__ void __launch_bounds__(PRE_THREADS, GSR_PV_MINBLOCKS)
preprocess_fwd_views_kernel(gsr_inputs in, FwdViewsDev fv) {
  __shared__ FwdViewShared sv[PV_MAX];
  const int tid = threadIdx.x, nv = fv.n;
#pragma unroll
  for (int v = 0; v < PV_MAX; v++) {
    if (v < nv) {
      if (tid == 0) sv[v].d = fv.v[v];
      if (tid < 16) { sv[v].cam.view[tid] = fv.v[v].viewmatrix[tid]; sv[v].cam.proj[tid] = fv.v[v].projmatrix[tid]; }
      if (tid < 3) sv[v].cam.campos[tid] = fv.v[v].campos[tid];
    }
  }
  __syncthreads();
  const int idx = (int)(blockIdx.x * PRE_THREADS + tid);
  if (idx >= in.P) return;

  // view-independent attributes, read once
  const float px = in.means3D[3 * idx], py = in.means3D[3 * idx + 1], pz = in.means3D[3 * idx + 2];
  float c3[6];
  if (in.cov3D_precomp) {
#pragma unroll
    for (int k = 0; k < 6; k++) c3[k] = in.cov3D_precomp[6 * (size_t)idx + k];
  } else {
    const float3 s_in = make_float3(in.scales[3 * idx], in.scales[3 * idx + 1], in.scales[3 * idx + 2]);
    const float4 q_in = reinterpret_cast<const float4*>(in.rotations)[idx];
    covariance3d(s_in, in.scale_modifier, q_in, c3);
  }
  const float op = in.opacities[idx];
  const float thr = -__logf(255.0f * op) - 1e-4f;
#ifdef GSR_PV_SH_CACHE
  float shv[48];  // the SH row, read at the first view that renders the Gaussian and kept across the loop
  bool have_sh = false;
#endif

  for (int v = 0; v < nv; v++) {
    const FwdViewShared& S = sv[v];
    const int W = S.d.W, H = S.d.H;
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    uint32_t touched = 0;
    int radius = 0;
    ushort4 rect = make_ushort4(0, 0, 0, 0);
    const float depth = affine_row(S.cam.view, 2, px, py, pz);
    if (depth > 0.2f) {
      const float hx = affine_row(S.cam.proj, 0, px, py, pz);
      const float hy = affine_row(S.cam.proj, 1, px, py, pz);
      const float hw = affine_row(S.cam.proj, 3, px, py, pz);
      const float p_w = frcp(add(hw, 0.0000001f));
      const float ndc_x = mul(hx, p_w), ndc_y = mul(hy, p_w);
      if (!in.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) S.d.cov3D[6 * (size_t)idx + k] = c3[k];
      }
      const float fy = fdiv((float)H, mul(2.0f, S.d.tanfovy)), fx = fdiv((float)W, mul(2.0f, S.d.tanfovx));
      const float3 cov = covariance2d(px, py, pz, fx, fy, S.d.tanfovx, S.d.tanfovy, c3, S.cam.view);
      const float det = sub(mul(cov.x, cov.z), mul(cov.y, cov.y));
      if (det != 0.0f) {
        const float det_inv = frcp(det);
        const float mid = mul(add(cov.x, cov.z), 0.5f);
        const float disc = fsqrt(fmaxf(sub(mul(mid, mid), det), 0.1f));
        const float lam = fmaxf(add(mid, disc), sub(mid, disc));
        const float my_radius = ceilf(mul(fsqrt(lam), 3.0f));
        const float pix_x = (float)__dmul_rn(__fma_rn(__dadd_rn((double)ndc_x, 1.0), (double)W, -1.0), 0.5);
        const float pix_y = (float)__dmul_rn(__fma_rn(__dadd_rn((double)ndc_y, 1.0), (double)H, -1.0), 0.5);
        const int ir = (int)my_radius;
        const float fr = (float)ir;
        const int x0 = min(gx, max(0, (int)mul(sub(pix_x, fr), 0.0625f)));
        const int y0 = min(gy, max(0, (int)mul(sub(pix_y, fr), 0.0625f)));
        const int x1 = min(gx, max(0, (int)mul(add(add(add(pix_x, fr), 16.0f), -1.0f), 0.0625f)));
        const int y1 = min(gy, max(0, (int)mul(add(add(add(pix_y, fr), 16.0f), -1.0f), 0.0625f)));
        const uint32_t area = (uint32_t)(x1 - x0) * (uint32_t)(y1 - y0);
        if (area != 0) {
          float rgb[3];
          uint8_t cbits = 0;
          if (in.colors_precomp) {
            rgb[0] = in.colors_precomp[3 * (size_t)idx];
            rgb[1] = in.colors_precomp[3 * (size_t)idx + 1];
            rgb[2] = in.colors_precomp[3 * (size_t)idx + 2];
          } else if (in.M == 16) {
#ifdef GSR_PV_SH_CACHE
            if (!have_sh) {  // the SH row is read at the first view that renders this Gaussian and kept
#else
            // (variant, slower: the SH row fetched again for every view that renders the Gaussian, from L1 / L2 after
            // the first time; fewer registers, three resident blocks)
            float shv[48];
            {
#endif
              const float4* sp = reinterpret_cast<const float4*>(in.shs) + (size_t)idx * 12;
              const int nload = in.D >= 3 ? 12 : (in.D == 2 ? 7 : (in.D == 1 ? 3 : 1));
#pragma unroll
              for (int k = 0; k < 12; k++) {
                float4 t = k < nload ? __ldg(sp + k) : make_float4(0.f, 0.f, 0.f, 0.f);
                shv[4 * k] = t.x; shv[4 * k + 1] = t.y; shv[4 * k + 2] = t.z; shv[4 * k + 3] = t.w;
              }
#ifdef GSR_PV_SH_CACHE
              have_sh = true;
#endif
            }
            sh_to_rgb(in.D, px, py, pz, S.cam.campos, [&](int i) { return shv[i]; }, rgb, cbits);
          } else {
            const float* sp = in.shs + (size_t)idx * in.M * 3;
            sh_to_rgb(in.D, px, py, pz, S.cam.campos, [&](int i) { return __ldg(sp + i); }, rgb, cbits);
          }
          S.d.depths[idx] = depth;
          S.d.means2D[idx] = make_float2(pix_x, pix_y);
          S.d.conic_opacity[idx] = make_float4(mul(cov.z, det_inv), mul(det_inv, -cov.y), mul(cov.x, det_inv), op);
          S.d.rgb_thr[idx] = make_float4(rgb[0], rgb[1], rgb[2], thr);
          S.d.clamped[idx] = cbits;
          radius = ir;
          rect = make_ushort4((unsigned short)x0, (unsigned short)y0, (unsigned short)x1, (unsigned short)y1);
          touched = area;
        }
      }
    }
    S.d.radii[idx] = radius;
    S.d.tiles_touched[idx] = touched;
    S.d.rects[idx] = rect;
    S.d.dkeys[idx] = touched ? __float_as_uint(depth) : 0xffffffffu;
  }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float d = affine_row(view, 2, means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
  present[idx] = d > 0.2f;
}

}  // namespace

int launch_preprocess_fwd(const gsr_inputs& in, Geom& g, int32_t* radii, Scratch& s, uint32_t* num_rendered_dev,
                          cudaStream_t st) {
  const int nblocks = (in.P + PRE_THREADS - 1) / PRE_THREADS;
  GSR_CUDA_CHECK(cudaMemsetAsync(s.scan_status, 0, s.scan_zero_bytes, st));
  GSR_CUDA_CHECK(cudaMemsetAsync(g.giant_count, 0, sizeof(uint32_t), st));
  (void)num_rendered_dev;  // written by the depth-order scan (launch_sorted_scan)
  if (in.shs_rest != nullptr || in.raw_params != 0)
    preprocess_fwd_kernel<true><<<nblocks, PRE_THREADS, 0, st>>>(in, g, radii, s.flags, s.dkeys_a);
  else
    preprocess_fwd_kernel<false><<<nblocks, PRE_THREADS, 0, st>>>(in, g, radii, s.flags, s.dkeys_a);
  GSR_LAUNCH_CHECK("preprocess_fwd_kernel");
  return GSR_OK;
}

int launch_preprocess_fwd_views(const gsr_inputs& in, int n_views, const gsr_view_fwd* views, cudaStream_t st) {
  FwdViewsDev fv;
  fv.n = n_views;
  for (int v = 0; v < n_views; v++) {
    const gsr_view_fwd& r = views[v];
    Geom g(r.geom, in.P);
    GSR_CUDA_CHECK(cudaMemsetAsync(g.giant_count, 0, sizeof(uint32_t), st));
    FwdViewDev& d = fv.v[v];
    d.depths = g.depths; d.means2D = g.means2D; d.conic_opacity = g.conic_opacity; d.rgb_thr = g.rgb_thr;
    d.cov3D = g.cov3D; d.clamped = g.clamped; d.tiles_touched = g.tiles_touched; d.rects = g.rects;
    d.dkeys = g.offsets_sorted; d.radii = r.radii;
    d.viewmatrix = r.viewmatrix; d.projmatrix = r.projmatrix; d.campos = r.campos;
    d.tanfovx = r.tanfovx; d.tanfovy = r.tanfovy; d.W = r.W; d.H = r.H;
  }
  const int nblocks = (in.P + PRE_THREADS - 1) / PRE_THREADS;
  preprocess_fwd_views_kernel<<<nblocks, PRE_THREADS, 0, st>>>(in, fv);
  GSR_LAUNCH_CHECK("preprocess_fwd_views_kernel");
  return GSR_OK;
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st) {
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, view, present);
  GSR_LAUNCH_CHECK("mark_visible_kernel");
  return GSR_OK;
}

}  // namespace gsr
