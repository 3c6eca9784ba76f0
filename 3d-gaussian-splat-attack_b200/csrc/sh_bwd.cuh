// sh_bwd.cuh -- SH colour backward (DGR/cuda_rasterizer/backward.cu:20-139 computeColorFromSH backward) and small
// helpers shared by the per-view and the multi-view per-Gaussian backward kernels.
#pragma once
#include "common.cuh"

namespace gsr {
namespace {

constexpr float C0 = 0.28209479177387814f;
constexpr float C1 = 0.4886025119029199f;
__constant__ float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                            -1.0925484305920792f, 0.5462742152960396f};
__constant__ float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                            0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                            -0.5900435899266435f};

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 operator*(float s, V3 v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// dL/dsh (written through `put`) and dL/d(view direction) for SH degree <= 3.
template <typename Get, typename Put>
__device__ __forceinline__ V3 sh_backward(int deg, float x, float y, float z, V3 dRGB, Get sh, Put put) {
  V3 gx{0, 0, 0}, gy{0, 0, 0}, gz{0, 0, 0};
  put(0, C0);
  if (deg > 0) {
    put(1, -C1 * y); put(2, C1 * z); put(3, -C1 * x);
    gx = -C1 * sh(3); gy = -C1 * sh(1); gz = C1 * sh(2);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      put(4, C2[0] * xy); put(5, C2[1] * yz); put(6, C2[2] * (2.f * zz - xx - yy));
      put(7, C2[3] * xz); put(8, C2[4] * (xx - yy));
      gx = gx + (C2[0] * y) * sh(4) + (C2[2] * 2.f * -x) * sh(6) + (C2[3] * z) * sh(7) + (C2[4] * 2.f * x) * sh(8);
      gy = gy + (C2[0] * x) * sh(4) + (C2[1] * z) * sh(5) + (C2[2] * 2.f * -y) * sh(6) + (C2[4] * 2.f * -y) * sh(8);
      gz = gz + (C2[1] * y) * sh(5) + (C2[2] * 4.f * z) * sh(6) + (C2[3] * x) * sh(7);
      if (deg > 2) {
        put(9, C3[0] * y * (3.f * xx - yy)); put(10, C3[1] * xy * z);
        put(11, C3[2] * y * (4.f * zz - xx - yy)); put(12, C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
        put(13, C3[4] * x * (4.f * zz - xx - yy)); put(14, C3[5] * z * (xx - yy));
        put(15, C3[6] * x * (xx - 3.f * yy));
        gx = gx + (C3[0] * 6.f * xy) * sh(9) + (C3[1] * yz) * sh(10) + (C3[2] * -2.f * xy) * sh(11) +
             (C3[3] * -6.f * xz) * sh(12) + (C3[4] * (-3.f * xx + 4.f * zz - yy)) * sh(13) +
             (C3[5] * 2.f * xz) * sh(14) + (C3[6] * 3.f * (xx - yy)) * sh(15);
        gy = gy + (C3[0] * 3.f * (xx - yy)) * sh(9) + (C3[1] * xz) * sh(10) +
             (C3[2] * (-3.f * yy + 4.f * zz - xx)) * sh(11) + (C3[3] * -6.f * yz) * sh(12) +
             (C3[4] * -2.f * xy) * sh(13) + (C3[5] * -2.f * yz) * sh(14) + (C3[6] * -6.f * xy) * sh(15);
        gz = gz + (C3[1] * xy) * sh(10) + (C3[2] * 8.f * yz) * sh(11) + (C3[3] * 3.f * (2.f * zz - xx - yy)) * sh(12) +
             (C3[4] * 8.f * xz) * sh(13) + (C3[5] * (xx - yy)) * sh(14);
      }
    }
  }
  return V3{dot(gx, dRGB), dot(gy, dRGB), dot(gz, dRGB)};
}

__device__ __forceinline__ void red_add4(float4* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

}  // namespace
}  // namespace gsr
