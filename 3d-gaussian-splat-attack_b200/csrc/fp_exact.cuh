// fp_exact.cuh -- single-precision operations emitted as inline PTX with explicit control over the
// rounding qualifier.  mul/add/sub carry NO rounding qualifier (ptxas may fuse them into FFMA, as it
// may in the reference build); fma_ is fma.rn (never split); div/rcp/sqrt are the IEEE .rn forms.
#pragma once

namespace fpx {

__device__ __forceinline__ float mul(float a, float b) {
  float d;
  asm("mul.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ float add(float a, float b) {
  float d;
  asm("add.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ float sub(float a, float b) {
  float d;
  asm("sub.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ float fma_(float a, float b, float c) {
  float d;
  asm("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float fdiv(float a, float b) {
  float d;
  asm("div.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ float frcp(float a) {
  float d;
  asm("rcp.rn.f32 %0, %1;" : "=f"(d) : "f"(a));
  return d;
}
__device__ __forceinline__ float fsqrt(float a) {
  float d;
  asm("sqrt.rn.f32 %0, %1;" : "=f"(d) : "f"(a));
  return d;
}

}  // namespace fpx
