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

// Packed pairs of f32 (sm_100a FFMA2 / FMUL2 / FADD2: two IEEE round-to-nearest results per issue slot, each
// component identical to the scalar instruction's).  A pair lives in a 64-bit register.
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// acc <- a * b + acc, accumulator tied to one register pair (no copy at control-flow joins)
__device__ __forceinline__ void fma2_acc(unsigned long long& acc, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
// (lo, hi) <- (a_lo, a_hi) * w + (lo, hi) on scalar registers: the pairing is left to ptxas
__device__ __forceinline__ void fma2_bcast(float& lo, float& hi, float a_lo, float a_hi, float w) {
  asm("{ .reg .b64 a, b, c;\n\t"
      "mov.b64 a, {%2, %3};\n\t"
      "mov.b64 b, {%4, %4};\n\t"
      "mov.b64 c, {%0, %1};\n\t"
      "fma.rn.f32x2 c, a, b, c;\n\t"
      "mov.b64 {%0, %1}, c; }" : "+f"(lo), "+f"(hi) : "f"(a_lo), "f"(a_hi), "f"(w));
}
// Component-wise pair arithmetic on scalar registers (one FADD2 / FMUL2 / FFMA2 issue slot for two results).  Each
// component is the IEEE round-to-nearest result of the scalar add / mul / fma, i.e. what add() / mul() / fma_() above
// give when ptxas does not contract them -- use only where the scalar sequence is known NOT to be contracted, or
// where the contraction is spelled out (fma2*).
__device__ __forceinline__ void add2(float& r0, float& r1, float a0, float a1, float b0, float b1) {
  asm("{ .reg .b64 a, b, c;\n\t"
      "mov.b64 a, {%2, %3};\n\t"
      "mov.b64 b, {%4, %5};\n\t"
      "add.rn.f32x2 c, a, b;\n\t"
      "mov.b64 {%0, %1}, c; }" : "=f"(r0), "=f"(r1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ void mul2(float& r0, float& r1, float a0, float a1, float b0, float b1) {
  asm("{ .reg .b64 a, b, c;\n\t"
      "mov.b64 a, {%2, %3};\n\t"
      "mov.b64 b, {%4, %5};\n\t"
      "mul.rn.f32x2 c, a, b;\n\t"
      "mov.b64 {%0, %1}, c; }" : "=f"(r0), "=f"(r1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
// (lo, hi) <- (a0, a1) * (b0, b1) + (lo, hi)
__device__ __forceinline__ void fma2_pair(float& lo, float& hi, float a0, float a1, float b0, float b1) {
  asm("{ .reg .b64 a, b, c;\n\t"
      "mov.b64 a, {%2, %3};\n\t"
      "mov.b64 b, {%4, %5};\n\t"
      "mov.b64 c, {%0, %1};\n\t"
      "fma.rn.f32x2 c, a, b, c;\n\t"
      "mov.b64 {%0, %1}, c; }" : "+f"(lo), "+f"(hi) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
// (r0, r1) <- (a0, a1) * w
__device__ __forceinline__ void mul2_bcast(float& r0, float& r1, float a0, float a1, float w) {
  asm("{ .reg .b64 a, b, c;\n\t"
      "mov.b64 a, {%2, %3};\n\t"
      "mov.b64 b, {%4, %4};\n\t"
      "mul.rn.f32x2 c, a, b;\n\t"
      "mov.b64 {%0, %1}, c; }" : "=f"(r0), "=f"(r1) : "f"(a0), "f"(a1), "f"(w));
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

}  // namespace fpx
