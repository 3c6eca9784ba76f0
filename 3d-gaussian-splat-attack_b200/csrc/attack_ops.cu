// attack_ops.cu -- the elementwise steps either side of the rasterizer in an attack iteration (SURVEY.md 8(f)
// N1 / N4), as HBM-bound CUDA kernels on the flat attribute / gradient buffers.
//
//   PGD updates        attack.py:25-173 (gaussian_*_linf_attack, gaussian_*_l2_attack): the reference runs 6-9
//                      torch elementwise kernels (and two norms) per attribute per iteration; here one kernel
//                      per attribute (plus one sum-of-squares reduction for the L2 variants).
//   image hand-off     attack.py:438-443,538-543: clamp(x,0,1)*255 -> uint8 HWC; detectors/detectron2_detector.py:78:
//                      clip(x*255 + 0.5, 0, 255) and its backward.
// The L-inf update and the image kernels round exactly like the torch sequences they replace (separate
// multiply / add, no contraction); the L2 update differs from torch only in the summation order of the norms.
#include "common.cuh"

namespace gsr {
namespace {

constexpr int EW_THREADS = 256;
inline int ew_blocks(int64_t n_items) {
  const int64_t b = (n_items + EW_THREADS - 1) / EW_THREADS;
  return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

__device__ __forceinline__ float sign_of(float g) { return g > 0.f ? 1.f : (g < 0.f ? -1.f : g); }  // nan -> nan, 0 -> 0

// p <- clamp((p + -(alpha sign(g))) - p0, -eps, eps) + p0        (attack.py:25-31 and siblings)
__device__ __forceinline__ float linf_one(float p, float g, float p0, float alpha, float eps) {
  const float eta = -__fmul_rn(alpha, sign_of(g));
  const float d = __fsub_rn(__fadd_rn(p, eta), p0);
  return __fadd_rn(fminf(fmaxf(d, -eps), eps), p0);
}

__global__ void __launch_bounds__(EW_THREADS)
pgd_linf_kernel(float* __restrict__ p, const float* __restrict__ g, const float* __restrict__ p0, int64_t n,
                float alpha, float eps) {
  // 16-byte accesses only when all three pointers allow them (slices of a flat buffer may start anywhere)
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(p0)) & 15) == 0;
  const int64_t n4 = vec ? n >> 2 : 0, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 a = reinterpret_cast<float4*>(p)[i];
    const float4 b = reinterpret_cast<const float4*>(g)[i], c = reinterpret_cast<const float4*>(p0)[i];
    a.x = linf_one(a.x, b.x, c.x, alpha, eps); a.y = linf_one(a.y, b.y, c.y, alpha, eps);
    a.z = linf_one(a.z, b.z, c.z, alpha, eps); a.w = linf_one(a.w, b.w, c.w, alpha, eps);
    reinterpret_cast<float4*>(p)[i] = a;
  }
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = linf_one(p[i], g[i], p0[i], alpha, eps);
}

// out[0] += sum g^2 (double accumulator; one atomic per block)
__global__ void __launch_bounds__(EW_THREADS)
sumsq_kernel(const float* __restrict__ g, int64_t n, double* __restrict__ out) {
  __shared__ double sh[EW_THREADS / 32];
  double acc = 0.0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = g[i];
    acc += (double)v * (double)v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < EW_THREADS / 32; w++) t += sh[w];
    atomicAdd(out, t);
  }
}

// One row (= one Gaussian's slice of the attribute, `width` floats) per thread:
//   p <- p - alpha g / ||g||        (zero step when ||g|| == 0)
//   d  = p - p0 ; rows with ||d|| > eps are scaled by eps / (||d|| + 1e-7)   (Tensor.renorm(2, 0, eps))
//   p <- p0 + d                                                    (attack.py:57-75 and siblings)
__global__ void __launch_bounds__(EW_THREADS)
pgd_l2_kernel(float* __restrict__ p, const float* __restrict__ g, const float* __restrict__ p0, int64_t rows, int width,
              float alpha, float eps, const double* __restrict__ sumsq) {
  const float norm = (float)sqrt(*sumsq);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    float* pr = p + r * width;
    const float* gr = g + r * width;
    const float* qr = p0 + r * width;
    float ss = 0.f;
    for (int k = 0; k < width; k++) {
      const float eta = norm > 0.f ? -__fmul_rn(alpha, __fdiv_rn(gr[k], norm)) : 0.f;
      const float d = __fsub_rn(__fadd_rn(pr[k], eta), qr[k]);
      pr[k] = d;  // parked; rescaled below
      ss = fmaf(d, d, ss);
    }
    const float dn = sqrtf(ss);
    const float s = dn > eps ? eps / (dn + 1e-7f) : 1.f;
    for (int k = 0; k < width; k++) pr[k] = __fadd_rn(qr[k], __fmul_rn(pr[k], s));
  }
}

// uint8 HWC <- trunc(clamp(x, 0, 1) * 255) of a float CHW image
__global__ void __launch_bounds__(EW_THREADS)
image_u8_hwc_kernel(const float* __restrict__ chw, uint8_t* __restrict__ hwc, int64_t HW) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += stride) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float v = __fmul_rn(fminf(fmaxf(chw[c * HW + i], 0.f), 1.f), 255.f);
      hwc[3 * i + c] = (uint8_t)v;  // nan clamps to 0 like torch's clamp(min, max) then byte()
    }
  }
}

// y = clip(x * 255 + 0.5, 0, 255)   and   dL/dx = 255 dL/dy where 0 <= x*255+0.5 <= 255 (torch clamp backward)
__global__ void __launch_bounds__(EW_THREADS)
detector_scale_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ dy,
                      float* __restrict__ dx, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float t = __fadd_rn(__fmul_rn(x[i], 255.f), 0.5f);
    if (y) y[i] = fminf(fmaxf(t, 0.f), 255.f);
    if (dx) dx[i] = (t >= 0.f && t <= 255.f) ? __fmul_rn(dy[i], 255.f) : 0.f;
  }
}

}  // namespace

int launch_pgd_linf(float* p, const float* g, const float* p0, int64_t n, float alpha, float eps, cudaStream_t st) {
  if (n == 0) return GSR_OK;
  pgd_linf_kernel<<<ew_blocks((n + 3) / 4), EW_THREADS, 0, st>>>(p, g, p0, n, alpha, eps);
  GSR_LAUNCH_CHECK("pgd_linf_kernel");
  return GSR_OK;
}

int launch_pgd_l2(float* p, const float* g, const float* p0, int64_t rows, int width, float alpha, float eps,
                  double* sumsq_scratch, cudaStream_t st) {
  if (rows == 0 || width == 0) return GSR_OK;
  GSR_CUDA_CHECK(cudaMemsetAsync(sumsq_scratch, 0, sizeof(double), st));
  sumsq_kernel<<<ew_blocks(rows * width / 4 + 1), EW_THREADS, 0, st>>>(g, rows * width, sumsq_scratch);
  GSR_LAUNCH_CHECK("sumsq_kernel");
  pgd_l2_kernel<<<ew_blocks(rows), EW_THREADS, 0, st>>>(p, g, p0, rows, width, alpha, eps, sumsq_scratch);
  GSR_LAUNCH_CHECK("pgd_l2_kernel");
  return GSR_OK;
}

int launch_image_u8_hwc(const float* chw, uint8_t* hwc, int W, int H, cudaStream_t st) {
  const int64_t HW = (int64_t)W * H;
  if (HW == 0) return GSR_OK;
  image_u8_hwc_kernel<<<ew_blocks(HW), EW_THREADS, 0, st>>>(chw, hwc, HW);
  GSR_LAUNCH_CHECK("image_u8_hwc_kernel");
  return GSR_OK;
}

int launch_detector_scale(const float* x, float* y, const float* dy, float* dx, int64_t n, cudaStream_t st) {
  if (n == 0) return GSR_OK;
  detector_scale_kernel<<<ew_blocks(n), EW_THREADS, 0, st>>>(x, y, dy, dx, n);
  GSR_LAUNCH_CHECK("detector_scale_kernel");
  return GSR_OK;
}

}  // namespace gsr
