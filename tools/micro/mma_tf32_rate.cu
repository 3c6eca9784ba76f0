// Micro-benchmark: issue rate of the legacy mma.sync.m16n8k8 TF32 path on sm_100a (used to decide whether the
// object-channel accumulation of blend_fwd<16> can move to it).  nvcc -arch=sm_100a -o mma_rate mma_tf32_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, int iters, long long* cyc) {
  float c[8][4];
  for (int i = 0; i < 8; i++) for (int j = 0; j < 4; j++) c[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; i++) for (int j = 0; j < 4; j++) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 4 << 20); cudaMalloc(&cyc, 8);
  for (int warps = 4; warps <= 32; warps *= 2) {
    const int iters = 2000;
    k<<<148, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    k<<<148, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    // per SM sub-partition: warps/4 warps, each 8*iters MMAs
    printf("warps/SM %2d: %lld cycles, %.2f cycles per MMA per sub-partition\n", warps, h, (double)h / (8.0 * iters * warps / 4));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
