// binning.cu -- tile-key duplication and tile-range identification.
//
// Replaces duplicateWithKeys (DGR/cuda_rasterizer/rasterizer_impl.cu:70-111) and the
// cudaMemset + identifyTileRanges pair (rasterizer_impl.cu:116-138, :310-318).
// Integer-only; both are HBM-bound.  Keys are (tile_id << 32) | float_bits(depth), values the
// Gaussian index, emitted per Gaussian in row-major tile order starting at point_offsets[idx-1]:
// that emission order is part of the result (the sort is stable, so ties keep it).
#include "common.cuh"

namespace gsr {
namespace {

constexpr int DUP_THREADS = 256;
constexpr uint32_t DUP_SOLO_MAX = 6;  // rects up to this many tiles are written by their own lane

// One lane per Gaussian for small rects; rects larger than DUP_SOLO_MAX are expanded by the whole
// warp (coalesced 32-wide stores), which removes the reference's load imbalance on large splats.
__global__ void __launch_bounds__(DUP_THREADS)
duplicate_keys_kernel(int P, const float* __restrict__ depths, const uint32_t* __restrict__ offsets,
                      const ushort4* __restrict__ rects, const int32_t* __restrict__ radii, int gx,
                      uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int idx = blockIdx.x * DUP_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t area = 0, off = 0, dbits = 0, x0 = 0, y0 = 0, w = 0;
  if (idx < P && radii[idx] > 0) {
    const ushort4 r = rects[idx];
    x0 = r.x; y0 = r.y; w = (uint32_t)r.z - r.x;
    area = w * ((uint32_t)r.w - r.y);
    off = idx == 0 ? 0u : offsets[idx - 1];
    dbits = __float_as_uint(depths[idx]);
  }
  if (area > 0 && area <= DUP_SOLO_MAX) {
    uint32_t x = 0, y = 0;
    for (uint32_t t = 0; t < area; t++) {
      keys[off + t] = ((uint64_t)((y0 + y) * gx + x0 + x) << 32) | dbits;
      vals[off + t] = (uint32_t)idx;
      if (++x == w) { x = 0; y++; }
    }
  }
  uint32_t big = __ballot_sync(0xffffffffu, area > DUP_SOLO_MAX);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t a = __shfl_sync(0xffffffffu, area, src), o = __shfl_sync(0xffffffffu, off, src);
    const uint32_t bx = __shfl_sync(0xffffffffu, x0, src), by = __shfl_sync(0xffffffffu, y0, src);
    const uint32_t bw = __shfl_sync(0xffffffffu, w, src), db = __shfl_sync(0xffffffffu, dbits, src);
    const uint32_t id = (uint32_t)(blockIdx.x * DUP_THREADS + (threadIdx.x & ~31) + src);
    for (uint32_t t = lane; t < a; t += 32) {
      const uint32_t y = t / bw, x = t - y * bw;
      keys[o + t] = ((uint64_t)((by + y) * gx + bx + x) << 32) | db;
      vals[o + t] = id;
    }
  }
}

// ranges[tile] = [first, one-past-last) index of that tile's run in the sorted key list.
__global__ void tile_ranges_kernel(int64_t L, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const uint32_t cur = (uint32_t)(keys[i] >> 32);
  if (i == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = (uint32_t)i;
      ranges[cur].x = (uint32_t)i;
    }
  }
  if (i == L - 1) ranges[cur].y = (uint32_t)L;
}

}  // namespace

int launch_duplicate_keys(int P, const Geom& g, const int32_t* radii, int gx, uint64_t* keys, uint32_t* vals,
                          cudaStream_t st) {
  duplicate_keys_kernel<<<(P + DUP_THREADS - 1) / DUP_THREADS, DUP_THREADS, 0, st>>>(
      P, g.depths, g.point_offsets, g.rects, radii, gx, keys, vals);
  GSR_LAUNCH_CHECK("duplicate_keys_kernel");
  return GSR_OK;
}

int launch_tile_ranges(int64_t L, const uint64_t* keys_sorted, uint2* ranges, int tiles, cudaStream_t st) {
  GSR_CUDA_CHECK(cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), st));
  if (L > 0) {
    tile_ranges_kernel<<<(unsigned)((L + 255) / 256), 256, 0, st>>>(L, keys_sorted, ranges);
    GSR_LAUNCH_CHECK("tile_ranges_kernel");
  }
  return GSR_OK;
}

uint32_t higher_msb(uint32_t n) {
  // smallest b with (n >> b) == 0, found by bisection (getHigherMsb, rasterizer_impl.cu:35-50)
  uint32_t msb = 16, step = 16;
  while (step > 1) {
    step >>= 1;
    msb = (n >> msb) ? msb + step : msb - step;
  }
  return (n >> msb) ? msb + 1 : msb;
}

}  // namespace gsr
