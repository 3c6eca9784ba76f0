// binning.cu -- depth-ordered tile-pair emission, prefix sum in depth order, tile ranges.
//
// Replaces duplicateWithKeys (DGR/cuda_rasterizer/rasterizer_impl.cu:70-111) and the
// cudaMemset + identifyTileRanges pair (rasterizer_impl.cu:116-138, :310-318).  Integer-only, HBM-bound.
// The reference emits 64-bit (tile << 32 | depth bits) keys in Gaussian-index order and sorts them all;
// here the Gaussians are already ordered by (depth bits, index) (geom.perm, a 32-bit sort of P items), pairs
// are emitted in that order with a 32-bit tile id as the only key, and one stable sort by tile id
// (2 passes at 1080p instead of 6) yields the same per-tile order.
#include "common.cuh"

namespace gsr {
namespace {

constexpr int DUP_THREADS = 256;
constexpr uint32_t DUP_SOLO_MAX = 6;  // rects up to this many tiles are written by their own lane

__device__ __forceinline__ unsigned long long pack_status(uint32_t flag, uint32_t v) {
  return ((unsigned long long)flag << 32) | v;
}

// offsets_sorted[i] = sum_{k<=i} tiles_touched[perm[k]]; single pass, decoupled look-back over 1024-item blocks
__global__ void __launch_bounds__(256)
sorted_scan_kernel(int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ tiles_touched,
                   uint32_t* __restrict__ offsets_sorted, unsigned long long* status, uint32_t* ticket) {
  __shared__ uint32_t s_tile, s_warp[8], s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int base = (int)tile * SCAN_ITEMS + tid * 4;
  uint32_t v[4], sum = 0;
  if (base + 3 < P) {
    const uint4 pi = *reinterpret_cast<const uint4*>(perm + base);
    v[0] = tiles_touched[pi.x]; v[1] = tiles_touched[pi.y]; v[2] = tiles_touched[pi.z]; v[3] = tiles_touched[pi.w];
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = base + k < P ? tiles_touched[perm[base + k]] : 0u;
  }
#pragma unroll
  for (int k = 0; k < 4; k++) { sum += v[k]; v[k] = sum; }
  uint32_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < 8 ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    if (lane < 8) s_warp[lane] = w;
  }
  __syncthreads();
  const uint32_t block_total = s_warp[7];
  if (tid == 0) {
    volatile unsigned long long* st = status;
    uint32_t excl = 0;
    if (tile == 0) {
      st[0] = pack_status(2, block_total);
    } else {
      st[tile] = pack_status(1, block_total);
      __threadfence();
      int64_t j = (int64_t)tile - 1;
      while (true) {
        const unsigned long long w = st[j];
        const uint32_t flag = (uint32_t)(w >> 32);
        if (flag == 0) continue;
        excl += (uint32_t)w;
        if (flag == 2) break;
        j--;
      }
      st[tile] = pack_status(2, excl + block_total);
    }
    s_prefix = excl;
  }
  __syncthreads();
  const uint32_t off = s_prefix + (warp ? s_warp[warp - 1] : 0u) + (incl - sum);
  if (base + 3 < P) {
    *reinterpret_cast<uint4*>(offsets_sorted + base) = make_uint4(off + v[0], off + v[1], off + v[2], off + v[3]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) if (base + k < P) offsets_sorted[base + k] = off + v[k];
  }
}

// One lane per Gaussian (in depth order) for small rects; rects larger than DUP_SOLO_MAX are expanded by the
// whole warp (coalesced 32-wide stores), which removes the reference's load imbalance on large splats.
// Tiles of one Gaussian are emitted in row-major order, as the reference does.
__global__ void __launch_bounds__(DUP_THREADS)
duplicate_tiles_kernel(int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ offsets_sorted,
                       const uint32_t* __restrict__ tiles_touched, const ushort4* __restrict__ rects, int gx,
                       uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int i = blockIdx.x * DUP_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t area = 0, off = 0, x0 = 0, y0 = 0, w = 0, id = 0;
  if (i < P) {
    id = perm[i];
    area = tiles_touched[id];
    if (area > 0) {
      const ushort4 r = rects[id];
      x0 = r.x; y0 = r.y; w = (uint32_t)r.z - r.x;
      off = i == 0 ? 0u : offsets_sorted[i - 1];
    }
  }
  if (area > 0 && area <= DUP_SOLO_MAX) {
    uint32_t x = 0, y = 0;
    for (uint32_t t = 0; t < area; t++) {
      keys[off + t] = (y0 + y) * gx + x0 + x;
      vals[off + t] = id;
      if (++x == w) { x = 0; y++; }
    }
  }
  uint32_t big = __ballot_sync(0xffffffffu, area > DUP_SOLO_MAX);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t a = __shfl_sync(0xffffffffu, area, src), o = __shfl_sync(0xffffffffu, off, src);
    const uint32_t bx = __shfl_sync(0xffffffffu, x0, src), by = __shfl_sync(0xffffffffu, y0, src);
    const uint32_t bw = __shfl_sync(0xffffffffu, w, src), bid = __shfl_sync(0xffffffffu, id, src);
    for (uint32_t t = lane; t < a; t += 32) {
      const uint32_t y = t / bw, x = t - y * bw;
      keys[o + t] = (by + y) * gx + bx + x;
      vals[o + t] = bid;
    }
  }
}

// ranges[tile] = [first, one-past-last) index of that tile's run in the sorted tile-id list.
__global__ void tile_ranges_kernel(int64_t L, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const uint32_t cur = keys[i];
  if (i == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = keys[i - 1];
    if (cur != prev) {
      ranges[prev].y = (uint32_t)i;
      ranges[cur].x = (uint32_t)i;
    }
  }
  if (i == L - 1) ranges[cur].y = (uint32_t)L;
}

// debug only: the reference's sorted 64-bit keys, rebuilt from the sorted tile ids and the point list
__global__ void rebuild_keys_kernel(int64_t L, const uint32_t* __restrict__ tile_keys,
                                    const uint32_t* __restrict__ point_list, const float* __restrict__ depths,
                                    uint64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < L) out[i] = ((uint64_t)tile_keys[i] << 32) | __float_as_uint(depths[point_list[i]]);
}

}  // namespace

int launch_sorted_scan(int P, const Geom& g, Scratch& s, cudaStream_t st) {
  sorted_scan_kernel<<<(P + SCAN_ITEMS - 1) / SCAN_ITEMS, 256, 0, st>>>(P, g.perm, g.tiles_touched, g.offsets_sorted,
                                                                        s.scan2_status, s.scan_ticket + 1);
  GSR_LAUNCH_CHECK("sorted_scan_kernel");
  return GSR_OK;
}

int launch_duplicate_tiles(int P, const Geom& g, int gx, uint32_t* tile_keys, uint32_t* vals, cudaStream_t st) {
  duplicate_tiles_kernel<<<(P + DUP_THREADS - 1) / DUP_THREADS, DUP_THREADS, 0, st>>>(
      P, g.perm, g.offsets_sorted, g.tiles_touched, g.rects, gx, tile_keys, vals);
  GSR_LAUNCH_CHECK("duplicate_tiles_kernel");
  return GSR_OK;
}

int launch_rebuild_keys(int64_t L, const uint32_t* tile_keys, const uint32_t* point_list, const float* depths,
                        uint64_t* keys_out, cudaStream_t st) {
  if (L > 0) {
    rebuild_keys_kernel<<<(unsigned)((L + 255) / 256), 256, 0, st>>>(L, tile_keys, point_list, depths, keys_out);
    GSR_LAUNCH_CHECK("rebuild_keys_kernel");
  }
  return GSR_OK;
}

int launch_tile_ranges(int64_t L, const uint32_t* keys_sorted, uint2* ranges, int tiles, cudaStream_t st) {
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
