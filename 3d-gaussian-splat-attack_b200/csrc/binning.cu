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
constexpr uint32_t DUP_MED_MAX = 32;  // up to this many: four rects at a time, a quarter warp each; above: the whole warp

__device__ __forceinline__ unsigned long long pack_status(uint32_t flag, uint32_t v) {
  return ((unsigned long long)flag << 32) | v;
}

// offsets[i] = sum_{k<=i} tiles_touched[perm ? perm[k] : k]; single pass over 1024-item blocks, decoupled
// look-back done by a whole warp (32 predecessors per round trip).  The last block also stores the grand
// total (= num_rendered, rasterizer_impl.cu:281) when `total_out` is given.
__global__ void __launch_bounds__(256)
scan_tiles_kernel(int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ tiles_touched,
                  uint32_t* __restrict__ offsets, unsigned long long* status, uint32_t* ticket,
                  uint32_t* __restrict__ total_out, volatile uint32_t* total_host, int nblocks,
                  uint32_t* __restrict__ giant_count, uint4* __restrict__ giants, uint32_t* __restrict__ giants_w,
                  const ushort4* __restrict__ rects) {
  __shared__ uint32_t s_tile, s_warp[8], s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);  // blocks are numbered in start order -> look-back cannot deadlock
  __syncthreads();
  const uint32_t tile = s_tile;
  constexpr int V = SCAN_ITEMS / 256;  // consecutive items per thread (a multiple of 4: 16-byte loads)
  const int base = (int)tile * SCAN_ITEMS + tid * V;
  uint32_t v[V], sum = 0;
  if (base + V - 1 < P) {
#pragma unroll
    for (int q = 0; q < V / 4; q++) {
      if (perm) {
        const uint4 pi = *reinterpret_cast<const uint4*>(perm + base + 4 * q);
        v[4 * q] = tiles_touched[pi.x]; v[4 * q + 1] = tiles_touched[pi.y];
        v[4 * q + 2] = tiles_touched[pi.z]; v[4 * q + 3] = tiles_touched[pi.w];
      } else {
        const uint4 t = *reinterpret_cast<const uint4*>(tiles_touched + base + 4 * q);
        v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
      }
    }
  } else {
#pragma unroll
    for (int k = 0; k < V; k++) v[k] = base + k < P ? tiles_touched[perm ? perm[base + k] : (uint32_t)(base + k)] : 0u;
  }
  uint32_t raw[V];
#pragma unroll
  for (int k = 0; k < V; k++) raw[k] = v[k];
#pragma unroll
  for (int k = 0; k < V; k++) { sum += v[k]; v[k] = sum; }
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
    const uint32_t block_total = __shfl_sync(0xffffffffu, w, 7);
    volatile unsigned long long* st = status;
    uint32_t excl = 0;
    if (tile == 0) {
      if (lane == 0) st[0] = pack_status(2, block_total);
    } else {
      if (lane == 0) st[tile] = pack_status(1, block_total);
      int64_t j = (int64_t)tile - 1 - lane;  // lane 0 looks at the nearest predecessor
      while (true) {
        const unsigned long long w2 = j >= 0 ? st[j] : pack_status(2, 0);  // before block 0: prefix 0
        const uint32_t flag = (uint32_t)(w2 >> 32);
        const uint32_t ready = __ballot_sync(0xffffffffu, flag != 0);
        const uint32_t pref = __ballot_sync(0xffffffffu, flag == 2);
        const uint32_t need = pref ? ((2u << (__ffs(pref) - 1)) - 1u) : 0xffffffffu;  // lanes up to the first prefix
        if ((ready & need) != need) continue;  // a needed predecessor has not published yet: look again
        uint32_t part = (need >> lane) & 1u ? (uint32_t)w2 : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        excl += part;
        if (pref) break;
        j -= 32;
      }
      if (lane == 0) st[tile] = pack_status(2, excl + block_total);
    }
    if (lane == 0) {
      s_prefix = excl;
      if (total_out && (int)tile == nblocks - 1) {
        *total_out = excl + block_total;
        // the count also goes to pinned (device-mapped) host memory with a plain store instead of a D2H memcpy: a
        // memcpy would queue on the copy engine behind any bulk transfer another stream has in flight
        // (pipeline.HostStepPipeline drains 236 MB of gradients while the next step's forward needs this word)
        if (total_host) { *total_host = excl + block_total; __threadfence_system(); }
      }
    }
  }
  __syncthreads();
  const uint32_t off = s_prefix + (warp ? s_warp[warp - 1] : 0u) + (incl - sum);
  if (giants) {  // complete records of the rects that all blocks of the emission kernel expand together
#pragma unroll
    for (int k = 0; k < V; k++)
      if (raw[k] > GIANT_TILES) {
        const uint32_t slot = atomicAdd(giant_count, 1u);
        if (slot < (uint32_t)GIANT_CAP) {
          const uint32_t gid = perm[base + k];
          const ushort4 r = rects[gid];
          giants[slot] = make_uint4(gid, off + v[k] - raw[k], raw[k], (uint32_t)r.x | ((uint32_t)r.y << 16));
          giants_w[slot] = (uint32_t)r.z - r.x;
        }
      }
  }
  if (base + V - 1 < P) {
#pragma unroll
    for (int q = 0; q < V / 4; q++)
      *reinterpret_cast<uint4*>(offsets + base + 4 * q) =
          make_uint4(off + v[4 * q], off + v[4 * q + 1], off + v[4 * q + 2], off + v[4 * q + 3]);
  } else {
#pragma unroll
    for (int k = 0; k < V; k++) if (base + k < P) offsets[base + k] = off + v[k];
  }
}

// Emits the (tile id, Gaussian) pairs in depth order.  One lane per Gaussian for small rects (<= 6 tiles), four
// medium rects (<= 32 tiles: 40 % of the pairs of C2, 43 % of C4) at a time on a quarter warp each, larger ones
// expanded by the whole warp (coalesced 32-wide stores): no load imbalance on large splats as in the reference.  Tiles of one Gaussian are emitted in row-major order, as the reference does.
// Blocks stride over 256-Gaussian chunks and also build the digit histograms of the tile-id sort that follows
// (hist[p * 256 + digit of pass p]), so that sort needs no histogram pass of its own.
__global__ void __launch_bounds__(DUP_THREADS)
duplicate_tiles_kernel(int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ offsets_sorted,
                       const uint32_t* __restrict__ tiles_touched, const ushort4* __restrict__ rects, int gx,
                       uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ hist,
                       uint32_t mask0, int shift1, uint32_t mask1, uint32_t cap,
                       const uint32_t* __restrict__ giant_count, const uint4* __restrict__ giants,
                       const uint32_t* __restrict__ giants_w) {
  // hist != nullptr: the tile sort has at most two digits, (key & mask0) and ((key >> shift1) & mask1), both
  // <= 8 bits wide (launch_duplicate_tiles): sh[] has two 256-bin rows.  Pairs beyond `cap` (the capacity of a
  // speculative launch that was sized before the count was known) are neither written nor counted.
  __shared__ uint32_t sh[2 * 256];
  for (int k = threadIdx.x; k < 2 * 256; k += DUP_THREADS) sh[k] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  auto emit = [&](uint32_t pos, uint32_t key, uint32_t gid) {
    if (pos < cap) {
      keys[pos] = key;
      vals[pos] = gid;
      if (hist) {
        atomicAdd(&sh[key & mask0], 1u);
        atomicAdd(&sh[256 + ((key >> shift1) & mask1)], 1u);
      }
    }
  };
  // rects above GIANT_TILES were listed by the scan: they are left out of the per-warp paths below and expanded by
  // all blocks together at the end (unless the list overflowed: then every rect is handled where it is met)
  const uint32_t n_giants = __ldg(giant_count);
  const uint32_t giant_min = n_giants <= (uint32_t)GIANT_CAP ? GIANT_TILES : 0xffffffffu;
  const int nchunks = (P + DUP_THREADS - 1) / DUP_THREADS;
  for (int chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
    const int i = chunk * DUP_THREADS + threadIdx.x;
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
        emit(off + t, (y0 + y) * gx + x0 + x, id);
        if (++x == w) { x = 0; y++; }
      }
    }
    // medium rects (DUP_SOLO_MAX < area <= DUP_MED_MAX): four at a time, each expanded by a quarter warp
    uint32_t med = __ballot_sync(0xffffffffu, area > DUP_SOLO_MAX && area <= DUP_MED_MAX);
    const int sub = lane >> 3, sl = lane & 7;
    while (med) {
      // the sub-th lowest set bit of `med` is this quarter's Gaussian (none left: src = -1)
      uint32_t m = med;
      int src = -1;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int f = __ffs(m) - 1;
        if (q == sub) src = f;
        m &= m - 1;
      }
      med = m;
      const int s0 = src < 0 ? 0 : src;
      const uint32_t a_src = __shfl_sync(0xffffffffu, area, s0), o = __shfl_sync(0xffffffffu, off, s0);
      const uint32_t a = src < 0 ? 0u : a_src;
      const uint32_t bx = __shfl_sync(0xffffffffu, x0, s0), by = __shfl_sync(0xffffffffu, y0, s0);
      const uint32_t bw = __shfl_sync(0xffffffffu, w, s0), bid = __shfl_sync(0xffffffffu, id, s0);
      const float rw = 1.0f / (float)bw;
      const uint32_t amax = __reduce_max_sync(0xffffffffu, a);
      for (uint32_t t0 = 0; t0 < amax; t0 += 8) {
        const uint32_t t = t0 + sl;
        const uint32_t y = (uint32_t)(((float)t + 0.5f) * rw), x = t - y * bw;  // exact: t, bw <= 32
        if (t < a) emit(o + t, (by + y) * gx + bx + x, bid);
      }
    }
    // large rects: the whole warp expands one at a time (coalesced 32-wide stores)
    uint32_t big = __ballot_sync(0xffffffffu, area > DUP_MED_MAX && area <= giant_min);
    while (big) {
      const int src = __ffs(big) - 1;
      big &= big - 1;
      const uint32_t a = __shfl_sync(0xffffffffu, area, src), o = __shfl_sync(0xffffffffu, off, src);
      const uint32_t bx = __shfl_sync(0xffffffffu, x0, src), by = __shfl_sync(0xffffffffu, y0, src);
      const uint32_t bw = __shfl_sync(0xffffffffu, w, src), bid = __shfl_sync(0xffffffffu, id, src);
      for (uint32_t t0 = 0; t0 < a; t0 += 32) {
        const uint32_t t = t0 + lane;
        const uint32_t y = t / bw, x = t - y * bw;
        if (t < a) emit(o + t, (by + y) * gx + bx + x, bid);
      }
    }
  }
  if (giant_min != 0xffffffffu) {
    // one listed rect per block at a time, expanded by all 256 threads (coalesced stores)
    for (uint32_t gi = blockIdx.x; gi < n_giants; gi += gridDim.x) {
      const uint4 rec = giants[gi];  // (Gaussian, first output slot, tiles, x0 | y0 << 16)
      const uint32_t bw = giants_w[gi], bx = rec.w & 0xffffu, by = rec.w >> 16;
      for (uint32_t t = threadIdx.x; t < rec.z; t += DUP_THREADS) {
        const uint32_t y = t / bw, x = t - y * bw;
        emit(rec.y + t, (by + y) * gx + bx + x, rec.x);
      }
    }
  }
  __syncthreads();
  if (hist) {
    for (int k = threadIdx.x; k < 2 * 256; k += DUP_THREADS)
      if (sh[k]) atomicAdd(&hist[k], sh[k]);
  }
}

// ranges[tile] = [first, one-past-last) index of that tile's run in the sorted tile-id list; four list entries per
// thread (one 16-byte load).  The list length may live on the device (n_dev, bounded by n_cap).
__global__ void __launch_bounds__(256)
tile_ranges_kernel(int64_t n_cap, const uint32_t* __restrict__ n_dev, const uint32_t* __restrict__ keys,
                   uint2* __restrict__ ranges) {
  int64_t n = n_cap;
  if (n_dev) { const int64_t v = (int64_t)__ldg(n_dev); n = v < n_cap ? v : n_cap; }
  const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i0 >= n) return;
  uint32_t k[5];
  k[0] = i0 ? keys[i0 - 1] : 0u;
  if (i0 + 3 < n) {
    const uint4 v = *reinterpret_cast<const uint4*>(keys + i0);
    k[1] = v.x; k[2] = v.y; k[3] = v.z; k[4] = v.w;
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++) k[j + 1] = i0 + j < n ? keys[i0 + j] : 0u;
  }
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int64_t i = i0 + j;
    if (i >= n) break;
    const uint32_t cur = k[j + 1], prev = k[j];
    if (i == 0) {
      ranges[cur].x = 0;
    } else if (cur != prev) {
      ranges[prev].y = (uint32_t)i;
      ranges[cur].x = (uint32_t)i;
    }
    if (i == n - 1) ranges[cur].y = (uint32_t)n;
  }
}

// debug only: the reference's sorted 64-bit keys, rebuilt from the sorted tile ids and the point list
__global__ void rebuild_keys_kernel(int64_t L, const uint32_t* __restrict__ tile_keys,
                                    const uint32_t* __restrict__ point_list, const float* __restrict__ depths,
                                    uint64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < L) out[i] = ((uint64_t)tile_keys[i] << 32) | __float_as_uint(depths[point_list[i]]);
}

}  // namespace

int launch_sorted_scan(int P, const Geom& g, Scratch& s, uint32_t* num_rendered_dev, uint32_t* num_rendered_host,
                       cudaStream_t st) {
  const int nb = (P + SCAN_ITEMS - 1) / SCAN_ITEMS;
  scan_tiles_kernel<<<nb, 256, 0, st>>>(P, g.perm, g.tiles_touched, g.offsets_sorted, s.scan2_status, s.scan_ticket + 1,
                                        num_rendered_dev, num_rendered_host, nb, g.giant_count, g.giants, g.giants_w,
                                        g.rects);
  GSR_LAUNCH_CHECK("scan_tiles_kernel");
  return GSR_OK;
}

// The reference's point_offsets (InclusiveSum in index order, rasterizer_impl.cu:277); parity tests only.
int launch_index_scan(int P, const Geom& g, Scratch& s, uint32_t* out, cudaStream_t st) {
  const int nb = (P + SCAN_ITEMS - 1) / SCAN_ITEMS;
  GSR_CUDA_CHECK(cudaMemsetAsync(s.scan_status, 0, (size_t)(nb + 1) * sizeof(unsigned long long), st));
  GSR_CUDA_CHECK(cudaMemsetAsync(s.scan_ticket, 0, sizeof(uint32_t), st));
  scan_tiles_kernel<<<nb, 256, 0, st>>>(P, nullptr, g.tiles_touched, out, s.scan_status, s.scan_ticket, nullptr, nullptr, nb,
                                        nullptr, nullptr, nullptr, nullptr);
  GSR_LAUNCH_CHECK("scan_tiles_kernel");
  return GSR_OK;
}

// hist != nullptr (only when the tile sort's plan has at most two digits): the sort's digit histograms are built
// here.  cap = capacity of tile_keys / vals in pairs.
int launch_duplicate_tiles(int P, const Geom& g, int gx, uint32_t* tile_keys, uint32_t* vals, uint32_t* hist,
                           const SortPlan& plan, int64_t cap, cudaStream_t st) {
  const int nchunks = (P + DUP_THREADS - 1) / DUP_THREADS;
  const int blocks = nchunks < 148 * 6 ? nchunks : 148 * 6;
  if (hist && plan.passes > 2) { set_error("duplicate_tiles: fused histograms need <= 2 digits"); return GSR_ERR_INVALID; }
  const uint32_t mask0 = (1u << plan.bits[0]) - 1u;
  const int shift1 = plan.passes > 1 ? plan.shift[1] : 0;
  const uint32_t mask1 = plan.passes > 1 ? (1u << plan.bits[1]) - 1u : 0u;
  const uint32_t cap32 = cap > 0xffffffffll ? 0xffffffffu : (uint32_t)cap;
  duplicate_tiles_kernel<<<blocks, DUP_THREADS, 0, st>>>(P, g.perm, g.offsets_sorted, g.tiles_touched, g.rects, gx,
                                                         tile_keys, vals, hist, mask0, shift1, mask1, cap32,
                                                         g.giant_count, g.giants, g.giants_w);
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

int launch_tile_ranges(int64_t L_cap, const uint32_t* n_dev, const uint32_t* keys_sorted, uint2* ranges, int tiles,
                       cudaStream_t st) {
  GSR_CUDA_CHECK(cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), st));
  if (L_cap > 0) {
    tile_ranges_kernel<<<(unsigned)((L_cap + 1023) / 1024), 256, 0, st>>>(L_cap, n_dev, keys_sorted, ranges);
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
