// radix_sort.cu -- hand-written stable onesweep LSD radix sort of (u64 key, u32 value) pairs.
//
// Replaces cub::DeviceRadixSort::SortPairs (DGR/cuda_rasterizer/rasterizer_impl.cu:303-308):
// ascending, stable, on key bits [0, end_bit).  No CUB / Thrust.
//
//   1. one histogram kernel reads the keys once and builds the 256-bin digit histogram of every
//      pass (shared-memory atomics, then one global atomic per bin per block);
//   2. a one-block kernel turns each histogram into exclusive digit offsets;
//   3. one scatter kernel per 8-bit digit: each block takes a 4096-pair tile (dynamic ticket),
//      ranks its keys with warp-level __match_any_sync multi-split into per-warp shared-memory
//      digit counters (stable: item order inside a warp is preserved), chains tile prefixes with a
//      decoupled look-back over per-digit status words, stages the tile in shared memory in
//      sorted order and writes it out in coalesced per-digit runs.
// HBM traffic: 8 B/pair (histogram) + 24 B/pair/pass.
#include <cstdlib>
#include "common.cuh"

namespace gsr {
namespace {

// Look-back status word = 2 flag bits + a count.  32-bit words (30-bit counts) serve n < 2^30; larger sorts
// (up to the 2^32 pairs the reference's u32 offsets allow) use 64-bit words.
template <typename W> struct St;
template <> struct St<uint32_t> {
  static constexpr uint32_t AGG = 1u << 30, PREFIX = 1u << 31, MASK = (1u << 30) - 1;
};
template <> struct St<unsigned long long> {
  static constexpr unsigned long long AGG = 1ull << 62, PREFIX = 1ull << 63, MASK = (1ull << 62) - 1;
};
constexpr int64_t SORT_SMALL_LIMIT = (int64_t)1 << 30;
constexpr int NWARPS = SORT_THREADS / 32;

template <typename KeyT>
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const KeyT* __restrict__ keys, int64_t n, int passes, int end_bit,
                      uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[SORT_MAX_PASSES * 256];
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const KeyT k = keys[i];
    for (int p = 0; p < passes; p++) {
      const int shift = 8 * p, nb = min(8, end_bit - shift);
      atomicAdd(&sh[p * 256 + (uint32_t)((k >> shift) & ((1u << nb) - 1))], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// exclusive scan of each pass's 256 bins, in place; one warp per pass
__global__ void sort_scan_hist_kernel(uint32_t* hist, int passes) {
  const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (p >= passes) return;
  uint32_t* h = hist + p * 256;
  uint32_t v[8], sum = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) { v[k] = h[lane * 8 + k]; sum += v[k]; }
  uint32_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  uint32_t run = incl - sum;
#pragma unroll
  for (int k = 0; k < 8; k++) { h[lane * 8 + k] = run; run += v[k]; }
}

template <typename KeyT>
struct PassSmem {
  KeyT keys[SORT_TILE];
  uint32_t vals[SORT_TILE];
  uint32_t warp_hist[NWARPS][256];
  uint32_t digit_off[256];   // exclusive offset of each digit inside the sorted tile
  int32_t global_base[256];  // destination of sorted tile slot s with digit d is global_base[d] + s
  uint32_t scan_tmp[NWARPS];
  uint32_t tile;
};

#ifndef GSR_SORT_MINBLOCKS
#define GSR_SORT_MINBLOCKS 3
#endif
template <typename KeyT, typename W>
__global__ void __launch_bounds__(SORT_THREADS, GSR_SORT_MINBLOCKS)
onesweep_pass_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, int64_t n, int shift,
                     uint32_t digit_mask, const uint32_t* __restrict__ digit_base, W* status,
                     uint32_t* ticket) {
  constexpr W ST_AGG = St<W>::AGG, ST_PREFIX = St<W>::PREFIX, ST_MASK = St<W>::MASK;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PassSmem<KeyT>& S = *reinterpret_cast<PassSmem<KeyT>*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t ntiles = (uint32_t)((n + SORT_TILE - 1) / SORT_TILE);
  // a block keeps taking tiles (tickets are handed out in order, so look-back cannot deadlock) until none is left;
  // launched with one block per tile this loop runs once, launched with fewer it is a persistent kernel with a
  // bounded footprint per SM (views.ViewPipeline runs the sorts underneath the blending kernels)
  for (;;) {
  __syncthreads();
  if (tid == 0) S.tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < NWARPS; w++) S.warp_hist[w][tid] = 0;
  __syncthreads();
  const uint32_t tile = S.tile;
  if (tile >= ntiles) break;
  const int64_t tile_base = (int64_t)tile * SORT_TILE;
  const int tile_n = (int)min((int64_t)SORT_TILE, n - tile_base);

  // warp-striped load: warp w owns tile items [w*512, (w+1)*512), lane-interleaved
  KeyT key[SORT_ITEMS];
  uint32_t val[SORT_ITEMS], rank[SORT_ITEMS];
  const int wbase = warp * (32 * SORT_ITEMS) + lane;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    const int t = wbase + i * 32;
    if (t < tile_n) { key[i] = keys_in[tile_base + t]; val[i] = vals_in[tile_base + t]; }
    else { key[i] = (KeyT)~(KeyT)0; val[i] = 0; }
  }

  // stable per-warp ranking
  uint32_t* wh = S.warp_hist[warp];
  const uint32_t lt_mask = (1u << lane) - 1;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    const bool valid = wbase + i * 32 < tile_n;
    const uint32_t d = valid ? (uint32_t)(key[i] >> shift) & digit_mask : 0xffffffffu;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t pre = 0;
    if (valid && lane == leader) { pre = wh[d]; wh[d] = pre + __popc(peers); }
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank[i] = pre + __popc(peers & lt_mask);
    __syncwarp();
  }
  __syncthreads();

  // thread t owns digit t: prefix over warps, publish the tile aggregate, look back
  uint32_t count = 0;
#pragma unroll
  for (int w = 0; w < NWARPS; w++) {
    const uint32_t c = S.warp_hist[w][tid];
    S.warp_hist[w][tid] = count;
    count += c;
  }
  volatile W* st = status;
  if (tile == 0) st[tid] = ST_PREFIX | (W)count;
  else st[(size_t)tile * 256 + tid] = ST_AGG | (W)count;
  // exclusive scan of `count` over the 256 digits
  uint32_t incl = count;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) S.scan_tmp[warp] = incl;
  __syncthreads();
  uint32_t wpre = 0;
#pragma unroll
  for (int w = 0; w < NWARPS; w++) wpre += w < warp ? S.scan_tmp[w] : 0u;
  const uint32_t doff = wpre + incl - count;
  S.digit_off[tid] = doff;
  uint32_t excl = 0;
  if (tile != 0) {
    // decoupled look-back, four predecessors per round trip (independent loads); a status word carries both
    // flag and value, so no fence is needed around it
    int64_t j = (int64_t)tile - 1;
    while (true) {
      W w[4];
#pragma unroll
      for (int k = 0; k < 4; k++) { w[k] = ST_PREFIX; if (j - k >= 0) w[k] = st[(size_t)(j - k) * 256 + tid]; }
      bool done = false, stop = false;
      int adv = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        if (!done && !stop) {
          if ((w[k] & (ST_AGG | ST_PREFIX)) == 0) stop = true;
          else { excl += (uint32_t)(w[k] & ST_MASK); adv = k + 1; done = (w[k] & ST_PREFIX) != 0; }
        }
      }
      if (done) break;
      j -= adv;
    }
    st[(size_t)tile * 256 + tid] = ST_PREFIX | ((W)(excl + count) & ST_MASK);
  }
  S.global_base[tid] = (int32_t)(digit_base[tid] + excl) - (int32_t)doff;
  __syncthreads();

  // stage in sorted order
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; i++) {
    if (wbase + i * 32 < tile_n) {
      const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
      const uint32_t pos = S.digit_off[d] + wh[d] + rank[i];
      S.keys[pos] = key[i];
      S.vals[pos] = val[i];
    }
  }
  __syncthreads();
  // coalesced per-digit runs
  for (int s = tid; s < tile_n; s += SORT_THREADS) {
    const KeyT k = S.keys[s];
    const uint32_t d = (uint32_t)(k >> shift) & digit_mask;
    const int64_t dst = (int64_t)(uint32_t)(S.global_base[d] + s);
    keys_out[dst] = k;
    vals_out[dst] = S.vals[s];
  }
  }  // next tile
}

int sort_max_blocks() {  // GSR_SORT_MAX_BLOCKS=n caps the grid of a sort pass (0 / unset: one block per tile)
  static int v = -1;
  if (v < 0) { const char* e = getenv("GSR_SORT_MAX_BLOCKS"); v = e ? atoi(e) : 0; if (v < 0) v = 0; }
  return v;
}

}  // namespace

// Sorts in place across two buffer pairs: pass p reads pair (p & 1) and writes the other; the
// result is in pair (passes & 1).  Returns the pass count through *passes_out.
template <typename KeyT>
static int sort_pairs_pingpong(KeyT* keys0, uint32_t* vals0, KeyT* keys1, uint32_t* vals1, int64_t n, int end_bit,
                               void* temp, int* passes_out, bool hist_ready, cudaStream_t st) {
  const int passes = (end_bit + 7) / 8;
  if (passes_out) *passes_out = passes;
  if (passes < 1 || passes > (int)sizeof(KeyT) || passes > SORT_MAX_PASSES) {
    set_error("sort: end_bit %d out of range", end_bit); return GSR_ERR_INVALID;
  }
  if (n >= ((int64_t)1 << 32)) { set_error("sort: %lld pairs exceed the 2^32 limit", (long long)n); return GSR_ERR_INVALID; }
  if (n == 0) return GSR_OK;
  SortTemp T(temp, n);
  if (!hist_ready) {  // otherwise the caller zeroed `temp` and the producer of keys0 filled T.hist (sort_prepare)
    GSR_CUDA_CHECK(cudaMemsetAsync(temp, 0, T.zero_bytes(passes), st));
    const int hist_blocks = (int)min((int64_t)148 * 8, (n + 2047) / 2048);
    sort_histogram_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys0, n, passes, end_bit, T.hist);
    GSR_LAUNCH_CHECK("sort_histogram_kernel");
  }
  sort_scan_hist_kernel<<<1, 32 * SORT_MAX_PASSES, 0, st>>>(T.hist, passes);
  GSR_LAUNCH_CHECK("sort_scan_hist_kernel");
  const int64_t tiles = sort_tiles(n);
  const bool big = n >= SORT_SMALL_LIMIT;
  const int cap = sort_max_blocks();
  const unsigned grid = (unsigned)(cap > 0 && tiles > cap ? cap : tiles);
  if (big) GSR_CUDA_CHECK(cudaFuncSetAttribute(onesweep_pass_kernel<KeyT, unsigned long long>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PassSmem<KeyT>)));
  else GSR_CUDA_CHECK(cudaFuncSetAttribute(onesweep_pass_kernel<KeyT, uint32_t>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PassSmem<KeyT>)));
  for (int p = 0; p < passes; p++) {
    const int shift = 8 * p, nb = end_bit - shift < 8 ? end_bit - shift : 8;
    const KeyT* kin = (p & 1) ? keys1 : keys0;
    KeyT* kout = (p & 1) ? keys0 : keys1;
    const uint32_t* vin = (p & 1) ? vals1 : vals0;
    uint32_t* vout = (p & 1) ? vals0 : vals1;
    char* stp = (char*)T.status + (size_t)p * T.pass_stride;
    if (big)
      onesweep_pass_kernel<KeyT, unsigned long long><<<grid, SORT_THREADS, sizeof(PassSmem<KeyT>), st>>>(
          kin, kout, vin, vout, n, shift, (1u << nb) - 1, T.hist + p * 256, (unsigned long long*)stp, T.tickets + p);
    else
      onesweep_pass_kernel<KeyT, uint32_t><<<grid, SORT_THREADS, sizeof(PassSmem<KeyT>), st>>>(
          kin, kout, vin, vout, n, shift, (1u << nb) - 1, T.hist + p * 256, (uint32_t*)stp, T.tickets + p);
    GSR_LAUNCH_CHECK("onesweep_pass_kernel");
  }
  return GSR_OK;
}

int launch_sort_pairs_pingpong(uint64_t* keys0, uint32_t* vals0, uint64_t* keys1, uint32_t* vals1, int64_t n,
                               int end_bit, void* temp, int* passes_out, cudaStream_t st) {
  return sort_pairs_pingpong<uint64_t>(keys0, vals0, keys1, vals1, n, end_bit, temp, passes_out, false, st);
}
int launch_sort_pairs_pingpong_u32(uint32_t* keys0, uint32_t* vals0, uint32_t* keys1, uint32_t* vals1, int64_t n,
                                   int end_bit, void* temp, int* passes_out, bool hist_ready, cudaStream_t st) {
  return sort_pairs_pingpong<uint32_t>(keys0, vals0, keys1, vals1, n, end_bit, temp, passes_out, hist_ready, st);
}

// Zeroes the sort's temp block and returns where the producer of the keys must accumulate the digit
// histograms (pass p, digit d -> hist[p * 256 + d]) when it wants to skip the sort's own histogram pass.
int sort_prepare(void* temp, int64_t n, int passes, uint32_t** hist_out, cudaStream_t st) {
  SortTemp T(temp, n);
  GSR_CUDA_CHECK(cudaMemsetAsync(temp, 0, T.zero_bytes(passes), st));
  *hist_out = T.hist;
  return GSR_OK;
}

}  // namespace gsr
