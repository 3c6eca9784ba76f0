// radix_sort.cu -- hand-written stable onesweep LSD radix sort of (key, u32 value) pairs.
//
// Replaces cub::DeviceRadixSort::SortPairs (DGR/cuda_rasterizer/rasterizer_impl.cu:303-308):
// ascending, stable, on key bits [0, end_bit).  No CUB / Thrust.
//
//   1. digit histograms of every pass: built by the producer of the keys where that is cheaper
//      (duplicate_tiles_kernel), else by sort_histogram_kernel (one read of the keys, few long-running blocks);
//   2. one scatter kernel per digit (SortPlan: ceil(end_bit / 8) digits of nearly equal width, e.g. 7 + 6
//      bits for the 13-bit tile ids of a 1080p image: fewer bins mean longer coalesced runs per digit in
//      the scatter).  Each block takes a tile of 256 x ITEMS pairs (dynamic ticket), ranks its keys per warp
//      through a shared-memory peer table (stable: item order inside a warp is preserved), stages the keys in
//      sorted order, requests the values, chains tile prefixes with a two-level decoupled look-back over
//      per-digit status words and writes the tile out in per-digit runs.  Every block redoes the exclusive
//      scan of the 256-bin histogram itself (no kernel launch between the passes).
// The pair count may live on the device (n_dev): the grid is then sized from a capacity and surplus blocks
// leave at once -- the forward pass sorts the (tile, Gaussian) list without the host knowing its length.
// HBM traffic: 4 B/pair for a separate histogram of u32 keys + 16 B/pair/pass (u32 keys), 24 B/pair/pass (u64).
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
constexpr int NWARPS = SORT_THREADS / 32;

// make EXTRA=-DGSR_SORT_PROFILE: cycles per phase of sort_tile summed over tiles (thread 0 of each block), read
// with gsr_debug_sort_profile: [0] tiles, [1] load, [2] rank, [5] count + scan + stage keys + request values,
// [7] in-group part of the look-back (thread 0's digit), [4] rest of the look-back + stage values, [6] write-out
__device__ unsigned long long g_sort_prof[8];
#ifdef GSR_SORT_PROFILE
#define SORT_TICK(i) do { if (threadIdx.x == 0) { const long long _t = clock64(); atomicAdd(&g_sort_prof[i], (unsigned long long)(_t - _t0)); _t0 = _t; } } while (0)
#else
#define SORT_TICK(i) do { } while (0)
#endif

__device__ __forceinline__ int64_t device_count(int64_t n_cap, const uint32_t* n_dev) {
  if (!n_dev) return n_cap;
  const int64_t v = (int64_t)__ldg(n_dev);
  return v < n_cap ? v : n_cap;
}

struct DigitLayout {  // digit p of a key = (key >> shift[p]) & mask[p]
  int shift[SORT_MAX_PASSES];
  uint32_t mask[SORT_MAX_PASSES];
};

// Digit histograms of every pass from one read of the keys.  Few, long-running blocks: each block ends with one
// global reduction per populated bin, and reductions on the same word serialise in L2 (with one block per 2048 keys
// the 1024 bins of a 1M-key depth sort took ~370 reductions each: 15 us for a 4 MB read).  Four 16-byte loads are in
// flight per thread.
template <typename KeyT>
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const KeyT* __restrict__ keys, int64_t n_cap, const uint32_t* __restrict__ n_dev, int passes,
                      DigitLayout dl, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[SORT_MAX_PASSES * 256];
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int64_t n = device_count(n_cap, n_dev);
  constexpr int VEC = 16 / (int)sizeof(KeyT), UNR = 4;  // keys per 16-byte load, loads in flight
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * VEC;
  for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC; i0 < n; i0 += stride * UNR) {
    KeyT k[UNR][VEC];
    int m[UNR];
#pragma unroll
    for (int u = 0; u < UNR; u++) {
      const int64_t i = i0 + u * stride;
      m[u] = i >= n ? 0 : (i + VEC <= n ? VEC : (int)(n - i));
      if (m[u] == VEC) {  // the key array is 16-byte aligned (caller's buffer carving)
        const uint4 v = *reinterpret_cast<const uint4*>(keys + i);
        if (sizeof(KeyT) == 4) { k[u][0] = (KeyT)v.x; k[u][1] = (KeyT)v.y; k[u][VEC > 2 ? 2 : 0] = (KeyT)v.z; k[u][VEC > 2 ? 3 : 1] = (KeyT)v.w; }
        else { k[u][0] = (KeyT)(((unsigned long long)v.y << 32) | v.x); k[u][1] = (KeyT)(((unsigned long long)v.w << 32) | v.z); }
      } else {
        for (int j = 0; j < VEC; j++) k[u][j] = j < m[u] ? keys[i + j] : (KeyT)0;
      }
    }
    for (int p = 0; p < passes; p++) {
      const int sft = dl.shift[p];
      const uint32_t msk = dl.mask[p];
#pragma unroll
      for (int u = 0; u < UNR; u++)
#pragma unroll
        for (int j = 0; j < VEC; j++)
          if (j < m[u]) atomicAdd(&sh[p * 256 + ((uint32_t)(k[u][j] >> sft) & msk)], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

template <typename KeyT, int ITEMS>
struct PassSmem {
  KeyT keys[SORT_THREADS * ITEMS];
  uint32_t vals[SORT_THREADS * ITEMS];
  uint32_t peer_tab[NWARPS][2][256];  // ranking: lanes of the warp that hold each digit (zero between uses)
  uint32_t warp_hist[NWARPS][256];  // ranking: running per-warp digit counts; afterwards: where in the sorted tile
                                    // this warp's items of each digit begin
  int32_t global_base[256];         // destination of sorted tile slot s with digit d is global_base[d] + s
  uint32_t scan_tmp[2 * NWARPS];
  uint32_t tile;
};

// One tile: load -> rank -> publish / look back -> stage sorted -> write out.  FULL = all 256 x ITEMS slots
// hold a pair (no bounds tests anywhere on that path).
template <typename KeyT, typename W, int ITEMS, bool FULL>
__device__ __forceinline__ void sort_tile(PassSmem<KeyT, ITEMS>& S, const KeyT* __restrict__ keys_in,
                                          KeyT* __restrict__ keys_out, const uint32_t* __restrict__ vals_in,
                                          uint32_t* __restrict__ vals_out, uint32_t tile, int tile_n, int shift,
                                          uint32_t digit_mask, const uint32_t* __restrict__ digit_hist, W* status, W* gstatus) {
  constexpr W ST_AGG = St<W>::AGG, ST_PREFIX = St<W>::PREFIX, ST_MASK = St<W>::MASK;
  constexpr int TILE = SORT_THREADS * ITEMS;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t tile_base = (size_t)tile * TILE;
#ifdef GSR_SORT_PROFILE
  long long _t0 = clock64();
  if (tid == 0) atomicAdd(&g_sort_prof[0], 1ull);
#endif

  // warp-striped load: warp w owns tile items [w * 32 * ITEMS, (w + 1) * 32 * ITEMS), lane-interleaved
  const int wbase = warp * (32 * ITEMS) + lane;
  const KeyT* kp = keys_in + tile_base + wbase;
  const uint32_t* vp = vals_in + tile_base + wbase;
  KeyT key[ITEMS];
  uint32_t rk[ITEMS];  // digit << 16 | rank among the warp's items of that digit; later the slot in the sorted tile
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    if (FULL || wbase + i * 32 < tile_n) key[i] = kp[i * 32];
    else key[i] = (KeyT)~(KeyT)0;
  }

#ifdef GSR_SORT_PROFILE
  { uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) x ^= (uint32_t)key[i];
    if (x == 0x9e3779b9u) g_sort_prof[0] = 1; }  // waits for the loads
#endif
  SORT_TICK(1);
  // Stable per-warp ranking.  The lanes that hold the same digit are found through shared memory: every lane ORs its
  // lane bit into a per-warp table entry of its digit, then reads the entry back -- 2 instructions where one ballot
  // per digit bit costs ~28 issue slots per item (the kernel is issue-bound in this phase) and MATCH.ANY ~150 cycles
  // of an SM sub-partition for random digits.  The lowest lane of each group advances the digit's running count and
  // clears the entry; two tables alternate so that the next item's ORs cannot meet a pending clear.
  uint32_t* wh = S.warp_hist[warp];
  const uint32_t lane_bit = 1u << lane, lt_mask = lane_bit - 1u;
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const bool valid = FULL || wbase + i * 32 < tile_n;
    const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
    uint32_t* tb = S.peer_tab[warp][i & 1];
    if (valid) atomicOr(&tb[d], lane_bit);
    __syncwarp();
    const uint32_t peers = tb[d], pre = wh[d];
    __syncwarp();
    const uint32_t below = peers & lt_mask;
    if (valid && below == 0) { tb[d] = 0; wh[d] = pre + __popc(peers); }
    rk[i] = (d << 16) | (pre + __popc(below));
  }
  __syncthreads();
  SORT_TICK(2);

  // thread t owns digit t: counts of the eight warps, tile aggregate published for the tiles behind
  uint32_t count = 0;
#pragma unroll
  for (int w = 0; w < NWARPS; w++) count += S.warp_hist[w][tid];
  volatile W* st = status;    // [tile][256]   tile aggregates
  volatile W* gst = gstatus;  // [group][256]  per group of SORT_LB_GROUP tiles: total (AGG), then inclusive prefix (PREFIX)
  st[(size_t)tile * 256 + tid] = ST_AGG | (W)count;
  // two exclusive scans over the 256 digits at once: of `count` -> where each digit begins inside the sorted tile,
  // and of the whole input's digit histogram -> where each digit begins in the output (every block redoes this
  // 256-element scan instead of a separate kernel launch between the passes)
  const uint32_t total = digit_hist[tid];
  uint32_t incl = count, gincl = total;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o), g = __shfl_up_sync(0xffffffffu, gincl, o);
    if (lane >= o) { incl += t; gincl += g; }
  }
  if (lane == 31) { S.scan_tmp[warp] = incl; S.scan_tmp[NWARPS + warp] = gincl; }
  __syncthreads();  // also: every warp has read the counts it needs before they are overwritten below
  uint32_t wpre = 0, gpre = 0;
#pragma unroll
  for (int w = 0; w < NWARPS; w++) {
    wpre += w < warp ? S.scan_tmp[w] : 0u;
    gpre += w < warp ? S.scan_tmp[NWARPS + w] : 0u;
  }
  const uint32_t doff = wpre + incl - count, dbase = gpre + gincl - total;
  uint32_t run = doff;
#pragma unroll
  for (int w = 0; w < NWARPS; w++) { const uint32_t c = S.warp_hist[w][tid]; S.warp_hist[w][tid] = run; run += c; }
  __syncthreads();  // every digit's offsets are in place
  // stage the keys in sorted order; the values are only fetched now -- into the registers the keys leave -- and
  // their latency passes under the look-back below (keys + values + ranks in registers at once would cost a
  // resident block per SM)
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    if (FULL || wbase + i * 32 < tile_n) {
      rk[i] = wh[rk[i] >> 16] + (rk[i] & 0xffffu);
      S.keys[rk[i]] = key[i];
    }
  }
  uint32_t val[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    if (vals_in == nullptr) val[i] = (uint32_t)tile_base + wbase + i * 32;  // first pass over values 0, 1, 2, ...
    else val[i] = (FULL || wbase + i * 32 < tile_n) ? vp[i * 32] : 0u;
  }
  SORT_TICK(5);
  // Two-level decoupled look-back.  All resident tiles reach this point at about the same time, so a plain chain
  // (sum predecessors until one has published an inclusive prefix) needs ~sqrt(tiles in flight / 2) dependent round
  // trips to L2 -- measured: 10k of the 22k cycles a tile took.  Here a tile sums the aggregates of the tiles before it
  // in its own group of 8, the last tile of a group publishes the group total, and the walk over earlier groups (8 per
  // round trip) stops at the first group whose inclusive prefix is known.  A status word carries flag and value
  // together, so no fence is needed around it; a tile only ever waits for tiles with a smaller ticket, which have
  // started.  It runs AFTER the keys are staged and the values requested: what it waits for passes under that work.
  uint32_t excl = 0;
  {
    constexpr int LBW = SORT_LB_GROUP;  // words fetched per round trip
    const uint32_t grp = tile / SORT_LB_GROUP, first = grp * SORT_LB_GROUP;
    // both walks start with loads that depend on nothing: the aggregates of the tiles before this one in its group
    // and the newest LBW group words travel in the same round trip
    W a[LBW], g[LBW];
    int64_t j = (int64_t)grp - 1;
#pragma unroll
    for (int k = 0; k < LBW; k++) { a[k] = ST_AGG; if (first + k < tile) a[k] = st[(size_t)(first + k) * 256 + tid]; }
#pragma unroll
    for (int k = 0; k < LBW; k++) { g[k] = ST_PREFIX; if (j - k >= 0) g[k] = gst[(size_t)(j - k) * 256 + tid]; }
    while (true) {  // in-group: all predecessors must have published their aggregate
      bool ok = true;
#pragma unroll
      for (int k = 0; k < LBW; k++) ok = ok && (a[k] & ST_AGG) != 0;
      if (ok) break;
#pragma unroll
      for (int k = 0; k < LBW; k++) if (first + k < tile) a[k] = st[(size_t)(first + k) * 256 + tid];
    }
#pragma unroll
    for (int k = 0; k < LBW; k++) excl += (uint32_t)(a[k] & ST_MASK);
    SORT_TICK(7);  // in-group part, thread 0's digit only
    const bool leader = tile - first == SORT_LB_GROUP - 1;
    if (leader) gst[(size_t)grp * 256 + tid] = (grp == 0 ? ST_PREFIX : ST_AGG) | ((W)(excl + count) & ST_MASK);
    if (grp != 0) {
      uint32_t gsum = 0;
      while (true) {
        bool done = false, stop = false;
        int adv = 0;
#pragma unroll
        for (int k = 0; k < LBW; k++) {
          if (!done && !stop) {
            if ((g[k] & (ST_AGG | ST_PREFIX)) == 0) stop = true;
            else { gsum += (uint32_t)(g[k] & ST_MASK); adv = k + 1; done = (g[k] & ST_PREFIX) != 0; }
          }
        }
        if (done) break;
        j -= adv;
#pragma unroll
        for (int k = 0; k < LBW; k++) { g[k] = ST_PREFIX; if (j - k >= 0) g[k] = gst[(size_t)(j - k) * 256 + tid]; }
      }
      excl += gsum;
      if (leader) gst[(size_t)grp * 256 + tid] = ST_PREFIX | ((W)(excl + count) & ST_MASK);
    }
  }
  S.global_base[tid] = (int32_t)(dbase + excl) - (int32_t)doff;
#pragma unroll
  for (int i = 0; i < ITEMS; i++)
    if (FULL || wbase + i * 32 < tile_n) S.vals[rk[i]] = val[i];
  __syncthreads();
  SORT_TICK(4);

  // per-digit runs land contiguously in the output
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    const int s = tid + k * SORT_THREADS;
    if (FULL || s < tile_n) {
      const KeyT kk = S.keys[s];
      const uint32_t dst = (uint32_t)(S.global_base[(uint32_t)(kk >> shift) & digit_mask] + s);
      keys_out[dst] = kk;
      vals_out[dst] = S.vals[s];
    }
  }
  SORT_TICK(6);
}

#ifndef GSR_SORT_MINBLOCKS
#define GSR_SORT_MINBLOCKS 3
#endif
template <typename KeyT, typename W, int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS, GSR_SORT_MINBLOCKS)
onesweep_pass_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out, int64_t n_cap,
                     const uint32_t* __restrict__ n_dev, int shift, uint32_t digit_mask,
                     const uint32_t* __restrict__ digit_hist, W* status, W* gstatus, uint32_t* ticket) {
  constexpr int TILE = SORT_THREADS * ITEMS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PassSmem<KeyT, ITEMS>& S = *reinterpret_cast<PassSmem<KeyT, ITEMS>*>(smem_raw);
  const int tid = threadIdx.x;
  const int64_t n = device_count(n_cap, n_dev);
  const uint32_t ntiles = (uint32_t)((n + TILE - 1) / TILE);
  // a block keeps taking tiles (tickets are handed out in order, so look-back cannot deadlock) until none is left:
  // launched with one block per tile of the capacity this loop runs once (or not at all for surplus blocks),
  // launched with fewer it is a persistent kernel
  bool first_tile = true;
  for (;;) {
    __syncthreads();
    if (tid == 0) S.tile = atomicAdd(ticket, 1u);
    uint4* z = reinterpret_cast<uint4*>(&S.warp_hist[0][0]);  // 8 KB = 512 x 16 B
    z[tid] = make_uint4(0, 0, 0, 0);
    z[tid + SORT_THREADS] = make_uint4(0, 0, 0, 0);
    if (first_tile) {  // the ranking leaves the peer tables zeroed
      uint4* zp = reinterpret_cast<uint4*>(&S.peer_tab[0][0][0]);  // 16 KB
#pragma unroll
      for (int k = 0; k < 4; k++) zp[tid + k * SORT_THREADS] = make_uint4(0, 0, 0, 0);
      first_tile = false;
    }
    __syncthreads();
    const uint32_t tile = S.tile;
    if (tile >= ntiles) break;
    const int64_t left = n - (int64_t)tile * TILE;
    if (left >= TILE)
      sort_tile<KeyT, W, ITEMS, true>(S, keys_in, keys_out, vals_in, vals_out, tile, TILE, shift, digit_mask, digit_hist, status, gstatus);
    else
      sort_tile<KeyT, W, ITEMS, false>(S, keys_in, keys_out, vals_in, vals_out, tile, (int)left, shift, digit_mask, digit_hist, status, gstatus);
  }
}

int sort_max_blocks() {  // GSR_SORT_MAX_BLOCKS=n caps the grid of a sort pass (0 / unset: one block per tile)
  static int v = -1;
  if (v < 0) { const char* e = getenv("GSR_SORT_MAX_BLOCKS"); v = e ? atoi(e) : 0; if (v < 0) v = 0; }
  return v;
}

template <typename KeyT, typename W, int ITEMS>
int launch_pass(const KeyT* kin, KeyT* kout, const uint32_t* vin, uint32_t* vout, int64_t n_cap, const uint32_t* n_dev,
                int shift, uint32_t mask, const uint32_t* digit_hist, void* status, void* gstatus, uint32_t* ticket, unsigned grid,
                cudaStream_t st) {
  constexpr size_t smem = sizeof(PassSmem<KeyT, ITEMS>);
  static bool attr_set = false;  // one per instantiation
  if (!attr_set) {
    GSR_CUDA_CHECK(cudaFuncSetAttribute(onesweep_pass_kernel<KeyT, W, ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  onesweep_pass_kernel<KeyT, W, ITEMS><<<grid, SORT_THREADS, smem, st>>>(kin, kout, vin, vout, n_cap, n_dev, shift, mask,
                                                                         digit_hist, (W*)status, (W*)gstatus, ticket);
  GSR_LAUNCH_CHECK("onesweep_pass_kernel");
  return GSR_OK;
}

}  // namespace

int sort_profile_read(uint64_t* out_host, int reset) {
  if (cudaMemcpyFromSymbol(out_host, g_sort_prof, 8 * sizeof(uint64_t)) != cudaSuccess) return GSR_ERR_CUDA;
  if (reset) {
    const unsigned long long z[8] = {0};
    if (cudaMemcpyToSymbol(g_sort_prof, z, sizeof(z)) != cudaSuccess) return GSR_ERR_CUDA;
  }
  return GSR_OK;
}

SortPlan make_sort_plan(int64_t n_cap, int end_bit, int key_bytes) {
  SortPlan p;
  p.end_bit = end_bit;
  p.passes = (end_bit + 7) / 8;
  if (p.passes < 1) p.passes = 1;
  int done = 0;
  for (int i = 0; i < p.passes && i < SORT_MAX_PASSES; i++) {  // nearly equal digit widths, wider digits first
    const int left = p.passes - i;
    const int w = (end_bit - done + left - 1) / left;
    p.shift[i] = done;
    p.bits[i] = w;
    done += w;
  }
  // small sorts are latency-bound: half-size tiles double the blocks in flight; 64-bit keys keep the big tile
  p.items = (key_bytes == 4 && n_cap <= SORT_SMALL_TILE_LIMIT) ? SORT_ITEMS_SMALL : SORT_ITEMS;
  return p;
}

// Sorts across two buffer pairs: pass p reads pair (p & 1) and writes the other; the result is in pair
// (passes & 1).  n_dev (optional) holds the actual pair count on the device, n_cap bounds it.
template <typename KeyT>
static int sort_pairs_pingpong(KeyT* keys0, uint32_t* vals0, KeyT* keys1, uint32_t* vals1, int64_t n_cap,
                               const uint32_t* n_dev, const SortPlan& plan, void* temp, bool hist_ready, bool iota_vals,
                               cudaStream_t st) {
  const int passes = plan.passes;
  if (passes < 1 || passes > (int)sizeof(KeyT) || passes > SORT_MAX_PASSES) {
    set_error("sort: end_bit %d out of range", plan.end_bit); return GSR_ERR_INVALID;
  }
  if (n_cap >= ((int64_t)1 << 32)) { set_error("sort: %lld pairs exceed the 2^32 limit", (long long)n_cap); return GSR_ERR_INVALID; }
  if (n_cap == 0) return GSR_OK;
  SortTemp T(temp, n_cap, passes, plan.items);
  if (!hist_ready) {  // otherwise the caller zeroed `temp` and the producer of keys0 filled T.hist (sort_prepare)
    GSR_CUDA_CHECK(cudaMemsetAsync(temp, 0, T.zero_bytes(passes), st));
    DigitLayout dl;
    for (int p = 0; p < SORT_MAX_PASSES; p++) {
      dl.shift[p] = p < passes ? plan.shift[p] : 0;
      dl.mask[p] = p < passes ? (1u << plan.bits[p]) - 1u : 0u;
    }
    const int64_t per_block = 256 * 16;  // one sweep of a block = 256 threads x 4 loads x 4 keys
    const int hist_blocks = (int)((n_cap + per_block - 1) / per_block < 148 * 2 ? (n_cap + per_block - 1) / per_block : 148 * 2);
    sort_histogram_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys0, n_cap, n_dev, passes, dl, T.hist);
    GSR_LAUNCH_CHECK("sort_histogram_kernel");
  }
  const int64_t tiles = sort_tiles(n_cap, plan.items);
  const bool big = n_cap >= SORT_SMALL_LIMIT;
  const int cap = sort_max_blocks();
  const unsigned grid = (unsigned)(cap > 0 && tiles > cap ? cap : tiles);
  for (int p = 0; p < passes; p++) {
    const KeyT* kin = (p & 1) ? keys1 : keys0;
    KeyT* kout = (p & 1) ? keys0 : keys1;
    const uint32_t* vin = (p & 1) ? vals1 : vals0;
    if (p == 0 && iota_vals) vin = nullptr;  // the values are the indices 0 .. n-1: generated, not read
    uint32_t* vout = (p & 1) ? vals0 : vals1;
    char* stp = (char*)T.status + (size_t)p * T.pass_stride;
    const uint32_t mask = (1u << plan.bits[p]) - 1u;
    int r;
#define GSR_PASS(WT, IT) launch_pass<KeyT, WT, IT>(kin, kout, vin, vout, n_cap, n_dev, plan.shift[p], mask, T.hist + p * 256, stp, stp + T.group_off, T.tickets + p, grid, st)
    if (plan.items == SORT_ITEMS_SMALL && sizeof(KeyT) == 4) r = big ? GSR_PASS(unsigned long long, SORT_ITEMS_SMALL) : GSR_PASS(uint32_t, SORT_ITEMS_SMALL);
    else r = big ? GSR_PASS(unsigned long long, SORT_ITEMS) : GSR_PASS(uint32_t, SORT_ITEMS);
#undef GSR_PASS
    if (r != GSR_OK) return r;
  }
  return GSR_OK;
}

int launch_sort_pairs_pingpong(uint64_t* keys0, uint32_t* vals0, uint64_t* keys1, uint32_t* vals1, int64_t n,
                               const SortPlan& plan, void* temp, cudaStream_t st) {
  return sort_pairs_pingpong<uint64_t>(keys0, vals0, keys1, vals1, n, nullptr, plan, temp, false, false, st);
}
int launch_sort_pairs_pingpong_u32(uint32_t* keys0, uint32_t* vals0, uint32_t* keys1, uint32_t* vals1, int64_t n_cap,
                                   const uint32_t* n_dev, const SortPlan& plan, void* temp, bool hist_ready, bool iota_vals,
                                   cudaStream_t st) {
  return sort_pairs_pingpong<uint32_t>(keys0, vals0, keys1, vals1, n_cap, n_dev, plan, temp, hist_ready, iota_vals, st);
}

// Zeroes the sort's temp block and returns where the producer of the keys must accumulate the digit
// histograms (pass p, digit d -> hist[p * 256 + d]) when it wants to skip the sort's own histogram pass.
int sort_prepare(void* temp, int64_t n_cap, const SortPlan& plan, uint32_t** hist_out, cudaStream_t st) {
  SortTemp T(temp, n_cap, plan.passes, plan.items);
  GSR_CUDA_CHECK(cudaMemsetAsync(temp, 0, T.zero_bytes(plan.passes), st));
  *hist_out = T.hist;
  return GSR_OK;
}

}  // namespace gsr
