// static_scratch.cu -- gsr_sort_pairs_u32_static: the (u32 key, u32 value) radix sort of radix_sort.cu for callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The sort's temp block (digit histograms, tile tickets, look-back status words, the ping-pong key / value buffers)
// lives in ONE file-scope __device__ array of GSR_STATIC_SORT_SCRATCH_BYTES: every GPU that loads the library holds it
// until the process exits.  In host code the array's name is only a handle; its device address is looked up per
// device with cudaGetSymbolAddress (that device current) and cached.  Every call on a device shares the array: the
// sort zeroes the status prefix it needs at the start of each call (cudaMemsetAsync on the caller's stream), and two
// calls that overlap on the GPU (different streams, no ordering between them) race.
// No cudaMalloc / cudaFree / allocating library routine on this path.
#include <mutex>
#include "common.cuh"

namespace {

#define GSR_TRY(x) do { int _r = (x); if (_r != GSR_OK) return _r; } while (0)

static __device__ __align__(256) unsigned char g_static_sort_scratch[GSR_STATIC_SORT_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_addr_mutex;
void* g_scratch_addr[MAX_DEVICES];  // device address of g_static_sort_scratch on each device (NULL: not looked up yet)

// Device address of the scratch array on the CURRENT device (looked up once per device).
int static_scratch_address(void** out) {
  int dev = -1;
  GSR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) { gsr::set_error("static scratch: device %d out of range", dev); return GSR_ERR_INVALID; }
  std::lock_guard<std::mutex> lock(g_addr_mutex);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    GSR_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_sort_scratch));
    if (!p || ((uintptr_t)p & 255)) { gsr::set_error("static scratch: bad symbol address %p", p); return GSR_ERR_CUDA; }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return GSR_OK;
}

}  // namespace

extern "C" int gsr_sort_pairs_u32_static(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in,
                                         uint32_t* vals_out, int64_t n, int32_t end_bit, void* stream) {
  using namespace gsr;
  cudaStream_t st = (cudaStream_t)stream;
  if (n < 0 || end_bit < 1 || end_bit > 32) { set_error("static sort: bad n/end_bit"); return GSR_ERR_INVALID; }
  if (n > (int64_t)GSR_STATIC_SORT_MAX_PAIRS) {
    set_error("static sort: %lld pairs exceed the built-in scratch (at most %lld; use gsr_sort_pairs_u32 with a caller buffer)",
              (long long)n, (long long)GSR_STATIC_SORT_MAX_PAIRS);
    return GSR_ERR_CAPACITY;
  }
  if (n == 0) return GSR_OK;
  if (!keys_in || !keys_out || !vals_in || !vals_out) { set_error("static sort: NULL buffer"); return GSR_ERR_INVALID; }
  const SortPlan plan = make_sort_plan(n, end_bit, 4);
  const size_t st_bytes = sort_temp_bytes(n, 32, 4);
  if (st_bytes + 2 * align_up((size_t)n * 4) > (size_t)GSR_STATIC_SORT_SCRATCH_BYTES) {
    set_error("static sort: %lld pairs need more than the %lld-byte built-in scratch", (long long)n,
              (long long)GSR_STATIC_SORT_SCRATCH_BYTES);
    return GSR_ERR_CAPACITY;
  }
  void* base = nullptr;
  GSR_TRY(static_scratch_address(&base));
  Carver c(base);
  void* st_temp = c.take<char>(st_bytes);
  uint32_t* ktmp = c.take<uint32_t>((size_t)n);
  uint32_t* vtmp = c.take<uint32_t>((size_t)n);
  const bool odd = plan.passes & 1;  // pass p reads pair (p & 1): the result lands in pair (passes & 1)
  uint32_t* k0 = odd ? ktmp : keys_out;
  uint32_t* v0 = odd ? vtmp : vals_out;
  uint32_t* k1 = odd ? keys_out : ktmp;
  uint32_t* v1 = odd ? vals_out : vtmp;
  if (keys_in != k0) GSR_CUDA_CHECK(cudaMemcpyAsync(k0, keys_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
  if (vals_in != v0) GSR_CUDA_CHECK(cudaMemcpyAsync(v0, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
  return launch_sort_pairs_pingpong_u32(k0, v0, k1, v1, n, nullptr, plan, st_temp, false, false, st);
}
