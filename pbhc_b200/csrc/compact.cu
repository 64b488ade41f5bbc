// compact.cu - env ids of a reset mask: `reset_buf.nonzero().flatten()` (legged_robot_base.py:311,315) on the device.
// The reference reads the mask back to the host every step to build env_ids; here the ids and their count stay on the
// device (or are read back as 4 + 4 * count bytes when the simulator setter wants a host-visible index tensor).
//   pass 1: a block counts the nonzero flags of its 4096-item span            -> scratch[block]
//   pass 2: a block sums scratch[0 .. block) (its first output slot), ranks its own nonzero flags with warp ballots
//           and writes the ids in ascending order; the last block writes the total.
// No atomics, so the output order does not depend on scheduling.
#include <mutex>

#include "common.cuh"

namespace pbhc {
namespace {

constexpr int kThreads = 256;
constexpr int kItems = 16;                 // consecutive flags per thread
constexpr int kSpan = kThreads * kItems;   // flags per block

// flags of thread t of block b: [b * kSpan + t * kItems, + kItems) - consecutive per thread so that ranks are ascending in
// (warp, lane, item) order.  Returns the bit mask of the nonzero ones.
__device__ __forceinline__ uint32_t load_mask(const int64_t* __restrict__ flags, int64_t n, int64_t first) {
  uint32_t m = 0;
  if (first + kItems <= n) {
    const longlong2* p = reinterpret_cast<const longlong2*>(flags + first);  // first is a multiple of 16 items = 128 B
#pragma unroll
    for (int i = 0; i < kItems / 2; ++i) {
      const longlong2 v = __ldg(p + i);
      m |= (v.x != 0 ? 1u : 0u) << (2 * i);
      m |= (v.y != 0 ? 1u : 0u) << (2 * i + 1);
    }
  } else {
    for (int i = 0; i < kItems; ++i)
      if (first + i < n && flags[first + i] != 0) m |= 1u << i;
  }
  return m;
}

__global__ void __launch_bounds__(kThreads) compact_count_kernel(const int64_t* __restrict__ flags, int64_t n, int32_t* __restrict__ block_count) {
  __shared__ int warp_total[kThreads / 32];
  const int64_t first = (int64_t)blockIdx.x * kSpan + (int64_t)threadIdx.x * kItems;
  int c = __popc(load_mask(flags, n, first));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) warp_total[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += warp_total[w];
    block_count[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kThreads) compact_write_kernel(const int64_t* __restrict__ flags, int64_t n, const int32_t* __restrict__ block_count,
                                                                 int32_t* __restrict__ ids, int32_t* __restrict__ count) {
  __shared__ int warp_total[kThreads / 32];
  __shared__ int block_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // first output slot of this block = sum of the counts of the blocks before it
  int before = 0;
  for (int b = threadIdx.x; b < (int)blockIdx.x; b += kThreads) before += block_count[b];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  if (lane == 0) warp_total[warp] = before;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += warp_total[w];
    block_base = t;
    if (blockIdx.x == gridDim.x - 1) *count = t + block_count[blockIdx.x];
  }
  __syncthreads();
  const int64_t first = (int64_t)blockIdx.x * kSpan + (int64_t)threadIdx.x * kItems;
  const uint32_t m = load_mask(flags, n, first);
  const int c = __popc(m);
  // exclusive scan of c over the block: lanes, then warps
  int incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  __syncthreads();  // warp_total is reused
  if (lane == 31) warp_total[warp] = incl;
  __syncthreads();
  int slot = block_base + incl - c;
  for (int w = 0; w < warp; ++w) slot += warp_total[w];
  uint32_t rest = m;
  while (rest) {
    const int i = __ffs(rest) - 1;
    rest &= rest - 1;
    ids[slot++] = (int32_t)(first + i);
  }
}

int compact_launch(const int64_t* flags, int64_t n, int32_t* ids, int32_t* count, int32_t* scratch, cudaStream_t st) {
  const int blocks = (int)((n + kSpan - 1) / kSpan);
  compact_count_kernel<<<blocks, kThreads, 0, st>>>(flags, n, scratch);
  compact_write_kernel<<<blocks, kThreads, 0, st>>>(flags, n, scratch, ids, count);
  return check_launch("pbhc_compact_ids");
}

int compact_check(const int64_t* flags, int64_t n, const int32_t* ids, const int32_t* count) {
  if (!flags || !ids || !count) { set_error("pbhc_compact_ids: flags, ids and count must not be NULL"); return PBHC_E_BADARG; }
  if (n <= 0 || n > 0x7fffffffLL) { set_error("pbhc_compact_ids: n = %lld outside [1, 2^31)", (long long)n); return PBHC_E_RANGE; }
  if (reinterpret_cast<uintptr_t>(flags) & 15) { set_error("pbhc_compact_ids: flags must be 16-byte aligned"); return PBHC_E_ALIGN; }
  return PBHC_OK;
}

// ---- the variant that allocates nothing: the per-block counts live in a file-scope device array --------------------------------
// 4096 block counts = 16 KB that every GPU that loads the library holds until exit; enough for kStaticMaxN flags.
constexpr int kStaticBlocks = 4096;
constexpr int64_t kStaticMaxN = (int64_t)kStaticBlocks * kSpan;  // 16 777 216 = PBHC_COMPACT_STATIC_MAX_N
static_assert(kStaticMaxN == PBHC_COMPACT_STATIC_MAX_N, "header and kernel disagree on the static scratch size");
static __device__ __align__(256) int32_t g_static_block_count[kStaticBlocks];

// In host code `g_static_block_count` is only a stand-in for the symbol, not a device address: the address is asked for
// once per device, with that device current, and cached.
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
int32_t* g_static_ptr[kMaxDevices] = {};

int static_scratch(int32_t** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("pbhc_compact_ids_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return PBHC_E_CUDA;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev]) {
    int32_t* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_block_count);
    if (e != cudaSuccess || !p) {
      set_error("pbhc_compact_ids_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      (void)cudaGetLastError();
      return PBHC_E_CUDA;
    }
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return PBHC_OK;
}

}  // namespace
}  // namespace pbhc

extern "C" int pbhc_compact_ids_static(const int64_t* flags, int64_t n, int32_t* ids, int32_t* count, void* stream) {
  using namespace pbhc;
  if (const int rc = compact_check(flags, n, ids, count)) return rc;
  if (n > kStaticMaxN) {
    set_error("pbhc_compact_ids_static: n = %lld exceeds PBHC_COMPACT_STATIC_MAX_N = %lld (use pbhc_compact_ids with caller scratch)",
              (long long)n, (long long)kStaticMaxN);
    return PBHC_E_RANGE;
  }
  int32_t* scratch = nullptr;
  if (const int rc = static_scratch(&scratch)) return rc;
  // nothing in the array has to start at zero: pass 1 overwrites the counts of every block pass 2 reads
  return compact_launch(flags, n, ids, count, scratch, static_cast<cudaStream_t>(stream));
}

extern "C" int64_t pbhc_compact_scratch_ints(int64_t n) { return n <= 0 ? 0 : (n + pbhc::kSpan - 1) / pbhc::kSpan; }

extern "C" int pbhc_compact_ids(const int64_t* flags, int64_t n, int32_t* ids, int32_t* count, int32_t* scratch, void* stream) {
  using namespace pbhc;
  if (const int rc = compact_check(flags, n, ids, count)) return rc;
  if (!scratch) { set_error("pbhc_compact_ids: scratch must hold pbhc_compact_scratch_ints(n) int32"); return PBHC_E_BADARG; }
  return compact_launch(flags, n, ids, count, scratch, static_cast<cudaStream_t>(stream));
}
