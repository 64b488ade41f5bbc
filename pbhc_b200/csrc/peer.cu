// peer.cu - the cross-rank exchange of the path over NVLink peer memory instead of a library collective.
// The path's only collectives are SUM all-reduces of a few hundred bytes (the 64-double statistics vector of a step, the
// (n, sum, sum of squares) of the advantages): an NCCL all-reduce of that size is pure launch + protocol latency
// (~20-30 us at 8 ranks).  Here every rank owns a small buffer in symmetric memory that all its peers have mapped
// (NVSwitch gives every GPU a direct path to every peer): one block per rank
//   1. stores its n values into slot [parity][rank] of EVERY rank's buffer (peer stores over NVLink),
//   2. fences, then raises its flag on every rank's signal pad (release, system scope),
//   3. waits until the flags of all ranks on its own pad carry this call's sequence number (acquire),
//   4. adds the `world` slots of its own buffer in rank order - the same order on every rank, so the result is
//      bit-identical everywhere and independent of arrival order.
// Two slot sets alternate by the parity of the sequence number: a rank can only enter call t+2 after every peer
// published call t+1, i.e. after every peer finished reading call t.  The sequence number lives in device memory and
// is advanced by the kernel itself, so the launch is CUDA-graph capturable and replayable.
#include "common.cuh"

namespace pbhc {
namespace {

constexpr int kPeerThreads = 128;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_kernel(double* __restrict__ data, int n, const uint64_t* __restrict__ bufs,
                                                                      const uint64_t* __restrict__ sigs, uint32_t* __restrict__ counter, int rank,
                                                                      int world, int sig_base) {
  __shared__ uint32_t s_seq;
  if (threadIdx.x == 0) s_seq = *counter + 1u;
  __syncthreads();
  const uint32_t seq = s_seq;
  const int par = (int)(seq & 1u);
  for (int i = threadIdx.x; i < n; i += kPeerThreads) {
    const double v = data[i];
    for (int p = 0; p < world; ++p) reinterpret_cast<double*>(bufs[p])[(int64_t)(par * world + rank) * n + i] = v;
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int p = (int)threadIdx.x;
    st_release_sys(reinterpret_cast<uint32_t*>(sigs[p]) + sig_base + rank, seq);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(sigs[rank]) + sig_base + p;
    while ((int32_t)(ld_acquire_sys(mine) - seq) < 0) { }
  }
  __syncthreads();
  const volatile double* own = reinterpret_cast<const volatile double*>(bufs[rank]) + (int64_t)par * world * n;
  for (int i = threadIdx.x; i < n; i += kPeerThreads) {
    double s = 0.0;
    for (int p = 0; p < world; ++p) s += own[(int64_t)p * n + i];
    data[i] = s;
  }
  if (threadIdx.x == 0) *counter = seq;
}

}  // namespace
}  // namespace pbhc

extern "C" int pbhc_peer_allreduce_sum(double* data, int32_t n, const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                                       uint32_t* counter, int32_t rank, int32_t world, int32_t signal_base, void* stream) {
  using namespace pbhc;
  if (!data || !peer_buffers || !peer_signal_pads || !counter) { set_error("pbhc_peer_allreduce_sum: null argument"); return PBHC_E_BADARG; }
  if (n <= 0 || n > PBHC_PEER_MAX_DOUBLES || world < 1 || world > kPeerThreads || rank < 0 || rank >= world || signal_base < 0) {
    set_error("pbhc_peer_allreduce_sum: n = %d (1..%d), rank %d of %d", n, PBHC_PEER_MAX_DOUBLES, rank, world); return PBHC_E_RANGE;
  }
  peer_allreduce_kernel<<<1, kPeerThreads, 0, static_cast<cudaStream_t>(stream)>>>(data, n, peer_buffers, peer_signal_pads, counter, rank, world, signal_base);
  return check_launch("pbhc_peer_allreduce_sum");
}
