// rollout.cu - RolloutStorage data movement: the minibatch shuffle-gather of every stored key in
// ONE launch, and the time-out bootstrap of the stored rewards.
//
// Reference: agents/modules/data_utils.py:116-152 (mini_batch_generator: `getattr(self, key)[:].flatten(0, 1)[indices]`
// for every stored key, indices = randperm(T*N)) and agents/ppo/ppo.py:266-270 (rewards += gamma * values * time_outs).
// Pure HBM-bound byte moves.  A warp owns a tile of 32 destination rows: each lane reads one index (one coalesced
// 256-byte read per tile, shared by all keys), then for every key the 32 destination rows are ONE dense span that the
// warp writes front to back with the widest vector the key's alignment allows (fully coalesced stores whatever the
// row width, 4-byte `rewards` rows included); the source row of each vector comes from a lane shuffle of the indices.
#include "common.cuh"

namespace pbhc {

struct GatherKeys {
  const void* src[PBHC_MAX_GATHER_KEYS];
  void* dst[PBHC_MAX_GATHER_KEYS];
  uint32_t row_vecs[PBHC_MAX_GATHER_KEYS];  // row length in vectors of 1 << vec_log2[k] bytes
  uint8_t vec_log2[PBHC_MAX_GATHER_KEYS];  // 0, 2, 3, 4 = flat span of 1/4/8/16-byte vectors; 5 = shifted 16-byte row copy (row_vecs in words)
  uint8_t kind[PBHC_MAX_GATHER_KEYS];      // staged kernel: 0 = through registers, 1 = bulk rows, 2 = 4-byte row fill, 3 = 4-byte flat fill
  uint8_t rpc[PBHC_MAX_GATHER_KEYS];       // staged kernel: rows per stage
  uint32_t row_bytes[PBHC_MAX_GATHER_KEYS];
  int n_keys;
};

template <typename V>
__device__ __forceinline__ void gather_tile(const V* __restrict__ src, V* __restrict__ dst, uint32_t W, int64_t row0, int rows, int64_t my_src,
                                            int lane) {
  // destination span: rows [row0, row0+rows) x W vectors, dense; vector j of the span lives in row j / W
  V* d = dst + row0 * W;
  const uint32_t total = (uint32_t)rows * W;
  uint32_t row = lane / W, col = lane - row * W;
  const uint32_t drow = 32 / W, dcol = 32 - drow * W;  // advancing j by 32 = drow rows + dcol columns (+1 row on wrap)
#pragma unroll 4
  for (uint32_t j = lane; j < ((total + 31u) & ~31u); j += 32) {
    const int64_t s = __shfl_sync(0xffffffffu, my_src, row < 32 ? row : 31);
    if (j < total && s >= 0) d[j] = __ldg(src + s * W + col);
    row += drow;
    col += dcol;
    if (col >= W) { col -= W; ++row; }
  }
}

// Wide rows whose byte length is a multiple of 4 but not of 16 (critic_obs: 463 floats): source and destination row
// starts then sit at different 16-byte phases.  The row is moved as 16-byte ALIGNED loads of the source chunks that
// cover it and 16-byte aligned stores of the destination body; each output vector is cut out of two neighbouring source
// chunks at the (warp-uniform) word phase P.  Head and tail words (<= 3 each) go as 4-byte copies.
template <int P>
__device__ __forceinline__ void shifted_body(const uint4* __restrict__ sa, uint4* __restrict__ d, uint32_t nb, int lane) {
#pragma unroll 4
  for (uint32_t v = lane; v < nb; v += 32) {
    const uint4 A = __ldg(sa + v);
    uint4 o = A;
    if (P != 0) {
      const uint4 B = __ldg(sa + v + 1);  // holds at least one word of this vector, so it lies inside the allocation
      if (P == 1) o = make_uint4(A.y, A.z, A.w, B.x);
      if (P == 2) o = make_uint4(A.z, A.w, B.x, B.y);
      if (P == 3) o = make_uint4(A.w, B.x, B.y, B.z);
    }
    d[v] = o;
  }
}

__device__ __noinline__ void copy_row_shifted(const uint32_t* __restrict__ src_row, uint32_t* __restrict__ dst_row, uint32_t W, int lane) {
  uint32_t head = (uint32_t)((16 - ((uintptr_t)dst_row & 15)) & 15) >> 2;  // words until the destination is 16-byte aligned
  if (head > W) head = W;
  if (lane < (int)head) dst_row[lane] = __ldg(src_row + lane);
  const uint32_t nb = (W - head) >> 2;
  const uint32_t* s = src_row + head;
  const uint32_t p = (uint32_t)((uintptr_t)s >> 2) & 3;
  const uint4* sa = (const uint4*)(s - p);
  uint4* d = (uint4*)(dst_row + head);
  switch (p) {
    case 0: shifted_body<0>(sa, d, nb, lane); break;
    case 1: shifted_body<1>(sa, d, nb, lane); break;
    case 2: shifted_body<2>(sa, d, nb, lane); break;
    default: shifted_body<3>(sa, d, nb, lane); break;
  }
  const uint32_t done = head + 4 * nb;
  if (lane < (int)(W - done)) dst_row[done + lane] = __ldg(src_row + done + lane);
}

__device__ __forceinline__ void key_through_registers(const GatherKeys& K, int k, int64_t row0, int rows, int64_t s, int lane) {
  const uint32_t W = K.row_vecs[k];
  switch (K.vec_log2[k]) {
    case 5:
      for (int r = 0; r < rows; ++r) {
        const int64_t sr = __shfl_sync(0xffffffffu, s, r);
        if (sr >= 0) copy_row_shifted((const uint32_t*)K.src[k] + sr * W, (uint32_t*)K.dst[k] + (row0 + r) * W, W, lane);
      }
      break;
    case 4: gather_tile((const uint4*)K.src[k], (uint4*)K.dst[k], W, row0, rows, s, lane); break;
    case 3: gather_tile((const uint2*)K.src[k], (uint2*)K.dst[k], W, row0, rows, s, lane); break;
    case 2: gather_tile((const uint32_t*)K.src[k], (uint32_t*)K.dst[k], W, row0, rows, s, lane); break;
    default: gather_tile((const uint8_t*)K.src[k], (uint8_t*)K.dst[k], W, row0, rows, s, lane); break;
  }
}

__global__ void __launch_bounds__(256, 4) gather_keys_kernel(const __grid_constant__ GatherKeys K, const int64_t* __restrict__ idx, int64_t n_rows,
                                                          int64_t src_rows, int* __restrict__ bad) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t n_tiles = (n_rows + 31) >> 5;
  for (int64_t t = warp; t < n_tiles; t += n_warps) {
    const int64_t row0 = t << 5;
    const int rows = (int)(n_rows - row0 < 32 ? n_rows - row0 : 32);
    int64_t s = -1;
    if (lane < rows) {
      s = __ldg(idx + row0 + lane);
      if (s < 0 || s >= src_rows) { *bad = 1; s = -1; }  // torch raises IndexError; the caller checks the flag
    }
    for (int k = 0; k < K.n_keys; ++k) key_through_registers(K, k, row0, rows, s, lane);
  }
}

// ---- shared-memory staged variant (large jobs) ----------------------------------------------------------------------
// The register kernel above is bound by load latency (ncu: ~37 warps stalled on long scoreboard per issue, 18% issue
// utilisation): a random row gather needs far more bytes in flight than registers can hold.  Here every warp owns a ring
// of kStages x kStage bytes of shared memory.  A stage is filled with a dense piece of the destination span - by one TMA
// bulk copy per row (16-byte aligned rows; one lane issues one row) or by 4-byte cp.async (rows at other phases) - and
// drained by ONE bulk store shared -> global; fills of the next stage are issued before the previous one is drained, so
// a warp keeps up to two stages of reads and one of writes in flight without holding a register.
constexpr int kStage = 8192, kStages = 3, kGatherWarps = 8;
constexpr int kStagedSmem = kGatherWarps * kStages * kStage + kGatherWarps * kStages * 8;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldgsts4(uint32_t s, const void* g, uint64_t pol) {
  asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;" ::"r"(s), "l"(g), "l"(pol) : "memory");
}
// L2 policies: the wide rows stream through once (evict first); the narrow keys' source arrays (4 bytes per row, a few
// MB each) are re-read sector by sector by eight different rows (evict last).  Measured neutral at 786432 rows (82.3%
// of the HBM roofline with and without): DRAM reads stay ~27% above the algorithmic bytes (64-byte DRAM bursts at the
// random row edges), the kernel runs at 93% of the measured copy bandwidth in DRAM traffic.
__device__ __forceinline__ uint64_t policy_stream() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t policy_keep() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra WAIT_DONE;\nbra WAIT_LOOP;\nWAIT_DONE:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

struct StagePipe {
  uint32_t ring, bars;  // shared addresses of this warp's stages / mbarriers
  uint32_t cnt = 0, parity = 0;
  bool pend = false, pend_bulk = false;
  uint32_t pend_stage = 0, pend_bytes = 0;
  void* pend_dst = nullptr;
  int lane;
  uint64_t stream;  // L2 evict-first policy of the stores

  // waits until the stage of the next piece is no longer read by the store that used it kStages pieces ago
  __device__ __forceinline__ uint32_t acquire() {
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kStages - 2) : "memory");
    __syncwarp();
    return cnt % kStages;
  }
  // drains the piece issued before the current one: wait for its fill, then one bulk store
  __device__ __forceinline__ void drain(bool last) {
    if (!pend) return;
    if (pend_bulk) {
      mbar_wait(bars + 8 * pend_stage, (parity >> pend_stage) & 1);
      parity ^= 1u << pend_stage;
    } else {
      if (last) asm volatile("cp.async.wait_group 0;" ::: "memory");
      else asm volatile("cp.async.wait_group 1;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // cp.async wrote through the generic proxy
    }
    __syncwarp();
    if (lane == 0) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(pend_dst), "r"(ring + pend_stage * kStage),
                   "r"(pend_bytes), "l"(stream)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    pend = false;
  }
  // the piece whose fills were just issued becomes the pending one (every piece commits one cp.async group, empty for bulk)
  __device__ __forceinline__ void issued(bool bulk, uint32_t stage, void* dst, uint32_t bytes) {
    asm volatile("cp.async.commit_group;" ::: "memory");
    drain(false);
    pend = true; pend_bulk = bulk; pend_stage = stage; pend_dst = dst; pend_bytes = bytes;
    ++cnt;
  }
};

__global__ void __launch_bounds__(kGatherWarps * 32, 1) gather_keys_staged_kernel(const __grid_constant__ GatherKeys K, const int64_t* __restrict__ idx,
                                                                                   int64_t n_rows, int64_t src_rows, int* __restrict__ bad) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  StagePipe P;
  P.lane = lane;
  const uint64_t pol_stream = policy_stream(), pol_keep = policy_keep();
  P.stream = pol_stream;
  P.ring = smem_addr(smem) + wib * kStages * kStage;
  P.bars = smem_addr(smem) + kGatherWarps * kStages * kStage + wib * kStages * 8;
  if (lane < kStages) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(P.bars + 8 * lane) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const int64_t warp = (int64_t)blockIdx.x * kGatherWarps + wib;
  const int64_t n_warps = (int64_t)gridDim.x * kGatherWarps;
  const int64_t n_tiles = (n_rows + 31) >> 5;
  for (int64_t t = warp; t < n_tiles; t += n_warps) {
    const int64_t row0 = t << 5;
    const int rows = (int)(n_rows - row0 < 32 ? n_rows - row0 : 32);
    int64_t s = -1;
    if (lane < rows) {
      s = __ldg(idx + row0 + lane);
      if (s < 0 || s >= src_rows) { *bad = 1; s = -1; }
    }
    // keys that cannot be staged (1-byte granular rows, rows larger than a stage) and the ragged last tile go through
    // registers, first, while the fills of the previous tile are still in flight
    for (int k = 0; k < K.n_keys; ++k)
      if (K.kind[k] == 0 || rows < 32) key_through_registers(K, k, row0, rows, s, lane);
    if (rows < 32) continue;
    for (int k = 0; k < K.n_keys; ++k) {
      const uint32_t kind = K.kind[k];
      if (kind == 0) continue;
      const uint32_t rb = K.row_bytes[k], rpc = K.rpc[k];
      const uint8_t* src = (const uint8_t*)K.src[k];
      uint8_t* dst = (uint8_t*)K.dst[k] + row0 * rb;
      if (kind == 3) {  // rows shorter than a warp: the whole 32-row span is one piece, word j of it from row j / W
        const uint32_t W = rb >> 2, total = 32 * W;
        const uint32_t stage = P.acquire(), sm = P.ring + stage * kStage;
        uint32_t row = lane / W, col = lane - row * W;
        const uint32_t drow = 32 / W, dcol = 32 - drow * W;
#pragma unroll 2
        for (uint32_t j = lane; j < total; j += 32) {  // total is a multiple of 32
          const int64_t sr = __shfl_sync(0xffffffffu, s, row < 32 ? row : 31);
          if (sr >= 0) ldgsts4(sm + 4 * j, src + sr * rb + 4 * col, pol_keep);
          row += drow;
          col += dcol;
          if (col >= W) { col -= W; ++row; }
        }
        P.issued(false, stage, dst, total * 4);
        continue;
      }
      for (uint32_t r0 = 0; r0 < 32; r0 += rpc) {
        const uint32_t nr = 32 - r0 < rpc ? 32 - r0 : rpc;
        const uint32_t stage = P.acquire(), sm = P.ring + stage * kStage;
        if (kind == 1) {  // one bulk copy per row, lane i moves row r0 + i
          const int64_t sr = __shfl_sync(0xffffffffu, s, (r0 + lane) & 31);
          const bool ok = lane < (int)nr && sr >= 0;
          const uint32_t n_ok = __popc(__ballot_sync(0xffffffffu, ok));
          const uint32_t bar = P.bars + 8 * stage;
          if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(n_ok * rb) : "memory");
          __syncwarp();
          if (ok)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(sm + lane * rb),
                         "l"(src + sr * rb), "r"(rb), "r"(bar), "l"(pol_stream)
                         : "memory");
          P.issued(true, stage, dst + r0 * rb, nr * rb);
        } else {  // 4-byte fills, row by row (source and destination rows sit at different 16-byte phases)
          const uint32_t W = rb >> 2;
          for (uint32_t r = 0; r < nr; ++r) {
            const int64_t sr = __shfl_sync(0xffffffffu, s, r0 + r);
            if (sr < 0) continue;
            const uint8_t* g = src + sr * rb + 4 * lane;
            uint32_t d = sm + r * rb + 4 * lane;
#pragma unroll 4
            for (uint32_t i = lane; i < W; i += 32, g += 128, d += 128) ldgsts4(d, g, pol_stream);
          }
          P.issued(false, stage, dst + r0 * rb, nr * rb);
        }
      }
    }
  }
  P.drain(true);
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  __syncwarp();
}

__global__ void __launch_bounds__(256) timeout_bootstrap_kernel(float* __restrict__ rewards, const float* __restrict__ values,
                                                                const uint8_t* __restrict__ time_outs, int64_t n, int R, float gamma) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * R) return;
  const int64_t e = i / R;
  // rewards + (gamma * values) * time_outs (ppo.py:268), the two products and the sum rounded separately like eager torch
  rewards[i] = rewards[i] + gamma * values[i] * (time_outs[e] ? 1.0f : 0.0f);
}

}  // namespace pbhc

extern "C" int pbhc_gather_keys(const PbhcGatherKey* keys, int32_t n_keys, const int64_t* idx, int64_t n_rows, int64_t src_rows,
                                int32_t* bad_index_flag, void* stream) {
  using namespace pbhc;
  if (!keys || !idx || !bad_index_flag || n_keys <= 0 || src_rows < 0 || n_rows < 0) {
    set_error("pbhc_gather_keys: null / negative argument"); return PBHC_E_BADARG;
  }
  if (n_keys > PBHC_MAX_GATHER_KEYS) { set_error("pbhc_gather_keys: %d keys > PBHC_MAX_GATHER_KEYS", n_keys); return PBHC_E_RANGE; }
  GatherKeys K{};
  K.n_keys = n_keys;
  for (int k = 0; k < n_keys; ++k) {
    const PbhcGatherKey& g = keys[k];
    if (!g.src || !g.dst || g.row_bytes <= 0) { set_error("pbhc_gather_keys: key %d null / empty", k); return PBHC_E_BADARG; }
    if (g.row_bytes > (1LL << 26)) { set_error("pbhc_gather_keys: key %d row of %lld bytes", k, (long long)g.row_bytes); return PBHC_E_RANGE; }
    const uintptr_t a = (uintptr_t)g.src | (uintptr_t)g.dst | (uintptr_t)g.row_bytes;
    int lg = (a & 15) == 0 ? 4 : (a & 7) == 0 ? 3 : (a & 3) == 0 ? 2 : 0;
    uint32_t vecs = (uint32_t)(g.row_bytes >> lg);
    if ((lg == 2 || lg == 3) && g.row_bytes >= 128) { vecs = (uint32_t)(g.row_bytes >> 2); lg = 5; }  // shifted 16-byte row copy
    K.src[k] = g.src; K.dst[k] = g.dst; K.vec_log2[k] = (uint8_t)lg; K.row_vecs[k] = vecs;
    // staging class: rows of a multiple of 4 bytes that fit a stage (4 of them for the 4-byte row fill, whose pieces
    // must start and end on 16 bytes of the destination)
    K.row_bytes[k] = (uint32_t)g.row_bytes;
    K.kind[k] = 0; K.rpc[k] = 32;
    if ((a & 3) == 0 && ((uintptr_t)g.dst & 15) == 0) {
      if (g.row_bytes < 128) K.kind[k] = 3;
      else if ((a & 15) == 0 && g.row_bytes <= kStage) { K.kind[k] = 1; K.rpc[k] = (uint8_t)(kStage / g.row_bytes > 32 ? 32 : kStage / g.row_bytes); }
      else if (4 * g.row_bytes <= kStage) { K.kind[k] = 2; K.rpc[k] = (uint8_t)((kStage / g.row_bytes > 32 ? 32 : kStage / g.row_bytes) & ~3); }
    }
  }
  if (n_rows == 0) return PBHC_OK;
  const int64_t tiles = (n_rows + 31) >> 5;
  static int sms = 0;
  if (!sms) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaFuncSetAttribute(gather_keys_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStagedSmem) != cudaSuccess) {
      set_error("pbhc_gather_keys: device setup failed: %s", cudaGetErrorString(cudaGetLastError())); return PBHC_E_CUDA;
    }
    sms = n;
  }
  // jobs of at least two tiles per resident warp (B200: 75776 rows): one 8-warp block per SM with the shared-memory ring
  // (measured: 77% of the HBM roofline at 98304 rows and 83% at 786432 rows of 3677 B, against 55% / 63% through registers)
  if (tiles >= 2LL * sms * kGatherWarps) {
    gather_keys_staged_kernel<<<sms, kGatherWarps * 32, kStagedSmem, (cudaStream_t)stream>>>(K, idx, n_rows, src_rows, bad_index_flag);
    return check_launch("pbhc_gather_keys");
  }
  static int resident = 0;  // persistent grid: every resident block of every SM, tiles strided over the warps
  if (!resident) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gather_keys_kernel, 256, 0) != cudaSuccess || per_sm < 1) {
      set_error("pbhc_gather_keys: occupancy query failed"); return PBHC_E_CUDA;
    }
    resident = per_sm * sms;
  }
  int64_t blocks = (tiles + 7) / 8;
  if (blocks > resident) blocks = resident;
  gather_keys_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(K, idx, n_rows, src_rows, bad_index_flag);
  return check_launch("pbhc_gather_keys");
}

extern "C" int pbhc_gather_rows(const void* src, int64_t src_rows, int64_t row_bytes, const int64_t* idx, int64_t n_rows, void* dst,
                                int32_t* bad_index_flag, void* stream) {
  PbhcGatherKey k{src, dst, row_bytes};
  return pbhc_gather_keys(&k, 1, idx, n_rows, src_rows, bad_index_flag, stream);
}

extern "C" int pbhc_timeout_bootstrap(float* rewards, const float* values, const uint8_t* time_outs, int64_t n_envs, int32_t R,
                                      float gamma, void* stream) {
  using namespace pbhc;
  if (!rewards || !values || !time_outs || n_envs < 0 || R <= 0) { set_error("pbhc_timeout_bootstrap: null / negative argument"); return PBHC_E_BADARG; }
  if (n_envs == 0) return PBHC_OK;
  const int64_t n = n_envs * R;
  timeout_bootstrap_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rewards, values, time_outs, n_envs, R, gamma);
  return check_launch("pbhc_timeout_bootstrap");
}
