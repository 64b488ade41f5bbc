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
  uint8_t vec_log2[PBHC_MAX_GATHER_KEYS];
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

__global__ void __launch_bounds__(256) gather_keys_kernel(const __grid_constant__ GatherKeys K, const int64_t* __restrict__ idx, int64_t n_rows,
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
    for (int k = 0; k < K.n_keys; ++k) {
      const uint32_t W = K.row_vecs[k];
      switch (K.vec_log2[k]) {
        case 4: gather_tile((const uint4*)K.src[k], (uint4*)K.dst[k], W, row0, rows, s, lane); break;
        case 3: gather_tile((const uint2*)K.src[k], (uint2*)K.dst[k], W, row0, rows, s, lane); break;
        case 2: gather_tile((const uint32_t*)K.src[k], (uint32_t*)K.dst[k], W, row0, rows, s, lane); break;
        default: gather_tile((const uint8_t*)K.src[k], (uint8_t*)K.dst[k], W, row0, rows, s, lane); break;
      }
    }
  }
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
    const int lg = (a & 15) == 0 ? 4 : (a & 7) == 0 ? 3 : (a & 3) == 0 ? 2 : 0;
    K.src[k] = g.src; K.dst[k] = g.dst; K.vec_log2[k] = (uint8_t)lg; K.row_vecs[k] = (uint32_t)(g.row_bytes >> lg);
  }
  if (n_rows == 0) return PBHC_OK;
  const int64_t tiles = (n_rows + 31) >> 5;
  int64_t blocks = (tiles + 7) / 8;
  if (blocks > 148LL * 8) blocks = 148LL * 8;  // persistent: 8 blocks of 8 warps per SM, tiles strided over them
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
