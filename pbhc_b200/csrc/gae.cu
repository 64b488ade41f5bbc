// gae.cu - K4: GAE reverse-time scan over the rollout buffer + advantage statistics + normalise.
// Thread per (env, reward head): [T,N,R] buffers are read/written fully coalesced; the time loop is
// the only sequential axis.  Reference: agents/ppo/ppo.py:338-355, agents/mh_ppo/mh_ppo.py:376-395.
#include "common.cuh"

namespace pbhc {

constexpr int kGaeThreads = 256;
constexpr int kTB = 8;  // time steps whose loads are issued together (and whose head sums share one barrier pair)

__device__ __forceinline__ double block_sum(double v, double* sm) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sm[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (l < (blockDim.x + 31) / 32) ? sm[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  return v;  // valid in thread 0
}

__global__ void __launch_bounds__(kGaeThreads) gae_kernel(
    const float* __restrict__ rewards, const float* __restrict__ values, const uint8_t* __restrict__ dones,
    const float* __restrict__ last_values, int T, int64_t N, int R, int E, float gamma, float lam,
    float* __restrict__ returns, float* __restrict__ adv_raw, double* __restrict__ stats3) {
  __shared__ float sm_adv[kTB][kGaeThreads];
  __shared__ double sm_red[8];
  const int local = threadIdx.x;
  const int e = local / R, r = local - e * R;
  const int64_t n0 = (int64_t)blockIdx.x * E;
  const int64_t n = n0 + e;
  const bool active = (e < E) && (n < N);
  const int64_t NR = N * R;
  const int64_t col = active ? n * R + r : 0;
  float next_v = active ? last_values[col] : 0.0f;
  float adv = 0.0f;
  double s1 = 0.0, s2 = 0.0;
  for (int t1 = T; t1 > 0; t1 -= kTB) {
    float rw[kTB], vl[kTB];
    uint8_t dn[kTB];
#pragma unroll
    for (int k = 0; k < kTB; ++k) {
      const int t = t1 - 1 - k;
      if (active && t >= 0) {
        rw[k] = rewards[t * NR + col];
        vl[k] = values[t * NR + col];
        dn[k] = dones[t * N + n];
      }
    }
#pragma unroll
    for (int k = 0; k < kTB; ++k) {
      const int t = t1 - 1 - k;
      if (t < 0) break;  // uniform
      float a = 0.0f;
      if (active) {
        const float nt = 1.0f - (dn[k] ? 1.0f : 0.0f);
        const float delta = rw[k] + nt * gamma * next_v - vl[k];
        adv = delta + nt * gamma * lam * adv;
        const float ret = adv + vl[k];
        returns[t * NR + col] = ret;
        a = ret - vl[k];
        next_v = vl[k];
      }
      if (R == 1) {
        if (active) { adv_raw[t * N + n] = a; s1 += (double)a; s2 += (double)a * (double)a; }
      } else {
        sm_adv[k][local] = a;
      }
    }
    if (R > 1) {
      // sum over the reward heads (mh_ppo.py:391): thread (step k, env ee) adds the R values of its env in head order
      __syncthreads();
      for (int idx = local; idx < kTB * E; idx += kGaeThreads) {
        const int k = idx / E, ee = idx - k * E, t = t1 - 1 - k;
        if (t >= 0 && n0 + ee < N) {
          float sum = 0.0f;
          for (int i = 0; i < R; ++i) sum += sm_adv[k][ee * R + i];
          adv_raw[t * N + n0 + ee] = sum;
          s1 += (double)sum; s2 += (double)sum * (double)sum;
        }
      }
      __syncthreads();
    }
  }
  s1 = block_sum(s1, sm_red);
  s2 = block_sum(s2, sm_red);
  if (threadIdx.x == 0) {
    const int64_t cnt = (N - n0 < E ? N - n0 : E) * (int64_t)T;
    atomicAdd(stats3 + 0, (double)cnt);
    atomicAdd(stats3 + 1, s1);
    atomicAdd(stats3 + 2, s2);
  }
}

__global__ void __launch_bounds__(256) adv_normalize_kernel(float* __restrict__ adv, int64_t n, const double* __restrict__ st) {
  const double cnt = st[0], mean_d = st[1] / cnt;
  double var = (st[2] - st[1] * mean_d) / (cnt - 1.0);  // unbiased (torch.std default)
  if (var < 0.0) var = 0.0;
  const float mean = (float)mean_d, denom = (float)sqrt(var) + 1e-8f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    adv[i] = (adv[i] - mean) / denom;
}

}  // namespace pbhc

extern "C" int pbhc_gae(const float* rewards, const float* values, const uint8_t* dones, const float* last_values,
                        int32_t T, int64_t N, int32_t R, float gamma, float lam, float* returns, float* adv_raw,
                        double* stats3, void* stream) {
  using namespace pbhc;
  if (!rewards || !values || !dones || !last_values || !returns || !adv_raw || !stats3) { set_error("pbhc_gae: null argument"); return PBHC_E_BADARG; }
  if (T <= 0 || N < 0 || R <= 0 || R > kGaeThreads) { set_error("pbhc_gae: bad shape T=%d N=%lld R=%d", T, (long long)N, R); return PBHC_E_BADARG; }
  if (N == 0) return PBHC_OK;
  const int E = kGaeThreads / R;
  const int64_t blocks = (N + E - 1) / E;
  if (blocks > 0x7fffffffLL) { set_error("pbhc_gae: too many envs"); return PBHC_E_RANGE; }
  gae_kernel<<<(unsigned)blocks, kGaeThreads, 0, (cudaStream_t)stream>>>(rewards, values, dones, last_values, T, N, R, E, gamma, lam, returns, adv_raw, stats3);
  return check_launch("pbhc_gae");
}

extern "C" int pbhc_adv_normalize(float* adv, int64_t n, const double* stats3_global, void* stream) {
  using namespace pbhc;
  if (!adv || !stats3_global || n < 0) { set_error("pbhc_adv_normalize: null argument"); return PBHC_E_BADARG; }
  if (n == 0) return PBHC_OK;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  adv_normalize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(adv, n, stats3_global);
  return check_launch("pbhc_adv_normalize");
}
