// track.cu - K3: the fused motion-tracking env step (lookup + differences + termination + rewards + reset +
// observations + history) as ONE persistent kernel.
//
// Mapping.  One warp owns one env per iteration of a grid-stride loop; lanes run over bodies (J <= 64), dofs
// (D <= 32), reward terms and flat row elements.  Work that is scalar per env (frame index / blend, root-frame
// velocities, heading quaternion) is hoisted into a "pre" phase where lane i computes it for the i-th env of the
// warp's next 32 envs, so its cost is amortised over the lanes instead of being replicated 32x.  The observation
// rows are assembled in per-warp shared memory in their final layout (each producer writes its value to every
// group that contains it, the history ring is scattered in with a per-launch map) and leave with 128/64-bit
// coalesced stores; simulator state enters with 128-bit loads.  The 16 per-env sums share one transposing
// butterfly (16 shuffles instead of 80) whose per-value summation order equals the plain xor butterfly's.
// Cross-env sums (adaptive-sigma errors, log means, reset statistics) are accumulated per lane in fp64 registers
// and reduced deterministically - no atomics, no host sync.  Lane-indexed configuration lives in shared memory.
// Arithmetic is the reference's unfused fp32 op sequence (-fmad=false): see DESIGN.md "arithmetic parity".
// Reference: legged_robot_base.py:297-338, 346-380, 408-489, 715-793, 944-1087; motion_tracking.py:330-358,
// 548-747, 1030-1328; helpers.py:128-152; history_handler.py:33-48 (paths under humanoidverse/envs|utils).

#ifndef __CUDACC_RTC__
#include <cstdlib>
#include <string>
#endif
#include "track_common.cuh"
#ifdef PBHC_SPEC
#include "track_spec.h"  // spec::cfg_bytes: the PbhcTrackCfg of one env config as a constant (csrc/jit.cu compiles this file once per config)
#endif

namespace pbhc {

#ifndef F_TW
#define F_TW 8   // warps (envs in flight) per block
#endif
#ifndef F_BPS
#define F_BPS 3  // resident blocks per SM the register budget is sized for
#endif
constexpr int kTW = F_TW;
constexpr int kTrackMaxBlocks = 148 * 8;
constexpr int kChunk = 16;                  // envs per pre phase (lanes used by it)
constexpr int kFutChunk = 3;                // future steps whose dof / key-body rows are fetched together (general_tracking window)

struct TrackParams {
  PbhcTrackCfg c;
  PbhcTableDesc tab;
  PbhcTrackIO io;
  int64_t n;
  int row_off[PBHC_MAX_OBS_GROUPS + 1];  // smem offsets of the obs rows; [n_groups] = history push staging
  int rows_len;                          // floats of `rows` (obs rows + push staging)
  int push_len;                          // floats pushed into the history per step
  int push_off[PBHC_MAX_HIST_KEYS];      // offset of key k inside the push staging
  int hist_base[PBHC_MAX_HIST_KEYS];     // offset of key k inside the global history row
  int hist_stage;                      // 1: the history ring is staged through the (then idle) simulator-state buffer
  int sm_body, sm_cf, sm_rows, sm_rv, sm_pre, sm_per_warp;  // per-warp shared-memory offsets (floats)
  int o_hmap, o_pmap, o_warp, warp_bytes;                    // byte offsets inside the dynamic shared memory
  int ob_body, ob_cf, ob_rows, ob_rv, ob_pre;                // byte offsets of the per-warp regions
  int e_pos, e_vel, e_ang, e_dp, e_dv, e_c;                  // float offsets inside a table record
  int body_vec, cf_vec, hist_vec, obs_vec[PBHC_MAX_OBS_GROUPS];  // widest vector access each row alignment allows
  int dump;                                                  // float offset (in `rows`) of the scratch region absent destinations write to
  int group_direct[PBHC_MAX_OBS_GROUPS];  // 1: the group holds only future-target sources, written straight to global memory
  int has_air_time;                      // the feet_air_time reward is active (only then the reference advances feet_air_time)
  int hist_any_mul;                      // some history destination has mul != 1
  int n_hdst_max;                        // max destinations of one history element
  int rew_extra;                         // bit (id - R_RADIAL_BODY_VEL): that R15 reward term is active
  int need_rootq;                        // relyaw sources or the feet-heading rewards are in use: the loop reads the root quaternion
  int fut_win;                           // the staging buffer has room for kFutChunk steps of future-window rows behind the per-step table
};

// block-level tables (built once per block from the launch parameters)
struct BlockTables {
  double scal[PBHC_SCALARS];
  float lim_lo[PBHC_MAX_DOF], lim_hi[PBHC_MAX_DOF], lim_vel[PBHC_MAX_DOF], lim_tq[PBHC_MAX_DOF];
  int term_gates;  // bit 0 / 1 / 2: the dof_pos / dof_vel / torque termination applies in this step (legged_robot_base.py:455-473)
  float r_scale[PBHC_MAX_REWARDS], r_pen[PBHC_MAX_REWARDS], r_sigma[PBHC_MAX_REWARDS];
  int r_kind[PBHC_MAX_REWARDS], r_arg[PBHC_MAX_REWARDS];
  float sum_div[32];                 // divisor applied to the k-th value of the 16-way reduction ([16..31]: second reduction of general_tracking)
  int sum_slot[32];                  // rv slot of the k-th value
  signed char key_slot[64];          // general_tracking: position of body j in the key-body list (-1: not a key body)
  int f_g[6], f_off[6], f_row[6];    // future-target source i: its direct group (-1: none) and offset there; its smem row offset (-1: none)
  alignas(16) int sdst[PBHC_NUM_OBS_SRC][8];     // (read as int4) {count, smem row offsets of source k in the obs groups that show it and in the push staging}
  float obs_scale[PBHC_NUM_OBS_SRC], obs_noise[PBHC_NUM_OBS_SRC];
  uint64_t noise_mask;               // bit k: source k carries observation noise
  int hist_head_w[PBHC_MAX_HIST_KEYS];
};

// k-th value of the 16-way reduction
enum Sum16 { U_UP_SQ, U_LO_SQ, U_VR_SQ, U_FEET_SQ, U_ROT_SQ, U_VEL_SQ, U_ANG_SQ, U_UP_N, U_LO_N, U_VR_N, U_DJA2, U_DJV2, U_TQ2,
             U_DACT2, U_DV2, U_ACC2 };
// general_tracking: first reduction = body sums (the dof sums U_DJA2.. keep their index in a second reduction)
enum Sum16G { GU_UP_SQ, GU_LO_SQ, GU_VR_SQ, GU_FEET_SQ, GU_UP_N, GU_LO_N, GU_VR_N, GU_KEY_N, GU_LUP_N, GU_LLO_N, GU_LVR_N, GU_LKEY_N,
              GU_LKEY_POS, GU_LKEY_ROT, GU_KEY_VEL, GU_KEY_ANG };
constexpr int V_SPARE = 30;  // rv slot nobody reads

__device__ void build_tables(const TrackParams& P, BlockTables& T, int* hmap, int* pmap) {
  const PbhcTrackCfg& c = P.c;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < PBHC_SCALARS; i += nt) T.scal[i] = P.io.scalars[i];
  __syncthreads();
  const uint64_t push_count = (uint64_t)T.scal[PBHC_S_PUSH_COUNT];
  if (tid < c.D) {  // soft limits of this launch (legged_robot_base.py:974-1002); the curriculum values are launch constants
    float lo = c.soft_lo[tid], hi = c.soft_hi[tid];
    if (c.cur_dof_pos.enable) {
      const float m = (c.hard_lo[tid] + c.hard_hi[tid]) / 2.0f, r = c.hard_hi[tid] - c.hard_lo[tid];
      const float cv = (float)T.scal[PBHC_S_SOFT_DOF_POS];
      lo = m - 0.5f * r * cv; hi = m + 0.5f * r * cv;
    }
    T.lim_lo[tid] = lo; T.lim_hi[tid] = hi;
    const float vs = c.cur_dof_vel.enable ? (float)T.scal[PBHC_S_SOFT_DOF_VEL] : c.soft_dof_vel_limit;
    T.lim_vel[tid] = c.vel_limit[tid] * vs;
    T.lim_tq[tid] = c.torque_limit[tid] * (c.cur_torque.enable ? (float)T.scal[PBHC_S_SOFT_TORQUE] : c.soft_torque_limit);
  }
  if (tid == 0) {  // `if torch.rand(1) < p` (legged_robot_base.py:455, 464, 473): one draw per step and check, the same for every env
    int gates = P.io.inject_term_gates;
    if (gates < 0) {
      const Rng rng = {make_uint2((uint32_t)c.seed, (uint32_t)(c.seed >> 32)), 0xffffffffu, 0xffffffffu, (uint32_t)push_count};
      const uint4 r = rng.draw(3);
      gates = (u01(r.x) < c.term_dof_pos_prob ? 1 : 0) | (u01(r.y) < c.term_dof_vel_prob ? 2 : 0) | (u01(r.z) < c.term_torque_prob ? 4 : 0);
    }
    T.term_gates = gates;
  }
  if (tid < PBHC_MAX_REWARDS) {
    int kind = K_RAW, arg = V_ONE;
    float sig = 1.0f;
    if (tid < c.n_rewards) {
      switch (c.reward_id[tid]) {
        case R_CONTACT_MASK: arg = V_CONTACT; break;
        case R_CONTACT_MASK_V2: arg = V_CONTACT_V2; break;
        case R_MAX_JOINT_POS: kind = K_EXP; arg = G_MAX_JOINT; break;
        case R_BODY_POS_EXT: kind = K_EXP2; arg = G_LOWER; break;
        case R_VR3: kind = K_EXP; arg = G_VR3; break;
        case R_FEET_POS: kind = K_EXP; arg = G_FEET; break;
        case R_BODY_ROT: kind = K_EXP; arg = G_ROT; break;
        case R_BODY_ANG_VEL: kind = K_EXP; arg = G_ANG_VEL; break;
        case R_BODY_VEL: kind = K_EXP; arg = G_VEL; break;
        case R_JOINT_POS: kind = K_EXP; arg = G_JOINT_POS; break;
        case R_JOINT_VEL: kind = K_EXP; arg = G_JOINT_VEL; break;
        case R_TORQUES: arg = V_TQ2; break;
        case R_ACTION_RATE: arg = V_DACT2; break;
        case R_DOF_VEL: arg = V_DV2; break;
        case R_DOF_ACC: arg = V_ACC2; break;
        case R_ORIENTATION: arg = V_ORIENT; break;
        case R_ALIVE: arg = V_ALIVE; break;
        case R_FEET_AIR_TIME: arg = V_AIR; break;
        case R_FEET_FORCES: arg = V_FORCE; break;
        case R_STUMBLE: arg = V_STUMBLE; break;
        case R_SLIPPAGE: arg = V_SLIP; break;
        case R_FOOT_SLIP: arg = V_FSLIP; break;
        case R_LIM_DOF_POS: arg = V_LIM_POS; break;
        case R_LIM_DOF_VEL: arg = V_LIM_VEL; break;
        case R_LIM_TORQUE: arg = V_LIM_TQ; break;
        case R_COLLISION: arg = V_COLLIDE; break;
        case R_ROOT_VEL: kind = K_EXP; arg = G_ROOT_VEL; break;
        case R_ANCHOR_POS: kind = K_EXP; arg = G_ANCHOR_POS; break;
        case R_ANCHOR_ROT: kind = K_EXP; arg = G_ANCHOR_ROT; break;
        case R_LKEY_POS: kind = K_EXP; arg = G_LKEY_POS; break;
        case R_LKEY_ROT: kind = K_EXP; arg = G_LKEY_ROT; break;
        case R_KEY_VEL: kind = K_EXP; arg = G_KEY_VEL; break;
        case R_KEY_ANG: kind = K_EXP; arg = G_KEY_ANG; break;
        case R_RADIAL_BODY_VEL: arg = V_RAD_BODY; break;
        case R_RADIAL_JOINT_VEL: arg = V_RAD_JOINT; break;
        case R_FEET_HEADING: arg = V_FEET_HEAD; break;
        case R_FEET_ORI: arg = V_FEET_ORI; break;
        case R_FEET_HEADING_CONTACT: arg = V_FEET_HEAD_C; break;
        case R_FEET_ORI_CONTACT: arg = V_FEET_ORI_C; break;
        default: break;
      }
      if (kind != K_RAW) sig = (float)T.scal[PBHC_S_SIGMA + arg];
    }
    T.r_kind[tid] = kind; T.r_arg[tid] = arg; T.r_sigma[tid] = sig;
    T.r_scale[tid] = tid < c.n_rewards ? c.reward_scale[tid] : 0.0f;
    T.r_pen[tid] = (tid < c.n_rewards && c.reward_is_penalty[tid]) ? (float)T.scal[PBHC_S_PENALTY] : 1.0f;
  }
  if (tid < 32) {
    float div = 1.0f; int slot = V_SPARE;
    const int u = tid & 15;
    const bool dof_table = c.general ? tid >= 16 : tid < 16;   // where the dof sums are reduced
    if (dof_table || !c.general) {
      switch (u) {
        case U_UP_SQ: if (!c.general) { div = (float)c.n_upper; slot = V_SIGMA_ERR + G_UPPER; } break;
        case U_LO_SQ: if (!c.general) { div = (float)c.n_lower; slot = V_SIGMA_ERR + G_LOWER; } break;
        case U_VR_SQ: if (!c.general) { div = (float)c.n_vr3; slot = V_SIGMA_ERR + G_VR3; } break;
        case U_FEET_SQ: if (!c.general) { div = (float)c.n_feet; slot = V_SIGMA_ERR + G_FEET; } break;
        case U_ROT_SQ: if (!c.general) { div = (float)c.J; slot = V_SIGMA_ERR + G_ROT; } break;
        case U_VEL_SQ: if (!c.general) { div = (float)c.J; slot = V_SIGMA_ERR + G_VEL; } break;
        case U_ANG_SQ: if (!c.general) { div = (float)c.J; slot = V_SIGMA_ERR + G_ANG_VEL; } break;
        case U_UP_N: if (!c.general) slot = V_UP_N; break;
        case U_LO_N: if (!c.general) slot = V_LO_N; break;
        case U_VR_N: if (!c.general) slot = V_VR_N; break;
        case U_DJA2: div = (float)c.D; slot = V_SIGMA_ERR + G_JOINT_POS; break;
        case U_DJV2: div = (float)c.D; slot = V_SIGMA_ERR + G_JOINT_VEL; break;
        case U_TQ2: slot = V_TQ2; break;
        case U_DACT2: slot = V_DACT2; break;
        case U_DV2: slot = V_DV2; break;
        default: slot = V_ACC2; break;
      }
      if (!c.general && tid >= 16) { div = 1.0f; slot = V_SPARE; }
    } else {
      const float nk = (float)c.n_key;
      switch (u) {
        case GU_UP_SQ: div = (float)c.n_upper; slot = V_SIGMA_ERR + G_UPPER; break;
        case GU_LO_SQ: div = (float)c.n_lower; slot = V_SIGMA_ERR + G_LOWER; break;
        case GU_VR_SQ: div = (float)c.n_vr3; slot = V_SIGMA_ERR + G_VR3; break;
        case GU_FEET_SQ: div = (float)c.n_feet; slot = V_SIGMA_ERR + G_FEET; break;
        case GU_UP_N: slot = V_UP_N; break;
        case GU_LO_N: slot = V_LO_N; break;
        case GU_VR_N: slot = V_VR_N; break;
        case GU_KEY_N: slot = V_KEY_N; break;
        case GU_LUP_N: slot = V_LUP_N; break;
        case GU_LLO_N: slot = V_LLO_N; break;
        case GU_LVR_N: slot = V_LVR_N; break;
        case GU_LKEY_N: slot = V_LKEY_N; break;
        case GU_LKEY_POS: div = nk; slot = V_SIGMA_ERR + G_LKEY_POS; break;
        case GU_LKEY_ROT: div = nk; slot = V_SIGMA_ERR + G_LKEY_ROT; break;
        case GU_KEY_VEL: div = nk; slot = V_SIGMA_ERR + G_KEY_VEL; break;
        default: div = nk; slot = V_SIGMA_ERR + G_KEY_ANG; break;
      }
    }
    T.sum_div[tid] = div; T.sum_slot[tid] = slot;
  }
  if (tid < 6) {  // where the future-target sources go (at most one direct group and one shared-memory row each: checked by the launcher)
    const int k = S_FUT_HEIGHT + tid;
    int fg = -1, fo = 0, fr = -1;
    for (int g = 0; g < c.n_groups; ++g) {
      const int o = c.src_dst[k][g];
      if (o < 0) continue;
      if (P.group_direct[g]) { fg = g; fo = o; } else fr = P.row_off[g] + o;
    }
    T.f_g[tid] = fg; T.f_off[tid] = fo; T.f_row[tid] = fr;
  }
  if (tid < 64) {
    int ks = -1;
    for (int i = 0; i < c.n_key; ++i) if (c.key_ids[i] == tid) ks = i;
    T.key_slot[tid] = (signed char)ks;
  }
#if F_PUT
  if (tid < PBHC_NUM_OBS_SRC) {  // {count, valid destinations...} of source tid
    int cnt = 0;
    for (int g = 0; g < c.n_groups; ++g) if (c.src_dst[tid][g] >= 0 && !P.group_direct[g]) T.sdst[tid][++cnt] = P.row_off[g] + c.src_dst[tid][g];
    for (int h = 0; h < c.n_hist_keys; ++h) if (c.hist_src[h] == tid) { T.sdst[tid][++cnt] = P.row_off[c.n_groups] + P.push_off[h]; break; }
    T.sdst[tid][0] = cnt;
    for (int i = cnt + 1; i < 8; ++i) T.sdst[tid][i] = P.dump;
  }
#else
  for (int i = tid; i < PBHC_NUM_OBS_SRC * 8; i += nt) {
    const int k = i >> 3, g = i & 7;
    int off = P.dump;  // absent destination: the scratch region
    if (g < PBHC_MAX_OBS_GROUPS) { if (g < c.n_groups && c.src_dst[k][g] >= 0 && !P.group_direct[g]) off = P.row_off[g] + c.src_dst[k][g]; }
    else if (g == PBHC_MAX_OBS_GROUPS) for (int h = 0; h < c.n_hist_keys; ++h) if (c.hist_src[h] == k) off = P.row_off[c.n_groups] + P.push_off[h];
    T.sdst[k][g] = off;
  }
#endif
  if (tid < PBHC_NUM_OBS_SRC) {  // (noise_scales[key] * current_noise_curriculum_value is a double product in the reference, helpers.py:142)
    T.obs_scale[tid] = c.obs_scale[tid];
    T.obs_noise[tid] = c.cur_noise.enable ? (float)((double)c.obs_noise[tid] * T.scal[PBHC_S_NOISE_CUR]) : c.obs_noise[tid];
  }
  if (tid == 0) {
    uint64_t m = 0;
    for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k) if (c.obs_noise[k] != 0.0f) m |= 1ull << k;
    T.noise_mask = m;
  }
  // history ring: physical slot of logical slot s of key k is (head_k + s) % len_k, head_k = -push_count mod len_k
  // (HistoryHandler keeps newest-first order, history_handler.py:40-48).  hmap[d][i]: smem destination of physical
  // element i of the global history row in the d-th obs segment that shows it; pmap[i]: physical element that
  // receives push element i.
  if (tid < c.n_hist_keys) {
    const int len = c.hist_len[tid];
    T.hist_head_w[tid] = (int)((len - ((push_count + 1) % len)) % len);
  }
  for (int i = tid; i < c.hist_row; i += nt) {
    int k = 0;
    while (k + 1 < c.n_hist_keys && i >= P.hist_base[k + 1]) ++k;
    const int len = c.hist_len[k], dim = c.hist_dim[k];
    const int r = i - P.hist_base[k], ps = r / dim, cc = r - ps * dim;
    const int head = (int)((len - (push_count % len)) % len);
    const int s = (ps - head + len) % len;  // logical slot
    int nd = 0;
    const int H = c.hist_row;
    for (int e = 0; e < c.n_hist_dst && nd < kMaxHistDstPerKey; ++e) {
      const PbhcHistDst& h = c.hist_dst[e];
      if (h.key != k) continue;
      const int m = s < h.len ? ((P.row_off[h.group] + h.offset + s * dim + cc) | (P.hist_any_mul ? (e << 24) : 0)) : P.dump + (i & 63);
      hmap[nd < 2 ? 2 * i + nd : nd * H + i] = m;
      ++nd;
    }
    for (; nd < (P.n_hdst_max < 2 ? 2 : P.n_hdst_max); ++nd) hmap[nd < 2 ? 2 * i + nd : nd * H + i] = P.dump + (i & 63);
  }
  __syncthreads();
  for (int i = tid; i < P.push_len; i += nt) {
    int k = 0;
    while (k + 1 < c.n_hist_keys && i >= P.push_off[k + 1]) ++k;
    pmap[i] = P.hist_base[k] + T.hist_head_w[k] * c.hist_dim[k] + (i - P.push_off[k]);
  }
  __syncthreads();
}

// Two measured alternatives are kept behind macros (B200, 65536 envs, track kernel alone): the kernel is bound by
// instruction issue, not by LSU wavefronts, so the variants with fewer instructions win even though they make more
// shared-memory wavefronts.  F_PUT=1: only valid destinations are stored (predicated)            -> 0.356 ms vs 0.345 ms
//                            F_HIST=1: scalar conflict-free history scatter instead of 128-bit loads -> 0.358 ms vs 0.345 ms
#ifndef F_PUT
#define F_PUT 0
#endif
#ifndef F_HIST
#define F_HIST 0
#endif
#if F_PUT
// writes `v` of element `e` of one observation source into every row that shows it.  sd = T.sdst[source] =
// {count, dst0, dst1, ...}: only the valid destinations (count is the same for every lane and env: uniform branches)
__device__ __forceinline__ void put1(const int* sd, float* rows, int e, float v, bool = true) {
  const int4 d = *reinterpret_cast<const int4*>(sd);
  float* r = rows + e;
  if (d.x > 0) r[d.y] = v;
  if (d.x > 1) r[d.z] = v;
  if (d.x > 2) r[d.w] = v;
  if (d.x > 3) { r[sd[4]] = v; if (d.x > 4) r[sd[5]] = v; }
}
__device__ __forceinline__ void put3(const int* sd, float* rows, int e, float x, float y, float z, bool = true) {
  const int4 d = *reinterpret_cast<const int4*>(sd);
  float* r = rows + e;
  if (d.x > 0) { r[d.y] = x; r[d.y + 1] = y; r[d.y + 2] = z; }
  if (d.x > 1) { r[d.z] = x; r[d.z + 1] = y; r[d.z + 2] = z; }
  if (d.x > 2) { r[d.w] = x; r[d.w + 1] = y; r[d.w + 2] = z; }
  for (int i = 3; i < d.x; ++i) { const int o = sd[i + 1]; r[o] = x; r[o + 1] = y; r[o + 2] = z; }
}

#else
// (wide: the config has more than two observation groups - a launch constant, so the branch is uniform)
__device__ __forceinline__ void put1(const int* sd, float* rows, int e, float v, bool wide = true) {
  const int4 d = *reinterpret_cast<const int4*>(sd);
  const int push = sd[4];
  float* r = rows + e;
  r[d.x] = v; r[d.y] = v; r[push] = v;
  if (wide) { r[d.z] = v; r[d.w] = v; }
}
__device__ __forceinline__ void put3(const int* sd, float* rows, int e, float x, float y, float z, bool wide = true) {
  const int4 d = *reinterpret_cast<const int4*>(sd);
  const int push = sd[4];
  float* r = rows + e;
  r[d.x] = x; r[d.x + 1] = y; r[d.x + 2] = z;
  r[d.y] = x; r[d.y + 1] = y; r[d.y + 2] = z;
  r[push] = x; r[push + 1] = y; r[push + 2] = z;
  if (wide) {
    r[d.z] = x; r[d.z + 1] = y; r[d.z + 2] = z;
    r[d.w] = x; r[d.w + 1] = y; r[d.w + 2] = z;
  }
}

#endif
// One element (three for emit3) of observation source k to every row that shows it: (value + U(-1,1)*noise) * scale with an
// independent draw per destination, like parse_observation() does per obs group and for the history (helpers.py:128-152;
// torch.rand_like there, Philox keyed by (seed, env, step, source, element) here).  Noise-free sources take the plain path.
struct NoiseKey { uint64_t seed; int64_t env; uint32_t step; bool wide; };  // (+ the launch constant put1 / put3 branch on)
__device__ __forceinline__ float noisy(float raw, uint32_t bits, float ns, float sc) { return (raw + (u01(bits) * 2.0f - 1.0f) * ns) * sc; }
__device__ __noinline__ void emit_noisy(const BlockTables& T, float* rows, int k, int e, float raw, const NoiseKey& nk) {
  const Rng rng = {make_uint2((uint32_t)nk.seed, (uint32_t)(nk.seed >> 32)), (uint32_t)nk.env, (uint32_t)((uint64_t)nk.env >> 32), nk.step};
  const uint32_t st = 4096u + ((uint32_t)k * 256u + (uint32_t)e) * 2u;
  const uint4 a = rng.draw(st);
  const uint32_t b = rng.draw(st + 1u).x;
  const float ns = T.obs_noise[k], sc = T.obs_scale[k];
  const int* sd = T.sdst[k];
  float* r = rows + e;
#if F_PUT
  const uint32_t bits[5] = {a.x, a.y, a.z, a.w, b};
  for (int i = 0; i < sd[0]; ++i) r[sd[i + 1]] = noisy(raw, bits[i], ns, sc);
#else
  r[sd[0]] = noisy(raw, a.x, ns, sc); r[sd[1]] = noisy(raw, a.y, ns, sc); r[sd[2]] = noisy(raw, a.z, ns, sc);
  r[sd[3]] = noisy(raw, a.w, ns, sc); r[sd[4]] = noisy(raw, b, ns, sc);
#endif
}
template <bool kNoise>
__device__ __forceinline__ void emit1(const BlockTables& T, float* rows, int k, int e, float raw, const NoiseKey& nk) {
  if (kNoise && ((T.noise_mask >> k) & 1ull)) emit_noisy(T, rows, k, e, raw, nk);
  else put1(T.sdst[k], rows, e, raw * T.obs_scale[k], nk.wide);
}
template <bool kNoise>
__device__ __forceinline__ void emit3(const BlockTables& T, float* rows, int k, int e, float x, float y, float z, const NoiseKey& nk) {
  if (kNoise && ((T.noise_mask >> k) & 1ull)) { emit_noisy(T, rows, k, e, x, nk); emit_noisy(T, rows, k, e + 1, y, nk); emit_noisy(T, rows, k, e + 2, z, nk); }
  else { const float sc = T.obs_scale[k]; put3(T.sdst[k], rows, e, x * sc, y * sc, z * sc, nk.wide); }
}

// global -> shared copy of `count` floats with the widest vector the row alignment allows (vec = 4, 2 or 1)
__device__ __forceinline__ void warp_load(float* dst, const float* src, int count, int vec, int lane) {
  if (vec == 4) {
    const float4* s = reinterpret_cast<const float4*>(src);
    float4* d = reinterpret_cast<float4*>(dst);
    for (int i = lane; i < (count >> 2); i += 32) d[i] = s[i];
    for (int i = (count & ~3) + lane; i < count; i += 32) dst[i] = src[i];
  } else if (vec == 2) {
    const float2* s = reinterpret_cast<const float2*>(src);
    float2* d = reinterpret_cast<float2*>(dst);
    for (int i = lane; i < (count >> 1); i += 32) d[i] = s[i];
    if ((count & 1) && lane == 0) dst[count - 1] = src[count - 1];
  } else {
    for (int i = lane; i < count; i += 32) dst[i] = src[i];
  }
}

// Asynchronous global -> shared copies (LDGSTS): the warp queues the next block of input while it computes on the current
// one, so no lane ever waits on HBM latency for the simulator state or the history ring.
__device__ __forceinline__ void cp_async16(uint32_t s, const void* g) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(g) : "memory"); }
__device__ __forceinline__ void cp_async8(uint32_t s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(g) : "memory"); }
__device__ __forceinline__ void cp_async4(uint32_t s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(g) : "memory"); }
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void warp_load_async(float* dst, const float* src, int count, int vec, int lane) {
  const uint32_t sd = (uint32_t)__cvta_generic_to_shared(dst);  // one address conversion per copy, byte offsets from there
  const char* gs = reinterpret_cast<const char*>(src);
  if (vec == 4) {
    const int nb = (count >> 2) << 4;
#pragma unroll 1
    for (int o = lane << 4; o < nb; o += 512) cp_async16(sd + o, gs + o);
#pragma unroll 1
    for (int i = (count & ~3) + lane; i < count; i += 32) cp_async4(sd + 4 * i, gs + 4 * i);
  } else if (vec == 2) {
    const int nb = (count >> 1) << 3;
#pragma unroll 1
    for (int o = lane << 3; o < nb; o += 256) cp_async8(sd + o, gs + o);
    if ((count & 1) && lane == 0) cp_async4(sd + 4 * (count - 1), gs + 4 * (count - 1));
  } else {
#pragma unroll 1
    for (int o = lane << 2; o < 4 * count; o += 128) cp_async4(sd + o, gs + o);
  }
}

// shared -> global copy of one observation row, clipped to +-clip (legged_robot_base.py:326-328)
__device__ __forceinline__ void warp_store_clipped(float* dst, const float* src, int count, int vec, float clip, int lane) {
  if (vec == 4) {
    const float4* s = reinterpret_cast<const float4*>(src);
    float4* d = reinterpret_cast<float4*>(dst);
    for (int i = lane; i < (count >> 2); i += 32) {
      float4 v = s[i];
      v.x = fminf(fmaxf(v.x, -clip), clip); v.y = fminf(fmaxf(v.y, -clip), clip);
      v.z = fminf(fmaxf(v.z, -clip), clip); v.w = fminf(fmaxf(v.w, -clip), clip);
      d[i] = v;
    }
  } else if (vec == 2) {
    const float2* s = reinterpret_cast<const float2*>(src);
    float2* d = reinterpret_cast<float2*>(dst);
    for (int i = lane; i < (count >> 1); i += 32) {
      float2 v = s[i];
      v.x = fminf(fmaxf(v.x, -clip), clip); v.y = fminf(fmaxf(v.y, -clip), clip);
      d[i] = v;
    }
  } else {
    for (int i = lane; i < count; i += 32) dst[i] = fminf(fmaxf(src[i], -clip), clip);
  }
}


// Scatters one env's history ring (global, physical slot order) into the observation rows.  hmap holds, per physical
// element, its first two destinations interleaved ({m0, m1}, -1 = none) followed by planes for destinations 2, 3.
__device__ __forceinline__ void hist_scatter(const float* gh, float* rows, const int* hmap, const PbhcTrackCfg& c, int n_dst,
                                             int any_mul, int vec, int lane, bool zero) {
  const int H = c.hist_row;
#if F_HIST
  (void)vec;  // consecutive lanes take consecutive elements: coalesced loads, conflict-free shared stores
  const int2* m2 = reinterpret_cast<const int2*>(hmap);
  if (!any_mul) {
#pragma unroll 4
    for (int i = lane; i < H; i += 32) {
      const float v = zero ? 0.0f : gh[i];
      const int2 m = m2[i];  // absent destinations point at the scratch region
      rows[m.x] = v; rows[m.y] = v;
    }
  } else {
    for (int i = lane; i < H; i += 32) {
      const float v = zero ? 0.0f : gh[i];
      const int2 m = m2[i];
      rows[m.x & 0xffffff] = v * c.hist_dst[m.x >> 24].mul;
      rows[m.y & 0xffffff] = v * c.hist_dst[m.y >> 24].mul;
    }
  }
#else
  if (!any_mul && vec == 4) {
    const float4* g4 = reinterpret_cast<const float4*>(gh);
    const int4* m4 = reinterpret_cast<const int4*>(hmap);
    for (int i = lane; i < (H >> 2); i += 32) {
      const float4 v = zero ? make_float4(0.f, 0.f, 0.f, 0.f) : g4[i];
      const int4 a = m4[2 * i], b = m4[2 * i + 1];
      rows[a.x] = v.x; rows[a.y] = v.x; rows[a.z] = v.y; rows[a.w] = v.y;
      rows[b.x] = v.z; rows[b.y] = v.z; rows[b.z] = v.w; rows[b.w] = v.w;
    }
  } else {
    for (int i = lane; i < H; i += 32) {
      const float v = zero ? 0.0f : gh[i];
      const int m0 = hmap[2 * i], m1 = hmap[2 * i + 1];
      rows[m0 & 0xffffff] = any_mul ? v * c.hist_dst[m0 >> 24].mul : v;
      rows[m1 & 0xffffff] = any_mul ? v * c.hist_dst[m1 >> 24].mul : v;
    }
  }
#endif
  for (int d = 2; d < n_dst; ++d)
    for (int i = lane; i < H; i += 32) {
      const int m = hmap[d * H + i];
      if (m >= 0) rows[m & 0xffffff] = zero ? 0.0f : (any_mul ? gh[i] * c.hist_dst[m >> 24].mul : gh[i]);
    }
}

// out-of-line copies of the larger helpers: the general_tracking kernel calls slerp from eight places and is bound by
// instruction fetch (instruction-cache hit rate 72 %, 2.2 stall cycles per instruction on `no_instruction`): one shared body
static __device__ __noinline__ Q4 slerp_nl(Q4 a, Q4 b, float t) { return slerp(a, b, t); }

// radial_velocity_potential (motion_tracking.py:78-94) from the three sums over the flattened vectors: dot = cur . ref,
// cc = |cur|^2, rr = |ref|^2.  F.cosine_similarity divides each vector by max(norm, 1e-8) before the product; here the sum of
// products is divided once (same value up to rounding).  sigma_theta = 0.75, alpha_norm = 0.4 (the function's defaults).
static __device__ __noinline__ float radial_potential(float dot, float cc, float rr) {
  const float nc = sqrtf(cc), nr = sqrtf(rr);
  const float cs = dot / (fmaxf(nc, 1e-8f) * fmaxf(nr, 1e-8f));
  const float pc = expf(-(1.0f - cs) / 0.75f);
  const float ratio = nc / nr;
  const float pn = ratio * expf(0.4f * (1.0f - powf(ratio, 2.5f)));
  return pc * pn;
}

__host__ __device__ inline int vec_of(int64_t row_floats, const void* base) {
  const uintptr_t a = (uintptr_t)base;
  if ((row_floats & 3) == 0 && (a & 15) == 0) return 4;
  if ((row_floats & 1) == 0 && (a & 7) == 0) return 2;
  return 1;
}

__device__ __forceinline__ int up4(int v) { return (v + 3) & ~3; }

template <bool kNoise, bool kGeneral>
__device__ __forceinline__ void track_body(const TrackParams& P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BlockTables& T = *reinterpret_cast<BlockTables*>(smem_raw);
#ifdef PBHC_SPEC
  // the config is a compile-time constant: flags fold, loops over J / D / future steps have known trip counts
  const PbhcTrackCfg& c = *reinterpret_cast<const PbhcTrackCfg*>(spec::cfg_bytes);
#else
  const PbhcTrackCfg& c = P.c;
#endif
  const PbhcTrackIO& io = P.io;
  const PbhcTableDesc& tab = P.tab;
  int* hmap = reinterpret_cast<int*>(smem_raw + P.o_hmap);
  int* pmap = reinterpret_cast<int*>(smem_raw + P.o_pmap);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int J = c.J, Jb = c.Jb, D = c.D, L = c.L, Cm = tab.contact_size;
  unsigned char* wp = smem_raw + P.o_warp + warp * P.warp_bytes;
  float* rv = reinterpret_cast<float*>(wp);                       // fixed-size tables first: compile-time offsets from wp
  float* pre = rv + kRv;
  float* sm_body = pre + kPre * kChunk;
  float* sm_cf = reinterpret_cast<float*>(wp + P.ob_cf);
  float* rows = reinterpret_cast<float*>(wp + P.ob_rows);
  build_tables(P, T, hmap, pmap);

  const float far_thr = (float)T.scal[PBHC_S_FAR_THR];
  const uint32_t push_count = (uint32_t)(uint64_t)T.scal[PBHC_S_PUSH_COUNT];
  const int e_pos = P.e_pos, e_vel = P.e_vel, e_ang = P.e_ang, e_dp = P.e_dp, e_dv = P.e_dv, e_c = P.e_c;
  const int body_vec = P.body_vec, cf_vec = P.cf_vec, hist_vec = P.hist_vec;
  uint64_t feet_mask = 0;
  for (int f = 0; f < c.n_feet; ++f) feet_mask |= 1ull << c.feet_ids[f];
  const int my_foot = lane < c.n_feet ? (lane == 0 ? c.feet_ids[0] : c.feet_ids[1]) : 0;
  const bool ld = lane < D;
  const float lim_lo = ld ? T.lim_lo[lane] : 0.f, lim_hi = ld ? T.lim_hi[lane] : 0.f;
  const float lim_v = ld ? T.lim_vel[lane] : 0.f, lim_t = ld ? T.lim_tq[lane] : 0.f;
  double acc_reg = 0.0, acc_reg2 = 0.0;  // slot = lane (and lane + 32 for general_tracking)

  const int stride = (int)gridDim.x * kTW, n_envs = (int)P.n;  // P.n < 2^31 / kChunk (checked by the launcher): 32-bit env ids
  const int hist_stage = P.hist_stage;
  auto fetch_sim_state = [&](int64_t m) {  // queue env m's rigid-body state and contact forces into the staging buffers
    warp_load_async(sm_body, io.body_state + m * io.body_env_stride, Jb * 13, body_vec, lane);
    warp_load_async(sm_cf, io.contact_forces + m * (Jb * 3), Jb * 3, cf_vec, lane);
  };
  if ((int)blockIdx.x * kTW + warp < n_envs) fetch_sim_state((int)blockIdx.x * kTW + warp);
  for (int chunk = (int)blockIdx.x * kTW + warp; chunk < n_envs; chunk += kChunk * stride) {
    // ---- pre phase: lane i = i-th env of this chunk (legged_robot_base.py:346-353, motion_tracking.py:548-558, 678-739) ----
    __syncwarp();
    {
      const int64_t m = chunk + lane * stride;
      if (lane < kChunk && m < n_envs) {
        const int64_t ep = io.episode_length_buf[m] + 1;  // legged_robot_base.py:299
        const float start = io.env_f32[m * PBHC_ENVF_STRIDE];
        const int clip = resolve_clip(tab, io.clip_of_motion, io.motion_ids[m]);
        const float t = (float)(ep + 1) * c.dt + start;  // kick_motion_res, motion_tracking.py:554
        const FrameBlend fb = frame_blend(tab, clip, t);
        const float* root = io.root_state + m * 13;
        const Q4 rq = {root[3], root[4], root[5], root[6]};
        const V3 blv = quat_rotate_inverse(rq, {root[7], root[8], root[9]});
        const V3 bav = quat_rotate_inverse(rq, {root[10], root[11], root[12]});
        const V3 pg = quat_rotate_inverse(rq, {0.0f, 0.0f, -1.0f});
        const Q4 hinv = heading_quat_inv_fast(rq);  // (within 1.2e-7 of the exact value like the reference's atan2 / sin / cos sequence, a fraction of its code)
        pre[(P_F0) * kChunk + lane] = __int_as_float((int)fb.f0);
        pre[(P_F1) * kChunk + lane] = __int_as_float((int)fb.f1);
        pre[(P_BLEND) * kChunk + lane] = fb.blend;
        pre[(P_PHASE) * kChunk + lane] = t / tab.clip_length[clip];  // motion_tracking.py:738-739
        pre[(P_HZ) * kChunk + lane] = hinv.z; pre[(P_HW) * kChunk + lane] = hinv.w;
        pre[(P_RPX) * kChunk + lane] = root[0]; pre[(P_RPY) * kChunk + lane] = root[1]; pre[(P_RPZ) * kChunk + lane] = root[2];
        pre[(P_BLV + 0) * kChunk + lane] = blv.x; pre[(P_BLV + 1) * kChunk + lane] = blv.y; pre[(P_BLV + 2) * kChunk + lane] = blv.z;
        pre[(P_BAV + 0) * kChunk + lane] = bav.x; pre[(P_BAV + 1) * kChunk + lane] = bav.y; pre[(P_BAV + 2) * kChunk + lane] = bav.z;
        pre[(P_PG + 0) * kChunk + lane] = pg.x; pre[(P_PG + 1) * kChunk + lane] = pg.y; pre[(P_PG + 2) * kChunk + lane] = pg.z;
        pre[(P_CLIP) * kChunk + lane] = __int_as_float(clip);
        pre[(P_GRAVN) * kChunk + lane] = norm2(pg.x, pg.y);
        pre[(P_ORIENT) * kChunk + lane] = pg.x * pg.x + pg.y * pg.y;
        if constexpr (kGeneral) {  // self.rpy, legged_robot_base.py:349
          float roll, pitch;
          euler_roll_pitch(rq, roll, pitch);
          pre[(P_ROLL) * kChunk + lane] = roll; pre[(P_PITCH) * kChunk + lane] = pitch;
        }
      }
    }
    __syncwarp();
    for (int e = 0; e < kChunk; ++e) {
      const int n32 = chunk + e * stride;
      if (n32 >= n_envs) break;
      const int64_t n = n32;  // sign-extended 32-bit: every n * row_size below is one IMAD.WIDE
      const NoiseKey nk = {P.c.seed, n + io.env_offset, push_count, c.n_groups > 2};
      cp_async_wait();  // this env's simulator state, queued while the previous env was being written out
      __syncwarp();
      // ---- A. per-env scalars ------------------------------------------------------------------------------
      if (lane == 0) {  // per-env scalars: loaded with the first wave of loads, parked in shared memory until phases D-G
        const float4* ge = reinterpret_cast<const float4*>(io.env_f32 + n * PBHC_ENVF_STRIDE);
        *reinterpret_cast<float4*>(rv + V_ENVF) = ge[0];
        *reinterpret_cast<float4*>(rv + V_ENVF + 4) = ge[1];
        *reinterpret_cast<long long*>(rv + V_EP) = io.episode_length_buf[n];
      }
      const int clip = __float_as_int(pre[(P_CLIP) * kChunk + e]);
      const float b = pre[(P_BLEND) * kChunk + e], omb = 1.0f - b;
      const float* r0 = tab.records + (int64_t)__float_as_int(pre[(P_F0) * kChunk + e]) * tab.rec_stride;
      const float* r1 = tab.records + (int64_t)__float_as_int(pre[(P_F1) * kChunk + e]) * tab.rec_stride;
      const float hz = pre[(P_HZ) * kChunk + e], hw = pre[(P_HW) * kChunk + e];
      const V3 rp = {pre[(P_RPX) * kChunk + e], pre[(P_RPY) * kChunk + e], pre[(P_RPZ) * kChunk + e]};
      const float org[3] = {io.env_origins[n * 3], io.env_origins[n * 3 + 1], io.env_origins[n * 3 + 2]};
      // history ring -> its places in the observation rows (HistoryHandler.query, history_handler.py:46-48)
      if (c.hist_row > 0 && !hist_stage) hist_scatter(io.hist + n * c.hist_row, rows, hmap, c, P.n_hdst_max, P.hist_any_mul, hist_vec, lane, false);
      __syncwarp();
      // relyaw = rpy[:, 2] - ref_init_rpy[0, 2] and its inverse heading quaternion (motion_tracking.py:686-687, rotations.py:309-322);
      // the root quaternion itself for the feet-heading rewards.  Rare configs: every lane computes the same values
      Q4 rootq = {0.f, 0.f, 0.f, 1.f};
      float relyaw = 0.f, ry_z = 0.f, ry_w = 1.f;
      if (P.need_rootq) {
        const float* root = io.root_state + n * 13;
        rootq = {root[3], root[4], root[5], root[6]};
        if (c.src_used_mask & ((1ull << S_RELYAW) | (1ull << S_LOCAL_REF_RELYAW))) {
          relyaw = euler_yaw(rootq) - c.ref_init_yaw;
          const float half = relyaw * 0.5f;
          ry_z = -sinf(half); ry_w = cosf(half);
        }
      }
      // ---- general_tracking: anchor quantities (general_tracking.py:738-803), the same values in every lane -----------
      Q4 g_raq = {0.f, 0.f, 0.f, 1.f}, g_dori = {0.f, 0.f, 0.f, 1.f};  // robot anchor rotation; delta_ori_w
      V3 g_ap = {0.f, 0.f, 0.f}, g_dpos = {0.f, 0.f, 0.f};              // reference anchor position; delta_pos_w
      bool t_refz = false, t_ori = false;
      if constexpr (kGeneral) {
        const int a = c.anchor_index;
        const float* bs = sm_body + a * 13;
        const V3 rap = {bs[0], bs[1], bs[2]};
        g_raq = {bs[3], bs[4], bs[5], bs[6]};
        const float4 qa = __ldg(reinterpret_cast<const float4*>(r0) + a), qb = __ldg(reinterpret_cast<const float4*>(r1) + a);
        const Q4 ar = slerp_nl({qa.x, qa.y, qa.z, qa.w}, {qb.x, qb.y, qb.z, qb.w}, b);
        g_ap = {(omb * __ldg(r0 + e_pos + 3 * a) + b * __ldg(r1 + e_pos + 3 * a)) + org[0],
                (omb * __ldg(r0 + e_pos + 3 * a + 1) + b * __ldg(r1 + e_pos + 3 * a + 1)) + org[1],
                (omb * __ldg(r0 + e_pos + 3 * a + 2) + b * __ldg(r1 + e_pos + 3 * a + 2)) + org[2]};
        // delta_pos_w aliases robot_anchor_pos_w_repeat in the reference (:752-753): the z overwrite is seen by
        // _obs_local_body_pos too, so g_dpos stands for both
        g_dpos = {rap.x, rap.y, g_ap.z};
        g_dori = yaw_quat_fast(quat_mul(g_raq, quat_conj(ar)));
        float m6[6];
        mat_cols2(quat_mul(quat_conj(g_raq), ar), m6);                                     // _obs_anchor_ref_rot (:785-792)
        const V3 arp = quat_apply(quat_conj(g_raq), {g_ap.x - rap.x, g_ap.y - rap.y, g_ap.z - rap.z});  // _obs_anchor_ref_pos
        if (lane < 6) {
          const float mv = lane == 0 ? m6[0] : lane == 1 ? m6[1] : lane == 2 ? m6[2] : lane == 3 ? m6[3] : lane == 4 ? m6[4] : m6[5];
          emit1<kNoise>(T, rows, S_ANCHOR_ROT, lane, mv, nk);
          if (lane < 3) emit1<kNoise>(T, rows, S_ANCHOR_POS, lane, lane == 0 ? arp.x : lane == 1 ? arp.y : arp.z, nk);
        }
        const V3 dpa = {g_ap.x - rap.x, g_ap.y - rap.y, g_ap.z - rap.z};                   // dif_anchor_body_pos
        const float ang = quat_angle(quat_mul(ar, quat_conj(g_raq)));                       // dif_global_body_rot[anchor]
        // dif_root_velocity (:656-657): reference root velocity in the reference root frame minus base_lin_vel
        const float4 ra0 = __ldg(reinterpret_cast<const float4*>(r0)), rb0 = __ldg(reinterpret_cast<const float4*>(r1));
        const Q4 rrq = slerp_nl({ra0.x, ra0.y, ra0.z, ra0.w}, {rb0.x, rb0.y, rb0.z, rb0.w}, b);
        const V3 rrv = {omb * __ldg(r0 + e_vel) + b * __ldg(r1 + e_vel), omb * __ldg(r0 + e_vel + 1) + b * __ldg(r1 + e_vel + 1),
                        omb * __ldg(r0 + e_vel + 2) + b * __ldg(r1 + e_vel + 2)};
        const V3 rl = quat_rotate_inverse(rrq, rrv);
        const V3 drv = {rl.x - pre[(P_BLV + 0) * kChunk + e], rl.y - pre[(P_BLV + 1) * kChunk + e], rl.z - pre[(P_BLV + 2) * kChunk + e]};
        const float apz = g_ap.z - rap.z;                                                    // dif_anchor_pos_z (:797)
        const float aori = quat_rotate_inverse(ar, {0.0f, 0.0f, -1.0f}).z - quat_rotate_inverse(g_raq, {0.0f, 0.0f, -1.0f}).z;
        t_refz = c.term_ref_pos_z && fabsf(apz) > c.term_ref_pos_z_thr;                      // :241-249
        t_ori = c.term_ref_ori && fabsf(aori) > c.term_ref_ori_thr;
        if (lane == 0) {
          rv[V_SIGMA_ERR + G_ANCHOR_POS] = ((dpa.x * dpa.x + dpa.y * dpa.y) + dpa.z * dpa.z) / 3.0f;
          rv[V_SIGMA_ERR + G_ANCHOR_ROT] = ang * ang;
          rv[V_SIGMA_ERR + G_ROOT_VEL] = ((drv.x * drv.x + drv.y * drv.y) + drv.z * drv.z) / 3.0f;
        }
      }
      // ---- B. bodies: reference (lerp / slerp of the two frames), extended state, differences, heading-frame features
      float sums[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) sums[i] = 0.0f;
      int far = 0, n_collide = 0, bodyz = 0, hit = 0;
      float rb_dot = 0.f, rb_cc = 0.f, rb_rr = 0.f;  // teleop_radial_body_velocity_extend
#pragma unroll 1
      for (int j0 = 0; j0 < J; j0 += 32) {  // (one copy of the body section: the specialised kernel knows J and would duplicate it for a 33rd body)
        const int j = j0 + lane;
        bool collide = false;
        if (j < J) {
          V3 p, v, w; Q4 q;
          if (j < Jb) {
            const float* bs = sm_body + j * 13;
            p = {bs[0], bs[1], bs[2]}; q = {bs[3], bs[4], bs[5], bs[6]}; v = {bs[7], bs[8], bs[9]}; w = {bs[10], bs[11], bs[12]};
          } else {  // motion_tracking.py:619-643
            const int x = j - Jb;
            const float* bs = sm_body + c.ext_parent[x] * 13;
            const Q4 qp = {bs[3], bs[4], bs[5], bs[6]};
            const V3 off = {c.ext_pos[x][0], c.ext_pos[x][1], c.ext_pos[x][2]};
            const Q4 qe = {c.ext_rot_xyzw[x][0], c.ext_rot_xyzw[x][1], c.ext_rot_xyzw[x][2], c.ext_rot_xyzw[x][3]};
            const V3 r = quat_rotate(qe, quat_rotate(qp, off));
            p = {r.x + bs[0], r.y + bs[1], r.z + bs[2]};
            q = quat_mul(qp, qe);
            w = {bs[10], bs[11], bs[12]};
            v = {bs[7] + (w.y * off.z - w.z * off.y), bs[8] + (w.z * off.x - w.x * off.z), bs[9] + (w.x * off.y - w.y * off.x)};
          }
          const float4 qa = __ldg(reinterpret_cast<const float4*>(r0) + j);
          const float4 qb = __ldg(reinterpret_cast<const float4*>(r1) + j);
          const Q4 rr = slerp_nl({qa.x, qa.y, qa.z, qa.w}, {qb.x, qb.y, qb.z, qb.w}, b);
          float rpz[3], rvv[3], rw[3];
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            rpz[k] = (omb * __ldg(r0 + e_pos + 3 * j + k) + b * __ldg(r1 + e_pos + 3 * j + k)) + org[k];
            rvv[k] = omb * __ldg(r0 + e_vel + 3 * j + k) + b * __ldg(r1 + e_vel + 3 * j + k);
            rw[k] = omb * __ldg(r0 + e_ang + 3 * j + k) + b * __ldg(r1 + e_ang + 3 * j + k);
          }
          if (io.ref_body_pos_extend) {
            float* o = io.ref_body_pos_extend + n * (3 * J) + 3 * j;
            o[0] = rpz[0]; o[1] = rpz[1]; o[2] = rpz[2];
          }
          if (io.ref_body_rot_extend) reinterpret_cast<float4*>(io.ref_body_rot_extend)[n * J + j] = make_float4(rr.x, rr.y, rr.z, rr.w);
          const V3 dp = {rpz[0] - p.x, rpz[1] - p.y, rpz[2] - p.z};
          const float dq0 = rr.x - q.x, dq1 = rr.y - q.y, dq2 = rr.z - q.z, dq3 = rr.w - q.w;
          const V3 dv = {rvv[0] - v.x, rvv[1] - v.y, rvv[2] - v.z};
          const V3 dw = {rw[0] - w.x, rw[1] - w.y, rw[2] - w.z};
          const float m_pos = ((dp.x * dp.x + dp.y * dp.y) + dp.z * dp.z) / 3.0f;
          const float nrm = norm3(dp.x, dp.y, dp.z);
          const uint64_t bit = 1ull << j;
          if constexpr (!kGeneral) {
            if (c.upper_mask & bit) { sums[U_UP_SQ] += m_pos; sums[U_UP_N] += nrm; }
            if (c.lower_mask & bit) { sums[U_LO_SQ] += m_pos; sums[U_LO_N] += nrm; }
            if (c.vr3_mask & bit) { sums[U_VR_SQ] += m_pos; sums[U_VR_N] += nrm; }
            if (feet_mask & bit) sums[U_FEET_SQ] += m_pos;
            sums[U_ROT_SQ] += (((dq0 * dq0 + dq1 * dq1) + dq2 * dq2) + dq3 * dq3) / 4.0f;
            sums[U_VEL_SQ] += ((dv.x * dv.x + dv.y * dv.y) + dv.z * dv.z) / 3.0f;
            sums[U_ANG_SQ] += ((dw.x * dw.x + dw.y * dw.y) + dw.z * dw.z) / 3.0f;
          } else {  // general_tracking.py:760-783, 303-335, 1131-1222
            const V3 rel = quat_apply(g_dori, {rpz[0] - g_ap.x, rpz[1] - g_ap.y, rpz[2] - g_ap.z});
            const V3 dl2 = {(g_dpos.x + rel.x) - p.x, (g_dpos.y + rel.y) - p.y, (g_dpos.z + rel.z) - p.z};  // dif_local_body_pos
            const float nl = norm3(dl2.x, dl2.y, dl2.z);
            if (c.upper_mask & bit) { sums[GU_UP_SQ] += m_pos; sums[GU_UP_N] += nrm; sums[GU_LUP_N] += nl; }
            if (c.lower_mask & bit) { sums[GU_LO_SQ] += m_pos; sums[GU_LO_N] += nrm; sums[GU_LLO_N] += nl; }
            if (c.vr3_mask & bit) { sums[GU_VR_SQ] += m_pos; sums[GU_VR_N] += nrm; sums[GU_LVR_N] += nl; }
            if (feet_mask & bit) sums[GU_FEET_SQ] += m_pos;
            if (c.key_mask & bit) {
              sums[GU_KEY_N] += nrm; sums[GU_LKEY_N] += nl;
              sums[GU_LKEY_POS] += ((dl2.x * dl2.x + dl2.y * dl2.y) + dl2.z * dl2.z) / 3.0f;
              const float la = quat_angle(quat_mul(quat_mul(g_dori, rr), quat_conj(q)));  // dif_local_body_rot
              sums[GU_LKEY_ROT] += la * la;
              sums[GU_KEY_VEL] += ((dv.x * dv.x + dv.y * dv.y) + dv.z * dv.z) / 3.0f;
              sums[GU_KEY_ANG] += ((dw.x * dw.x + dw.y * dw.y) + dw.z * dw.z) / 3.0f;
              const int ks = T.key_slot[j];
              const V3 lp = quat_apply(quat_conj(g_raq), {p.x - g_dpos.x, p.y - g_dpos.y, p.z - g_dpos.z});  // _obs_local_body_pos
              emit3<kNoise>(T, rows, S_LKEY_POS, 3 * ks, lp.x, lp.y, lp.z, nk);
              float m6[6];
              mat_cols2(quat_mul(quat_conj(g_raq), q), m6);                                  // _obs_local_body_rot
              emit3<kNoise>(T, rows, S_LKEY_ROT, 6 * ks, m6[0], m6[1], m6[2], nk);
              emit3<kNoise>(T, rows, S_LKEY_ROT, 6 * ks + 3, m6[3], m6[4], m6[5], nk);
            }
            if ((c.body_z_mask & bit) && fabsf(dl2.z) > c.term_body_z_thr) bodyz = 1;        // :251-255
            (void)dq0; (void)dq1; (void)dq2; (void)dq3;
          }
          if (nrm > far_thr) far = 1;
          if (j < Jb && (c.penalised_mask & bit)) {
            const float* fc = sm_cf + 3 * j;
            collide = __fmaf_rn(fc[2], fc[2], __fmaf_rn(fc[1], fc[1], fc[0] * fc[0])) > 0x1.47ae16p-7f;  // |f| > 0.1 on the squared norm (0x1.47ae16p-7: the largest fp32 whose root is <= 0.1f): no square root of the exact zeros of bodies in the air
          }
          if (c.term_contact && j < Jb && (c.term_contact_mask & bit)) {  // |f| > 1 (legged_robot_base.py:435)
            const float* fc = sm_cf + 3 * j;
            if (__fmaf_rn(fc[2], fc[2], __fmaf_rn(fc[1], fc[1], fc[0] * fc[0])) > 0x1.000002p0f) hit = 1;
          }
          if (!kGeneral || (c.src_used_mask & ((1ull << S_DIF_LOCAL) | (1ull << S_LOCAL_REF)))) {
            const V3 dl = quat_rotate_z(hz, hw, dp);  // motion_tracking.py:694-697
            emit3<kNoise>(T, rows, S_DIF_LOCAL, 3 * j, dl.x, dl.y, dl.z, nk);
            const V3 lr = quat_rotate_z(hz, hw, {rpz[0] - rp.x, rpz[1] - rp.y, rpz[2] - rp.z});  // :705-707
            emit3<kNoise>(T, rows, S_LOCAL_REF, 3 * j, lr.x, lr.y, lr.z, nk);
          }
          if (c.src_used_mask & ((1ull << S_LOCAL_REF_VEL) | (1ull << S_GLOBAL_REF_VEL))) {
            const V3 lv = quat_rotate_z(hz, hw, {rvv[0], rvv[1], rvv[2]});
            emit3<kNoise>(T, rows, S_LOCAL_REF_VEL, 3 * j, lv.x, lv.y, lv.z, nk);
            emit3<kNoise>(T, rows, S_GLOBAL_REF_VEL, 3 * j, rvv[0], rvv[1], rvv[2], nk);
          }
          if (c.src_used_mask & (1ull << S_LOCAL_REF_RELYAW)) {  // the reference VELOCITIES in the relyaw frame (motion_tracking.py:720-721)
            const V3 ly = quat_rotate_z(ry_z, ry_w, {rvv[0], rvv[1], rvv[2]});
            emit3<kNoise>(T, rows, S_LOCAL_REF_RELYAW, 3 * j, ly.x, ly.y, ly.z, nk);
          }
          if ((c.src_used_mask & (1ull << S_VR3PT)) && (c.vr3_mask & bit)) {
            // _obs_vr_3point_pos (:724-731): `heading_inv_rot.repeat(3, 1)` pairs row 3 n + k with the heading of env (3 n + k) mod N
            for (int k = 0; k < 3; ++k)
              if (c.vr3_ids[k] == j) {
                const int64_t src = (3 * n + k) % P.n;
                const float2 h = reinterpret_cast<const float2*>(io.heading_scratch)[src];
                const V3 o = quat_rotate_z(h.x, h.y, {rpz[0] - rp.x, rpz[1] - rp.y, rpz[2] - rp.z});
                emit3<kNoise>(T, rows, S_VR3PT, 3 * k, o.x, o.y, o.z, nk);
              }
          }
          if (!kGeneral && (P.rew_extra & 1)) {  // cur = body velocity, ref = dif + cur (motion_tracking.py:1238-1244)
            const V3 rf = {dv.x + v.x, dv.y + v.y, dv.z + v.z};
            rb_dot += (v.x * rf.x + v.y * rf.y) + v.z * rf.z;
            rb_cc += (v.x * v.x + v.y * v.y) + v.z * v.z;
            rb_rr += (rf.x * rf.x + rf.y * rf.y) + rf.z * rf.z;
          }
        }
        n_collide += __popc(__ballot_sync(0xffffffffu, collide));
      }
      const bool any_far = __any_sync(0xffffffffu, far);
      if (!kGeneral && (P.rew_extra & 1)) {
        rb_dot = warp_sum(rb_dot); rb_cc = warp_sum(rb_cc); rb_rr = warp_sum(rb_rr);
        if (lane == 0) rv[V_RAD_BODY] = radial_potential(rb_dot, rb_cc, rb_rr);
      }
      // ---- C. dofs (lane = dof) ---------------------------------------------------------------------------
      float dpos = 0.f, dvel = 0.f, act = 0.f, tq = 0.f, lact = 0.f, ldv = 0.f, ddef = 0.f, kp = 1.f, kd = 1.f;
      float dja = 0.f, djv = 0.f;
      if (ld) {
        const float2 ds = reinterpret_cast<const float2*>(io.dof_state)[n * D + lane];
        dpos = ds.x; dvel = ds.y;
        act = io.actions[n * D + lane]; tq = io.torques[n * D + lane];
        lact = io.last_actions[n * D + lane]; ldv = io.last_dof_vel[n * D + lane];
        ddef = io.default_dof_pos[n * io.default_dof_stride + lane];
        kp = io.kp_scale[n * D + lane]; kd = io.kd_scale[n * D + lane];
        dja = (omb * __ldg(r0 + e_dp + lane) + b * __ldg(r1 + e_dp + lane)) - dpos;   // motion_tracking.py:661-665
        djv = (omb * __ldg(r0 + e_dv + lane) + b * __ldg(r1 + e_dv + lane)) - dvel;
      }
      const float max_dja = warp_max(fabsf(dja));
      if constexpr (!kGeneral) {
        if (P.rew_extra & 2) {  // teleop_radial_joint_velocity (motion_tracking.py:1286-1293): cur = dof_vel, ref = dif + cur
          const float rf = djv + dvel;
          const float d0 = warp_sum(ld ? dvel * rf : 0.f), d1 = warp_sum(ld ? dvel * dvel : 0.f), d2 = warp_sum(ld ? rf * rf : 0.f);
          if (lane == 0) rv[V_RAD_JOINT] = radial_potential(d0, d1, d2);
        }
        if (c.n_future > 0 && (c.src_used_mask & ((1ull << S_FUT_REF_DOF_POS) | (1ull << S_FUT_REF_DOF_VEL)))) {
          // kick_motion_res_multiplestep (motion_tracking.py:562-579, 594-599): reference dofs at (episode_length + 1 + i) dt + start
          const int64_t ep1 = *reinterpret_cast<const long long*>(rv + V_EP) + 1;
          const float st = rv[V_ENVF];
#pragma unroll 1
          for (int s2 = 0; s2 < c.n_future; ++s2) {
            const FrameBlend fb = frame_blend(tab, clip, (float)(ep1 + 1 + s2) * c.dt + st);
            const float* q0 = tab.records + fb.f0 * tab.rec_stride;
            const float* q1 = tab.records + fb.f1 * tab.rec_stride;
            const float bb = fb.blend, ob = 1.0f - bb;
            if (ld) {
              emit1<kNoise>(T, rows, S_FUT_REF_DOF_POS, s2 * D + lane, ob * __ldg(q0 + e_dp + lane) + bb * __ldg(q1 + e_dp + lane), nk);
              emit1<kNoise>(T, rows, S_FUT_REF_DOF_VEL, s2 * D + lane, ob * __ldg(q0 + e_dv + lane) + bb * __ldg(q1 + e_dv + lane), nk);
            }
          }
        }
      }
      if constexpr (kGeneral) {  // the body sums take the whole first reduction; the dof sums get their own
        const float tot = warp_sum16(sums, lane);
        const int idx = (lane >> 1) & 15;
        rv[T.sum_slot[idx]] = tot / T.sum_div[idx];
#pragma unroll
        for (int i = 0; i < 16; ++i) sums[i] = 0.0f;
      }
      sums[U_DJA2] = dja * dja; sums[U_DJV2] = djv * djv; sums[U_TQ2] = tq * tq;
      const float dact = lact - act;
      sums[U_DACT2] = dact * dact; sums[U_DV2] = dvel * dvel;
      const float dacc = (ldv - dvel) / c.dt;
      sums[U_ACC2] = ld ? dacc * dacc : 0.f;
      int dof_viol = 0;  // bit 0 / 1 / 2: this dof is beyond its position / velocity / torque termination limit (legged_robot_base.py:449-476)
      if (ld) {
        if (c.term_dof_pos && (dpos < c.term_pos_lo[lane] || dpos > c.term_pos_hi[lane])) dof_viol |= 1;
        if (c.term_dof_vel && fabsf(dvel) > c.vel_limit[lane] * c.term_dof_vel_scale) dof_viol |= 2;
        if (c.term_torque && fabsf(tq) > c.torque_limit[lane] * c.term_torque_scale) dof_viol |= 4;
      }
      float l_pos = 0.f, l_vel = 0.f, l_tq = 0.f;
      if (ld) {  // legged_robot_base.py:974-1002
        l_pos = -fminf(dpos - lim_lo, 0.0f) + fmaxf(dpos - lim_hi, 0.0f);
        l_vel = fminf(fmaxf(fabsf(dvel) - lim_v, 0.0f), 1.0f);
        l_tq = c.cur_torque.enable ? fminf(fmaxf(fabsf(tq) - lim_t, 0.0f), 1.0f) : fmaxf(fabsf(tq) - lim_t, 0.0f);
      }
      l_pos = warp_sum(l_pos); l_vel = warp_sum(l_vel); l_tq = warp_sum(l_tq);
      // dof-wise observation features and the last_* state are written now (values of an env that does not reset) so
      // that nothing of this phase stays live across the next ones; the rare reset branch overwrites them
      auto emit_dof = [&](float p_, float v_, float a_, float kp_, float kd_, float def_) {
        emit1<kNoise>(T, rows, S_DOF_POS, lane, p_ - def_, nk);
        emit1<kNoise>(T, rows, S_DOF_VEL, lane, v_, nk);
        emit1<kNoise>(T, rows, S_ACTIONS, lane, a_, nk);
        emit1<kNoise>(T, rows, S_DR_KP, lane, kp_, nk);
        emit1<kNoise>(T, rows, S_DR_KD, lane, kd_, nk);
        io.last_actions[n * D + lane] = a_;   // _post_compute_observations_callback, legged_robot_base.py:398-405
        io.last_dof_pos[n * D + lane] = p_;
        io.last_dof_vel[n * D + lane] = v_;
      };
      if (ld) {
        emit_dof(dpos, dvel, act, kp, kd, ddef);
        if (c.src_used_mask & ((1ull << S_DIF_JOINT) | (1ull << S_DIF_JOINT_VEL))) {
          emit1<kNoise>(T, rows, S_DIF_JOINT, lane, dja, nk);
          emit1<kNoise>(T, rows, S_DIF_JOINT_VEL, lane, djv, nk);
        }
      }
      {
        const float tot = warp_sum16(sums, lane);
        const int idx = (lane >> 1) & 15;
        rv[T.sum_slot[(kGeneral ? 16 : 0) + idx]] = tot / T.sum_div[(kGeneral ? 16 : 0) + idx];
        if (idx == U_DJA2) rv[V_JOINT_N] = sqrtf(tot);
      }
      // ---- D. feet (lane = foot): contacts, feet_air_time, contact-force / slip terms -----------------------------
      // (the per-env scalars were parked in shared memory at the top of the iteration: nothing of them stays in registers
      // across the body / dof phases)
      int64_t ep = *reinterpret_cast<const long long*>(rv + V_EP) + 1;
      const int64_t last_ep = ep;
      float* envf = io.env_f32 + n * PBHC_ENVF_STRIDE;
      const float4 ef0 = *reinterpret_cast<const float4*>(rv + V_ENVF);      // start, motion_len, air_time[2]
      const float4 ef1 = *reinterpret_cast<const float4*>(rv + V_ENVF + 4);  // last_contacts[2], end_time_ratio, pad
      float start = ef0.x, mlen = ef0.y, etr = ef1.z;
      float x_air = 0.f, x_force = 0.f, x_slip = 0.f, x_fslip = 0.f, x_cm = 0.f, contact_f = 0.f, air_f = 0.f;
      bool x_stumble = false;
      float cm_f = 0.f;  // contacts_filt of this lane's foot (general_tracking observes it)
      if (lane < c.n_feet) {
        const float* fc = sm_cf + 3 * my_foot;
        const float* bs = sm_body + 13 * my_foot;
        const float last_c = lane == 0 ? ef1.x : ef1.y;
        air_f = lane == 0 ? ef0.z : ef0.w;
        const float fn = norm3_z(fc[0], fc[1], fc[2]);
        contact_f = fn > 1.0f ? 1.0f : 0.0f;                                   // legged_robot_base.py:352-353
        const float filt = (contact_f != 0.0f || last_c != 0.0f) ? 1.0f : 0.0f;
        cm_f = filt;
        if (Cm) x_cm = fabsf(filt - (omb * __ldg(r0 + e_c + lane) + b * __ldg(r1 + e_c + lane)));
        // feet_air_time, motion_tracking.py:1307-1319 (uses last_contacts of the previous step)
        const bool fz = fc[2] > 1.0f || last_c != 0.0f;
        if (P.has_air_time) {
          const float first = (air_f > 0.0f && fz) ? 1.0f : 0.0f;
          air_f += c.dt;
          x_air = (air_f - c.desired_feet_air_time) * first;
          air_f *= fz ? 0.0f : 1.0f;
        }
        x_force = fmaxf(fn - c.max_contact_force, 0.0f);                        // :1321-1323
        x_stumble = norm2_z(fc[0], fc[1]) > 5.0f * fabsf(fc[2]);                  // :1325-1328
        x_slip = norm3_z(bs[7], bs[8], bs[9]) * (fn > 1.0f ? 1.0f : 0.0f);        // legged_robot_base.py:1004-1010
        x_fslip = (fn > 1.0f ? 1.0f : 0.0f) * norm2_z(bs[7], bs[8]);              // :1012-1017
      }
      if (P.rew_extra & 0x3c) {  // feet_heading_alignment(_contact), penalty_feet_ori(_contact): legged_robot_base.py:1031-1082
        float hd = 0.f, ori = 0.f;
        if (lane < c.n_feet) {
          const float* bs = sm_body + 13 * my_foot;
          const Q4 fq = {bs[3], bs[4], bs[5], bs[6]};
          const V3 ff = quat_apply(fq, {1.0f, 0.0f, 0.0f}), rf = quat_apply(rootq, {1.0f, 0.0f, 0.0f});
          hd = fabsf(wrap_to_pi(atan2f(ff.y, ff.x) - atan2f(rf.y, rf.x)));
          const V3 g = quat_rotate_inverse(fq, {0.0f, 0.0f, -1.0f});
          ori = sqrtf(g.x * g.x + g.y * g.y);
        }
        const float hc = hd * cm_f, oc = ori * cm_f;
        const float s_hd = hd + __shfl_xor_sync(0xffffffffu, hd, 1), s_ori = ori + __shfl_xor_sync(0xffffffffu, ori, 1);
        const float s_hc = hc + __shfl_xor_sync(0xffffffffu, hc, 1), s_oc = oc + __shfl_xor_sync(0xffffffffu, oc, 1);
        if (lane == 0) { rv[V_FEET_HEAD] = s_hd; rv[V_FEET_ORI] = s_ori; rv[V_FEET_HEAD_C] = s_hc; rv[V_FEET_ORI_C] = s_oc; }
      }
      x_air += __shfl_xor_sync(0xffffffffu, x_air, 1); x_force += __shfl_xor_sync(0xffffffffu, x_force, 1);
      x_slip += __shfl_xor_sync(0xffffffffu, x_slip, 1); x_fslip += __shfl_xor_sync(0xffffffffu, x_fslip, 1);
      x_cm += __shfl_xor_sync(0xffffffffu, x_cm, 1);
      const bool stumble = __any_sync(0xffffffffu, x_stumble);
      float air1 = __shfl_sync(0xffffffffu, air_f, 1), contact1 = __shfl_sync(0xffffffffu, contact_f, 1);  // lane 0 packs the state
      // ---- general_tracking: future reference window (_get_future_motion_targets, general_tracking.py:500-565) ------
      if constexpr (kGeneral) {
        if (c.n_future > 0) {
          // destinations of a future-target source: a group made only of such sources is written straight to global memory
          // (scaled and clipped here), any other group through its shared-memory row
          struct FDst { float* gp; int row; float sc; };
          auto fdst = [&](int k) {
            const int i = k - S_FUT_HEIGHT, g = T.f_g[i];
            FDst d;
            d.gp = g >= 0 ? io.obs[g] + n * c.group_dim[g] + T.f_off[i] : nullptr;
            d.row = T.f_row[i]; d.sc = T.obs_scale[k];
            return d;
          };
          const float clipv = c.clip_obs;
          auto fput = [&](const FDst& d, int el, float val) {
            const float sv = val * d.sc;
            if (d.gp) d.gp[el] = fminf(fmaxf(sv, -clipv), clipv);
            if (d.row >= 0) rows[d.row + el] = sv;
          };
          __syncwarp();  // the simulator-state buffer is idle from here on: it holds the per-step table
          float* fut = sm_body;
          const int F = c.n_future, K = c.n_key, a = c.anchor_index;
          const float t0 = (float)ep * c.dt + start;                 // episode_length_buf * dt + start (:509), no +1
          // where the window's values sit in a record: the packed table, or the compact per-frame records built from it
          // (pbhc_future_table_build: the same bits, contiguous - ~3 lines per frame instead of ~10 scattered sectors)
          const bool ft = io.fut_table != nullptr;
          const float* f_rec = ft ? io.fut_table : tab.records;
          const int64_t f_stride = ft ? io.fut_stride : tab.rec_stride;
          const int o_aq = ft ? 4 : 4 * a, o_rv = ft ? 8 : e_vel, o_ra = ft ? 11 : e_ang, o_rp = ft ? 14 : e_pos, o_ap = ft ? 17 : e_pos + 3 * a;
          const int o_dp = ft ? 20 : e_dp, o_key = ft ? 20 + D : e_pos, f_lines = ft ? io.fut_stride : 0;
          if (lane < F) {  // lane = future step: frame pair, root-frame quantities, anchor pose
            const float ts = (float)c.future_steps[lane] * c.dt + t0;
            const FrameBlend fb = frame_blend(tab, clip, ts);
            const float* q0 = f_rec + fb.f0 * f_stride;
            const float* q1 = f_rec + fb.f1 * f_stride;
            const float bb = fb.blend, ob = 1.0f - bb;
            // the passes below read this step's dof positions and key-body positions: on a library larger than L2 every one of
            // those reads would be a dependent HBM access - ask for the lines now (positions: J x 12 B, dofs: D x 4 B per frame)
            if (ft) { for (int o = 32; o < f_lines; o += 32) { prefetch_l2(q0 + o); prefetch_l2(q1 + o); } }
            else {
              for (int o = 0; o < 3 * J; o += 32) { prefetch_l2(q0 + e_pos + o); prefetch_l2(q1 + e_pos + o); }
              for (int o = 0; o < D; o += 32) { prefetch_l2(q0 + e_dp + o); prefetch_l2(q1 + e_dp + o); }
              prefetch_l2(q0 + e_dp + D - 1); prefetch_l2(q1 + e_dp + D - 1);
            }
            const float4 a0 = __ldg(reinterpret_cast<const float4*>(q0)), a1 = __ldg(reinterpret_cast<const float4*>(q1));
            const Q4 rq = slerp_nl({a0.x, a0.y, a0.z, a0.w}, {a1.x, a1.y, a1.z, a1.w}, bb);
            const V3 rvl = quat_rotate_inverse(rq, {ob * __ldg(q0 + o_rv) + bb * __ldg(q1 + o_rv), ob * __ldg(q0 + o_rv + 1) + bb * __ldg(q1 + o_rv + 1),
                                                    ob * __ldg(q0 + o_rv + 2) + bb * __ldg(q1 + o_rv + 2)});
            const V3 ral = quat_rotate_inverse(rq, {ob * __ldg(q0 + o_ra) + bb * __ldg(q1 + o_ra), ob * __ldg(q0 + o_ra + 1) + bb * __ldg(q1 + o_ra + 1),
                                                    ob * __ldg(q0 + o_ra + 2) + bb * __ldg(q1 + o_ra + 2)});
            const float rz = (ob * __ldg(q0 + o_rp + 2) + bb * __ldg(q1 + o_rp + 2)) + org[2];
            float roll, pitch;
            euler_roll_pitch(rq, roll, pitch);
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(q0 + o_aq)), b1 = __ldg(reinterpret_cast<const float4*>(q1 + o_aq));
            const Q4 aq = slerp_nl({b0.x, b0.y, b0.z, b0.w}, {b1.x, b1.y, b1.z, b1.w}, bb);
            fut[F_F0 * 32 + lane] = __int_as_float((int)fb.f0); fut[F_F1 * 32 + lane] = __int_as_float((int)fb.f1);
            fut[F_BLEND * 32 + lane] = bb;
            fut[F_APX * 32 + lane] = (ob * __ldg(q0 + o_ap) + bb * __ldg(q1 + o_ap)) + org[0];
            fut[F_APY * 32 + lane] = (ob * __ldg(q0 + o_ap + 1) + bb * __ldg(q1 + o_ap + 1)) + org[1];
            fut[F_APZ * 32 + lane] = (ob * __ldg(q0 + o_ap + 2) + bb * __ldg(q1 + o_ap + 2)) + org[2];
            fut[F_AQX * 32 + lane] = aq.x; fut[F_AQY * 32 + lane] = aq.y; fut[F_AQZ * 32 + lane] = aq.z; fut[F_AQW * 32 + lane] = aq.w;
            fput(fdst(S_FUT_HEIGHT), lane, rz);
            { const FDst d = fdst(S_FUT_RP); fput(d, 2 * lane, roll); fput(d, 2 * lane + 1, pitch); }
            { const FDst d = fdst(S_FUT_LIN); fput(d, 3 * lane, rvl.x); fput(d, 3 * lane + 1, rvl.y); fput(d, 3 * lane + 2, rvl.z); }
            fput(fdst(S_FUT_YAW), lane, ral.z);
            if (lane == 0) {  // obs_next_step_mimic_buf (:553-563): the step-0 slices
              emit1<kNoise>(T, rows, S_NEXT_STEP, 0, rz, nk); emit1<kNoise>(T, rows, S_NEXT_STEP, 1, roll, nk);
              emit1<kNoise>(T, rows, S_NEXT_STEP, 2, pitch, nk); emit3<kNoise>(T, rows, S_NEXT_STEP, 3, rvl.x, rvl.y, rvl.z, nk);
              emit1<kNoise>(T, rows, S_NEXT_STEP, 6, ral.z, nk);
            }
          }
          __syncwarp();
          const FDst d_dof = fdst(S_FUT_DOF), d_key = fdst(S_FUT_KEY);
          // In a compact record the dof positions and the key-body positions of a frame are ONE contiguous run of D + 3K floats
          // (56 for the teacher config).  Reading them straight from the table makes every step of the two loops below a
          // dependent global load (ncu: 19 % of the kernel's stall samples sit on these two loops).  With the compact table the
          // runs of kFutChunk steps x 2 frames are fetched together by cp.async into the idle part of the staging buffer: the
          // table latency is exposed once per chunk instead of once per step, and the loops read shared memory.  Copy width:
          // 16 bytes when the run is a multiple of 4 floats, 8 bytes when it is even (the teacher config: 29 + 3 x 11 = 62 floats
          // = 31 pieces; the run starts 80 bytes into a 16-byte aligned record and the staged rows are 248 bytes apart).
          const int row_f = D + 3 * K;
          const int wf = (row_f & 3) == 0 ? 4 : ((row_f & 1) == 0 ? 2 : 0);  // floats per copy
          // Which pair of loops runs is decided on the host from the launch itself (fut_staged_ok: compact table bound, even run,
          // staging on).  The specialised kernel gets it as a constant and so carries ONE pair: with both compiled in, this
          // instruction-fetch-bound kernel is slower on either path (B200, one box, 1.283 ms without the staged loops at all,
          // 1.297 staged / 1.344 direct with both, 1.257 with the staged pair alone).
#ifdef PBHC_SPEC
          const bool staged = spec::fut_staged != 0;
#else
          const bool staged = ft && wf != 0 && P.fut_win;
#endif
          if (staged) {
            float* win = fut + kFut * 32;
            const uint32_t win_s = (uint32_t)__cvta_generic_to_shared(win);
            const int pieces = row_f / wf;
#pragma unroll 1
            for (int c0 = 0; c0 < F; c0 += kFutChunk) {
              const int ns = F - c0 < kFutChunk ? F - c0 : kFutChunk;
              __syncwarp();  // (the previous chunk has been read)
              for (int i = lane; i < 2 * ns * pieces; i += 32) {
                const int r = i / pieces, pc = i - r * pieces, st = c0 + (r >> 1);
                const int fr = __float_as_int(fut[((r & 1) ? F_F1 : F_F0) * 32 + st]);
                const uint32_t dsts = win_s + 4 * (r * row_f + wf * pc);
                const float* srcg = f_rec + (int64_t)fr * f_stride + o_dp + wf * pc;
                if (wf == 4) cp_async16(dsts, srcg); else cp_async8(dsts, srcg);
              }
              cp_async_wait();
              __syncwarp();
#pragma unroll 1
              for (int sl = 0; sl < ns; ++sl) {  // dof_pos: lane = dof
                const int st = c0 + sl;
                const float bb = fut[F_BLEND * 32 + st], ob = 1.0f - bb;
                if (ld) {
                  const float v = ob * win[(2 * sl) * row_f + lane] + bb * win[(2 * sl + 1) * row_f + lane];
                  fput(d_dof, st * D + lane, v);
                  if (st == 0) emit1<kNoise>(T, rows, S_NEXT_STEP, 7 + lane, v, nk);
                }
              }
#pragma unroll 1
              for (int sl0 = 0; sl0 < ns; sl0 += 2) {  // key bodies in the reference anchor frame: lane = (step, key)
                const int sl = sl0 + (lane >> 4), k = lane & 15, st = c0 + sl;
                if (sl < ns && k < K) {
                  const float* w0 = win + (2 * sl) * row_f + D + 3 * k;
                  const float* w1 = w0 + row_f;
                  const float bb = fut[F_BLEND * 32 + st], ob = 1.0f - bb;
                  const V3 d = {((ob * w0[0] + bb * w1[0]) + org[0]) - fut[F_APX * 32 + st],
                                ((ob * w0[1] + bb * w1[1]) + org[1]) - fut[F_APY * 32 + st],
                                ((ob * w0[2] + bb * w1[2]) + org[2]) - fut[F_APZ * 32 + st]};
                  const Q4 aq = {fut[F_AQX * 32 + st], fut[F_AQY * 32 + st], fut[F_AQZ * 32 + st], fut[F_AQW * 32 + st]};
                  const V3 lp = quat_apply(quat_conj(aq), d);
                  const int el = st * 3 * K + 3 * k;
                  fput(d_key, el, lp.x); fput(d_key, el + 1, lp.y); fput(d_key, el + 2, lp.z);
                  if (st == 0) emit3<kNoise>(T, rows, S_NEXT_STEP, 7 + D + 3 * k, lp.x, lp.y, lp.z, nk);
                }
              }
            }
          } else {
#pragma unroll 1
          for (int s = 0; s < F; ++s) {  // dof_pos of every future step: lane = dof (not unrolled: the specialised kernel knows F, and 20 copies of the body only cost instruction-cache misses)
            const float* q0 = f_rec + (int64_t)__float_as_int(fut[F_F0 * 32 + s]) * f_stride;
            const float* q1 = f_rec + (int64_t)__float_as_int(fut[F_F1 * 32 + s]) * f_stride;
            const float bb = fut[F_BLEND * 32 + s], ob = 1.0f - bb;
            if (ld) {
              const float v = ob * __ldg(q0 + o_dp + lane) + bb * __ldg(q1 + o_dp + lane);
              fput(d_dof, s * D + lane, v);
              if (s == 0) emit1<kNoise>(T, rows, S_NEXT_STEP, 7 + lane, v, nk);
            }
          }
#pragma unroll 1
          for (int s0 = 0; s0 < F; s0 += 2) {  // key bodies in the reference anchor frame: two steps per pass, lane = (step, key)
            const int s = s0 + (lane >> 4), k = lane & 15;
            if (s < F && k < K) {
              const int jo = o_key + 3 * (ft ? k : c.key_ids[k]);
              const float* q0 = f_rec + (int64_t)__float_as_int(fut[F_F0 * 32 + s]) * f_stride;
              const float* q1 = f_rec + (int64_t)__float_as_int(fut[F_F1 * 32 + s]) * f_stride;
              const float bb = fut[F_BLEND * 32 + s], ob = 1.0f - bb;
              const V3 d = {((ob * __ldg(q0 + jo) + bb * __ldg(q1 + jo)) + org[0]) - fut[F_APX * 32 + s],
                            ((ob * __ldg(q0 + jo + 1) + bb * __ldg(q1 + jo + 1)) + org[1]) - fut[F_APY * 32 + s],
                            ((ob * __ldg(q0 + jo + 2) + bb * __ldg(q1 + jo + 2)) + org[2]) - fut[F_APZ * 32 + s]};
              const Q4 aq = {fut[F_AQX * 32 + s], fut[F_AQY * 32 + s], fut[F_AQZ * 32 + s], fut[F_AQW * 32 + s]};
              const V3 lp = quat_apply(quat_conj(aq), d);
              const int el = s * 3 * K + 3 * k;
              fput(d_key, el, lp.x); fput(d_key, el + 1, lp.y); fput(d_key, el + 2, lp.z);
              if (s == 0) emit3<kNoise>(T, rows, S_NEXT_STEP, 7 + D + 3 * k, lp.x, lp.y, lp.z, nk);
            }
          }
          }
          __syncwarp();
        }
      }
      if (hist_stage) {  // the simulator-state buffer is idle from here on: the history ring arrives in it during E-H
        __syncwarp();
        warp_load_async(sm_body, io.hist + n * c.hist_row, c.hist_row, hist_vec, lane);
      }
      // ---- E. termination (legged_robot_base.py:408-489, motion_tracking.py:330-358) -------------------------------
      const bool t_grav = c.term_gravity && pre[(P_GRAVN) * kChunk + e] > c.term_gravity_thr;
      const bool t_far = c.term_motion_far && any_far;
      const bool t_out = ep > c.max_episode_length;
      const bool t_end = c.term_motion_end && ((float)ep * c.dt + start > mlen);
      const bool time_out = t_out || t_end;
      bool t_bodyz = false;
      if constexpr (kGeneral) t_bodyz = c.term_body_z && __any_sync(0xffffffffu, bodyz);
      bool t_contact = false, t_low = false, t_dpos = false, t_dvel = false, t_tq = false;
      if (c.term_contact || c.term_low_height || c.term_dof_pos || c.term_dof_vel || c.term_torque) {  // base class, legged_robot_base.py:433-476
        const int gates = T.term_gates;
        t_contact = c.term_contact && __any_sync(0xffffffffu, hit);
        t_low = c.term_low_height && rp.z < c.term_min_height;
        t_dpos = c.term_dof_pos && (gates & 1) && __any_sync(0xffffffffu, dof_viol & 1);
        t_dvel = c.term_dof_vel && (gates & 2) && __any_sync(0xffffffffu, dof_viol & 2);
        t_tq = c.term_torque && (gates & 4) && __any_sync(0xffffffffu, dof_viol & 4);
      }
      const bool terminated = t_grav || t_far || time_out || t_refz || t_ori || t_bodyz || t_contact || t_low || t_dpos || t_dvel || t_tq;
      const bool reset = terminated || (io.force_reset != nullptr && io.force_reset[n] != 0);  // forced: resample_motion (motion_tracking.py:385-389)
      if (lane == 0) {
        rv[V_SIGMA_ERR + G_MAX_JOINT] = max_dja;
        rv[V_T_GRAV] = t_grav; rv[V_T_FAR] = t_far; rv[V_T_OUT] = time_out; rv[V_T_END] = t_end;
        rv[V_RESET] = reset; rv[V_ONE] = 1.0f; rv[V_ALIVE] = 1.0f;
        if constexpr (kGeneral) { rv[V_T_REFZ] = t_refz; rv[V_T_ORI] = t_ori; rv[V_T_BODYZ] = t_bodyz; }
        rv[V_CONTACT] = 1.0f - x_cm / (float)c.n_feet; rv[V_CONTACT_V2] = 0.5f - x_cm / (float)c.n_feet;
        rv[V_ORIENT] = pre[(P_ORIENT) * kChunk + e];
        rv[V_AIR] = x_air; rv[V_FORCE] = x_force; rv[V_STUMBLE] = stumble ? 1.0f : 0.0f; rv[V_SLIP] = x_slip; rv[V_FSLIP] = x_fslip;
        rv[V_LIM_POS] = l_pos; rv[V_LIM_VEL] = l_vel; rv[V_LIM_TQ] = l_tq; rv[V_COLLIDE] = (float)n_collide;
      }
      __syncwarp();
      // ---- F. rewards (lane = reward term; legged_robot_base.py:715-761) ------------------------------------------
      float rew = 0.0f;
      if (lane < c.n_rewards) {
        const int kind = T.r_kind[lane];
        float r = rv[T.r_arg[lane]];
        if (kind != K_RAW) {
          r = expf(-r / T.r_sigma[lane]);
          if (kind == K_EXP2)  // teleop_body_position_extend, motion_tracking.py:1177-1199
            r = r * c.w_lower + expf(-rv[V_SIGMA_ERR + G_UPPER] / (float)T.scal[PBHC_S_SIGMA + G_UPPER]) * c.w_upper;
        }
        rew = r * T.r_scale[lane] * T.r_pen[lane];
      }
      const float term_rew = c.has_termination ? ((terminated && !time_out) ? 1.0f : 0.0f) * c.termination_scale : 0.0f;
      {
        const int R = c.n_rewards;
        float* es = io.episode_sums + n * c.n_episode_sums;
        if (c.use_vec_reward) {
          float v = rew;
          if (c.only_positive_rewards) v = fmaxf(v, 0.0f);
          if (lane == R - 1) v += term_rew;  // the reference's stale loop index, legged_robot_base.py:744
          if (lane < R) io.rew_buf[n * (R + 1) + lane] = v;
          if (lane == R) io.rew_buf[n * (R + 1) + R] = 0.0f;
        } else {
          float tot = warp_sum(lane < R ? rew : 0.0f);
          if (c.only_positive_rewards) tot = fmaxf(tot, 0.0f);
          if (lane == 0) io.rew_buf[n] = tot + term_rew;
        }
        float ev = 0.0f;
        if (lane < c.n_episode_sums) ev = es[lane] + (lane < R ? rew : term_rew);
        if (reset) {  // extras["episode"], legged_robot_base.py:509-513
          if (lane < c.n_episode_sums) io.last_episode_sums[n * c.n_episode_sums + lane] = ev;
          ev = 0.0f;
        }
        if (lane < c.n_episode_sums) es[lane] = ev;
      }
      // ---- G. reset (legged_robot_base.py:491-516, 647-686, 587-635; motion_tracking.py:274-296, 369-384, 445-543) ----
      int64_t delay = io.action_delay_idx[n];
      float air0 = air_f, contact0 = contact_f;  // meaningful in lane 0
      float root_z = rp.z;                       // simulator root height as the observations see it (live view: new value after a reset)
      if (reset) {
        const Rng rng = {make_uint2((uint32_t)P.c.seed, (uint32_t)(P.c.seed >> 32)), (uint32_t)(n + io.env_offset), (uint32_t)((uint64_t)(n + io.env_offset) >> 32), push_count};
        if (ld) io.actions[n * D + lane] = 0.0f;
        if (lane == 0 && io.last_episode_length_buf) io.last_episode_length_buf[n] = last_ep;  // extras["episode"]["end_epis_length"], legged_robot_base.py:515
        if (lane == 0 && io.refresh_buf) io.refresh_buf[n] = (int64_t)push_count + 1;             // need_to_refresh_envs, :493
        air0 = 0.0f; air1 = 0.0f; contact0 = 0.0f; contact1 = 0.0f;
        if (c.hist_row > 0) {  // HistoryHandler.reset: zero the ring (and what the observation rows already took from it)
          for (int i = lane; i < c.hist_row; i += 32) io.hist[n * c.hist_row + i] = 0.0f;
          hist_scatter(io.hist + n * c.hist_row, rows, hmap, c, P.n_hdst_max, P.hist_any_mul, hist_vec, lane, true);
        }
        ep = 0;
        float kp2 = 1.f, kd2 = 1.f, dpos2 = 0.f, dvel2 = 0.f;  // (nothing of the dof phase is kept live for this rare branch)
        if (ld) {  // _episodic_domain_randomization
          kp2 = io.kp_scale[n * D + lane]; kd2 = io.kd_scale[n * D + lane];
          if (c.randomize_pd_gain) {
            kp2 = io.inject_kp ? io.inject_kp[n * D + lane] : (c.kp_hi - c.kp_lo) * rng.uniform(16 + lane) + c.kp_lo;
            kd2 = io.inject_kd ? io.inject_kd[n * D + lane] : (c.kd_hi - c.kd_lo) * rng.uniform(48 + lane) + c.kd_lo;
            io.kp_scale[n * D + lane] = kp2; io.kd_scale[n * D + lane] = kd2;
          }
          if (c.randomize_rfi_lim)
            io.rfi_lim_scale[n * D + lane] = io.inject_rfi ? io.inject_rfi[n * D + lane] : (c.rfi_hi - c.rfi_lo) * rng.uniform(80 + lane) + c.rfi_lo;
          if (c.use_rao)
            io.rao_scale[n * D + lane] = io.inject_rao ? io.inject_rao[n * D + lane] : (c.rao_lim - (-c.rao_lim)) * rng.uniform(112 + lane) + (-c.rao_lim);
        }
        if (c.randomize_ctrl_delay) {
          delay = io.inject_delay ? io.inject_delay[n] : c.delay_lo + (int64_t)(rng.draw(1).x % (uint32_t)(c.delay_hi - c.delay_lo + 1));
          if (lane == 0) io.action_delay_idx[n] = delay;
        }
        etr = ((float)last_ep * c.dt + start) / mlen;                       // motion_tracking.py:281-282
        mlen = tab.clip_length[clip];                                       // _resample_motion_times
        start = io.inject_start_times ? io.inject_start_times[n] : (c.start_time_zero ? 0.0f : rng.uniform(0) * mlen);
        // _reset_dofs / _reset_root_states from the motion state at the new time
        const FrameBlend f2 = frame_blend(tab, clip, (float)(ep + 1) * c.dt + start);
        const float b2 = f2.blend, omb2 = 1.0f - b2;
        const float* q0 = tab.records + f2.f0 * tab.rec_stride;
        const float* q1 = tab.records + f2.f1 * tab.rec_stride;
        if (ld) {
          if constexpr (kGeneral) {  // general_tracking._reset_dofs (:464-483): motion at episode_length*dt + start (no +1), uniform noise
            const FrameBlend f3 = frame_blend(tab, clip, (float)ep * c.dt + start);
            const float b3 = f3.blend, omb3 = 1.0f - b3;
            const float* p0 = tab.records + f3.f0 * tab.rec_stride;
            const float* p1 = tab.records + f3.f1 * tab.rec_stride;
            dpos2 = omb3 * p0[e_dp + lane] + b3 * p1[e_dp + lane];
            dvel2 = omb3 * p0[e_dv + lane] + b3 * p1[e_dv + lane];
            if (c.reset_noise[0] != 0.0f) dpos2 = dpos2 + rng.uniform(144 + lane) * c.reset_noise[0];
            if (c.reset_noise[1] != 0.0f) dvel2 = dvel2 + rng.uniform(176 + lane) * c.reset_noise[1];
          } else {
            dpos2 = omb2 * q0[e_dp + lane] + b2 * q1[e_dp + lane];
            dvel2 = omb2 * q0[e_dv + lane] + b2 * q1[e_dv + lane];
            if (c.reset_noise[0] != 0.0f) dpos2 = dpos2 + rng.normal(144 + lane) * c.reset_noise[0];
            if (c.reset_noise[1] != 0.0f) dvel2 = dvel2 + rng.normal(176 + lane) * c.reset_noise[1];
          }
          reinterpret_cast<float2*>(io.dof_state)[n * D + lane] = make_float2(dpos2, dvel2);
        }
        float v = 0.0f;
        if (lane < 13) {
          if (lane >= 3 && lane < 7) {
            const float4 a = reinterpret_cast<const float4*>(q0)[0], d4 = reinterpret_cast<const float4*>(q1)[0];
            const Q4 rr = slerp_nl({a.x, a.y, a.z, a.w}, {d4.x, d4.y, d4.z, d4.w}, b2);
            Q4 nq = {0.0f, 0.0f, 0.0f, 1.0f};  // small_random_quaternions(n, 0)
            if (c.reset_noise[3] != 0.0f) {
              V3 ax = {rng.normal(2), rng.normal(3), rng.normal(4)};
              const float an = norm3(ax.x, ax.y, ax.z);
              const float ang = c.reset_noise[3] * rng.uniform(5);
              const float sh = sinf(ang / 2.0f);
              nq = {sh * (ax.x / an), sh * (ax.y / an), sh * (ax.z / an), cosf(ang / 2.0f)};
            }
            const Q4 o = quat_mul(nq, rr);
            v = lane == 3 ? o.x : (lane == 4 ? o.y : (lane == 5 ? o.z : o.w));
          } else {
            const int el = lane < 3 ? e_pos + lane : (lane < 10 ? e_vel + (lane - 7) : e_ang + (lane - 10));
            v = omb2 * q0[el] + b2 * q1[el];
            if (lane < 3) v = v + org[lane];
            const float ns = lane < 3 ? c.reset_noise[2] : (lane < 10 ? c.reset_noise[4] : c.reset_noise[5]);
            if (ns != 0.0f) v = v + rng.normal(8 + lane) * ns;
          }
          io.root_state[n * 13 + lane] = v;
        }
        if constexpr (kGeneral) root_z = __shfl_sync(0xffffffffu, v, 2);
        if (ld) emit_dof(dpos2, dvel2, 0.0f, kp2, kd2, io.default_dof_pos[n * io.default_dof_stride + lane]);
      }
      // ---- H. observation features into the rows (legged_robot_base.py:763-793, 1114-1215; motion_tracking.py:944-1015)
      if (lane < 16) {
        int k, ce = 0;
        float val;
        if (lane < 9) { k = lane / 3; ce = lane - 3 * k; val = pre[(P_BLV + lane) * kChunk + e]; }
        else if (lane < 12) { k = S_DR_COM; ce = lane - 9; val = io.base_com_bias[n * 3 + ce]; }
        else if (lane == 12) { k = S_PHASE; val = pre[(P_PHASE) * kChunk + e]; }   // stale for envs that reset, like the reference
        else if (lane == 13) { k = S_DR_FRICTION; val = io.friction[n]; }
        else if (lane == 14) { k = S_DR_DELAY; val = (float)delay; }
        else { k = S_BASE_POS_Z; val = rp.z; }
        emit1<kNoise>(T, rows, k, ce, val, nk);
      }
      for (int i = lane; i < L; i += 32) emit1<kNoise>(T, rows, S_DR_MASS, i, io.link_mass_scale[n * L + i], nk);
      if ((c.src_used_mask & (1ull << S_RELYAW)) && lane == 0) emit1<kNoise>(T, rows, S_RELYAW, 0, relyaw, nk);  // stale for envs that reset, like the reference
      if constexpr (kGeneral) {  // getters of general_tracking.py:879-891, legged_robot_base.py:1193
        const float cmv = __shfl_sync(0xffffffffu, cm_f, (lane - 2) & 31);  // contacts_filt (zeroed by a reset, :664-666)
        if (lane < 6) {
          int k, ce = 0;
          float val;
          if (lane < 2) { k = S_ROLL_PITCH; ce = lane; val = pre[(P_ROLL + lane) * kChunk + e]; }  // stale for envs that reset
          else if (lane < 4) { k = S_CONTACT_MASK; ce = lane - 2; val = reset ? 0.0f : cmv; }
          else if (lane == 4) { k = S_ROOT_HEIGHT; val = root_z; }
          else { k = S_DR_BASE_MASS; val = io.base_mass_scale ? io.base_mass_scale[n] : 1.0f; }
          emit1<kNoise>(T, rows, k, ce, val, nk);
        }
      }
      if (lane == 0) { rv[V_ETR] = etr; }
      if (hist_stage) {  // history ring -> its places in the rows (zeros are already there for an env that reset)
        cp_async_wait();
        __syncwarp();
        if (!reset) hist_scatter(sm_body, rows, hmap, c, P.n_hdst_max, P.hist_any_mul, hist_vec, lane, false);
      }
      __syncwarp();
      if (n32 + stride < n_envs) fetch_sim_state(n + stride);
      // ---- I. rows -> global (clipped), history push, state write-back -----------------------------------------------
      for (int g = 0; g < c.n_groups; ++g) {
        if (kGeneral && P.group_direct[g]) continue;
        const int dim = c.group_dim[g];
        warp_store_clipped(io.obs[g] + n * dim, rows + P.row_off[g], dim, P.obs_vec[g], c.clip_obs, lane);
      }
      if (P.push_len > 0) {  // HistoryHandler.add (history_handler.py:40-44) as a ring-buffer push: one slot per key is written
        float* gh = io.hist + n * c.hist_row;
        const float* ps = rows + P.row_off[c.n_groups];
        for (int i = lane; i < P.push_len; i += 32) gh[pmap[i]] = ps[i];
      }
      if (lane == 0) {
        const int pw = push_step(c, io, n, __float_as_int(ef1.w), reset, P.c.seed, push_count);
        reinterpret_cast<float4*>(envf)[0] = make_float4(start, mlen, air0, c.n_feet > 1 ? air1 : 0.0f);
        reinterpret_cast<float4*>(envf)[1] = make_float4(contact0, c.n_feet > 1 ? contact1 : 0.0f, etr, __int_as_float(pw));
        io.episode_length_buf[n] = ep;
        io.reset_buf[n] = reset ? 1 : 0;
        io.time_out_buf[n] = time_out ? 1 : 0;
      }
      if (lane < kAccLo) {  // per-lane fp64 accumulation of the cross-env sums (slot = lane, layout PBHC_A_*)
        double a = (double)rv[lane];
        if (lane == V_RESET_EPLEN) a = reset ? (double)last_ep : 0.0;
        if (lane == V_ETR_SQ) { const double x = (double)rv[V_ETR]; a = x * x; }
        acc_reg += a;
      }
      if constexpr (kGeneral) if (lane < 8) acc_reg2 += (double)rv[32 + lane];
      if (lane >= 8 && lane < 13) {  // PBHC_A_TERM_CONTACT .. PBHC_A_TERM_TORQUE
        const int slot = 32 + lane;
        acc_reg2 += slot == PBHC_A_TERM_CONTACT ? (double)t_contact : slot == PBHC_A_TERM_LOW_HEIGHT ? (double)t_low : slot == PBHC_A_TERM_DOF_POS ? (double)t_dpos :
                    slot == PBHC_A_TERM_DOF_VEL ? (double)t_dvel : (double)t_tq;
      }
    }
  }
  // block partials: every warp parks its 64 sums in its (now idle) pre-phase buffer, 64 threads add them up in warp order
  __syncwarp();
  double* wacc = reinterpret_cast<double*>(pre);
  wacc[lane] = lane < kAccLo ? acc_reg : 0.0;
  wacc[32 + lane] = ((kGeneral && lane < 8) || (lane >= 8 && lane < 13)) ? acc_reg2 : 0.0;
  __syncthreads();
  if (threadIdx.x < PBHC_ACC) {
    double v = 0.0;
    for (int w = 0; w < kTW; ++w) v += reinterpret_cast<const double*>(smem_raw + P.o_warp + w * P.warp_bytes + kRv * 4)[threadIdx.x];
    io.partials[(int64_t)blockIdx.x * PBHC_ACC + threadIdx.x] = v;
  }
}

#ifdef __CUDACC_RTC__
}  // namespace pbhc
extern "C" __global__ void __launch_bounds__(pbhc::kTW * 32, 2) pbhc_track_spec(const __grid_constant__ pbhc::TrackParams P) {
  pbhc::track_body<PBHC_NOISE != 0, true>(P);
}
#else
// every env's inverse heading quaternion (z, w) BEFORE the step: what the vr_3point_pos observation of another env reads
__global__ void heading_kernel(const float* __restrict__ root_state, int64_t n, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* r = root_state + i * 13;
  const Q4 h = heading_quat_inv_fast({r[3], r[4], r[5], r[6]});
  reinterpret_cast<float2*>(out)[i] = make_float2(h.z, h.w);
}

// configs only the warp-per-env kernel serves: the observation sources and reward terms added with ABI 9
static bool needs_warp_kernel(const PbhcTrackCfg& c) {
  if (c.src_used_mask >> S_RELYAW) return true;
  for (int r = 0; r < c.n_rewards; ++r) if (c.reward_id[r] >= R_RADIAL_BODY_VEL) return true;
  return false;
}

template <bool kNoise, bool kGeneral>
__global__ void __launch_bounds__(kTW * 32, kGeneral ? 2 : F_BPS) track_kernel(const __grid_constant__ TrackParams P) { track_body<kNoise, kGeneral>(P); }

// deterministic reduction of the per-block partials: 32 groups of threads take every 32nd block (independent loads in
// flight instead of one serial chain), then a fixed-order sum over the groups
constexpr int kReduceGroups = 1024 / PBHC_ACC;
__global__ void __launch_bounds__(1024) track_reduce_kernel(const double* __restrict__ partials, int n_blocks, double* __restrict__ acc) {
  __shared__ double part[kReduceGroups][PBHC_ACC + 1];
  const int a = threadIdx.x & (PBHC_ACC - 1), g = threadIdx.x / PBHC_ACC;  // blockDim = kReduceGroups * PBHC_ACC
  // launched with programmatic stream serialization (launch_reduce): the block is resident while the step kernel drains;
  // the step kernel's writes are visible after this returns
  cudaGridDependencySynchronize();
  double v = 0.0;
#pragma unroll 4
  for (int b = g; b < n_blocks; b += kReduceGroups) v += partials[(int64_t)b * PBHC_ACC + a];
  part[g][a] = v;
  __syncthreads();
  if (g == 0) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kReduceGroups; ++k) t += part[k][a];
    acc[a] = t;
  }
}

__device__ double curriculum(double value, double avg, const PbhcCurriculum& c, double sign_down) {
  if (avg < c.down) value *= (1.0 + sign_down * c.degree);
  else if (avg > c.up) value *= (1.0 - sign_down * c.degree);
  return fmin(fmax(value, c.lo), c.hi);
}

// _update_adaptive_sigma (motion_tracking.py:1030-1048), _update_average_episode_length and the scalar curricula
// (legged_robot_base.py:875-939, motion_tracking.py:306-317), log_dict values (:397-418, legged_robot_base.py:425-429).
__global__ void track_finalize_kernel(const __grid_constant__ PbhcTrackCfg c, const double* __restrict__ acc, double* __restrict__ S) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const double n = acc[PBHC_A_NUM_ENVS];
  if (n <= 0.0) return;
  double* L = S + PBHC_S_LOG;
  // values logged before this step's updates
  L[PBHC_L_FAR_THR] = (double)(float)S[PBHC_S_FAR_THR];
  L[PBHC_L_PENALTY] = (double)(float)S[PBHC_S_PENALTY];
  L[PBHC_L_AVG_EPL] = S[PBHC_S_AVG_EPL];
  L[PBHC_L_SOFT_DOF_POS] = (double)(float)S[PBHC_S_SOFT_DOF_POS];
  L[PBHC_L_SOFT_DOF_VEL] = (double)(float)S[PBHC_S_SOFT_DOF_VEL];
  L[PBHC_L_SOFT_TORQUE] = (double)(float)S[PBHC_S_SOFT_TORQUE];
  L[PBHC_L_NOISE_CUR] = (double)(float)S[PBHC_S_NOISE_CUR];
  for (int k = 0; k < PBHC_NUM_SIGMA; ++k) { L[PBHC_L_ADP_SIGMA + k] = S[PBHC_S_SIGMA + k]; L[PBHC_L_ERROR_EMA + k] = S[PBHC_S_EMA + k]; }
  if (c.adaptive_enable) {
    uint32_t active = 0;
    for (int i = 0; i < c.n_rewards; ++i) {
      switch (c.reward_id[i]) {
        case R_MAX_JOINT_POS: active |= 1u << G_MAX_JOINT; break;
        case R_BODY_POS_EXT: active |= (1u << G_UPPER) | (1u << G_LOWER); break;
        case R_VR3: active |= 1u << G_VR3; break;
        case R_FEET_POS: active |= 1u << G_FEET; break;
        case R_BODY_ROT: active |= 1u << G_ROT; break;
        case R_BODY_VEL: active |= 1u << G_VEL; break;
        case R_BODY_ANG_VEL: active |= 1u << G_ANG_VEL; break;
        case R_JOINT_POS: active |= 1u << G_JOINT_POS; break;
        case R_JOINT_VEL: active |= 1u << G_JOINT_VEL; break;
        case R_ROOT_VEL: active |= 1u << G_ROOT_VEL; break;
        case R_ANCHOR_POS: active |= 1u << G_ANCHOR_POS; break;
        case R_ANCHOR_ROT: active |= 1u << G_ANCHOR_ROT; break;
        case R_LKEY_POS: active |= 1u << G_LKEY_POS; break;
        case R_LKEY_ROT: active |= 1u << G_LKEY_ROT; break;
        case R_KEY_VEL: active |= 1u << G_KEY_VEL; break;
        case R_KEY_ANG: active |= 1u << G_KEY_ANG; break;
        default: break;
      }
    }
    for (int k = 0; k < PBHC_NUM_SIGMA; ++k) {
      if (!(active & (1u << k))) continue;
      const double mean = (double)(float)(acc[PBHC_A_SIGMA_ERR + k] / n);  // error.mean() is an fp32 value
      const double ema = S[PBHC_S_EMA + k] * (1.0 - c.adaptive_alpha) + mean * c.adaptive_alpha;
      S[PBHC_S_EMA + k] = ema;
      const double sig = S[PBHC_S_SIGMA + k];
      if (c.adaptive_type == 2) S[PBHC_S_SIGMA + k] = fmin(ema * c.adaptive_scale, sig);
      else if (c.adaptive_type == 1) S[PBHC_S_SIGMA + k] = c.general ? ema : (fmin(ema, sig) + ema) / 2.0;  // general_tracking.py:985
      else S[PBHC_S_SIGMA + k] = fmin(ema, sig);
    }
  }
  L[PBHC_L_UPPER] = acc[PBHC_A_UPPER_NORM] / (n * c.n_upper);
  L[PBHC_L_LOWER] = acc[PBHC_A_LOWER_NORM] / (n * c.n_lower);
  L[PBHC_L_VR3] = acc[PBHC_A_VR3_NORM] / (n * c.n_vr3);
  L[PBHC_L_JOINT] = acc[PBHC_A_JOINT_NORM] / n;
  const double resets = acc[PBHC_A_RESET];
  const float reset_mean = (float)(resets / n) + 1e-15f;
  if (c.general) {  // general_tracking.py:303-335, 241-255
    L[PBHC_L_KEY] = acc[PBHC_A_KEY_NORM] / (n * c.n_key);
    L[PBHC_L_LOCAL_UPPER] = acc[PBHC_A_LOCAL_UPPER_NORM] / (n * c.n_upper);
    L[PBHC_L_LOCAL_LOWER] = acc[PBHC_A_LOCAL_LOWER_NORM] / (n * c.n_lower);
    L[PBHC_L_LOCAL_VR3] = acc[PBHC_A_LOCAL_VR3_NORM] / (n * c.n_vr3);
    L[PBHC_L_LOCAL_KEY] = acc[PBHC_A_LOCAL_KEY_NORM] / (n * c.n_key);
    L[PBHC_L_TERM_REF_POS_Z] = (double)((float)(acc[PBHC_A_TERM_REF_POS_Z] / n) / reset_mean);
    L[PBHC_L_TERM_REF_ORI] = (double)((float)(acc[PBHC_A_TERM_REF_ORI] / n) / reset_mean);
    L[PBHC_L_TERM_BODY_Z] = (double)((float)(acc[PBHC_A_TERM_BODY_Z] / n) / reset_mean);
  }
  if (c.term_contact) L[PBHC_L_TERM_CONTACT] = (double)((float)(acc[PBHC_A_TERM_CONTACT] / n) / reset_mean);        // legged_robot_base.py:425-429
  if (c.term_low_height) L[PBHC_L_TERM_LOW_HEIGHT] = (double)((float)(acc[PBHC_A_TERM_LOW_HEIGHT] / n) / reset_mean);
  if (c.term_dof_pos) L[PBHC_L_TERM_DOF_POS] = (double)((float)(acc[PBHC_A_TERM_DOF_POS] / n) / reset_mean);
  if (c.term_dof_vel) L[PBHC_L_TERM_DOF_VEL] = (double)((float)(acc[PBHC_A_TERM_DOF_VEL] / n) / reset_mean);
  if (c.term_torque) L[PBHC_L_TERM_TORQUE] = (double)((float)(acc[PBHC_A_TERM_TORQUE] / n) / reset_mean);
  L[PBHC_L_TERM_GRAVITY] = (double)((float)(acc[PBHC_A_TERM_GRAVITY] / n) / reset_mean);
  L[PBHC_L_TERM_FAR] = (double)((float)(acc[PBHC_A_TERM_FAR] / n) / reset_mean);
  L[PBHC_L_TERM_TIMEOUT] = (double)((float)(acc[PBHC_A_TERM_TIMEOUT] / n) / reset_mean);
  L[PBHC_L_TERM_MOTION_END] = (double)((float)(acc[PBHC_A_TERM_MOTION_END] / n) / reset_mean);
  if (resets > 0.0) {
    const float cur = (float)acc[PBHC_A_RESET_EPLEN] / (float)resets;
    const float avg = (float)S[PBHC_S_AVG_EPL] * (float)(1.0 - resets / c.num_compute_average_epl) + cur * (float)(resets / c.num_compute_average_epl);
    S[PBHC_S_AVG_EPL] = (double)avg;
    if (c.cur_penalty.enable) S[PBHC_S_PENALTY] = curriculum(S[PBHC_S_PENALTY], avg, c.cur_penalty, -1.0);
    if (c.cur_dof_pos.enable) S[PBHC_S_SOFT_DOF_POS] = curriculum(S[PBHC_S_SOFT_DOF_POS], avg, c.cur_dof_pos, 1.0);
    if (c.cur_dof_vel.enable) S[PBHC_S_SOFT_DOF_VEL] = curriculum(S[PBHC_S_SOFT_DOF_VEL], avg, c.cur_dof_vel, 1.0);
    if (c.cur_torque.enable) S[PBHC_S_SOFT_TORQUE] = curriculum(S[PBHC_S_SOFT_TORQUE], avg, c.cur_torque, 1.0);
    if (c.cur_far.enable) S[PBHC_S_FAR_THR] = curriculum(S[PBHC_S_FAR_THR], avg, c.cur_far, 1.0);
    if (c.cur_noise.enable) S[PBHC_S_NOISE_CUR] = curriculum(S[PBHC_S_NOISE_CUR], avg, c.cur_noise, -1.0);  // legged_robot_base.py:637-645
    const double mean = acc[PBHC_A_ETR_SUM] / n;
    L[PBHC_L_ETR_MEAN] = mean;
    const double var = n > 1.0 ? (acc[PBHC_A_ETR_SQ] - acc[PBHC_A_ETR_SUM] * mean) / (n - 1.0) : 0.0;
    L[PBHC_L_ETR_STD] = n > 1.0 ? sqrt(fmax(var, 0.0)) : nan("");  // torch.std() of one element is NaN (motion_tracking.py:414)
  }
  S[PBHC_S_PUSH_COUNT] += 1.0;
}


static int g_track_blocks_per_sm = 0;

// the deterministic reduction behind the step kernel, as a programmatic dependent launch: its launch latency overlaps the
// step kernel's tail instead of following its full drain (PBHC_REDUCE_PDL=0: a plain dependent launch)
static void launch_reduce(const double* partials, int blocks, double* acc, cudaStream_t stream) {
  static const bool pdl = [] { const char* e = getenv("PBHC_REDUCE_PDL"); return !(e && e[0] == '0'); }();
  if (!pdl) { track_reduce_kernel<<<1, kReduceGroups * PBHC_ACC, 0, stream>>>(partials, blocks, acc); return; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1); cfg.blockDim = dim3(kReduceGroups * PBHC_ACC); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, track_reduce_kernel, partials, blocks, acc) != cudaSuccess) {
    (void)cudaGetLastError();
    track_reduce_kernel<<<1, kReduceGroups * PBHC_ACC, 0, stream>>>(partials, blocks, acc);
  }
}
int jit_get(const char* what, const std::string& spec_header, bool noise, const void** kernel);   // jit.cu

// the future window of this launch can take the staged loops: compact table bound, its dof + key-body run copyable in 8- or
// 16-byte pieces, staging not switched off (PBHC_FUT_STAGE=0).  Used by pbhc_track_step AND pbhc_track_specialize, so the
// kernel that is warmed up is the kernel that is launched; a launch without the compact table gets the direct-read variant
// (forcing the staged loops onto the packed table is a misaligned cp.async - measured, it kills the context).
static bool fut_staged_ok(const PbhcTrackCfg& c, const PbhcTrackIO& io) {
  static const bool on = [] { const char* e = getenv("PBHC_FUT_STAGE"); return !(e && e[0] == '0'); }();
  return on && c.general && c.n_future > 0 && io.fut_table != nullptr && ((c.D + 3 * c.n_key) & 1) == 0;
}

// "track_spec.h" of the warp-per-env kernel: the env config as a constant byte image (seed zeroed: it stays a launch parameter)
// plus the launch facts the kernel is specialised on
static std::string track_spec_header(const PbhcTrackCfg& c, bool fut_staged) {
  PbhcTrackCfg z = c;
  z.seed = 0;
  const unsigned char* b = reinterpret_cast<const unsigned char*>(&z);
  std::string h = "// generated by pbhc_b200 (track.cu track_spec_header): the PbhcTrackCfg of one env config\nnamespace spec {\n"
                  "__device__ const unsigned char __align__(16) cfg_bytes[" + std::to_string(sizeof(z)) + "] = {";
  char buf[8];
  for (size_t i = 0; i < sizeof(z); ++i) { snprintf(buf, sizeof(buf), "%u,", (unsigned)b[i]); h += buf; if ((i & 63) == 63) h += "\n"; }
  h += "};\n";
  h += std::string("constexpr int fut_staged = ") + (fut_staged ? "1" : "0") + ";  // future window: staged loops (1) or direct table reads (0)\n";
  h += "}  // namespace spec\n";
  return h;
}

// the kernel specialised for this config (kernel == nullptr: compile into the on-disk cache only); cached per thread
static int jit_track(const PbhcTrackCfg& c, bool noise, bool fut_staged, const void** kernel) {
  static thread_local PbhcTrackCfg last;
  static thread_local bool last_noise = false, last_staged = false, have = false;
  static thread_local const void* last_fn = nullptr;
  static thread_local int last_rc = 0;
  PbhcTrackCfg z = c;
  z.seed = 0;
  if (kernel && have && last_noise == noise && last_staged == fut_staged && !memcmp(&last, &z, sizeof(z))) { *kernel = last_fn; return last_rc; }
  const void* fn = nullptr;
  const int rc = jit_get("track", track_spec_header(c, fut_staged), noise, kernel ? &fn : nullptr);
  if (kernel) { last = z; last_noise = noise; last_staged = fut_staged; have = true; last_fn = fn; last_rc = rc; *kernel = fn; }
  return rc;
}
int track_tile_launch(const PbhcTrackCfg& c, const PbhcTableDesc& tab, const PbhcTrackIO& io, int64_t n, cudaStream_t stream, int* blocks_out);  // track_tile.cu
int track_tile_prepare(const PbhcTrackCfg& c, const PbhcTableDesc& tab, const PbhcTrackIO& io);

}  // namespace pbhc

namespace pbhc {
// compact future-window records (include/pbhc_b200.h): one thread per (frame, float)
struct FutKeys { int v[PBHC_MAX_KEY]; };
__global__ void future_table_kernel(const float* __restrict__ rec, int64_t frames, int rec_stride, int J, int D, int a, int K, const __grid_constant__ FutKeys keys,
                                    float* __restrict__ out, int stride) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= frames * stride) return;
  const int64_t f = i / stride;
  const int o = (int)(i - f * stride);
  const float* r = rec + f * rec_stride;
  const int e_pos = 4 * J, e_vel = 7 * J, e_ang = 10 * J, e_dp = 13 * J;
  float v = 0.0f;
  if (o < 4) v = r[o];
  else if (o < 8) v = r[4 * a + (o - 4)];
  else if (o < 11) v = r[e_vel + (o - 8)];
  else if (o < 14) v = r[e_ang + (o - 11)];
  else if (o < 17) v = r[e_pos + (o - 14)];
  else if (o < 20) v = r[e_pos + 3 * a + (o - 17)];
  else if (o < 20 + D) v = r[e_dp + (o - 20)];
  else if (o < 20 + D + 3 * K) { const int q = o - 20 - D; v = r[e_pos + 3 * keys.v[q / 3] + q % 3]; }
  out[i] = v;
}
}  // namespace pbhc

extern "C" int32_t pbhc_future_table_stride(const PbhcTrackCfg* cfg) {
  if (!cfg || !cfg->general || cfg->n_future <= 0) return 0;
  return (20 + cfg->D + 3 * cfg->n_key + 3) & ~3;
}

extern "C" int pbhc_future_table_build(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, float* out, int32_t stride, void* stream) {
  using namespace pbhc;
  if (!cfg || !tab || !tab->records || !out) { set_error("pbhc_future_table_build: null argument"); return PBHC_E_BADARG; }
  if (stride != pbhc_future_table_stride(cfg) || stride <= 0 || cfg->n_key > PBHC_MAX_KEY || cfg->J != tab->num_joints || cfg->D != tab->num_dof) {
    set_error("pbhc_future_table_build: stride %d does not fit the config", stride); return PBHC_E_RANGE;
  }
  if (((uintptr_t)out & 15)) { set_error("pbhc_future_table_build: out must be 16-byte aligned"); return PBHC_E_ALIGN; }
  if (tab->total_frames == 0) return PBHC_OK;
  FutKeys keys;  // (the key ids travel in the kernel's argument block)
  for (int k = 0; k < PBHC_MAX_KEY; ++k) keys.v[k] = k < cfg->n_key ? cfg->key_ids[k] : 0;
  const int64_t total = tab->total_frames * stride;
  const int64_t blocks = (total + 255) / 256;
  future_table_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(tab->records, tab->total_frames, tab->rec_stride, cfg->J, cfg->D, cfg->anchor_index,
                                                                         cfg->n_key, keys, out, stride);
  return check_launch("pbhc_future_table_build");
}

extern "C" int pbhc_track_max_blocks(void) { return pbhc::kTrackMaxBlocks; }

namespace pbhc { int g_track_info[4] = {0, 0, 0, 0}; }  // kernel (0 warp-per-env, 1 tile), blocks, dynamic shared memory, resident blocks / SM
extern "C" int pbhc_track_launch_info(int32_t* out4) {
  if (!out4) return PBHC_E_BADARG;
  for (int i = 0; i < 4; ++i) out4[i] = pbhc::g_track_info[i];
  return PBHC_OK;
}

extern "C" int pbhc_track_specialize(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, const PbhcTrackIO* io) {
  if (!cfg || !tab || !io) { pbhc::set_error("pbhc_track_specialize: null argument"); return PBHC_E_BADARG; }
  if (cfg->general) {  // LeggedRobotGeneralTracking: the warp-per-env kernel, specialised for the config
    bool noise = false;
    for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k) noise |= cfg->obs_noise[k] != 0.0f;
    return pbhc::jit_track(*cfg, noise, pbhc::fut_staged_ok(*cfg, *io), nullptr) == 0 ? PBHC_OK : PBHC_E_CUDA;
  }
  if (pbhc::needs_warp_kernel(*cfg)) return PBHC_OK;  // served by the generic warp-per-env kernel
  return pbhc::track_tile_prepare(*cfg, *tab, *io);
}

extern "C" int pbhc_track_step(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, const PbhcTrackIO* io, int64_t n, void* stream) {
  using namespace pbhc;
  if (!cfg || !tab || !io || n < 0) { set_error("pbhc_track_step: null argument"); return PBHC_E_BADARG; }
  const PbhcTrackCfg& c = *cfg;
  if (c.J <= 0 || c.J > 64 || c.Jb > c.J || c.D <= 0 || c.D > PBHC_MAX_DOF || c.n_ext != c.J - c.Jb || c.n_ext > PBHC_MAX_EXT ||
      c.n_feet < 0 || c.n_feet > 2 || c.n_rewards < 0 || c.n_rewards >= 32 || c.n_episode_sums > 32 || c.n_groups < 0 ||
      c.n_groups > PBHC_MAX_OBS_GROUPS || c.n_hist_keys < 0 || c.n_hist_keys > PBHC_MAX_HIST_KEYS || c.n_hist_dst < 0 ||
      c.n_hist_dst > PBHC_MAX_HIST_DST || c.J != tab->num_joints || c.Jb != tab->num_bodies || c.D != tab->num_dof ||
      tab->total_frames > 0x7fffffffLL) {
    set_error("pbhc_track_step: configuration outside the supported envelope (J<=64, D<=32, feet<=2, rewards<32)"); return PBHC_E_RANGE;
  }
  const void* need[] = {io->body_state, io->root_state, io->dof_state, io->contact_forces, io->actions, io->torques, io->env_origins,
                        io->motion_ids, io->default_dof_pos, io->base_com_bias, io->friction, io->episode_length_buf, io->env_f32,
                        io->last_actions, io->last_dof_pos, io->last_dof_vel, io->kp_scale, io->kd_scale, io->rfi_lim_scale, io->rao_scale,
                        io->action_delay_idx, io->episode_sums, io->scalars, io->rew_buf, io->reset_buf, io->time_out_buf,
                        io->last_episode_sums, io->partials, io->acc};
  for (const void* p : need) if (!p) { set_error("pbhc_track_step: a required PbhcTrackIO pointer is NULL"); return PBHC_E_BADARG; }
  if (c.push_robots && (c.push_hi <= c.push_lo || c.push_lo < 1 || c.push_hi > 32767 || (float)c.push_hi / c.dt > 32767.0f)) {
    set_error("pbhc_track_step: push_interval_s [%d, %d) outside what the 16 + 15 bit push word holds", c.push_lo, c.push_hi); return PBHC_E_RANGE;
  }
  if (c.L < 0 || c.L > 48) { set_error("pbhc_track_step: L outside [0, 48]"); return PBHC_E_RANGE; }
  if ((c.L > 0 && !io->link_mass_scale) || (c.hist_row > 0 && !io->hist)) { set_error("pbhc_track_step: NULL link_mass_scale / hist"); return PBHC_E_BADARG; }
  for (int g = 0; g < c.n_groups; ++g) if (!io->obs[g] || c.group_dim[g] <= 0) { set_error("pbhc_track_step: NULL / empty obs group %d", g); return PBHC_E_BADARG; }
  if (((uintptr_t)tab->records & 15) || ((uintptr_t)io->dof_state & 7) || ((uintptr_t)io->env_f32 & 15) ||
      (io->ref_body_rot_extend && ((uintptr_t)io->ref_body_rot_extend & 15))) {
    set_error("pbhc_track_step: records / env_f32 / ref_body_rot_extend must be 16-B, dof_state 8-B aligned"); return PBHC_E_ALIGN;
  }
  if (c.general) {
    if (c.anchor_index < 0 || c.anchor_index >= c.Jb || c.n_key < 0 || c.n_key > PBHC_MAX_KEY || c.n_future < 0 || c.n_future > PBHC_MAX_FUTURE) {
      set_error("pbhc_track_step: general_tracking needs anchor < Jb, <= %d key bodies, <= %d future steps", PBHC_MAX_KEY, PBHC_MAX_FUTURE); return PBHC_E_RANGE;
    }
    for (int i = 0; i < c.n_key; ++i) if (c.key_ids[i] < 0 || c.key_ids[i] >= c.J) { set_error("pbhc_track_step: key body %d out of range", i); return PBHC_E_RANGE; }
    for (int k = S_FUT_HEIGHT; k <= S_FUT_KEY; ++k) if (c.obs_noise[k] != 0.0f) { set_error("pbhc_track_step: noise on future-target observations is not implemented"); return PBHC_E_RANGE; }
    if ((c.src_used_mask >> S_FUT_REF_DOF_POS) & 3ull) { set_error("pbhc_track_step: future_ref_dof_pos / _vel exist only in motion_tracking"); return PBHC_E_RANGE; }
    for (int r = 0; r < c.n_rewards; ++r)
      if (c.reward_id[r] == R_RADIAL_BODY_VEL || c.reward_id[r] == R_RADIAL_JOINT_VEL) { set_error("pbhc_track_step: the radial velocity rewards exist only in motion_tracking"); return PBHC_E_RANGE; }
    for (int r = 0; r < c.n_rewards; ++r)
      if (c.reward_id[r] == R_BODY_ROT || c.reward_id[r] == R_BODY_VEL || c.reward_id[r] == R_BODY_ANG_VEL) {
        set_error("pbhc_track_step: whole-body rotation / velocity rewards are not implemented for general_tracking"); return PBHC_E_RANGE;
      }
  } else {
    for (int k = S_ROLL_PITCH; k <= S_FUT_KEY; ++k)
      if ((c.src_used_mask >> k) & 1ull) { set_error("pbhc_track_step: observation source %d exists only in general_tracking", k); return PBHC_E_RANGE; }
    if (c.n_future < 0 || c.n_future > PBHC_MAX_FUTURE) { set_error("pbhc_track_step: future_ref_steps outside [0, %d]", PBHC_MAX_FUTURE); return PBHC_E_RANGE; }
    for (int r = 0; r < c.n_rewards; ++r) if (c.reward_id[r] >= R_ROOT_VEL && c.reward_id[r] <= R_KEY_ANG) { set_error("pbhc_track_step: reward %d exists only in general_tracking", c.reward_id[r]); return PBHC_E_RANGE; }
  }
  TrackParams P;
  P.c = c; P.tab = *tab; P.io = *io; P.n = n;
  P.rew_extra = 0;
  for (int r = 0; r < c.n_rewards; ++r) {
    if (c.reward_id[r] > R_FEET_ORI_CONTACT || c.reward_id[r] < 0) { set_error("pbhc_track_step: unknown reward id %d", c.reward_id[r]); return PBHC_E_RANGE; }
    if (c.reward_id[r] >= R_RADIAL_BODY_VEL) P.rew_extra |= 1 << (c.reward_id[r] - R_RADIAL_BODY_VEL);
  }
  if ((P.rew_extra & 0x3c) && c.n_feet != 2) { set_error("pbhc_track_step: the feet heading / orientation rewards take exactly two feet"); return PBHC_E_RANGE; }
  P.need_rootq = ((P.rew_extra & 0x3c) || (c.src_used_mask & ((1ull << S_RELYAW) | (1ull << S_LOCAL_REF_RELYAW)))) ? 1 : 0;
  if (c.src_used_mask & (1ull << S_VR3PT)) {
    if (c.n_vr3 != 3 || !io->heading_scratch || ((uintptr_t)io->heading_scratch & 7)) {
      set_error("pbhc_track_step: vr_3point_pos needs three motion_tracking_link bodies and an 8-byte aligned heading_scratch [N,2]"); return PBHC_E_BADARG;
    }
    for (int k = 0; k < 3; ++k) if (c.vr3_ids[k] < 0 || c.vr3_ids[k] >= c.J) { set_error("pbhc_track_step: vr3_ids[%d] out of range", k); return PBHC_E_RANGE; }
  }
  for (int k = S_FUT_REF_DOF_POS; k <= S_FUT_REF_DOF_VEL; ++k)
    if (((c.src_used_mask >> k) & 1ull) && c.obs_noise[k] != 0.0f && c.n_future * c.D > 256) { set_error("pbhc_track_step: noise on future_ref_dof_* needs future_ref_steps * D <= 256"); return PBHC_E_RANGE; }
  P.has_air_time = 0;
  for (int r = 0; r < c.n_rewards; ++r) if (c.reward_id[r] == R_FEET_AIR_TIME) P.has_air_time = 1;
  // a group made only of future-target sources never passes through shared memory
  for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) {
    bool fut = false, other = false;
    if (c.general && g < c.n_groups) {
      for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k) if (c.src_dst[k][g] >= 0) ((k >= S_FUT_HEIGHT && k <= S_FUT_KEY) ? fut : other) = true;
      for (int e = 0; e < c.n_hist_dst; ++e) if (c.hist_dst[e].group == g) other = true;
    }
    P.group_direct[g] = (fut && !other) ? 1 : 0;
  }
  for (int k = S_FUT_HEIGHT; c.general && k <= S_FUT_KEY; ++k) {
    int nd = 0, nr = 0;
    for (int g = 0; g < c.n_groups; ++g) if (c.src_dst[k][g] >= 0) (P.group_direct[g] ? nd : nr)++;
    if (nd > 1 || nr > 1) { set_error("pbhc_track_step: a future-target observation may appear in at most one pure and one mixed group"); return PBHC_E_RANGE; }
  }
  // history layout: global row = keys back to back ([len_k, dim_k] each); push staging = one slot per key
  int hrow = 0, push = 0;
  for (int k = 0; k < PBHC_MAX_HIST_KEYS; ++k) { P.hist_base[k] = hrow; P.push_off[k] = push; if (k < c.n_hist_keys) { hrow += c.hist_len[k] * c.hist_dim[k]; push += c.hist_dim[k]; } }
  if (hrow != c.hist_row) { set_error("pbhc_track_step: hist_row %d != sum len*dim %d", c.hist_row, hrow); return PBHC_E_BADARG; }
  P.push_len = push;
  auto up4h = [](int v) { return (v + 3) & ~3; };
  int off = 0;
  for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) { P.row_off[g] = off; if (g < c.n_groups && !P.group_direct[g]) off += up4h(c.group_dim[g]); }
  P.row_off[c.n_groups] = off;
  P.rows_len = off + up4h(push);
  // validate the destination tables against the row sizes
  for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k)
    for (int g = 0; g < c.n_groups; ++g)
      if (c.src_dst[k][g] >= 0 && c.src_dst[k][g] + src_dim(k, c.D, c.J, c.L, c.n_key, c.n_future) > c.group_dim[g]) { set_error("pbhc_track_step: src_dst[%d][%d] out of range", k, g); return PBHC_E_RANGE; }
  int per_key[PBHC_MAX_HIST_KEYS] = {0};
  P.hist_any_mul = 0; P.n_hdst_max = 0;
  for (int e = 0; e < c.n_hist_dst; ++e) {
    const PbhcHistDst& h = c.hist_dst[e];
    if (h.key < 0 || h.key >= c.n_hist_keys || h.group < 0 || h.group >= c.n_groups || h.len < 0 || h.len > c.hist_len[h.key] ||
        h.offset < 0 || h.offset + h.len * c.hist_dim[h.key] > c.group_dim[h.group]) { set_error("pbhc_track_step: hist_dst[%d] out of range", e); return PBHC_E_RANGE; }
    if (++per_key[h.key] > kMaxHistDstPerKey) { set_error("pbhc_track_step: a history key appears in more than %d segments", kMaxHistDstPerKey); return PBHC_E_RANGE; }
    if (per_key[h.key] > P.n_hdst_max) P.n_hdst_max = per_key[h.key];
    if (h.mul != 1.0f) P.hist_any_mul = 1;
  }
  for (int k = 0; k < c.n_hist_keys; ++k) if (c.hist_len[k] <= 0 || c.hist_dim[k] != src_dim(c.hist_src[k], c.D, c.J, c.L, c.n_key, c.n_future)) { set_error("pbhc_track_step: bad history key %d", k); return PBHC_E_BADARG; }
  if (n == 0) return PBHC_OK;
  if (c.src_used_mask & (1ull << S_VR3PT)) heading_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(io->root_state, n, io->heading_scratch);
  // LeggedRobotMotionTracking runs on the tile kernel (csrc/track_tile.cu); the warp-per-env kernel below serves
  // LeggedRobotGeneralTracking and the configs the tile kernel does not take (a body- or joint-difference source kept in
  // the history; PBHC_TRACK_KERNEL=warp forces it for A/B measurements)
  if (!c.general) {
    static const bool force_warp = [] { const char* e = getenv("PBHC_TRACK_KERNEL"); return e && e[0] == 'w'; }();
    bool tile_ok = !force_warp && c.n_feet >= 1 && !needs_warp_kernel(c);
    for (int k = 0; k < c.n_hist_keys; ++k) {
      const int s = c.hist_src[k];
      if (s == S_DIF_LOCAL || s == S_LOCAL_REF || s == S_LOCAL_REF_VEL || s == S_GLOBAL_REF_VEL || s == S_DIF_JOINT || s == S_DIF_JOINT_VEL) tile_ok = false;
    }
    if (tile_ok) {
      int blocks = 0;
      const int rc = track_tile_launch(c, *tab, *io, n, (cudaStream_t)stream, &blocks);
      if (rc < 0) return rc;
      if (rc == PBHC_OK) {
        launch_reduce(io->partials, blocks, io->acc, (cudaStream_t)stream);
        return check_launch("pbhc_track_step");
      }  // (rc > 0: tensors the tile kernel does not take, e.g. not 16-byte aligned - the kernel below runs)
    }
  }
  P.dump = P.rows_len;                                   // scratch for absent destinations: one source wide
  // (widest source: 3*J body features, 6*n_key key-body rotations, 7+D+3K next-step block, L link masses, 64 history lanes)
  int widest = 3 * c.J + 4;
  for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k) { const int w = src_dim(k, c.D, c.J, c.L, c.n_key, c.general ? 0 : c.n_future); if (w > widest) widest = w; }
  if (widest < 64) widest = 64;
  const int dump_len = up4h(widest);
  // per-warp layout: the fixed-size tables first, so that the kernel reaches rv / pre / the staging buffer with
  // compile-time offsets from one base register
  P.sm_rv = 0;
  P.sm_pre = kRv;
  P.sm_body = kRv + kPre * kChunk;
  P.sm_cf = P.sm_body + up4h(c.Jb * 13);
  int stage_len = up4h(c.Jb * 13) + up4h(c.Jb * 3);  // staging buffer = body state + contact forces (+ padding)
  // general_tracking runs two blocks per SM (register-bound): there is shared memory to spare, so the staging buffer is
  // widened to also take the (longer) history ring
  if (c.general && c.hist_row > stage_len && c.hist_row <= 1024) stage_len = up4h(c.hist_row);
  // ... and to take kFutChunk steps x 2 frames of the future window's dof / key-body rows behind its per-step table
  // (PBHC_FUT_STAGE=0: the window reads the table directly, for A/B runs)
  P.fut_win = 0;
  if (fut_staged_ok(c, *io)) {
    const int need = kFut * 32 + kFutChunk * 2 * (c.D + 3 * c.n_key);
    if (need > stage_len) stage_len = up4h(need);
    P.fut_win = 1;
  }
  {  // PBHC_TRACK_VERBOSE=1: which path the future window of this launch takes (once per process)
    static bool told = false;
    const char* v = getenv("PBHC_TRACK_VERBOSE");
    if (!told && v && v[0] == '1' && c.general && c.n_future > 0) {
      told = true;
      const int row_f = c.D + 3 * c.n_key, wf = (row_f & 3) == 0 ? 4 : ((row_f & 1) == 0 ? 2 : 0);
      fprintf(stderr, "[pbhc track] future window: run of %d floats per frame, compact table %s, staging %s -> %s\n", row_f,
              io->fut_table ? "yes" : "no", P.fut_win ? "on" : "off",
              (io->fut_table && wf && P.fut_win) ? (wf == 4 ? "staged, 16-byte copies" : "staged, 8-byte copies") : "direct table reads");
    }
  }
  P.sm_rows = P.sm_body + stage_len;
  P.sm_per_warp = P.sm_rows + up4h(P.rows_len) + dump_len;
  P.o_hmap = (int)((sizeof(BlockTables) + 15) & ~size_t(15));
  P.o_pmap = P.o_hmap + 4 * up4h((P.n_hdst_max < 2 ? 2 : P.n_hdst_max) * c.hist_row);
  P.o_warp = P.o_pmap + 4 * up4h(push);
  P.warp_bytes = 4 * P.sm_per_warp;
  P.hist_stage = (c.hist_row > 0 && c.hist_row <= stage_len) ? 1 : 0;
  P.ob_body = 4 * P.sm_body; P.ob_cf = 4 * P.sm_cf; P.ob_rows = 4 * P.sm_rows; P.ob_rv = 4 * P.sm_rv; P.ob_pre = 4 * P.sm_pre;
  P.e_pos = 4 * c.J; P.e_vel = 7 * c.J; P.e_ang = 10 * c.J; P.e_dp = 13 * c.J; P.e_dv = P.e_dp + c.D; P.e_c = P.e_dv + c.D;
  {
    const int bv = vec_of(io->body_env_stride, io->body_state);
    P.body_vec = (bv == 4 && ((c.Jb * 13) & 3) == 0) ? 4 : ((bv >= 2 && ((c.Jb * 13) & 1) == 0) ? 2 : 1);
    P.cf_vec = vec_of(c.Jb * 3, io->contact_forces);
    P.hist_vec = c.hist_row > 0 ? vec_of(c.hist_row, io->hist) : 1;
    for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) P.obs_vec[g] = g < c.n_groups ? vec_of(c.group_dim[g], io->obs[g]) : 1;
  }
  if (n > 0x7fffffffLL / (kChunk * 2)) { set_error("pbhc_track_step: too many envs for one launch"); return PBHC_E_RANGE; }
  const size_t shmem = (size_t)P.o_warp + (size_t)kTW * P.warp_bytes;
  if (shmem > 220 * 1024) { set_error("pbhc_track_step: shared memory %zu B too large", shmem); return PBHC_E_RANGE; }
  bool noise = false;
  for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k) noise |= c.obs_noise[k] != 0.0f;
  // two instantiations: the noise-free one (benchmark.yaml, evaluation) carries no Philox code in its observation stores
  if (c.general && c.n_future > 0 && kFut * 32 > P.sm_rows - P.sm_body) { set_error("pbhc_track_step: future table does not fit the staging buffer"); return PBHC_E_RANGE; }
  const int variant = (noise ? 1 : 0) + (c.general ? 2 : 0);
  const void* kern = variant == 0 ? (const void*)track_kernel<false, false> : variant == 1 ? (const void*)track_kernel<true, false> :
                     variant == 2 ? (const void*)track_kernel<false, true> : (const void*)track_kernel<true, true>;
  static thread_local size_t configured[4] = {0, 0, 0, 0};
  static thread_local int per_sm_of[4] = {0, 0, 0, 0};
  int which = 0;
  if (c.general) {  // the kernel specialised for this env config, when it can be had (NVRTC, cached on disk; PBHC_JIT=0 turns it off)
    const void* jk = nullptr;
    if (jit_track(c, noise, P.fut_win != 0, &jk) == 0 && jk) {
      static thread_local const void* jk_configured = nullptr;
      static thread_local size_t jk_shmem = 0;
      static thread_local int jk_per_sm = 0;
      if (jk != jk_configured || shmem > jk_shmem) {
        if (cudaFuncSetAttribute(jk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_track_step(attr)");
        (void)cudaFuncSetAttribute(jk, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        int per_sm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jk, kTW * 32, shmem);
        jk_configured = jk; jk_shmem = shmem; jk_per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
      }
      g_track_blocks_per_sm = jk_per_sm;
      int64_t blocks = (n + kTW - 1) / kTW;
      if (blocks > 148LL * jk_per_sm) blocks = 148LL * jk_per_sm;
      g_track_info[0] = 3; g_track_info[1] = (int)blocks; g_track_info[2] = (int)shmem; g_track_info[3] = jk_per_sm;
      void* args[] = {(void*)&P};
      if (cudaLaunchKernel(jk, dim3((unsigned)blocks), dim3(kTW * 32), args, shmem, (cudaStream_t)stream) != cudaSuccess) return check_launch("pbhc_track_step(spec)");
      launch_reduce(io->partials, (int)blocks, io->acc, (cudaStream_t)stream);
      return check_launch("pbhc_track_step");
    }
  }
  (void)which;
  if (shmem > configured[variant] || per_sm_of[variant] == 0) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_track_step(attr)");
    (void)cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured[variant] = shmem;
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kTW * 32, shmem);
    per_sm_of[variant] = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
  }
  g_track_blocks_per_sm = per_sm_of[variant];
  int64_t blocks = (n + kTW - 1) / kTW;
  const int64_t cap = 148LL * g_track_blocks_per_sm;
  if (blocks > cap) blocks = cap;
  g_track_info[0] = 0; g_track_info[1] = (int)blocks; g_track_info[2] = (int)shmem; g_track_info[3] = g_track_blocks_per_sm;
  {
    void* args[] = {(void*)&P};
    if (cudaLaunchKernel(kern, dim3((unsigned)blocks), dim3(kTW * 32), args, shmem, (cudaStream_t)stream) != cudaSuccess) return check_launch("pbhc_track_step(launch)");
  }
  launch_reduce(io->partials, (int)blocks, io->acc, (cudaStream_t)stream);
  return check_launch("pbhc_track_step");
}

extern "C" int pbhc_track_finalize(const PbhcTrackCfg* cfg, const double* acc, double* scalars, void* stream) {
  using namespace pbhc;
  if (!cfg || !acc || !scalars) { set_error("pbhc_track_finalize: null argument"); return PBHC_E_BADARG; }
  track_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*cfg, acc, scalars);
  return check_launch("pbhc_track_finalize");
}
#endif  // __CUDACC_RTC__
