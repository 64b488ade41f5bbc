// track_common.cuh - enums, per-env scalar table layout and small device helpers shared by the two K3 kernels
// (csrc/track.cu: warp-per-env kernel, LeggedRobotGeneralTracking; csrc/track_tile.cu: tile kernel, LeggedRobotMotionTracking).
#pragma once
#include "common.cuh"
#include "table.cuh"

namespace pbhc {

constexpr int kRv = 72;                     // per-env scalar table (see enum Rv)
constexpr int kMaxHistDstPerKey = 4;
constexpr int kPre = 24;                    // per-env values of the pre phase
constexpr int kFut = 12;                    // per-future-step values staged by the general_tracking future pass

enum ObsSrc { S_BASE_LIN_VEL, S_BASE_ANG_VEL, S_PROJ_GRAV, S_DOF_POS, S_DOF_VEL, S_ACTIONS, S_PHASE, S_DIF_LOCAL,
              S_LOCAL_REF, S_DR_COM, S_DR_MASS, S_DR_KP, S_DR_KD, S_DR_FRICTION, S_DR_DELAY, S_LOCAL_REF_VEL,
              S_GLOBAL_REF_VEL, S_DIF_JOINT, S_DIF_JOINT_VEL, S_BASE_POS_Z,
              // LeggedRobotGeneralTracking
              S_ROLL_PITCH, S_LKEY_POS, S_LKEY_ROT, S_ANCHOR_POS, S_ANCHOR_ROT, S_NEXT_STEP, S_CONTACT_MASK, S_ROOT_HEIGHT,
              S_DR_BASE_MASS, S_FUT_HEIGHT, S_FUT_RP, S_FUT_LIN, S_FUT_YAW, S_FUT_DOF, S_FUT_KEY,
              // both env classes (motion_tracking.py:686-731 / general_tracking.py:682-725): warp-per-env kernel only
              S_RELYAW, S_LOCAL_REF_RELYAW, S_VR3PT,
              // LeggedRobotMotionTracking with obs.future_ref_steps > 0 (motion_tracking.py:594-599)
              S_FUT_REF_DOF_POS, S_FUT_REF_DOF_VEL };
static_assert(S_FUT_REF_DOF_VEL + 1 == PBHC_NUM_OBS_SRC, "ObsSrc and PBHC_NUM_OBS_SRC disagree");
enum RewardId { R_CONTACT_MASK, R_MAX_JOINT_POS, R_BODY_POS_EXT, R_VR3, R_FEET_POS, R_BODY_ROT, R_BODY_ANG_VEL,
                R_BODY_VEL, R_JOINT_POS, R_JOINT_VEL, R_TORQUES, R_ACTION_RATE, R_FEET_AIR_TIME, R_FEET_FORCES,
                R_STUMBLE, R_SLIPPAGE, R_LIM_DOF_POS, R_LIM_DOF_VEL, R_LIM_TORQUE, R_COLLISION, R_DOF_VEL, R_DOF_ACC,
                R_FOOT_SLIP, R_ALIVE, R_ORIENTATION, R_CONTACT_MASK_V2,
                R_ROOT_VEL, R_ANCHOR_POS, R_ANCHOR_ROT, R_LKEY_POS, R_LKEY_ROT, R_KEY_VEL, R_KEY_ANG,
                // zero-scaled in the shipped configs: warp-per-env kernel only.  The two radial terms exist in
                // LeggedRobotMotionTracking only (motion_tracking.py:1238-1244, 1286-1293), the feet terms in the base class
                // (legged_robot_base.py:1031-1082)
                R_RADIAL_BODY_VEL, R_RADIAL_JOINT_VEL, R_FEET_HEADING, R_FEET_ORI, R_FEET_HEADING_CONTACT, R_FEET_ORI_CONTACT };
enum SigmaId { G_MAX_JOINT, G_UPPER, G_LOWER, G_VR3, G_FEET, G_ROT, G_VEL, G_ANG_VEL, G_JOINT_POS, G_JOINT_VEL,
               G_ROOT_VEL, G_ANCHOR_POS, G_ANCHOR_ROT, G_LKEY_POS, G_LKEY_ROT, G_KEY_VEL, G_KEY_ANG };
static_assert(G_KEY_ANG + 1 == PBHC_NUM_SIGMA, "SigmaId and PBHC_NUM_SIGMA disagree");
// per-env scalar table `rv`: slots 0..39 are laid out like `acc` (PBHC_A_*), the rest are raw reward values
enum Rv { V_SIGMA_ERR = 0, V_UP_N = 17, V_LO_N, V_VR_N, V_JOINT_N, V_T_GRAV, V_T_FAR, V_T_OUT, V_T_END, V_RESET, V_RESET_EPLEN,
          V_ETR, V_ETR_SQ, V_ONE, V_KEY_N = 32, V_LUP_N, V_LLO_N, V_LVR_N, V_LKEY_N, V_T_REFZ, V_T_ORI, V_T_BODYZ,
          V_CONTACT = 40, V_CONTACT_V2, V_TQ2, V_DACT2, V_DV2, V_ACC2, V_ORIENT, V_ALIVE, V_AIR, V_FORCE,
          V_STUMBLE, V_SLIP, V_FSLIP, V_LIM_POS, V_LIM_VEL, V_LIM_TQ, V_COLLIDE, V_RAD_BODY, V_RAD_JOINT, V_FEET_HEAD,
          V_ENVF = 60 /* 8 floats */, V_EP = 68 /* int64 */, V_FEET_ORI = 70, V_FEET_HEAD_C, V_FEET_ORI_C = 31 };
static_assert(V_FEET_HEAD == 59 && V_FEET_HEAD_C == 71, "rv layout");
static_assert(V_UP_N == PBHC_A_UPPER_NORM && V_ONE == PBHC_A_NUM_ENVS && V_KEY_N == PBHC_A_KEY_NORM && V_T_BODYZ == PBHC_A_TERM_BODY_Z &&
              V_RESET_EPLEN == PBHC_A_RESET_EPLEN && V_ETR_SQ == PBHC_A_ETR_SQ, "rv and acc layouts disagree");
constexpr int kAccLo = PBHC_A_NUM_ENVS + 1;  // slots [0, kAccLo) are accumulated by lane = slot; general_tracking adds [32, 40)
enum RewardKind { K_RAW, K_EXP, K_EXP2 };
// pre-phase values (per env)
enum Pre { P_F0, P_F1, P_BLEND, P_PHASE, P_HZ, P_HW, P_RPX, P_RPY, P_RPZ, P_BLV, P_BAV = P_BLV + 3, P_PG = P_BAV + 3,
           P_CLIP = P_PG + 3, P_GRAVN, P_ORIENT, P_ROLL, P_PITCH };
static_assert(P_PITCH + 1 <= kPre, "pre table too small");
// general_tracking: per-future-step values (lane s = step s) staged in the idle simulator-state buffer
enum Fut { F_F0, F_F1, F_BLEND, F_APX, F_APY, F_APZ, F_AQX, F_AQY, F_AQZ, F_AQW };
static_assert(F_AQW + 1 <= kFut, "future table too small");

__host__ __device__ inline int src_dim(int k, int D, int J, int L, int K = 0, int F = 0) {
  switch (k) {
    case S_ROLL_PITCH: case S_CONTACT_MASK: return 2;
    case S_LKEY_POS: return 3 * K;
    case S_LKEY_ROT: return 6 * K;
    case S_ANCHOR_ROT: return 6;
    case S_NEXT_STEP: return 7 + D + 3 * K;
    case S_ROOT_HEIGHT: case S_DR_BASE_MASS: case S_RELYAW: return 1;
    case S_VR3PT: return 9;
    case S_FUT_REF_DOF_POS: case S_FUT_REF_DOF_VEL: return F * D;
    case S_FUT_HEIGHT: case S_FUT_YAW: return F;
    case S_FUT_RP: return 2 * F;
    case S_FUT_LIN: return 3 * F;
    case S_FUT_DOF: return F * D;
    case S_FUT_KEY: return F * 3 * K;
    case S_DOF_POS: case S_DOF_VEL: case S_ACTIONS: case S_DR_KP: case S_DR_KD: case S_DIF_JOINT: case S_DIF_JOINT_VEL: return D;
    case S_DIF_LOCAL: case S_LOCAL_REF: case S_LOCAL_REF_VEL: case S_GLOBAL_REF_VEL: case S_LOCAL_REF_RELYAW: return 3 * J;
    case S_DR_MASS: return L;
    case S_PHASE: case S_DR_FRICTION: case S_DR_DELAY: case S_BASE_POS_Z: return 1;
    default: return 3;
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Sums 16 per-lane values across the warp with 16 shuffles.  Round r keeps, in each lane, the half of the values
// selected by one lane bit and adds the partner's copy of them; after the xor-16/8/4/2 rounds every lane holds one
// value index ((lane >> 1) & 15), the xor-1 round completes it.  For each value the additions are the same pairs in
// the same order as the plain xor butterfly (16, 8, 4, 2, 1), so the results are bit-identical to warp_sum().
__device__ __forceinline__ float warp_sum16(float (&v)[16], int lane) {
#pragma unroll
  for (int half = 8, bit = 16; half >= 1; half >>= 1, bit >>= 1) {
    const bool upper = (lane & bit) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = upper ? v[i] : v[i + half];
      const float keep = upper ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// Philox4x32-10 (common.cuh), one call per (env, step, stream)
static __device__ __noinline__ uint4 philox(uint4 ctr, uint2 key) { return philox4x32_10(ctr, key); }  // cold (resets, observation noise): out of line so the hot loop stays small
struct Rng {
  uint2 key; uint32_t env_lo, env_hi, step;
  __device__ __forceinline__ uint4 draw(uint32_t stream) const { return philox(make_uint4(env_lo, env_hi, step, stream), key); }
  __device__ __forceinline__ float uniform(uint32_t stream) const { return u01(draw(stream).x); }
  __device__ __noinline__ float normal(uint32_t stream) const {
    const uint4 r = draw(stream);
    const float u1 = fmaxf(u01(r.x), 1.0f / 16777216.0f), u2 = u01(r.y);
    return sqrtf(-2.0f * logf(u1)) * cosf(6.283185307179586f * u2);
  }
};

// _update_counters_each_step + the push part of _update_tasks_callback + _push_robots (legged_robot_base.py:143-148, 382-388,
// 1090-1106) for one env, run by one lane.  `word` = the env's push word (env_f32 slot 7): counter | threshold << 16 with
// threshold = (push_interval_s / dt).int(), so that the per-step part push_step() is an increment and a compare; the push itself - rare - is out of line:
// a pushed env draws its next interval and a root lin-vel x / y of U(-max, max), which is added to / stored in the root state
// unless the env also resets this step (the reset comes later in the reference's step and overwrites the root state).
static __device__ __noinline__ int push_cold(const PbhcTrackCfg& c, const PbhcTrackIO& io, int64_t n, bool reset, uint64_t seed, uint32_t step) {
  const int64_t g = n + io.env_offset;
  const Rng rng = {make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)), (uint32_t)g, (uint32_t)((uint64_t)g >> 32), step};
  const int iv = io.inject_push_interval ? (int)io.inject_push_interval[n] : c.push_lo + (int)(rng.draw(240).x % (uint32_t)(c.push_hi - c.push_lo));
  const uint4 r = rng.draw(241);
  const float lo = -c.max_push_vel, hi = c.max_push_vel;  // torch_rand_float: (upper - lower) * rand + lower
  const float vx = io.inject_push_vel ? io.inject_push_vel[2 * n] : __fadd_rn(__fmul_rn(hi - lo, u01(r.x)), lo);
  const float vy = io.inject_push_vel ? io.inject_push_vel[2 * n + 1] : __fadd_rn(__fmul_rn(hi - lo, u01(r.y)), lo);
  if (!reset) {
    float* rs = io.root_state + n * 13 + 7;
    rs[0] = c.push_fixed ? rs[0] + vx : vx; rs[1] = c.push_fixed ? rs[1] + vy : vy;
  }
  if (io.push_vel) { io.push_vel[2 * n] = vx; io.push_vel[2 * n + 1] = vy; }
  return (int)__fdiv_rn((float)iv, c.dt) << 16;  // counter 0 | (self.push_interval_s / self.dt).int()
}
// returns the env's new push word; stamps refresh_buf[n] when the env is pushed (the reset path stamps the envs it resets)
__device__ __forceinline__ int push_step(const PbhcTrackCfg& c, const PbhcTrackIO& io, int64_t n, int word, bool reset, uint64_t seed, uint32_t step) {
  if (!c.push_robots) return word;
  const bool pushed = ((word + 1) & 0xffff) == (word >> 16);
  if (pushed) {
    word = push_cold(c, io, n, reset, seed, step);
    if (io.refresh_buf) io.refresh_buf[n] = (int64_t)step + 1;
  } else word = (word & 0xffff0000) | ((word + 1) & 0xffff);
  return word;
}

// calc_heading_quat_inv (rotations.py:256-268, 296-306) = quat_from_angle_axis(-heading, z) (rotations.py:138-146)
__device__ __forceinline__ Q4 heading_quat_inv(Q4 q) {
  const V3 d = quat_rotate(q, {1.0f, 0.0f, 0.0f});
  const float heading = atan2f(d.y, d.x);
  const float theta = (-heading) / 2.0f;
  // normalize(axis) with axis = (0,0,1): x / norm.clamp(1e-9) = (0,0,1); xyz = axis * sin(theta); w = cos(theta)
  const float sz = 1.0f * sinf(theta), w = cosf(theta);
  const float n = fmaxf(norm4(0.0f * sinf(theta), 0.0f * sinf(theta), sz, w), 1e-9f);
  return {0.0f / n, 0.0f / n, sz / n, w / n};
}

// The same value without atan2f / sinf / cosf: with (c, s) = (cos, sin) of the heading, cos(h / 2) = sqrt((1 + c) / 2) >= 0 and
// sin(h / 2) = s / (2 cos(h / 2)) (or the mirrored pair where c < 0, which keeps the subtraction well conditioned).  Within
// 1.2e-7 of the exact value over all headings, like the sequence above (its normalisation divides by 1 +- 1 ulp).
__device__ __forceinline__ Q4 heading_quat_inv_fast(Q4 q) {
  const V3 d = quat_rotate(q, {1.0f, 0.0f, 0.0f});
  const float r = sqrtf(d.x * d.x + d.y * d.y);
  if (!(r > 1e-30f)) return {0.0f, 0.0f, 0.0f, 1.0f};  // atan2f(0, 0) = 0
  const float c = d.x / r, s = d.y / r;
  float sh, ch;
  if (c >= 0.0f) { ch = sqrtf((1.0f + c) * 0.5f); sh = s / (2.0f * ch); }
  else { const float a = sqrtf((1.0f - c) * 0.5f); sh = copysignf(a, s); ch = fabsf(s) / (2.0f * a); }
  return {0.0f, 0.0f, -sh, ch};
}

// yaw_quat (common.cuh) the same way: (0, 0, sin(yaw / 2), cos(yaw / 2)) from (cos, sin) of the yaw, no atan2f / sinf / cosf
__device__ __forceinline__ Q4 yaw_quat_fast(Q4 q) {
  const float dy = 2.0f * (q.w * q.z + q.x * q.y), dx = 1.0f - 2.0f * (q.y * q.y + q.z * q.z);
  const float r = sqrtf(dx * dx + dy * dy);
  if (!(r > 1e-30f)) return {0.0f, 0.0f, 0.0f, 1.0f};
  const float c = dx / r, s = dy / r;
  float sh, ch;
  if (c >= 0.0f) { ch = sqrtf((1.0f + c) * 0.5f); sh = s / (2.0f * ch); }
  else { const float a = sqrtf((1.0f - c) * 0.5f); sh = copysignf(a, s); ch = fabsf(s) / (2.0f * a); }
  return {0.0f, 0.0f, sh, ch};
}

// my_quat_rotate(q, v) for q = (0, 0, z, w) - the heading quaternion.  Same expression tree as quat_rotate_impl with
// the terms that multiply q.x = q.y = 0 dropped: they are exact zeros, so only the sign of a zero result can differ.
__device__ __forceinline__ V3 quat_rotate_z(float qz, float qw, V3 v) {
  const float s = 2.0f * (qw * qw) - 1.0f;
  const V3 a = {v.x * s, v.y * s, v.z * s};
  const V3 b = {(0.0f - qz * v.y) * qw * 2.0f, (qz * v.x) * qw * 2.0f, 0.0f};
  const float d = qz * v.z;
  return {a.x + b.x, a.y + b.y, a.z + qz * d * 2.0f};
}

}  // namespace pbhc
