// actuate.cu - the actuation half of the env step: action clip + control-delay queue
// (LeggedRobotBase._pre_physics_step, envs/legged_base_task/legged_robot_base.py:268-286) and the PD torque law with
// torque RFI / RAO randomisation and clipping (_compute_torques, :795-838), which the reference runs control_decimation
// (4) times per env step between physics substeps.  One thread per (env, dof); fp32 operations in the reference's order
// (the file is compiled without FMA contraction), so everything but the RFI random draw is bit-exact.
#include "common.cuh"

namespace pbhc {

__global__ void __launch_bounds__(256) pre_physics_kernel(const PbhcActuationCfg C, const float* __restrict__ actions_in, float* __restrict__ actions,
                                                          float* __restrict__ queue, const int64_t* __restrict__ delay_idx,
                                                          float* __restrict__ after_delay, unsigned long long* __restrict__ clip_count,
                                                          int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool at_limit = false;
  if (i < n * C.D) {
    const int64_t e = i / C.D;
    const int d = (int)(i - e * C.D);
    // torch.clip(actions, -clip, clip): min(max(x, lo), hi), NaN propagates
    const float x = actions_in[i];
    const float a = x != x ? x : fminf(fmaxf(x, -C.action_clip), C.action_clip);
    actions[i] = a;
    at_limit = fabsf(a) == C.action_clip;  // log_dict["action_clip_frac"] (:275-277)
    float out = a;
    if (C.Q > 0) {  // :279-282: queue[:, 1:] = queue[:, :-1]; queue[:, 0] = actions; after_delay = queue[env, delay_idx[env]]
      float* q = queue + (e * C.Q) * C.D + d;
      const int64_t pick = delay_idx[e];
      float prev = a;
      for (int k = 0; k < C.Q; ++k) {
        const float old = q[(int64_t)k * C.D];
        q[(int64_t)k * C.D] = prev;
        if (k == pick) out = prev;
        prev = old;
      }
    }
    after_delay[i] = out;
  }
  const unsigned cnt = __popc(__ballot_sync(0xffffffffu, at_limit));
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(clip_count, (unsigned long long)cnt);  // integer: order-independent
}

__global__ void __launch_bounds__(256) torques_kernel(const PbhcActuationCfg C, const float* __restrict__ actions, const float* __restrict__ dof_state,
                                                      const float* __restrict__ default_dof_pos, int default_stride,
                                                      const float* __restrict__ last_dof_vel, const float* __restrict__ kp_scale,
                                                      const float* __restrict__ kd_scale, const float* __restrict__ rfi_lim_scale,
                                                      const float* __restrict__ rao_scale, const float* __restrict__ rfi_uniform,
                                                      uint32_t call, float* __restrict__ torques, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * C.D) return;
  const int64_t e = i / C.D;
  const int d = (int)(i - e * C.D);
  const float a_s = actions[i] * (C.per_dof_action_scale ? C.action_scales[d] : C.action_scale);
  const float2 pv = __ldg(reinterpret_cast<const float2*>(dof_state) + i);  // isaacgym dof_state [N*D, 2] = (pos, vel)
  float t;
  if (C.control_type == 0) {         // "P" (:811)
    t = (kp_scale[i] * C.p_gains[d]) * ((a_s + default_dof_pos[e * default_stride + d]) - pv.x) - (kd_scale[i] * C.d_gains[d]) * pv.y;
  } else if (C.control_type == 1) {  // "V" (:813)
    t = (kp_scale[i] * C.p_gains[d]) * (a_s - pv.y) - ((kd_scale[i] * C.d_gains[d]) * (pv.y - last_dof_vel[i])) / C.sim_dt;
  } else {                           // "T" (:815)
    t = a_s;
  }
  const float lim = C.torque_limits[d];
  if (C.rfi) {  // :820  torques + (rand*2 - 1) * rfi_lim * _rfi_lim_scale * torque_limits
    float u;
    if (rfi_uniform) u = rfi_uniform[i];
    else {
      const int64_t ge = e + C.env_offset;
      const uint32_t step = call + (C.call_counter ? *C.call_counter : 0u);
      u = u01(philox4x32_10(make_uint4((uint32_t)ge, (uint32_t)(ge >> 32), step, (uint32_t)d), make_uint2((uint32_t)C.seed, (uint32_t)(C.seed >> 32))).x);
    }
    t = t + (((u * 2.0f - 1.0f) * C.rfi_lim) * rfi_lim_scale[i]) * lim;
  }
  if (C.rao) t = t + rao_scale[i] * lim;  // :832
  if (C.clip_torques) t = t != t ? t : fminf(fmaxf(t, -lim), lim);  // :835
  torques[i] = t;
}

__global__ void counter_bump_kernel(uint32_t* counter) { *counter += 1u; }  // (after the torque kernel, same stream: every block read the old value)

static int check_cfg(const PbhcActuationCfg* c, const char* who) {
  if (!c) { set_error("%s: null cfg", who); return PBHC_E_BADARG; }
  if (c->D <= 0 || c->D > PBHC_MAX_DOF || c->Q < 0 || c->Q > 64 || c->control_type < 0 || c->control_type > 2) {
    set_error("%s: D=%d Q=%d control_type=%d out of range", who, c->D, c->Q, c->control_type); return PBHC_E_RANGE;
  }
  return PBHC_OK;
}

}  // namespace pbhc

extern "C" int pbhc_pre_physics(const PbhcActuationCfg* cfg, const float* actions_in, float* actions, float* action_queue,
                                const int64_t* action_delay_idx, float* actions_after_delay, uint64_t* clip_count, int64_t n_envs,
                                void* stream) {
  using namespace pbhc;
  int rc = check_cfg(cfg, "pbhc_pre_physics");
  if (rc) return rc;
  if (!actions_in || !actions || !actions_after_delay || !clip_count || n_envs < 0 || (cfg->Q > 0 && (!action_queue || !action_delay_idx))) {
    set_error("pbhc_pre_physics: null argument"); return PBHC_E_BADARG;
  }
  if (n_envs == 0) return PBHC_OK;
  const int64_t n = n_envs * cfg->D;
  pre_physics_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*cfg, actions_in, actions, action_queue, action_delay_idx,
                                                                                      actions_after_delay, (unsigned long long*)clip_count, n_envs);
  return check_launch("pbhc_pre_physics");
}

extern "C" int pbhc_compute_torques(const PbhcActuationCfg* cfg, const float* actions_after_delay, const float* dof_state,
                                    const float* default_dof_pos, int32_t default_dof_stride, const float* last_dof_vel,
                                    const float* kp_scale, const float* kd_scale, const float* rfi_lim_scale, const float* rao_scale,
                                    const float* rfi_uniform, uint32_t call_index, float* torques, int64_t n_envs, void* stream) {
  using namespace pbhc;
  int rc = check_cfg(cfg, "pbhc_compute_torques");
  if (rc) return rc;
  if (!actions_after_delay || !dof_state || !torques || n_envs < 0 || (cfg->control_type != 2 && (!kp_scale || !kd_scale)) ||
      (cfg->control_type == 0 && !default_dof_pos) || (cfg->control_type == 1 && !last_dof_vel) || (cfg->rfi && !rfi_lim_scale) ||
      (cfg->rao && !rao_scale)) {
    set_error("pbhc_compute_torques: null argument"); return PBHC_E_BADARG;
  }
  if ((uintptr_t)dof_state & 7) { set_error("pbhc_compute_torques: dof_state must be 8-byte aligned"); return PBHC_E_ALIGN; }
  if (n_envs == 0) return PBHC_OK;
  const int64_t n = n_envs * cfg->D;
  torques_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*cfg, actions_after_delay, dof_state, default_dof_pos, default_dof_stride,
                                                                                  last_dof_vel, kp_scale, kd_scale, rfi_lim_scale, rao_scale, rfi_uniform,
                                                                                  call_index, torques, n_envs);
  if (cfg->call_counter && cfg->rfi && !rfi_uniform) counter_bump_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(cfg->call_counter);
  return check_launch("pbhc_compute_torques");
}
