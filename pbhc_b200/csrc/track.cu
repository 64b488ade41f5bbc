// track.cu - K3: the fused motion-tracking env step.  One warp owns one env per iteration of a
// persistent grid-stride loop: lanes run over bodies (J <= 64), dofs (D <= 32) and flat record /
// observation elements; everything per-env is staged once in per-warp shared memory (sim body state,
// the lerped/slerped reference record, the scaled observation features + history) so that every global
// access is a coalesced run of contiguous floats.  Cross-env sums (adaptive-sigma errors, log means,
// reset statistics) are accumulated per block and reduced deterministically - no atomics, no host sync.
// Reference: legged_robot_base.py:297-338, 346-380, 408-489, 715-793, 944-1087; motion_tracking.py:330-358,
// 548-747, 1030-1328; helpers.py:128-152; history_handler.py:33-48 (paths under humanoidverse/envs|utils).
#include "common.cuh"
#include "table.cuh"

namespace pbhc {

constexpr int kTW = 8;              // warps (envs in flight) per block
constexpr int kTrackMaxBlocks = 148 * 4;

struct TrackParams {
  PbhcTrackCfg c;
  PbhcTableDesc tab;
  PbhcTrackIO io;
  int64_t n;
  int sm_body, sm_ref, sm_cf, sm_feat, sm_per_warp;  // per-warp shared-memory offsets (floats)
  int src_off[PBHC_NUM_OBS_SRC + 1];                 // feature offsets of the observation sources; [last] = history
};

enum ObsSrc { S_BASE_LIN_VEL, S_BASE_ANG_VEL, S_PROJ_GRAV, S_DOF_POS, S_DOF_VEL, S_ACTIONS, S_PHASE, S_DIF_LOCAL,
              S_LOCAL_REF, S_DR_COM, S_DR_MASS, S_DR_KP, S_DR_KD, S_DR_FRICTION, S_DR_DELAY, S_LOCAL_REF_VEL,
              S_GLOBAL_REF_VEL, S_DIF_JOINT, S_DIF_JOINT_VEL, S_BASE_POS_Z };
enum RewardId { R_CONTACT_MASK, R_MAX_JOINT_POS, R_BODY_POS_EXT, R_VR3, R_FEET_POS, R_BODY_ROT, R_BODY_ANG_VEL,
                R_BODY_VEL, R_JOINT_POS, R_JOINT_VEL, R_TORQUES, R_ACTION_RATE, R_FEET_AIR_TIME, R_FEET_FORCES,
                R_STUMBLE, R_SLIPPAGE, R_LIM_DOF_POS, R_LIM_DOF_VEL, R_LIM_TORQUE, R_COLLISION, R_DOF_VEL, R_DOF_ACC,
                R_FOOT_SLIP, R_ALIVE, R_ORIENTATION, R_CONTACT_MASK_V2 };
enum SigmaId { G_MAX_JOINT, G_UPPER, G_LOWER, G_VR3, G_FEET, G_ROT, G_VEL, G_ANG_VEL, G_JOINT_POS, G_JOINT_VEL };

__host__ __device__ inline int src_dim(int k, int D, int J, int L) {
  switch (k) {
    case S_DOF_POS: case S_DOF_VEL: case S_ACTIONS: case S_DR_KP: case S_DR_KD: case S_DIF_JOINT: case S_DIF_JOINT_VEL: return D;
    case S_DIF_LOCAL: case S_LOCAL_REF: case S_LOCAL_REF_VEL: case S_GLOBAL_REF_VEL: return 3 * J;
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

// Philox4x32-10 (counter-based; one call per (env, step, stream))
__device__ __forceinline__ uint4 philox(uint4 ctr, uint2 key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u; key.y += 0xBB67AE85u;
  }
  return ctr;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x & 0xFFFFFFu) * (1.0f / 16777216.0f); }
struct Rng {
  uint2 key; uint32_t env_lo, env_hi, step;
  __device__ __forceinline__ uint4 draw(uint32_t stream) const { return philox(make_uint4(env_lo, env_hi, step, stream), key); }
  __device__ __forceinline__ float uniform(uint32_t stream) const { return u01(draw(stream).x); }
  __device__ __forceinline__ float normal(uint32_t stream) const {
    const uint4 r = draw(stream);
    const float u1 = fmaxf(u01(r.x), 1.0f / 16777216.0f), u2 = u01(r.y);
    return sqrtf(-2.0f * logf(u1)) * cosf(6.283185307179586f * u2);
  }
};

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

struct Sums {  // per-lane partial sums over bodies, reduced across the warp afterwards
  float up_sq, lo_sq, vr_sq, feet_sq, rot_sq, vel_sq, ang_sq, up_n, lo_n, vr_n, collide;
  int far;
};

__global__ void __launch_bounds__(kTW * 32) track_kernel(const __grid_constant__ TrackParams P) {
  extern __shared__ float smem[];
  __shared__ double sm_scal[PBHC_SCALARS];
  __shared__ double sm_acc[kTW][PBHC_ACC];
  const PbhcTrackCfg& c = P.c;
  const PbhcTrackIO& io = P.io;
  const PbhcTableDesc& tab = P.tab;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int J = c.J, Jb = c.Jb, D = c.D, L = c.L, Cm = tab.contact_size;
  float* sm_body = smem + warp * P.sm_per_warp + P.sm_body;
  float* sm_ref = smem + warp * P.sm_per_warp + P.sm_ref;
  float* sm_cf = smem + warp * P.sm_per_warp + P.sm_cf;
  float* feat = smem + warp * P.sm_per_warp + P.sm_feat;
  for (int i = threadIdx.x; i < PBHC_SCALARS; i += blockDim.x) sm_scal[i] = io.scalars[i];
  for (int i = threadIdx.x; i < kTW * PBHC_ACC; i += blockDim.x) (&sm_acc[0][0])[i] = 0.0;
  __syncthreads();
  double* acc = sm_acc[warp];
  const float far_thr = (float)sm_scal[PBHC_S_FAR_THR];
  const float penalty_scale = (float)sm_scal[PBHC_S_PENALTY];
  const uint64_t push_count = (uint64_t)sm_scal[PBHC_S_PUSH_COUNT];
  const int e_pos = 4 * J, e_vel = 7 * J, e_ang = 10 * J, e_dp = 13 * J, e_dv = e_dp + D, e_c = e_dv + D, e_end = e_c + Cm;
  const int hist_off = P.src_off[PBHC_NUM_OBS_SRC];

  for (int64_t n = (int64_t)blockIdx.x * kTW + warp; n < P.n; n += (int64_t)gridDim.x * kTW) {
    __syncwarp();
    // ---- A. per-env scalars ------------------------------------------------------------------
    int64_t ep = io.episode_length_buf[n] + 1;  // legged_robot_base.py:299
    const int64_t last_ep = ep;
    float* envf = io.env_f32 + n * PBHC_ENVF_STRIDE;
    float start = envf[0], mlen = envf[1];
    float air[PBHC_MAX_FEET], last_c[PBHC_MAX_FEET];
#pragma unroll
    for (int f = 0; f < 2; ++f) { air[f] = envf[2 + f]; last_c[f] = envf[4 + f]; }
    float etr = envf[6];
    const int clip = resolve_clip(tab, io.clip_of_motion, io.motion_ids[n]);
    const float t = (float)(ep + 1) * c.dt + start;  // kick_motion_res, motion_tracking.py:554
    const FrameBlend fb = frame_blend(tab, clip, t);
    const float b = fb.blend, omb = 1.0f - b;
    const float* r0 = tab.records + fb.f0 * tab.rec_stride;
    const float* r1 = tab.records + fb.f1 * tab.rec_stride;
    const float org[3] = {io.env_origins[n * 3], io.env_origins[n * 3 + 1], io.env_origins[n * 3 + 2]};
    // ---- C. stage sim body state, contact forces, reference record, history ------------------
    {
      const float* g = io.body_state + n * io.body_env_stride;
      for (int i = lane; i < Jb * 13; i += 32) sm_body[i] = g[i];
      const float* gc = io.contact_forces + n * (Jb * 3);
      for (int i = lane; i < Jb * 3; i += 32) sm_cf[i] = gc[i];
      for (int j = lane; j < J; j += 32) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(r0) + j);
        const float4 d = __ldg(reinterpret_cast<const float4*>(r1) + j);
        const Q4 s = slerp({a.x, a.y, a.z, a.w}, {d.x, d.y, d.z, d.w}, b);
        sm_ref[4 * j] = s.x; sm_ref[4 * j + 1] = s.y; sm_ref[4 * j + 2] = s.z; sm_ref[4 * j + 3] = s.w;
      }
#pragma unroll 4
      for (int e = e_pos + lane; e < e_end; e += 32) {
        float v = omb * __ldg(r0 + e) + b * __ldg(r1 + e);
        if (e < e_vel) v = v + org[(e - e_pos) % 3];
        sm_ref[e] = v;
      }
      // history, rotated into logical (newest-first) order
      const float* gh = io.hist + n * c.hist_row;
      int base = 0;
      for (int k = 0; k < c.n_hist_keys; ++k) {
        const int len = c.hist_len[k], dim = c.hist_dim[k];
        const int head = (int)((len - (push_count % len)) % len);
        for (int i = lane; i < len * dim; i += 32) {
          const int s = i / dim, cc = i - s * dim;
          int ps = head + s; if (ps >= len) ps -= len;
          feat[hist_off + base + i] = gh[base + ps * dim + cc];
        }
        base += len * dim;
      }
    }
    __syncwarp();
    // ---- D. root quantities (legged_robot_base.py:346-353) -----------------------------------
    float* root = io.root_state + n * 13;
    const V3 rp = {root[0], root[1], root[2]};
    const Q4 rq = {root[3], root[4], root[5], root[6]};
    const V3 blv = quat_rotate_inverse(rq, {root[7], root[8], root[9]});
    const V3 bav = quat_rotate_inverse(rq, {root[10], root[11], root[12]});
    const V3 pg = quat_rotate_inverse(rq, {0.0f, 0.0f, -1.0f});
    const Q4 hinv = heading_quat_inv(rq);
    // ---- E. bodies: extended state, differences, heading-frame features ----------------------
    Sums s = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    uint64_t feet_mask = 0;
    for (int f = 0; f < c.n_feet; ++f) feet_mask |= 1ull << c.feet_ids[f];
    for (int j = lane; j < J; j += 32) {
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
      const float* rr = sm_ref + 4 * j;
      const float* rpz = sm_ref + e_pos + 3 * j;
      const float* rv = sm_ref + e_vel + 3 * j;
      const float* rw = sm_ref + e_ang + 3 * j;
      const V3 dp = {rpz[0] - p.x, rpz[1] - p.y, rpz[2] - p.z};
      const float dq0 = rr[0] - q.x, dq1 = rr[1] - q.y, dq2 = rr[2] - q.z, dq3 = rr[3] - q.w;
      const V3 dv = {rv[0] - v.x, rv[1] - v.y, rv[2] - v.z};
      const V3 dw = {rw[0] - w.x, rw[1] - w.y, rw[2] - w.z};
      const float m_pos = ((dp.x * dp.x + dp.y * dp.y) + dp.z * dp.z) / 3.0f;
      const float nrm = norm3(dp.x, dp.y, dp.z);
      const uint64_t bit = 1ull << j;
      if (c.upper_mask & bit) { s.up_sq += m_pos; s.up_n += nrm; }
      if (c.lower_mask & bit) { s.lo_sq += m_pos; s.lo_n += nrm; }
      if (c.vr3_mask & bit) { s.vr_sq += m_pos; s.vr_n += nrm; }
      if (feet_mask & bit) s.feet_sq += m_pos;
      s.rot_sq += (((dq0 * dq0 + dq1 * dq1) + dq2 * dq2) + dq3 * dq3) / 4.0f;
      s.vel_sq += ((dv.x * dv.x + dv.y * dv.y) + dv.z * dv.z) / 3.0f;
      s.ang_sq += ((dw.x * dw.x + dw.y * dw.y) + dw.z * dw.z) / 3.0f;
      if (nrm > far_thr) s.far = 1;
      if (j < Jb && (c.penalised_mask & bit)) {
        const float* fc = sm_cf + 3 * j;
        if (norm3(fc[0], fc[1], fc[2]) > 0.1f) s.collide += 1.0f;
      }
      const V3 dl = quat_rotate(hinv, dp);  // motion_tracking.py:694-697
      float* f0 = feat + P.src_off[S_DIF_LOCAL] + 3 * j;
      const float sc0 = c.obs_scale[S_DIF_LOCAL];
      f0[0] = dl.x * sc0; f0[1] = dl.y * sc0; f0[2] = dl.z * sc0;
      const V3 lr = quat_rotate(hinv, {rpz[0] - rp.x, rpz[1] - rp.y, rpz[2] - rp.z});  // :705-707
      float* f1 = feat + P.src_off[S_LOCAL_REF] + 3 * j;
      const float sc1 = c.obs_scale[S_LOCAL_REF];
      f1[0] = lr.x * sc1; f1[1] = lr.y * sc1; f1[2] = lr.z * sc1;
      if (c.src_used_mask & ((1u << S_LOCAL_REF_VEL) | (1u << S_GLOBAL_REF_VEL))) {
        const V3 lv = quat_rotate(hinv, {rv[0], rv[1], rv[2]});
        float* f2 = feat + P.src_off[S_LOCAL_REF_VEL] + 3 * j;
        float* f3 = feat + P.src_off[S_GLOBAL_REF_VEL] + 3 * j;
        const float sc2 = c.obs_scale[S_LOCAL_REF_VEL], sc3 = c.obs_scale[S_GLOBAL_REF_VEL];
        f2[0] = lv.x * sc2; f2[1] = lv.y * sc2; f2[2] = lv.z * sc2;
        f3[0] = rv[0] * sc3; f3[1] = rv[1] * sc3; f3[2] = rv[2] * sc3;
      }
    }
    s.up_sq = warp_sum(s.up_sq); s.lo_sq = warp_sum(s.lo_sq); s.vr_sq = warp_sum(s.vr_sq); s.feet_sq = warp_sum(s.feet_sq);
    s.rot_sq = warp_sum(s.rot_sq); s.vel_sq = warp_sum(s.vel_sq); s.ang_sq = warp_sum(s.ang_sq);
    s.up_n = warp_sum(s.up_n); s.lo_n = warp_sum(s.lo_n); s.vr_n = warp_sum(s.vr_n); s.collide = warp_sum(s.collide);
    const bool far = __any_sync(0xffffffffu, s.far);
    // ---- F. dofs (lane = dof) ---------------------------------------------------------------------
    const bool ld = lane < D;
    float dpos = 0.f, dvel = 0.f, act = 0.f, tq = 0.f, lact = 0.f, ldv = 0.f, ddef = 0.f;
    if (ld) {
      const float2 ds = reinterpret_cast<const float2*>(io.dof_state)[n * D + lane];
      dpos = ds.x; dvel = ds.y;
      act = io.actions[n * D + lane]; tq = io.torques[n * D + lane];
      lact = io.last_actions[n * D + lane]; ldv = io.last_dof_vel[n * D + lane];
      ddef = io.default_dof_pos[n * io.default_dof_stride + lane];
    }
    const float dja = ld ? sm_ref[e_dp + lane] - dpos : 0.f;   // motion_tracking.py:661-665
    const float djv = ld ? sm_ref[e_dv + lane] - dvel : 0.f;
    const float max_dja = warp_max(fabsf(dja));
    const float sum_dja2 = warp_sum(dja * dja);
    const float sum_djv2 = warp_sum(djv * djv);
    const float sum_tq2 = warp_sum(tq * tq);
    const float dact = lact - act;
    const float sum_dact2 = warp_sum(dact * dact);
    const float sum_dv2 = warp_sum(dvel * dvel);
    const float dacc = (ldv - dvel) / c.dt;
    const float sum_acc2 = warp_sum(ld ? dacc * dacc : 0.f);
    float lim_pos = 0.f, lim_vel = 0.f, lim_tq = 0.f;
    if (ld) {
      float lo = c.soft_lo[lane], hi = c.soft_hi[lane];
      if (c.cur_dof_pos.enable) {  // legged_robot_base.py:977-981
        const float m = (c.hard_lo[lane] + c.hard_hi[lane]) / 2.0f, r = c.hard_hi[lane] - c.hard_lo[lane];
        const float cv = (float)sm_scal[PBHC_S_SOFT_DOF_POS];
        lo = m - 0.5f * r * cv; hi = m + 0.5f * r * cv;
      }
      lim_pos = -fminf(dpos - lo, 0.0f) + fmaxf(dpos - hi, 0.0f);
      const float vs = c.cur_dof_vel.enable ? (float)sm_scal[PBHC_S_SOFT_DOF_VEL] : c.soft_dof_vel_limit;
      lim_vel = fminf(fmaxf(fabsf(dvel) - c.vel_limit[lane] * vs, 0.0f), 1.0f);
      if (c.cur_torque.enable) lim_tq = fminf(fmaxf(fabsf(tq) - c.torque_limit[lane] * (float)sm_scal[PBHC_S_SOFT_TORQUE], 0.0f), 1.0f);
      else lim_tq = fmaxf(fabsf(tq) - c.torque_limit[lane] * c.soft_torque_limit, 0.0f);
    }
    lim_pos = warp_sum(lim_pos); lim_vel = warp_sum(lim_vel); lim_tq = warp_sum(lim_tq);
    // ---- G. feet (uniform across lanes) -------------------------------------------------------------
    float contacts[PBHC_MAX_FEET], filt[PBHC_MAX_FEET];
    float r_air = 0.f, r_force = 0.f, r_slip = 0.f, r_fslip = 0.f, cm_err = 0.f;
    bool stumble = false;
    for (int f = 0; f < c.n_feet; ++f) {
      const float* fc = sm_cf + 3 * c.feet_ids[f];
      const float* bs = sm_body + 13 * c.feet_ids[f];
      const float fn = norm3(fc[0], fc[1], fc[2]);
      contacts[f] = fn > 1.0f ? 1.0f : 0.0f;                                  // legged_robot_base.py:352-353
      filt[f] = (contacts[f] != 0.0f || last_c[f] != 0.0f) ? 1.0f : 0.0f;
      if (Cm) cm_err += fabsf(filt[f] - sm_ref[e_c + f]);
      // feet_air_time, motion_tracking.py:1307-1319 (uses last_contacts of the previous step)
      const bool cz = fc[2] > 1.0f;
      const bool fz = cz || last_c[f] != 0.0f;
      const float first = (air[f] > 0.0f && fz) ? 1.0f : 0.0f;
      air[f] += c.dt;
      r_air += (air[f] - c.desired_feet_air_time) * first;
      air[f] *= fz ? 0.0f : 1.0f;
      r_force += fmaxf(fn - c.max_contact_force, 0.0f);                        // :1321-1323
      if (norm2(fc[0], fc[1]) > 5.0f * fabsf(fc[2])) stumble = true;           // :1325-1328
      r_slip += norm3(bs[7], bs[8], bs[9]) * (fn > 1.0f ? 1.0f : 0.0f);        // legged_robot_base.py:1004-1010
      r_fslip += (fn > 1.0f ? 1.0f : 0.0f) * norm2(bs[7], bs[8]);              // :1012-1017
    }
    // ---- H. termination (legged_robot_base.py:408-489, motion_tracking.py:330-358) ---------------------
    const bool t_grav = c.term_gravity && norm2(pg.x, pg.y) > c.term_gravity_thr;
    const bool t_far = c.term_motion_far && far;
    const bool t_out = ep > c.max_episode_length;
    const bool t_end = c.term_motion_end && ((float)ep * c.dt + start > mlen);
    const bool time_out = t_out || t_end;
    const bool reset = t_grav || t_far || time_out;
    // ---- I. rewards (lane = reward term) ----------------------------------------------------------------
    const float e_up = s.up_sq / (float)c.n_upper, e_lo = s.lo_sq / (float)c.n_lower, e_vr = s.vr_sq / (float)c.n_vr3;
    const float e_feet = s.feet_sq / (float)c.n_feet, e_rot = s.rot_sq / (float)J, e_vel2 = s.vel_sq / (float)J, e_ang2 = s.ang_sq / (float)J;
    const float e_jp = sum_dja2 / (float)D, e_jv = sum_djv2 / (float)D;
    float rew = 0.0f;
    if (lane < c.n_rewards) {
      const double* sg = sm_scal + PBHC_S_SIGMA;
      float r = 0.0f;
      switch (c.reward_id[lane]) {
        case R_CONTACT_MASK: r = 1.0f - cm_err / (float)c.n_feet; break;
        case R_CONTACT_MASK_V2: r = 0.5f - cm_err / (float)c.n_feet; break;
        case R_MAX_JOINT_POS: r = expf(-max_dja / (float)sg[G_MAX_JOINT]); break;
        case R_BODY_POS_EXT: r = expf(-e_lo / (float)sg[G_LOWER]) * c.w_lower + expf(-e_up / (float)sg[G_UPPER]) * c.w_upper; break;
        case R_VR3: r = expf(-e_vr / (float)sg[G_VR3]); break;
        case R_FEET_POS: r = expf(-e_feet / (float)sg[G_FEET]); break;
        case R_BODY_ROT: r = expf(-e_rot / (float)sg[G_ROT]); break;
        case R_BODY_ANG_VEL: r = expf(-e_ang2 / (float)sg[G_ANG_VEL]); break;
        case R_BODY_VEL: r = expf(-e_vel2 / (float)sg[G_VEL]); break;
        case R_JOINT_POS: r = expf(-e_jp / (float)sg[G_JOINT_POS]); break;
        case R_JOINT_VEL: r = expf(-e_jv / (float)sg[G_JOINT_VEL]); break;
        case R_TORQUES: r = sum_tq2; break;
        case R_ACTION_RATE: r = sum_dact2; break;
        case R_DOF_VEL: r = sum_dv2; break;
        case R_DOF_ACC: r = sum_acc2; break;
        case R_ORIENTATION: r = pg.x * pg.x + pg.y * pg.y; break;
        case R_ALIVE: r = 1.0f; break;
        case R_FEET_AIR_TIME: r = r_air; break;
        case R_FEET_FORCES: r = r_force; break;
        case R_STUMBLE: r = stumble ? 1.0f : 0.0f; break;
        case R_SLIPPAGE: r = r_slip; break;
        case R_FOOT_SLIP: r = r_fslip; break;
        case R_LIM_DOF_POS: r = lim_pos; break;
        case R_LIM_DOF_VEL: r = lim_vel; break;
        case R_LIM_TORQUE: r = lim_tq; break;
        case R_COLLISION: r = s.collide; break;
        default: break;
      }
      rew = r * c.reward_scale[lane];
      if (c.reward_is_penalty[lane]) rew *= penalty_scale;
    }
    const float term_rew = c.has_termination ? ((reset && !time_out) ? 1.0f : 0.0f) * c.termination_scale : 0.0f;
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
    if (lane == 0) {
      acc[PBHC_A_SIGMA_ERR + G_MAX_JOINT] += max_dja; acc[PBHC_A_SIGMA_ERR + G_UPPER] += e_up;
      acc[PBHC_A_SIGMA_ERR + G_LOWER] += e_lo; acc[PBHC_A_SIGMA_ERR + G_VR3] += e_vr;
      acc[PBHC_A_SIGMA_ERR + G_FEET] += e_feet; acc[PBHC_A_SIGMA_ERR + G_ROT] += e_rot;
      acc[PBHC_A_SIGMA_ERR + G_VEL] += e_vel2; acc[PBHC_A_SIGMA_ERR + G_ANG_VEL] += e_ang2;
      acc[PBHC_A_SIGMA_ERR + G_JOINT_POS] += e_jp; acc[PBHC_A_SIGMA_ERR + G_JOINT_VEL] += e_jv;
      acc[PBHC_A_UPPER_NORM] += s.up_n; acc[PBHC_A_LOWER_NORM] += s.lo_n; acc[PBHC_A_VR3_NORM] += s.vr_n;
      acc[PBHC_A_JOINT_NORM] += sqrtf(sum_dja2);
      acc[PBHC_A_TERM_GRAVITY] += t_grav; acc[PBHC_A_TERM_FAR] += t_far;
      acc[PBHC_A_TERM_TIMEOUT] += time_out; acc[PBHC_A_TERM_MOTION_END] += t_end;
      acc[PBHC_A_NUM_ENVS] += 1.0;
    }
    // ---- J. reset (legged_robot_base.py:491-516, 647-686, 587-635; motion_tracking.py:274-296, 369-384, 445-543) --
    float kp = 1.f, kd = 1.f, delay_f;
    if (ld) { kp = io.kp_scale[n * D + lane]; kd = io.kd_scale[n * D + lane]; }
    int64_t delay = io.action_delay_idx[n];
    const float phase = t / tab.clip_length[clip];  // motion_tracking.py:738-739 (stale for envs that reset, like the reference)
    if (reset) {
      const Rng rng = {make_uint2((uint32_t)c.seed, (uint32_t)(c.seed >> 32)), (uint32_t)n, (uint32_t)(n >> 32), (uint32_t)push_count};
      act = 0.0f;
      if (ld) io.actions[n * D + lane] = 0.0f;
      for (int f = 0; f < c.n_feet; ++f) { air[f] = 0.0f; contacts[f] = 0.0f; filt[f] = 0.0f; }
      for (int i = lane; i < c.hist_row; i += 32) { feat[hist_off + i] = 0.0f; io.hist[n * c.hist_row + i] = 0.0f; }
      if (lane == 0) { acc[PBHC_A_RESET] += 1.0; acc[PBHC_A_RESET_EPLEN] += (double)last_ep; }
      ep = 0;
      if (ld) {  // _episodic_domain_randomization
        if (c.randomize_pd_gain) {
          kp = io.inject_kp ? io.inject_kp[n * D + lane] : (c.kp_hi - c.kp_lo) * rng.uniform(16 + lane) + c.kp_lo;
          kd = io.inject_kd ? io.inject_kd[n * D + lane] : (c.kd_hi - c.kd_lo) * rng.uniform(48 + lane) + c.kd_lo;
          io.kp_scale[n * D + lane] = kp; io.kd_scale[n * D + lane] = kd;
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
      start = io.inject_start_times ? io.inject_start_times[n] : rng.uniform(0) * mlen;
      // _reset_dofs / _reset_root_states from the motion state at the new time
      const FrameBlend f2 = frame_blend(tab, clip, (float)(ep + 1) * c.dt + start);
      const float b2 = f2.blend, omb2 = 1.0f - b2;
      const float* q0 = tab.records + f2.f0 * tab.rec_stride;
      const float* q1 = tab.records + f2.f1 * tab.rec_stride;
      if (ld) {
        dpos = omb2 * q0[e_dp + lane] + b2 * q1[e_dp + lane];
        dvel = omb2 * q0[e_dv + lane] + b2 * q1[e_dv + lane];
        if (c.reset_noise[0] != 0.0f) dpos = dpos + rng.normal(144 + lane) * c.reset_noise[0];
        if (c.reset_noise[1] != 0.0f) dvel = dvel + rng.normal(176 + lane) * c.reset_noise[1];
        reinterpret_cast<float2*>(io.dof_state)[n * D + lane] = make_float2(dpos, dvel);
      }
      if (lane < 13) {
        float v;
        if (lane >= 3 && lane < 7) {
          const float4 a = reinterpret_cast<const float4*>(q0)[0], d4 = reinterpret_cast<const float4*>(q1)[0];
          const Q4 rr = slerp({a.x, a.y, a.z, a.w}, {d4.x, d4.y, d4.z, d4.w}, b2);
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
          const int e = lane < 3 ? e_pos + lane : (lane < 10 ? e_vel + (lane - 7) : e_ang + (lane - 10));
          v = omb2 * q0[e] + b2 * q1[e];
          if (lane < 3) v = v + org[lane];
          const float ns = lane < 3 ? c.reset_noise[2] : (lane < 10 ? c.reset_noise[4] : c.reset_noise[5]);
          if (ns != 0.0f) v = v + rng.normal(8 + lane) * ns;
        }
        root[lane] = v;
      }
    }
    delay_f = (float)delay;
    // ---- K. observation features, groups, history push, state write-back -----------------------------------
    {
      const float* sc = c.obs_scale;
      if (lane < 3) {
        const float l3 = lane == 0 ? blv.x : (lane == 1 ? blv.y : blv.z);
        const float a3 = lane == 0 ? bav.x : (lane == 1 ? bav.y : bav.z);
        const float g3 = lane == 0 ? pg.x : (lane == 1 ? pg.y : pg.z);
        feat[P.src_off[S_BASE_LIN_VEL] + lane] = l3 * sc[S_BASE_LIN_VEL];
        feat[P.src_off[S_BASE_ANG_VEL] + lane] = a3 * sc[S_BASE_ANG_VEL];
        feat[P.src_off[S_PROJ_GRAV] + lane] = g3 * sc[S_PROJ_GRAV];
        feat[P.src_off[S_DR_COM] + lane] = io.base_com_bias[n * 3 + lane] * sc[S_DR_COM];
      }
      if (ld) {
        feat[P.src_off[S_DOF_POS] + lane] = (dpos - ddef) * sc[S_DOF_POS];
        feat[P.src_off[S_DOF_VEL] + lane] = dvel * sc[S_DOF_VEL];
        feat[P.src_off[S_ACTIONS] + lane] = act * sc[S_ACTIONS];
        feat[P.src_off[S_DR_KP] + lane] = kp * sc[S_DR_KP];
        feat[P.src_off[S_DR_KD] + lane] = kd * sc[S_DR_KD];
        feat[P.src_off[S_DIF_JOINT] + lane] = dja * sc[S_DIF_JOINT];
        feat[P.src_off[S_DIF_JOINT_VEL] + lane] = djv * sc[S_DIF_JOINT_VEL];
      }
      for (int i = lane; i < L; i += 32) feat[P.src_off[S_DR_MASS] + i] = io.link_mass_scale[n * L + i] * sc[S_DR_MASS];
      if (lane == 0) {
        feat[P.src_off[S_PHASE]] = phase * sc[S_PHASE];
        feat[P.src_off[S_DR_FRICTION]] = io.friction[n] * sc[S_DR_FRICTION];
        feat[P.src_off[S_DR_DELAY]] = delay_f * sc[S_DR_DELAY];
        feat[P.src_off[S_BASE_POS_Z]] = rp.z * sc[S_BASE_POS_Z];
      }
    }
    __syncwarp();
    for (int g = 0; g < c.n_groups; ++g) {
      const int dim = c.group_dim[g];
      const int32_t* map = c.group_map[g];
      const float* mul = c.group_mul[g];
      float* o = io.obs[g] + n * dim;
      for (int i = lane; i < dim; i += 32) {
        const float v = feat[__ldg(map + i)] * __ldg(mul + i);
        o[i] = fminf(fmaxf(v, -c.clip_obs), c.clip_obs);
      }
    }
    {  // HistoryHandler.add (history_handler.py:40-44) as a ring-buffer push: one slot per key is written
      float* gh = io.hist + n * c.hist_row;
      int base = 0;
      for (int k = 0; k < c.n_hist_keys; ++k) {
        const int len = c.hist_len[k], dim = c.hist_dim[k];
        const int head = (int)((len - ((push_count + 1) % len)) % len);
        const float* src = feat + P.src_off[c.hist_src[k]];
        for (int i = lane; i < dim; i += 32) gh[base + head * dim + i] = src[i];
        base += len * dim;
      }
    }
    if (io.ref_body_pos_extend) for (int i = lane; i < 3 * J; i += 32) io.ref_body_pos_extend[n * (3 * J) + i] = sm_ref[e_pos + i];
    if (io.ref_body_rot_extend) for (int i = lane; i < 4 * J; i += 32) io.ref_body_rot_extend[n * (4 * J) + i] = sm_ref[i];
    if (ld) {  // _post_compute_observations_callback, legged_robot_base.py:398-405
      io.last_actions[n * D + lane] = act;
      io.last_dof_pos[n * D + lane] = dpos;
      io.last_dof_vel[n * D + lane] = dvel;
    }
    if (lane == 0) {
      envf[0] = start; envf[1] = mlen; envf[6] = etr;
      for (int f = 0; f < 2; ++f) { envf[2 + f] = f < c.n_feet ? air[f] : 0.0f; envf[4 + f] = f < c.n_feet ? contacts[f] : 0.0f; }
      io.episode_length_buf[n] = ep;
      io.reset_buf[n] = reset ? 1 : 0;
      io.time_out_buf[n] = time_out ? 1 : 0;
      acc[PBHC_A_ETR_SUM] += (double)etr; acc[PBHC_A_ETR_SQ] += (double)etr * (double)etr;
    }
  }
  __syncthreads();
  if (threadIdx.x < PBHC_ACC) {
    double v = 0.0;
    for (int w = 0; w < kTW; ++w) v += sm_acc[w][threadIdx.x];
    io.partials[(int64_t)blockIdx.x * PBHC_ACC + threadIdx.x] = v;
  }
}

// deterministic reduction of the per-block partials
__global__ void track_reduce_kernel(const double* __restrict__ partials, int n_blocks, double* __restrict__ acc) {
  const int a = threadIdx.x;  // blockDim = PBHC_ACC
  double v = 0.0;
  for (int b = 0; b < n_blocks; ++b) v += partials[(int64_t)b * PBHC_ACC + a];
  acc[a] = v;
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
      else if (c.adaptive_type == 1) S[PBHC_S_SIGMA + k] = (fmin(ema, sig) + ema) / 2.0;
      else S[PBHC_S_SIGMA + k] = fmin(ema, sig);
    }
  }
  L[PBHC_L_UPPER] = acc[PBHC_A_UPPER_NORM] / (n * c.n_upper);
  L[PBHC_L_LOWER] = acc[PBHC_A_LOWER_NORM] / (n * c.n_lower);
  L[PBHC_L_VR3] = acc[PBHC_A_VR3_NORM] / (n * c.n_vr3);
  L[PBHC_L_JOINT] = acc[PBHC_A_JOINT_NORM] / n;
  const double resets = acc[PBHC_A_RESET];
  const float reset_mean = (float)(resets / n) + 1e-15f;
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
    const double mean = acc[PBHC_A_ETR_SUM] / n;
    L[PBHC_L_ETR_MEAN] = mean;
    const double var = n > 1.0 ? (acc[PBHC_A_ETR_SQ] - acc[PBHC_A_ETR_SUM] * mean) / (n - 1.0) : 0.0;
    L[PBHC_L_ETR_STD] = sqrt(fmax(var, 0.0));
  }
  S[PBHC_S_PUSH_COUNT] += 1.0;
}

static int g_track_blocks_per_sm = 0;

}  // namespace pbhc

extern "C" int pbhc_track_max_blocks(void) { return pbhc::kTrackMaxBlocks; }

extern "C" int pbhc_track_step(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, const PbhcTrackIO* io, int64_t n, void* stream) {
  using namespace pbhc;
  if (!cfg || !tab || !io || n < 0) { set_error("pbhc_track_step: null argument"); return PBHC_E_BADARG; }
  const PbhcTrackCfg& c = *cfg;
  if (c.J <= 0 || c.J > 64 || c.Jb > c.J || c.D <= 0 || c.D > PBHC_MAX_DOF || c.n_ext != c.J - c.Jb || c.n_ext > PBHC_MAX_EXT ||
      c.n_feet < 0 || c.n_feet > 2 || c.n_rewards < 0 || c.n_rewards >= 32 || c.n_episode_sums > 32 || c.n_groups > PBHC_MAX_OBS_GROUPS ||
      c.n_hist_keys > PBHC_MAX_HIST_KEYS || c.J != tab->num_joints || c.Jb != tab->num_bodies || c.D != tab->num_dof) {
    set_error("pbhc_track_step: configuration outside the supported envelope (J<=64, D<=32, feet<=2, rewards<32)"); return PBHC_E_RANGE;
  }
  const void* need[] = {io->body_state, io->root_state, io->dof_state, io->contact_forces, io->actions, io->torques, io->env_origins,
                        io->motion_ids, io->default_dof_pos, io->base_com_bias, io->friction, io->episode_length_buf, io->env_f32,
                        io->last_actions, io->last_dof_pos, io->last_dof_vel, io->kp_scale, io->kd_scale, io->rfi_lim_scale, io->rao_scale,
                        io->action_delay_idx, io->episode_sums, io->scalars, io->rew_buf, io->reset_buf, io->time_out_buf,
                        io->last_episode_sums, io->partials, io->acc};
  for (const void* p : need) if (!p) { set_error("pbhc_track_step: a required PbhcTrackIO pointer is NULL"); return PBHC_E_BADARG; }
  if ((c.L > 0 && !io->link_mass_scale) || (c.hist_row > 0 && !io->hist)) { set_error("pbhc_track_step: NULL link_mass_scale / hist"); return PBHC_E_BADARG; }
  for (int g = 0; g < c.n_groups; ++g) if (!io->obs[g] || !c.group_map[g] || !c.group_mul[g]) { set_error("pbhc_track_step: NULL obs group %d", g); return PBHC_E_BADARG; }
  if (((uintptr_t)tab->records & 15) || ((uintptr_t)io->dof_state & 7)) { set_error("pbhc_track_step: records must be 16-B, dof_state 8-B aligned"); return PBHC_E_ALIGN; }
  if (n == 0) return PBHC_OK;
  TrackParams P;
  P.c = c; P.tab = *tab; P.io = *io; P.n = n;
  int off = 0;
  for (int k = 0; k < PBHC_NUM_OBS_SRC; ++k) { P.src_off[k] = off; off += src_dim(k, c.D, c.J, c.L); }
  P.src_off[PBHC_NUM_OBS_SRC] = off;
  const int feat_len = off + c.hist_row;
  auto up4 = [](int v) { return (v + 3) & ~3; };
  P.sm_body = 0;
  P.sm_ref = up4(c.Jb * 13);
  P.sm_cf = P.sm_ref + up4(tab->rec_stride);
  P.sm_feat = P.sm_cf + up4(c.Jb * 3);
  P.sm_per_warp = P.sm_feat + up4(feat_len) + 1;  // +1: odd stride between warps
  const size_t shmem = (size_t)kTW * P.sm_per_warp * sizeof(float);
  if (shmem > 200 * 1024) { set_error("pbhc_track_step: shared memory %zu B too large", shmem); return PBHC_E_RANGE; }
  static thread_local size_t configured = 0;
  if (shmem > configured) {
    if (cudaFuncSetAttribute(track_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_track_step(attr)");
    configured = shmem;
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, track_kernel, kTW * 32, shmem);
    g_track_blocks_per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  }
  int64_t blocks = (n + kTW - 1) / kTW;
  const int64_t cap = 148LL * g_track_blocks_per_sm;
  if (blocks > cap) blocks = cap;
  track_kernel<<<(unsigned)blocks, kTW * 32, shmem, (cudaStream_t)stream>>>(P);
  track_reduce_kernel<<<1, PBHC_ACC, 0, (cudaStream_t)stream>>>(io->partials, (int)blocks, io->acc);
  return check_launch("pbhc_track_step");
}

extern "C" int pbhc_track_finalize(const PbhcTrackCfg* cfg, const double* acc, double* scalars, void* stream) {
  using namespace pbhc;
  if (!cfg || !acc || !scalars) { set_error("pbhc_track_finalize: null argument"); return PBHC_E_BADARG; }
  track_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*cfg, acc, scalars);
  return check_launch("pbhc_track_finalize");
}
