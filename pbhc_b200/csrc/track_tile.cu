// track_tile.cu - K3 for LeggedRobotMotionTracking (config A/B): the fused env step as a TILE kernel.
//
// A block owns a tile of kTE = 8 consecutive envs per iteration of a grid-stride loop.  Every per-env array of the step
// ([N, row] tensors) is one contiguous slab for such a tile, so the whole input of a tile (rigid-body state, contact
// forces, dof state, actions, gains, history ring, episode sums, per-env scalars: ~3.9 KB / env) arrives in shared
// memory by ~20 TMA bulk copies (cp.async.bulk + mbarrier, issued by the lanes of one warp from a descriptor table) and
// the whole output (observation rows assembled in their final layout, rewards, episode sums, state: ~4.6 KB / env)
// leaves by ~10 bulk stores; the next tile's loads are issued as soon as the inputs are dead, before the stores.
// No lane spends an instruction on moving simulator state or observation rows.
//
// Work is mapped as FLAT ITEMS over the tile so that all 32 lanes of a warp are busy: (env, body), (env, dof),
// (env, reward), (env, feature) ...; per-env reductions go through small per-item tables in shared memory that one
// thread per (env, sum) adds up in index order (deterministic, no shuffles, no atomics on floats).  Block-wide phases:
//   1 pre (frame pair / blend, root-frame velocities, gravity, heading) + history ring -> observation rows
//   2 dofs, extended bodies, feet                    3 bodies (+ dof sums)
//   4 body sums, per-env scalars / termination, scalar features      5 rewards + episode sums
//   6 per-env state, resets (warp per env, rare)      7 bulk stores, next tile's loads, cross-env sums (fp64)
//
// Arithmetic: this file is compiled with FMA contraction allowed.  The chains that feed bit-exact results - frame
// indices / blend, motion_end, the gravity and motion_far comparisons - are written with __fmul_rn / __fadd_rn
// (exact.cuh-style helpers below) and reproduce the reference's unfused fp32 sequence; everything else is checked
// against the reference within 1e-5 rel / 1e-6 abs.  See DESIGN.md "arithmetic parity".
// Reference: legged_robot_base.py:297-338, 346-380, 408-489, 715-793, 944-1087; motion_tracking.py:330-358,
// 548-747, 1030-1328; helpers.py:128-152; history_handler.py:33-48 (paths under humanoidverse/envs|utils).
#include "track_common.cuh"

namespace pbhc {

constexpr int kTE = 8;             // envs per tile = warps per block
constexpr int kNT = kTE * 32;
constexpr int kMaxIn = 24, kMaxOut = 16;
constexpr int kBodyRaw = 5;        // per-(env, body) values the body sums are made of
constexpr int kDofRaw = 9;         // per-(env, dof) values the dof sums are made of
constexpr int kRvT = 80;           // per-env scalar table: [0, 57) = enum Rv, then the per-foot values
constexpr int kSrcT = S_BASE_POS_Z + 1;  // observation sources of LeggedRobotMotionTracking
enum RvTile { V_FOOT = 57 /* 2 x 8 */, V_AIR0 = 73, V_AIR1, V_CONT0, V_CONT1 };
enum FootVal { FV_AIR, FV_FORCE, FV_SLIP, FV_FSLIP, FV_CM, FV_STUMBLE, FV_AIRF, FV_CONTACT };
enum BodyRaw { BR_MPOS, BR_NRM, BR_ROT, BR_VEL, BR_ANG };
enum DofRaw { DR_DJA, DR_DJV2, DR_TQ2, DR_DACT2, DR_DV2, DR_ACC2, DR_LPOS, DR_LVEL, DR_LTQ };
constexpr int kBodySums = 10;

struct TileDesc { const char* g; int row_bytes, g_stride, soff, pad; };  // one per-env array: global rows -> slab at byte offset soff

struct TileParams {
  PbhcTrackCfg c;
  PbhcTableDesc tab;
  PbhcTrackIO io;
  int n, n_tiles, n_in, n_out, bulk_ok, in_bytes;
  TileDesc in[kMaxIn], out[kMaxOut];
  int o_bar, o_flags, o_tab, o_hmap, o_hmul, o_rv, o_pre, o_rows;   // byte offsets of the fixed regions
  int i_body, i_cf, i_hist, i_dof, i_act, i_tq, i_lact, i_ldv, i_kp, i_kd, i_es, i_envf, i_root, i_org, i_ep, i_mid, i_com,
      i_fric, i_delay, i_mass;                                       // input slabs
  int q_rew, q_es, q_envf, q_ep, q_reset, q_lact, q_ldpos, q_ldvel;  // output slabs (the obs groups are at o_rows + 4 * row_base[g])
  int o_dpart, o_ext, o_bpart, o_rw;                                 // scratch: aliases of dead input slabs
  int row_base[PBHC_MAX_OBS_GROUPS];                                 // float offset of group g's slab inside the rows region
  int hist_base[PBHC_MAX_HIST_KEYS];
  int e_pos, e_vel, e_ang, e_dp, e_dv, e_c;
  int has_air_time, hist_any_mul, n_hd, inv_j, inv_d, inv_r, inv_f, n_cols, n_feat;
};

// ---- exact (unfused) arithmetic for the bit-exact chains --------------------------------------------------------------
__device__ __forceinline__ float xm(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xa(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xs(float a, float b) { return __fsub_rn(a, b); }

// motion_lib_base.py:503-513 with torch's promotions, unfused (table.cuh frame_blend under -fmad=false)
__device__ __forceinline__ FrameBlend frame_blend_x(const PbhcTableDesc& tab, int clip, float time) {
  const int nf = tab.clip_num_frames[clip];
  const float len = tab.clip_length[clip];
  const float dt = tab.clip_dt[clip];
  const float phase = clip01(__fdiv_rn(time, len));
  if (time < 0.0f) time = 0.0f;
  const int64_t idx0 = (int64_t)xm(phase, (float)(nf - 1));
  const int64_t idx1 = min(idx0 + 1, (int64_t)(nf - 1));
  const float blend = clip01(__fdiv_rn(xs(time, xm((float)idx0, dt)), dt));
  const int64_t start = tab.clip_frame_start[clip];
  return {idx0 + start, idx1 + start, blend};
}
// quat_rotate_impl of common.cuh, unfused
__device__ __forceinline__ V3 quat_rotate_x(Q4 q, V3 v, bool inverse) {
  const float s = xs(xm(2.0f, xm(q.w, q.w)), 1.0f);
  const V3 a = {xm(v.x, s), xm(v.y, s), xm(v.z, s)};
  const V3 cr = {xs(xm(q.y, v.z), xm(q.z, v.y)), xs(xm(q.z, v.x), xm(q.x, v.z)), xs(xm(q.x, v.y), xm(q.y, v.x))};
  const V3 b = {xm(xm(cr.x, q.w), 2.0f), xm(xm(cr.y, q.w), 2.0f), xm(xm(cr.z, q.w), 2.0f)};
  const float d = xa(xa(xm(q.x, v.x), xm(q.y, v.y)), xm(q.z, v.z));
  const V3 c = {xm(xm(q.x, d), 2.0f), xm(xm(q.y, d), 2.0f), xm(xm(q.z, d), 2.0f)};
  if (inverse) return {xa(xs(a.x, b.x), c.x), xa(xs(a.y, b.y), c.y), xa(xs(a.z, b.z), c.z)};
  return {xa(xa(a.x, b.x), c.x), xa(xa(a.y, b.y), c.y), xa(xa(a.z, b.z), c.z)};
}
// x / 3 correctly rounded in three instructions (product + FMA residual + FMA correction)
__device__ __forceinline__ float div3(float x) {
  const float q = x * 0.333333343f;
  return __fmaf_rn(__fmaf_rn(-3.0f, q, x), 0.333333343f, q);
}

// slerp (isaac_utils/rotations.py:210-232): the dot product and sin_half_theta decide the branches, so they keep the
// reference's unfused operation order; the interpolation itself may contract
__device__ __forceinline__ Q4 slerp_t(Q4 a, Q4 b, float t) {
  float c = xa(xa(xa(xm(a.x, b.x), xm(a.y, b.y)), xm(a.z, b.z)), xm(a.w, b.w));
  if (c < 0.0f) { b.x = -b.x; b.y = -b.y; b.z = -b.z; b.w = -b.w; }
  c = fabsf(c);
  const float h = acosf(c);
  const float s = sqrtf(xs(1.0f, xm(c, c)));
  const float ra = sinf((1.0f - t) * h) / s;
  const float rb = sinf(t * h) / s;
  Q4 q = {ra * a.x + rb * b.x, ra * a.y + rb * b.y, ra * a.z + rb * b.z, ra * a.w + rb * b.w};
  if (fabsf(s) < 0.001f) q = {0.5f * a.x + 0.5f * b.x, 0.5f * a.y + 0.5f * b.y, 0.5f * a.z + 0.5f * b.z, 0.5f * a.w + 0.5f * b.w};
  if (fabsf(c) >= 1.0f) q = a;
  return q;
}

// ---- block tables ---------------------------------------------------------------------------------------------------
struct SrcDst { int n; int base[PBHC_MAX_OBS_GROUPS]; int stride[PBHC_MAX_OBS_GROUPS]; int grp[PBHC_MAX_OBS_GROUPS]; int push; };
struct TileTables {
  float sigma[PBHC_NUM_SIGMA];
  float far_thr;
  uint32_t push_count;
  float lim_lo[PBHC_MAX_DOF], lim_hi[PBHC_MAX_DOF], lim_vel[PBHC_MAX_DOF], lim_tq[PBHC_MAX_DOF];
  float r_scale[PBHC_MAX_REWARDS], r_pen[PBHC_MAX_REWARDS], r_sigma[PBHC_MAX_REWARDS];
  int r_kind[PBHC_MAX_REWARDS], r_arg[PBHC_MAX_REWARDS];
  uint64_t bs_mask[kBodySums];
  int bs_raw[kBodySums], bs_slot[kBodySums];
  float bs_div[kBodySums];
  SrcDst sd[kSrcT];
  float obs_scale[kSrcT], obs_noise[kSrcT];
  uint32_t noise_mask;
  int gdim[PBHC_MAX_OBS_GROUPS];
  int feat_src[64], feat_el[64];   // scalar feature i of an env: source and element
};

struct NoiseKeyT { uint64_t seed; int64_t env; uint32_t step; };
struct Rows { float* rows; float* ghist; };   // observation slabs (shared) and the env's global history row

__device__ __forceinline__ float clampf(float v, float c) { return fminf(fmaxf(v, -c), c); }

// (value + U(-1,1) * noise) * scale with an independent draw per destination (helpers.py:128-152)
__device__ __noinline__ void emit_noisy_t(const TileTables& T, const Rows& R, int k, int e, int el, float raw, float clip, const NoiseKeyT& nk) {
  const Rng rng = {make_uint2((uint32_t)nk.seed, (uint32_t)(nk.seed >> 32)), (uint32_t)nk.env, (uint32_t)((uint64_t)nk.env >> 32), nk.step};
  const uint32_t st = 4096u + ((uint32_t)k * 256u + (uint32_t)el) * 2u;
  const uint4 a = rng.draw(st);
  const uint32_t b = rng.draw(st + 1u).x;
  const float ns = T.obs_noise[k], sc = T.obs_scale[k];
  const SrcDst& d = T.sd[k];
  for (int i = 0; i < d.n; ++i) {
    const int g = d.grp[i];
    const uint32_t bits = g == 0 ? a.x : g == 1 ? a.y : g == 2 ? a.z : a.w;
    R.rows[d.base[i] + e * d.stride[i] + el] = clampf((raw + (u01(bits) * 2.0f - 1.0f) * ns) * sc, clip);
  }
  if (d.push >= 0) R.ghist[d.push + el] = (raw + (u01(b) * 2.0f - 1.0f) * ns) * sc;
}
// one element of observation source k of env e (tile-local) to every row that shows it (clipped) and to the history ring
template <bool kNoise>
__device__ __forceinline__ void emit1(const TileTables& T, const Rows& R, int k, int e, int el, float raw, float clip, const NoiseKeyT& nk) {
  if (kNoise && ((T.noise_mask >> k) & 1u)) { emit_noisy_t(T, R, k, e, el, raw, clip, nk); return; }
  const SrcDst& d = T.sd[k];
  const float sv = raw * T.obs_scale[k], cv = clampf(sv, clip);
  for (int i = 0; i < d.n; ++i) R.rows[d.base[i] + e * d.stride[i] + el] = cv;
  if (d.push >= 0) R.ghist[d.push + el] = sv;
}
template <bool kNoise>
__device__ __forceinline__ void emit3(const TileTables& T, const Rows& R, int k, int e, int el, float x, float y, float z, float clip, const NoiseKeyT& nk) {
  if (kNoise && ((T.noise_mask >> k) & 1u)) {
    emit_noisy_t(T, R, k, e, el, x, clip, nk); emit_noisy_t(T, R, k, e, el + 1, y, clip, nk); emit_noisy_t(T, R, k, e, el + 2, z, clip, nk);
    return;
  }
  const SrcDst& d = T.sd[k];
  const float sc = T.obs_scale[k];
  const float sx = x * sc, sy = y * sc, sz = z * sc;
  const float cx = clampf(sx, clip), cy = clampf(sy, clip), cz = clampf(sz, clip);
  for (int i = 0; i < d.n; ++i) {
    float* r = R.rows + d.base[i] + e * d.stride[i] + el;
    r[0] = cx; r[1] = cy; r[2] = cz;
  }
  if (d.push >= 0) { float* h = R.ghist + d.push + el; h[0] = sx; h[1] = sy; h[2] = sz; }
}

// ---- TMA / mbarrier ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, int bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait_t(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nTWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra TWAIT_DONE;\nbra TWAIT_LOOP;\nTWAIT_DONE:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, int bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, uint32_t src, int bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ void build_tile_tables(const TileParams& P, TileTables& T, int* hmap, float* hmul, const double* scal_g, double* scal) {
  const PbhcTrackCfg& c = P.c;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < PBHC_SCALARS; i += nt) scal[i] = scal_g[i];
  __syncthreads();
  const uint64_t push_count = (uint64_t)scal[PBHC_S_PUSH_COUNT];
  if (tid < PBHC_NUM_SIGMA) T.sigma[tid] = (float)scal[PBHC_S_SIGMA + tid];
  if (tid == 0) { T.far_thr = (float)scal[PBHC_S_FAR_THR]; T.push_count = (uint32_t)push_count; }
  if (tid < c.D) {  // soft limits of this launch (legged_robot_base.py:974-1002), unfused like track.cu
    float lo = c.soft_lo[tid], hi = c.soft_hi[tid];
    if (c.cur_dof_pos.enable) {
      const float m = __fdiv_rn(xa(c.hard_lo[tid], c.hard_hi[tid]), 2.0f), r = xs(c.hard_hi[tid], c.hard_lo[tid]);
      const float cv = (float)scal[PBHC_S_SOFT_DOF_POS];
      lo = xs(m, xm(xm(0.5f, r), cv)); hi = xa(m, xm(xm(0.5f, r), cv));
    }
    T.lim_lo[tid] = lo; T.lim_hi[tid] = hi;
    const float vs = c.cur_dof_vel.enable ? (float)scal[PBHC_S_SOFT_DOF_VEL] : c.soft_dof_vel_limit;
    T.lim_vel[tid] = xm(c.vel_limit[tid], vs);
    T.lim_tq[tid] = xm(c.torque_limit[tid], c.cur_torque.enable ? (float)scal[PBHC_S_SOFT_TORQUE] : c.soft_torque_limit);
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
        default: break;
      }
      if (kind != K_RAW) sig = (float)scal[PBHC_S_SIGMA + arg];
    }
    T.r_kind[tid] = kind; T.r_arg[tid] = arg; T.r_sigma[tid] = sig;
    T.r_scale[tid] = tid < c.n_rewards ? c.reward_scale[tid] : 0.0f;
    T.r_pen[tid] = (tid < c.n_rewards && c.reward_is_penalty[tid]) ? (float)scal[PBHC_S_PENALTY] : 1.0f;
  }
  if (tid < kBodySums) {
    uint64_t feet_mask = 0;
    for (int f = 0; f < c.n_feet; ++f) feet_mask |= 1ull << c.feet_ids[f];
    const uint64_t all = c.J >= 64 ? ~0ull : ((1ull << c.J) - 1ull);
    uint64_t m = all; int raw = BR_MPOS, slot = 0; float div = 1.0f;
    switch (tid) {
      case 0: m = c.upper_mask; raw = BR_MPOS; div = (float)c.n_upper; slot = V_SIGMA_ERR + G_UPPER; break;
      case 1: m = c.lower_mask; raw = BR_MPOS; div = (float)c.n_lower; slot = V_SIGMA_ERR + G_LOWER; break;
      case 2: m = c.vr3_mask; raw = BR_MPOS; div = (float)c.n_vr3; slot = V_SIGMA_ERR + G_VR3; break;
      case 3: m = feet_mask; raw = BR_MPOS; div = (float)c.n_feet; slot = V_SIGMA_ERR + G_FEET; break;
      case 4: raw = BR_ROT; div = (float)c.J; slot = V_SIGMA_ERR + G_ROT; break;
      case 5: raw = BR_VEL; div = (float)c.J; slot = V_SIGMA_ERR + G_VEL; break;
      case 6: raw = BR_ANG; div = (float)c.J; slot = V_SIGMA_ERR + G_ANG_VEL; break;
      case 7: m = c.upper_mask; raw = BR_NRM; slot = V_UP_N; break;
      case 8: m = c.lower_mask; raw = BR_NRM; slot = V_LO_N; break;
      default: m = c.vr3_mask; raw = BR_NRM; slot = V_VR_N; break;
    }
    T.bs_mask[tid] = m & all; T.bs_raw[tid] = raw; T.bs_slot[tid] = slot; T.bs_div[tid] = div;
  }
  if (tid < kSrcT) {
    SrcDst d;
    d.n = 0; d.push = -1;
    for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) { d.base[g] = 0; d.stride[g] = 0; d.grp[g] = 0; }
    for (int g = 0; g < c.n_groups; ++g)
      if (c.src_dst[tid][g] >= 0) { d.base[d.n] = P.row_base[g] + c.src_dst[tid][g]; d.stride[d.n] = c.group_dim[g]; d.grp[d.n] = g; ++d.n; }
    for (int h = 0; h < c.n_hist_keys; ++h)
      if (c.hist_src[h] == tid) {  // ring push: one slot per key (HistoryHandler.add, history_handler.py:40-44)
        const int len = c.hist_len[h];
        d.push = P.hist_base[h] + (int)((len - ((push_count + 1) % len)) % len) * c.hist_dim[h];
        break;
      }
    T.sd[tid] = d;
    T.obs_scale[tid] = c.obs_scale[tid]; T.obs_noise[tid] = c.obs_noise[tid];
  }
  if (tid == 0) {
    uint32_t m = 0;
    for (int k = 0; k < kSrcT; ++k) if (c.obs_noise[k] != 0.0f) m |= 1u << k;
    T.noise_mask = m;
    for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) T.gdim[g] = g < c.n_groups ? c.group_dim[g] : 0;
    // scalar features of an env (legged_robot_base.py:1114-1215, motion_tracking.py:944-1015)
    int n = 0;
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_BASE_LIN_VEL; T.feat_el[n++] = i; }
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_BASE_ANG_VEL; T.feat_el[n++] = i; }
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_PROJ_GRAV; T.feat_el[n++] = i; }
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_DR_COM; T.feat_el[n++] = i; }
    T.feat_src[n] = S_PHASE; T.feat_el[n++] = 0;
    T.feat_src[n] = S_DR_FRICTION; T.feat_el[n++] = 0;
    T.feat_src[n] = S_DR_DELAY; T.feat_el[n++] = 0;
    T.feat_src[n] = S_BASE_POS_Z; T.feat_el[n++] = 0;
    for (int i = 0; i < c.L && n < 64; ++i) { T.feat_src[n] = S_DR_MASS; T.feat_el[n++] = i; }
  }
  // history ring: physical slot of logical slot s of key k is (head_k + s) % len_k, head_k = -push_count mod len_k
  // (HistoryHandler keeps newest-first order, history_handler.py:40-48).  hmap[d][i] = destination of physical element
  // i of the ring in the d-th observation segment that shows it: row offset | group << 24, -1 = none.
  for (int i = tid; i < c.hist_row; i += nt) {
    int k = 0;
    while (k + 1 < c.n_hist_keys && i >= P.hist_base[k + 1]) ++k;
    const int len = c.hist_len[k], dim = c.hist_dim[k];
    const int r = i - P.hist_base[k], ps = r / dim, cc = r - ps * dim;
    const int head = (int)((len - (push_count % len)) % len);
    const int s = (ps - head + len) % len;  // logical slot
    int nd = 0;
    for (int e = 0; e < c.n_hist_dst && nd < P.n_hd; ++e) {
      const PbhcHistDst& h = c.hist_dst[e];
      if (h.key != k) continue;
      hmap[nd * c.hist_row + i] = s < h.len ? ((P.row_base[h.group] + h.offset + s * dim + cc) | (h.group << 24)) : -1;
      if (P.hist_any_mul) hmul[nd * c.hist_row + i] = h.mul;
      ++nd;
    }
    for (; nd < P.n_hd; ++nd) { hmap[nd * c.hist_row + i] = -1; if (P.hist_any_mul) hmul[nd * c.hist_row + i] = 1.0f; }
  }
  __syncthreads();
}

template <bool kNoise>
__global__ void __launch_bounds__(kNT, 3) track_tile_kernel(const __grid_constant__ TileParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const PbhcTrackCfg& c = P.c;
  const PbhcTrackIO& io = P.io;
  const PbhcTableDesc& tab = P.tab;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  TileTables& T = *reinterpret_cast<TileTables*>(smem + P.o_tab);
  int* hmap = reinterpret_cast<int*>(smem + P.o_hmap);
  float* hmul = reinterpret_cast<float*>(smem + P.o_hmul);
  int* s_far = reinterpret_cast<int*>(smem + P.o_flags);          // [kTE]
  int* s_collide = s_far + kTE;                                   // [kTE]
  int* s_state = s_collide + kTE;                                 // [kTE] bit 0 reset, bit 1 time_out
  float* rv = reinterpret_cast<float*>(smem + P.o_rv);            // [kTE][kRvT]
  float* pre = reinterpret_cast<float*>(smem + P.o_pre);          // [kTE][kPre]
  float* rows = reinterpret_cast<float*>(smem + P.o_rows);
  const float* s_body = reinterpret_cast<const float*>(smem + P.i_body);
  const float* s_cf = reinterpret_cast<const float*>(smem + P.i_cf);
  const float* s_hist = reinterpret_cast<const float*>(smem + P.i_hist);
  const float2* s_dof = reinterpret_cast<const float2*>(smem + P.i_dof);
  const float* s_act = reinterpret_cast<const float*>(smem + P.i_act);
  const float* s_tq = reinterpret_cast<const float*>(smem + P.i_tq);
  const float* s_lact = reinterpret_cast<const float*>(smem + P.i_lact);
  const float* s_ldv = reinterpret_cast<const float*>(smem + P.i_ldv);
  const float* s_kp = reinterpret_cast<const float*>(smem + P.i_kp);
  const float* s_kd = reinterpret_cast<const float*>(smem + P.i_kd);
  const float* s_es = reinterpret_cast<const float*>(smem + P.i_es);
  const float* s_envf = reinterpret_cast<const float*>(smem + P.i_envf);
  const float* s_root = reinterpret_cast<const float*>(smem + P.i_root);
  const float* s_org = reinterpret_cast<const float*>(smem + P.i_org);
  const long long* s_ep = reinterpret_cast<const long long*>(smem + P.i_ep);
  const long long* s_mid = reinterpret_cast<const long long*>(smem + P.i_mid);
  const float* s_com = reinterpret_cast<const float*>(smem + P.i_com);
  const float* s_fric = reinterpret_cast<const float*>(smem + P.i_fric);
  const long long* s_delay = reinterpret_cast<const long long*>(smem + P.i_delay);
  const float* s_mass = reinterpret_cast<const float*>(smem + P.i_mass);
  float* q_rew = reinterpret_cast<float*>(smem + P.q_rew);
  float* q_es = reinterpret_cast<float*>(smem + P.q_es);
  float* q_envf = reinterpret_cast<float*>(smem + P.q_envf);
  long long* q_ep = reinterpret_cast<long long*>(smem + P.q_ep);
  long long* q_reset = reinterpret_cast<long long*>(smem + P.q_reset);
  float* q_lact = reinterpret_cast<float*>(smem + P.q_lact);
  float* q_ldpos = reinterpret_cast<float*>(smem + P.q_ldpos);
  float* q_ldvel = reinterpret_cast<float*>(smem + P.q_ldvel);
  float* dpart = reinterpret_cast<float*>(smem + P.o_dpart);     // [kTE][kDofRaw][D]   (in the dead history slab)
  float* s_ext = reinterpret_cast<float*>(smem + P.o_ext);       // [kTE][n_ext][13]    (in the dead history slab)
  float* bpart = reinterpret_cast<float*>(smem + P.o_bpart);     // [kTE][kBodyRaw][J]  (in the dead dof-input slabs)
  float* s_rw = reinterpret_cast<float*>(smem + P.o_rw);         // [kTE][32] raw reward terms (scalar reward mode)
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t bar = sbase + P.o_bar;

  if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); fence_async(); }
  for (int i = tid; i < kTE * kRvT; i += kNT) rv[i] = 0.0f;  // slots nobody writes are accumulated too
  build_tile_tables(P, T, hmap, hmul, io.scalars, reinterpret_cast<double*>(smem + P.o_rows));  // (the rows region is idle: scratch for the scalars)

  const int J = c.J, Jb = c.Jb, D = c.D, n_ext = c.n_ext, H = c.hist_row, n_envs = P.n;
  const int R = c.n_rewards;
  const float clipv = c.clip_obs, dt = c.dt;
  const int e_pos = P.e_pos, e_vel = P.e_vel, e_ang = P.e_ang, e_dp = P.e_dp, e_dv = P.e_dv, e_c = P.e_c;
  const float far_thr = T.far_thr;
  const uint32_t push_count = T.push_count;
  double acc_reg = 0.0;  // slot = tid (< kAccLo)

  auto issue_loads = [&](int tile0) {  // warp 0: one descriptor per lane
    if (lane == 0) mbar_expect_tx(bar, P.in_bytes);
    __syncwarp();
    for (int i = lane; i < P.n_in; i += 32) {
      const TileDesc d = P.in[i];
      if (d.g_stride == d.row_bytes) bulk_load(sbase + d.soff, d.g + (int64_t)tile0 * d.row_bytes, kTE * d.row_bytes, bar);
      else for (int e = 0; e < kTE; ++e) bulk_load(sbase + d.soff + e * d.row_bytes, d.g + (int64_t)(tile0 + e) * d.g_stride, d.row_bytes, bar);
    }
  };

  const int n_tiles = P.n_tiles;
  int tile = blockIdx.x;
  bool loaded = false;
  uint32_t phase = 0;
  if (P.bulk_ok && tile < n_tiles && (tile + 1) * kTE <= n_envs) { if (warp == 0) issue_loads(tile * kTE); loaded = true; }

  for (; tile < n_tiles; tile += gridDim.x) {
    const int tile0 = tile * kTE;
    const int cnt = min(kTE, n_envs - tile0);
    if (warp == 0) bulk_wait_read();  // the previous tile's bulk stores have left the output slabs
    if (loaded) { mbar_wait_t(bar, phase); phase ^= 1u; }
    else {  // ragged last tile / unaligned tensors: cooperative word copies
      for (int i = 0; i < P.n_in; ++i) {
        const TileDesc d = P.in[i];
        const int rw = d.row_bytes >> 2;
        uint32_t* dst = reinterpret_cast<uint32_t*>(smem + d.soff);
        for (int w = tid; w < cnt * rw; w += kNT) {
          const int e = w / rw, x = w - e * rw;
          dst[w] = reinterpret_cast<const uint32_t*>(d.g + (int64_t)(tile0 + e) * d.g_stride)[x];
        }
      }
    }
    if (tid < kTE) { s_far[tid] = 0; s_collide[tid] = 0; }
    __syncthreads();

    // ---- 1. pre (legged_robot_base.py:346-353, motion_tracking.py:548-558, 678-739) + history ring -> rows ------------------
    if (warp < 3 && lane < cnt) {
      const int e = lane;
      float* pe = pre + e * kPre;
      const float* root = s_root + e * 13;
      const Q4 rq = {root[3], root[4], root[5], root[6]};
      if (warp == 0) {
        const int64_t ep = s_ep[e] + 1;  // legged_robot_base.py:299
        const float start = s_envf[e * PBHC_ENVF_STRIDE];
        const int clip = resolve_clip(tab, io.clip_of_motion, s_mid[e]);
        const float t = xa(xm((float)(ep + 1), dt), start);  // kick_motion_res, motion_tracking.py:554
        const FrameBlend fb = frame_blend_x(tab, clip, t);
        pe[P_F0] = __int_as_float((int)fb.f0); pe[P_F1] = __int_as_float((int)fb.f1); pe[P_BLEND] = fb.blend;
        pe[P_PHASE] = __fdiv_rn(t, tab.clip_length[clip]);  // motion_tracking.py:738-739
        pe[P_CLIP] = __int_as_float(clip);
      } else if (warp == 1) {
        const V3 blv = quat_rotate_inverse(rq, {root[7], root[8], root[9]});
        const V3 bav = quat_rotate_inverse(rq, {root[10], root[11], root[12]});
        pe[P_BLV] = blv.x; pe[P_BLV + 1] = blv.y; pe[P_BLV + 2] = blv.z;
        pe[P_BAV] = bav.x; pe[P_BAV + 1] = bav.y; pe[P_BAV + 2] = bav.z;
        pe[P_RPX] = root[0]; pe[P_RPY] = root[1]; pe[P_RPZ] = root[2];
      } else {
        const V3 pg = quat_rotate_x(rq, {0.0f, 0.0f, -1.0f}, true);
        const Q4 hinv = heading_quat_inv(rq);
        pe[P_PG] = pg.x; pe[P_PG + 1] = pg.y; pe[P_PG + 2] = pg.z;
        pe[P_GRAVN] = norm2(pg.x, pg.y);
        pe[P_ORIENT] = pg.x * pg.x + pg.y * pg.y;
        pe[P_HZ] = hinv.z; pe[P_HW] = hinv.w;
      }
    }
    // HistoryHandler.query (history_handler.py:46-48): every physical ring element to its places in the rows
    for (int i = tid; i < H; i += kNT) {
      for (int d = 0; d < P.n_hd; ++d) {
        const int m = hmap[d * H + i];
        if (m < 0) continue;
        const float mul = P.hist_any_mul ? hmul[d * H + i] : 1.0f;
        float* dst = rows + (m & 0xffffff);
        const int es = T.gdim[m >> 24];
        for (int e = 0; e < cnt; ++e) dst[e * es] = clampf(s_hist[e * H + i] * mul, clipv);
      }
    }
    __syncthreads();

    // ---- 2. dofs (item = env x dof), extended bodies, feet ----------------------------------------------------------------------
    for (int it = tid; it < cnt * D; it += kNT) {
      const int e = (it * P.inv_d) >> 16, d = it - e * D;
      const int64_t n = tile0 + e;
      const float* pe = pre + e * kPre;
      const float b = pe[P_BLEND], omb = 1.0f - b;
      const float* r0 = tab.records + (int64_t)__float_as_int(pe[P_F0]) * tab.rec_stride;
      const float* r1 = tab.records + (int64_t)__float_as_int(pe[P_F1]) * tab.rec_stride;
      const float2 ds = s_dof[it];
      const float dpos = ds.x, dvel = ds.y;
      const float act = s_act[it], tq = s_tq[it], lact = s_lact[it], ldv = s_ldv[it];
      const float ddef = io.default_dof_pos[n * io.default_dof_stride + d];
      const float dja = (omb * __ldg(r0 + e_dp + d) + b * __ldg(r1 + e_dp + d)) - dpos;   // motion_tracking.py:661-665
      const float djv = (omb * __ldg(r0 + e_dv + d) + b * __ldg(r1 + e_dv + d)) - dvel;
      const float dact = lact - act, dacc = (ldv - dvel) / dt;
      const float lim_t = T.lim_tq[d];
      float* dp_ = dpart + e * (kDofRaw * D) + d;
      dp_[DR_DJA * D] = dja; dp_[DR_DJV2 * D] = djv * djv; dp_[DR_TQ2 * D] = tq * tq; dp_[DR_DACT2 * D] = dact * dact;
      dp_[DR_DV2 * D] = dvel * dvel; dp_[DR_ACC2 * D] = dacc * dacc;
      dp_[DR_LPOS * D] = -fminf(dpos - T.lim_lo[d], 0.0f) + fmaxf(dpos - T.lim_hi[d], 0.0f);  // legged_robot_base.py:974-1002
      dp_[DR_LVEL * D] = fminf(fmaxf(fabsf(dvel) - T.lim_vel[d], 0.0f), 1.0f);
      dp_[DR_LTQ * D] = c.cur_torque.enable ? fminf(fmaxf(fabsf(tq) - lim_t, 0.0f), 1.0f) : fmaxf(fabsf(tq) - lim_t, 0.0f);
      const Rows Rw = {rows, io.hist + n * H};
      const NoiseKeyT nk = {c.seed, n + io.env_offset, push_count};
      emit1<kNoise>(T, Rw, S_DOF_POS, e, d, dpos - ddef, clipv, nk);
      emit1<kNoise>(T, Rw, S_DOF_VEL, e, d, dvel, clipv, nk);
      emit1<kNoise>(T, Rw, S_ACTIONS, e, d, act, clipv, nk);
      emit1<kNoise>(T, Rw, S_DR_KP, e, d, s_kp[it], clipv, nk);
      emit1<kNoise>(T, Rw, S_DR_KD, e, d, s_kd[it], clipv, nk);
      if (c.src_used_mask & ((1ull << S_DIF_JOINT) | (1ull << S_DIF_JOINT_VEL))) {
        emit1<kNoise>(T, Rw, S_DIF_JOINT, e, d, dja, clipv, nk);
        emit1<kNoise>(T, Rw, S_DIF_JOINT_VEL, e, d, djv, clipv, nk);
      }
      q_lact[it] = act; q_ldpos[it] = dpos; q_ldvel[it] = dvel;  // _post_compute_observations_callback, legged_robot_base.py:398-405
    }
    if (warp == kTE - 2) {  // extended bodies (motion_tracking.py:619-643): position chain unfused (it feeds the motion_far comparison)
      for (int it = lane; it < cnt * n_ext; it += 32) {
        const int e = it / n_ext, x = it - e * n_ext;
        const float* bs = s_body + (e * Jb + c.ext_parent[x]) * 13;
        const Q4 qp = {bs[3], bs[4], bs[5], bs[6]};
        const V3 off = {c.ext_pos[x][0], c.ext_pos[x][1], c.ext_pos[x][2]};
        const Q4 qe = {c.ext_rot_xyzw[x][0], c.ext_rot_xyzw[x][1], c.ext_rot_xyzw[x][2], c.ext_rot_xyzw[x][3]};
        const V3 r = quat_rotate_x(qe, quat_rotate_x(qp, off, false), false);
        const Q4 q = quat_mul(qp, qe);
        float* o = s_ext + it * 13;
        o[0] = xa(r.x, bs[0]); o[1] = xa(r.y, bs[1]); o[2] = xa(r.z, bs[2]);
        o[3] = q.x; o[4] = q.y; o[5] = q.z; o[6] = q.w;
        const V3 w = {bs[10], bs[11], bs[12]};
        o[7] = bs[7] + (w.y * off.z - w.z * off.y); o[8] = bs[8] + (w.z * off.x - w.x * off.z); o[9] = bs[9] + (w.x * off.y - w.y * off.x);
        o[10] = w.x; o[11] = w.y; o[12] = w.z;
      }
    }
    if (warp == kTE - 1) {  // feet: contacts, feet_air_time, contact-force / slip terms (item = env x foot)
      for (int it = lane; it < cnt * c.n_feet; it += 32) {
        const int e = it / c.n_feet, f = it - e * c.n_feet;
        const int fb = c.feet_ids[f];
        const float* pe = pre + e * kPre;
        const float b = pe[P_BLEND], omb = 1.0f - b;
        const float* fc = s_cf + (e * Jb + fb) * 3;
        const float* bs = s_body + (e * Jb + fb) * 13;
        const float* ef = s_envf + e * PBHC_ENVF_STRIDE;
        const float last_c = ef[4 + f];
        float air_f = ef[2 + f];
        const float fn = norm3(fc[0], fc[1], fc[2]);
        const float contact_f = fn > 1.0f ? 1.0f : 0.0f;                       // legged_robot_base.py:352-353
        const float filt = (contact_f != 0.0f || last_c != 0.0f) ? 1.0f : 0.0f;
        float x_cm = 0.0f, x_air = 0.0f;
        if (tab.contact_size) {
          const float* r0 = tab.records + (int64_t)__float_as_int(pe[P_F0]) * tab.rec_stride;
          const float* r1 = tab.records + (int64_t)__float_as_int(pe[P_F1]) * tab.rec_stride;
          x_cm = fabsf(filt - (omb * __ldg(r0 + e_c + f) + b * __ldg(r1 + e_c + f)));
        }
        const bool fz = fc[2] > 1.0f || last_c != 0.0f;                         // motion_tracking.py:1307-1319
        if (P.has_air_time) {
          const float first = (air_f > 0.0f && fz) ? 1.0f : 0.0f;
          air_f = xa(air_f, dt);
          x_air = (air_f - c.desired_feet_air_time) * first;
          air_f *= fz ? 0.0f : 1.0f;
        }
        float* o = rv + e * kRvT + V_FOOT + f * 8;
        o[FV_AIR] = x_air;
        o[FV_FORCE] = fmaxf(fn - c.max_contact_force, 0.0f);                     // :1321-1323
        o[FV_STUMBLE] = norm2(fc[0], fc[1]) > 5.0f * fabsf(fc[2]) ? 1.0f : 0.0f;  // :1325-1328
        o[FV_SLIP] = norm3(bs[7], bs[8], bs[9]) * contact_f;                     // legged_robot_base.py:1004-1010
        o[FV_FSLIP] = contact_f * norm2(bs[7], bs[8]);                           // :1012-1017
        o[FV_CM] = x_cm; o[FV_AIRF] = air_f; o[FV_CONTACT] = contact_f;
      }
    }
    __syncthreads();

    // ---- 3. bodies (item = env x body): reference pose (lerp / slerp of the frame pair), differences, heading-frame features --
    if (warp == kTE - 1) {  // dof sums (item = env x value), added in dof order
      for (int it = lane; it < cnt * kDofRaw; it += 32) {
        const int e = it / kDofRaw, k = it - e * kDofRaw;
        const float* src = dpart + (e * kDofRaw + k) * D;
        float* r = rv + e * kRvT;
        if (k == DR_DJA) {
          float s = 0.0f, mx = 0.0f;
          for (int d = 0; d < D; ++d) { const float v = src[d]; s += v * v; mx = fmaxf(mx, fabsf(v)); }
          r[V_SIGMA_ERR + G_JOINT_POS] = s / (float)D; r[V_JOINT_N] = sqrtf(s); r[V_SIGMA_ERR + G_MAX_JOINT] = mx;
        } else {
          float s = 0.0f;
          for (int d = 0; d < D; ++d) s += src[d];
          if (k == DR_DJV2) r[V_SIGMA_ERR + G_JOINT_VEL] = s / (float)D;
          else r[k == DR_TQ2 ? V_TQ2 : k == DR_DACT2 ? V_DACT2 : k == DR_DV2 ? V_DV2 : k == DR_ACC2 ? V_ACC2 : k == DR_LPOS ? V_LIM_POS : k == DR_LVEL ? V_LIM_VEL : V_LIM_TQ] = s;
        }
      }
    }
    for (int it = tid; it < cnt * J; it += kNT) {
      const int e = (it * P.inv_j) >> 16, j = it - e * J;
      const int64_t n = tile0 + e;
      const float* pe = pre + e * kPre;
      const float b = pe[P_BLEND], omb = 1.0f - b;
      const float* r0 = tab.records + (int64_t)__float_as_int(pe[P_F0]) * tab.rec_stride;
      const float* r1 = tab.records + (int64_t)__float_as_int(pe[P_F1]) * tab.rec_stride;
      const float* bs = j < Jb ? s_body + (e * Jb + j) * 13 : s_ext + (e * n_ext + (j - Jb)) * 13;
      const float4 qa = __ldg(reinterpret_cast<const float4*>(r0) + j);
      const float4 qb = __ldg(reinterpret_cast<const float4*>(r1) + j);
      float rpz[3], rvv[3], rw[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        rpz[k] = xa(xa(xm(omb, __ldg(r0 + e_pos + 3 * j + k)), xm(b, __ldg(r1 + e_pos + 3 * j + k))), s_org[e * 3 + k]);
        rvv[k] = omb * __ldg(r0 + e_vel + 3 * j + k) + b * __ldg(r1 + e_vel + 3 * j + k);
        rw[k] = omb * __ldg(r0 + e_ang + 3 * j + k) + b * __ldg(r1 + e_ang + 3 * j + k);
      }
      const Q4 rr = slerp_t({qa.x, qa.y, qa.z, qa.w}, {qb.x, qb.y, qb.z, qb.w}, b);
      if (io.ref_body_pos_extend) {
        float* o = io.ref_body_pos_extend + n * (3 * J) + 3 * j;
        o[0] = rpz[0]; o[1] = rpz[1]; o[2] = rpz[2];
      }
      if (io.ref_body_rot_extend) reinterpret_cast<float4*>(io.ref_body_rot_extend)[n * J + j] = make_float4(rr.x, rr.y, rr.z, rr.w);
      const V3 dp = {xs(rpz[0], bs[0]), xs(rpz[1], bs[1]), xs(rpz[2], bs[2])};
      const float dq0 = rr.x - bs[3], dq1 = rr.y - bs[4], dq2 = rr.z - bs[5], dq3 = rr.w - bs[6];
      const V3 dv = {rvv[0] - bs[7], rvv[1] - bs[8], rvv[2] - bs[9]};
      const V3 dw = {rw[0] - bs[10], rw[1] - bs[11], rw[2] - bs[12]};
      const float nrm = norm3(dp.x, dp.y, dp.z);
      float* bp = bpart + e * (kBodyRaw * J) + j;
      bp[BR_MPOS * J] = div3((dp.x * dp.x + dp.y * dp.y) + dp.z * dp.z);
      bp[BR_NRM * J] = nrm;
      bp[BR_ROT * J] = (((dq0 * dq0 + dq1 * dq1) + dq2 * dq2) + dq3 * dq3) * 0.25f;
      bp[BR_VEL * J] = div3((dv.x * dv.x + dv.y * dv.y) + dv.z * dv.z);
      bp[BR_ANG * J] = div3((dw.x * dw.x + dw.y * dw.y) + dw.z * dw.z);
      if (nrm > far_thr) s_far[e] = 1;
      if (j < Jb && ((c.penalised_mask >> j) & 1ull)) {
        const float* fc = s_cf + (e * Jb + j) * 3;
        if (norm3(fc[0], fc[1], fc[2]) > 0.1f) atomicAdd(&s_collide[e], 1);
      }
      const Rows Rw = {rows, io.hist + n * H};
      const NoiseKeyT nk = {c.seed, n + io.env_offset, push_count};
      const float hz = pe[P_HZ], hw = pe[P_HW];
      {
        const V3 dl = quat_rotate_z(hz, hw, dp);  // motion_tracking.py:694-697
        emit3<kNoise>(T, Rw, S_DIF_LOCAL, e, 3 * j, dl.x, dl.y, dl.z, clipv, nk);
        const V3 lr = quat_rotate_z(hz, hw, {rpz[0] - pe[P_RPX], rpz[1] - pe[P_RPY], rpz[2] - pe[P_RPZ]});  // :705-707
        emit3<kNoise>(T, Rw, S_LOCAL_REF, e, 3 * j, lr.x, lr.y, lr.z, clipv, nk);
      }
      if (c.src_used_mask & ((1ull << S_LOCAL_REF_VEL) | (1ull << S_GLOBAL_REF_VEL))) {
        const V3 lv = quat_rotate_z(hz, hw, {rvv[0], rvv[1], rvv[2]});
        emit3<kNoise>(T, Rw, S_LOCAL_REF_VEL, e, 3 * j, lv.x, lv.y, lv.z, clipv, nk);
        emit3<kNoise>(T, Rw, S_GLOBAL_REF_VEL, e, 3 * j, rvv[0], rvv[1], rvv[2], clipv, nk);
      }
    }
    __syncthreads();

    // ---- 4. body sums (item = env x sum, bodies added in index order), per-env scalars + termination, scalar features ------------
    for (int it = tid; it < cnt * kBodySums; it += kNT) {
      const int e = it / kBodySums, k = it - e * kBodySums;
      const float* src = bpart + (e * kBodyRaw + T.bs_raw[k]) * J;
      uint64_t m = T.bs_mask[k];
      float s = 0.0f;
      while (m) { const int j = __ffsll((long long)m) - 1; m &= m - 1; s += src[j]; }
      rv[e * kRvT + T.bs_slot[k]] = s / T.bs_div[k];
    }
    if (warp == kTE - 1 && lane < cnt) {  // legged_robot_base.py:408-489, motion_tracking.py:330-358
      const int e = lane;
      float* r = rv + e * kRvT;
      const float* pe = pre + e * kPre;
      const float* f0 = r + V_FOOT;
      const float* f1 = r + V_FOOT + 8;
      const bool two = c.n_feet > 1;
      const int64_t ep = s_ep[e] + 1;
      const float start = s_envf[e * PBHC_ENVF_STRIDE], mlen = s_envf[e * PBHC_ENVF_STRIDE + 1];
      const bool t_grav = c.term_gravity && pe[P_GRAVN] > c.term_gravity_thr;
      const bool t_far = c.term_motion_far && s_far[e] != 0;
      const bool t_out = ep > c.max_episode_length;
      const bool t_end = c.term_motion_end && (xa(xm((float)ep, dt), start) > mlen);
      const bool time_out = t_out || t_end;
      const bool reset = t_grav || t_far || time_out;
      const float x_cm = f0[FV_CM] + (two ? f1[FV_CM] : 0.0f);
      r[V_T_GRAV] = t_grav; r[V_T_FAR] = t_far; r[V_T_OUT] = time_out; r[V_T_END] = t_end;
      r[V_RESET] = reset; r[V_ONE] = 1.0f; r[V_ALIVE] = 1.0f;
      r[V_CONTACT] = 1.0f - x_cm / (float)c.n_feet; r[V_CONTACT_V2] = 0.5f - x_cm / (float)c.n_feet;
      r[V_ORIENT] = pe[P_ORIENT];
      r[V_AIR] = f0[FV_AIR] + (two ? f1[FV_AIR] : 0.0f); r[V_FORCE] = f0[FV_FORCE] + (two ? f1[FV_FORCE] : 0.0f);
      r[V_STUMBLE] = (f0[FV_STUMBLE] != 0.0f || (two && f1[FV_STUMBLE] != 0.0f)) ? 1.0f : 0.0f;
      r[V_SLIP] = f0[FV_SLIP] + (two ? f1[FV_SLIP] : 0.0f); r[V_FSLIP] = f0[FV_FSLIP] + (two ? f1[FV_FSLIP] : 0.0f);
      r[V_COLLIDE] = (float)s_collide[e];
      r[V_AIR0] = f0[FV_AIRF]; r[V_AIR1] = two ? f1[FV_AIRF] : 0.0f; r[V_CONT0] = f0[FV_CONTACT]; r[V_CONT1] = two ? f1[FV_CONTACT] : 0.0f;
      s_state[e] = (reset ? 1 : 0) | (time_out ? 2 : 0);
    }
    for (int it = tid; it < cnt * P.n_feat; it += kNT) {  // scalar features (the reset branch re-emits them for its env)
      const int e = (it * P.inv_f) >> 16, i = it - e * P.n_feat;
      const int64_t n = tile0 + e;
      const float* pe = pre + e * kPre;
      const int k = T.feat_src[i], el = T.feat_el[i];
      float val;
      if (i < 9) val = pe[P_BLV + i];
      else if (k == S_DR_COM) val = s_com[e * 3 + el];
      else if (k == S_PHASE) val = pe[P_PHASE];        // stale for envs that reset, like the reference
      else if (k == S_DR_FRICTION) val = s_fric[e];
      else if (k == S_DR_DELAY) val = (float)s_delay[e];
      else if (k == S_BASE_POS_Z) val = pe[P_RPZ];
      else val = s_mass[e * c.L + el];
      const Rows Rw = {rows, io.hist + n * H};
      const NoiseKeyT nk = {c.seed, n + io.env_offset, push_count};
      emit1<kNoise>(T, Rw, k, e, el, val, clipv, nk);
    }
    __syncthreads();

    // ---- 5. rewards (item = env x term; legged_robot_base.py:715-761) + episode sums ----------------------------------------------
    for (int it = tid; it < cnt * P.n_cols; it += kNT) {
      const int e = (it * P.inv_r) >> 16, r = it - e * P.n_cols;
      const int64_t n = tile0 + e;
      const float* rve = rv + e * kRvT;
      const int st = s_state[e];
      const bool reset = st & 1, time_out = st & 2;
      float rew = 0.0f;
      if (r < R) {
        const int kind = T.r_kind[r];
        float v = rve[T.r_arg[r]];
        if (kind != K_RAW) {
          v = expf(-v / T.r_sigma[r]);
          if (kind == K_EXP2)  // teleop_body_position_extend, motion_tracking.py:1177-1199
            v = v * c.w_lower + expf(-rve[V_SIGMA_ERR + G_UPPER] / T.sigma[G_UPPER]) * c.w_upper;
        }
        rew = v * T.r_scale[r] * T.r_pen[r];
      }
      const float term_rew = c.has_termination ? ((reset && !time_out) ? 1.0f : 0.0f) * c.termination_scale : 0.0f;
      if (c.use_vec_reward) {
        float v = rew;
        if (c.only_positive_rewards) v = fmaxf(v, 0.0f);
        if (r == R - 1) v += term_rew;  // the reference's stale loop index, legged_robot_base.py:744
        if (r < R) q_rew[e * (R + 1) + r] = v;
        if (r == R) q_rew[e * (R + 1) + R] = 0.0f;
      } else if (r < R) s_rw[e * 32 + r] = rew;
      if (r < c.n_episode_sums) {
        float ev = s_es[e * c.n_episode_sums + r] + (r < R ? rew : term_rew);
        if (reset) { io.last_episode_sums[n * c.n_episode_sums + r] = ev; ev = 0.0f; }  // extras["episode"], legged_robot_base.py:509-513
        q_es[e * c.n_episode_sums + r] = ev;
      }
    }
    __syncthreads();

    // ---- 6. per-env state; reset (warp = env; legged_robot_base.py:491-516, 647-686, 587-635; motion_tracking.py:274-296, 369-384, 445-543)
    if (warp < cnt) {
      const int e = warp;
      const int64_t n = tile0 + e;
      float* rve = rv + e * kRvT;
      const float* pe = pre + e * kPre;
      const int st = s_state[e];
      const bool reset = st & 1, time_out = st & 2;
      const int64_t last_ep = s_ep[e] + 1;
      int64_t ep = last_ep;
      float start = s_envf[e * PBHC_ENVF_STRIDE], mlen = s_envf[e * PBHC_ENVF_STRIDE + 1], etr = s_envf[e * PBHC_ENVF_STRIDE + 6];
      float air0 = rve[V_AIR0], air1 = rve[V_AIR1], contact0 = rve[V_CONT0], contact1 = rve[V_CONT1];
      if (!c.use_vec_reward && lane == 0) {
        float tot = 0.0f;
        for (int r = 0; r < R; ++r) tot += s_rw[e * 32 + r];
        if (c.only_positive_rewards) tot = fmaxf(tot, 0.0f);
        const float term_rew = c.has_termination ? ((reset && !time_out) ? 1.0f : 0.0f) * c.termination_scale : 0.0f;
        q_rew[e] = tot + term_rew;
      }
      if (reset) {
        const int clip = __float_as_int(pe[P_CLIP]);
        const bool ld = lane < D;
        const Rows Rw = {rows, io.hist + n * H};
        const NoiseKeyT nk = {c.seed, n + io.env_offset, push_count};
        const Rng rng = {make_uint2((uint32_t)c.seed, (uint32_t)(c.seed >> 32)), (uint32_t)(n + io.env_offset), (uint32_t)((uint64_t)(n + io.env_offset) >> 32), push_count};
        if (ld) io.actions[n * D + lane] = 0.0f;
        air0 = 0.0f; air1 = 0.0f; contact0 = 0.0f; contact1 = 0.0f;
        // HistoryHandler.reset: zero the ring and what the observation rows already took from it
        for (int i = lane; i < H; i += 32) {
          io.hist[n * H + i] = 0.0f;
          for (int d = 0; d < P.n_hd; ++d) {
            const int m = hmap[d * H + i];
            if (m >= 0) rows[(m & 0xffffff) + e * T.gdim[m >> 24]] = 0.0f;
          }
        }
        __syncwarp();
        ep = 0;
        float kp2 = 1.f, kd2 = 1.f, dpos2 = 0.f, dvel2 = 0.f;
        if (ld) {  // _episodic_domain_randomization
          kp2 = io.kp_scale[n * D + lane]; kd2 = io.kd_scale[n * D + lane];
          if (c.randomize_pd_gain) {
            kp2 = io.inject_kp ? io.inject_kp[n * D + lane] : xa(xm(xs(c.kp_hi, c.kp_lo), rng.uniform(16 + lane)), c.kp_lo);
            kd2 = io.inject_kd ? io.inject_kd[n * D + lane] : xa(xm(xs(c.kd_hi, c.kd_lo), rng.uniform(48 + lane)), c.kd_lo);
            io.kp_scale[n * D + lane] = kp2; io.kd_scale[n * D + lane] = kd2;
          }
          if (c.randomize_rfi_lim)
            io.rfi_lim_scale[n * D + lane] = io.inject_rfi ? io.inject_rfi[n * D + lane] : xa(xm(xs(c.rfi_hi, c.rfi_lo), rng.uniform(80 + lane)), c.rfi_lo);
          if (c.use_rao)
            io.rao_scale[n * D + lane] = io.inject_rao ? io.inject_rao[n * D + lane] : xa(xm(xs(c.rao_lim, -c.rao_lim), rng.uniform(112 + lane)), -c.rao_lim);
        }
        int64_t delay = s_delay[e];
        if (c.randomize_ctrl_delay) {
          delay = io.inject_delay ? io.inject_delay[n] : c.delay_lo + (int64_t)(rng.draw(1).x % (uint32_t)(c.delay_hi - c.delay_lo + 1));
          if (lane == 0) io.action_delay_idx[n] = delay;
        }
        etr = __fdiv_rn(xa(xm((float)last_ep, dt), start), mlen);              // motion_tracking.py:281-282
        mlen = tab.clip_length[clip];                                       // _resample_motion_times
        start = io.inject_start_times ? io.inject_start_times[n] : xm(rng.uniform(0), mlen);
        // _reset_dofs / _reset_root_states from the motion state at the new time (unfused: these are simulator state)
        const FrameBlend f2 = frame_blend_x(tab, clip, xa(xm((float)(ep + 1), dt), start));
        const float b2 = f2.blend, omb2 = 1.0f - b2;
        const float* q0 = tab.records + f2.f0 * tab.rec_stride;
        const float* q1 = tab.records + f2.f1 * tab.rec_stride;
        if (ld) {
          dpos2 = xa(xm(omb2, q0[e_dp + lane]), xm(b2, q1[e_dp + lane]));
          dvel2 = xa(xm(omb2, q0[e_dv + lane]), xm(b2, q1[e_dv + lane]));
          if (c.reset_noise[0] != 0.0f) dpos2 = xa(dpos2, xm(rng.normal(144 + lane), c.reset_noise[0]));
          if (c.reset_noise[1] != 0.0f) dvel2 = xa(dvel2, xm(rng.normal(176 + lane), c.reset_noise[1]));
          reinterpret_cast<float2*>(io.dof_state)[n * D + lane] = make_float2(dpos2, dvel2);
        }
        if (lane < 13) {
          float v;
          if (lane >= 3 && lane < 7) {
            const float4 a = reinterpret_cast<const float4*>(q0)[0], d4 = reinterpret_cast<const float4*>(q1)[0];
            const Q4 rr = slerp_t({a.x, a.y, a.z, a.w}, {d4.x, d4.y, d4.z, d4.w}, b2);
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
            v = xa(xm(omb2, q0[el]), xm(b2, q1[el]));
            if (lane < 3) v = xa(v, s_org[e * 3 + lane]);
            const float ns = lane < 3 ? c.reset_noise[2] : (lane < 10 ? c.reset_noise[4] : c.reset_noise[5]);
            if (ns != 0.0f) v = xa(v, xm(rng.normal(8 + lane), ns));
          }
          io.root_state[n * 13 + lane] = v;
        }
        // the observations of an env that reset: dof-wise features from the new state, scalar features again (their push
        // slots were just zeroed; ref_motion_phase & co. stay stale like the reference's)
        if (ld) {
          const float ddef = io.default_dof_pos[n * io.default_dof_stride + lane];
          emit1<kNoise>(T, Rw, S_DOF_POS, e, lane, dpos2 - ddef, clipv, nk);
          emit1<kNoise>(T, Rw, S_DOF_VEL, e, lane, dvel2, clipv, nk);
          emit1<kNoise>(T, Rw, S_ACTIONS, e, lane, 0.0f, clipv, nk);
          emit1<kNoise>(T, Rw, S_DR_KP, e, lane, kp2, clipv, nk);
          emit1<kNoise>(T, Rw, S_DR_KD, e, lane, kd2, clipv, nk);
          q_lact[e * D + lane] = 0.0f; q_ldpos[e * D + lane] = dpos2; q_ldvel[e * D + lane] = dvel2;
        }
        for (int i = lane; i < P.n_feat; i += 32) {
          const int k = T.feat_src[i], el = T.feat_el[i];
          float val;
          if (i < 9) val = pe[P_BLV + i];
          else if (k == S_DR_COM) val = s_com[e * 3 + el];
          else if (k == S_PHASE) val = pe[P_PHASE];
          else if (k == S_DR_FRICTION) val = s_fric[e];
          else if (k == S_DR_DELAY) val = (float)delay;
          else if (k == S_BASE_POS_Z) val = pe[P_RPZ];
          else val = s_mass[e * c.L + el];
          emit1<kNoise>(T, Rw, k, e, el, val, clipv, nk);
        }
      }
      if (lane == 0) {
        float4* of = reinterpret_cast<float4*>(q_envf + e * PBHC_ENVF_STRIDE);
        of[0] = make_float4(start, mlen, air0, c.n_feet > 1 ? air1 : 0.0f);
        of[1] = make_float4(contact0, c.n_feet > 1 ? contact1 : 0.0f, etr, s_envf[e * PBHC_ENVF_STRIDE + 7]);
        q_ep[e] = ep;
        q_reset[e] = reset ? 1 : 0;
        io.time_out_buf[n] = time_out ? 1 : 0;
        rve[V_ETR] = etr;
        rve[V_RESET_EPLEN] = reset ? (float)last_ep : 0.0f;
      }
    }
    fence_async();
    __syncthreads();

    // ---- 7. outputs leave, the next tile's inputs are requested, cross-env sums -----------------------------------------------------
    const bool bulk_now = P.bulk_ok && cnt == kTE;
    if (bulk_now) {
      if (warp == 0) {
        for (int i = lane; i < P.n_out; i += 32) {
          const TileDesc d = P.out[i];
          bulk_store(const_cast<char*>(d.g) + (int64_t)tile0 * d.row_bytes, sbase + d.soff, kTE * d.row_bytes);
        }
        bulk_commit();
      }
    } else {
      for (int i = 0; i < P.n_out; ++i) {
        const TileDesc d = P.out[i];
        const int nw = (cnt * d.row_bytes) >> 2;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(smem + d.soff);
        uint32_t* dst = reinterpret_cast<uint32_t*>(const_cast<char*>(d.g) + (int64_t)tile0 * d.row_bytes);
        for (int w = tid; w < nw; w += kNT) dst[w] = src[w];
      }
    }
    const int next = tile + (int)gridDim.x;
    loaded = false;
    if (P.bulk_ok && next < n_tiles && (next + 1) * kTE <= n_envs) { if (warp == 0) issue_loads(next * kTE); loaded = true; }
    if (tid < kAccLo) {  // per-thread fp64 accumulation of the cross-env sums (slot = tid, layout PBHC_A_*), envs in index order
      for (int e = 0; e < cnt; ++e) {
        double a = (double)rv[e * kRvT + tid];
        if (tid == V_ETR_SQ) { const double x = (double)rv[e * kRvT + V_ETR]; a = x * x; }
        acc_reg += a;
      }
    }
  }
  if (warp == 0) bulk_wait_read();
  if (tid < PBHC_ACC) io.partials[(int64_t)blockIdx.x * PBHC_ACC + tid] = tid < kAccLo ? acc_reg : 0.0;
}

extern int g_track_info[4];

// host side: slab layout + descriptor tables + launch (called by pbhc_track_step for general == 0)
int track_tile_launch(const PbhcTrackCfg& c, const PbhcTableDesc& tab, const PbhcTrackIO& io, int64_t n, cudaStream_t stream, int* blocks_out) {
  static thread_local TileParams P;
  P.c = c; P.tab = tab; P.io = io;
  if (n > 0x7fffffffLL / 64) { set_error("pbhc_track_step: too many envs for one launch"); return PBHC_E_RANGE; }
  P.n = (int)n; P.n_tiles = (int)((n + kTE - 1) / kTE);
  const int D = c.D, J = c.J, Jb = c.Jb, H = c.hist_row, L = c.L;
  if (c.n_feet > 2) { set_error("pbhc_track_step: at most two feet"); return PBHC_E_RANGE; }
  P.has_air_time = 0;
  for (int r = 0; r < c.n_rewards; ++r) if (c.reward_id[r] == R_FEET_AIR_TIME) P.has_air_time = 1;
  int hrow = 0;
  for (int k = 0; k < PBHC_MAX_HIST_KEYS; ++k) { P.hist_base[k] = hrow; if (k < c.n_hist_keys) hrow += c.hist_len[k] * c.hist_dim[k]; }
  int per_key[PBHC_MAX_HIST_KEYS] = {0};
  P.hist_any_mul = 0; P.n_hd = 0;
  for (int e = 0; e < c.n_hist_dst; ++e) {
    const PbhcHistDst& h = c.hist_dst[e];
    if (++per_key[h.key] > P.n_hd) P.n_hd = per_key[h.key];
    if (h.mul != 1.0f) P.hist_any_mul = 1;
  }
  P.e_pos = 4 * J; P.e_vel = 7 * J; P.e_ang = 10 * J; P.e_dp = 13 * J; P.e_dv = P.e_dp + D; P.e_c = P.e_dv + D;
  auto inv16 = [](int d) { return d > 0 ? (65536 + d - 1) / d : 0; };
  P.n_cols = c.use_vec_reward ? (c.n_rewards + 1 > c.n_episode_sums ? c.n_rewards + 1 : c.n_episode_sums) : (c.n_rewards > c.n_episode_sums ? c.n_rewards : c.n_episode_sums);
  if (P.n_cols < 1) P.n_cols = 1;
  P.n_feat = 16 + (L < 48 ? L : 48);
  if (L > 48) { set_error("pbhc_track_step: more than 48 randomised link masses"); return PBHC_E_RANGE; }
  P.inv_j = inv16(J); P.inv_d = inv16(D); P.inv_r = inv16(P.n_cols); P.inv_f = inv16(P.n_feat);
  // ---- shared-memory layout ----
  int off = 0;
  auto take = [&](int bytes) { const int o = off; off += (bytes + 15) & ~15; return o; };
  P.o_bar = take(16);
  P.o_flags = take(3 * kTE * 4);
  P.o_tab = take((int)sizeof(TileTables));
  P.o_hmap = take(4 * (P.n_hd > 0 ? P.n_hd : 1) * (H > 0 ? H : 1));
  P.o_hmul = P.hist_any_mul ? take(4 * P.n_hd * H) : P.o_hmap;
  P.o_rv = take(4 * kTE * kRvT);
  P.o_pre = take(4 * kTE * kPre);
  P.o_rw = take(c.use_vec_reward ? 16 : 4 * kTE * 32);
  P.n_in = 0; P.n_out = 0; P.in_bytes = 0;
  bool ok = true;
  auto slab_in = [&](const void* g, int row_bytes, int64_t g_stride, int min_bytes = 0) {
    int bytes = kTE * row_bytes;
    const int o = take(bytes > min_bytes ? bytes : min_bytes);
    TileDesc d = {reinterpret_cast<const char*>(g), row_bytes, (int)g_stride, o, 0};
    P.in[P.n_in++] = d;
    P.in_bytes += bytes;
    const bool contig = g_stride == row_bytes;
    if (((uintptr_t)g & 15) || (!contig && ((row_bytes & 15) || (g_stride & 15)))) ok = false;
    return o;
  };
  auto slab_out = [&](void* g, int row_bytes) {
    const int o = take(kTE * row_bytes);
    TileDesc d = {reinterpret_cast<const char*>(g), row_bytes, row_bytes, o, 0};
    P.out[P.n_out++] = d;
    if ((uintptr_t)g & 15) ok = false;
    return o;
  };
  P.i_body = slab_in(io.body_state, Jb * 13 * 4, io.body_env_stride * 4);
  P.i_cf = slab_in(io.contact_forces, Jb * 12, Jb * 12);
  // the history slab doubles as scratch once it is scattered: dof partials + extended-body states
  const int scratch1 = 4 * kTE * (kDofRaw * D + c.n_ext * 13);
  if (H > 0) P.i_hist = slab_in(io.hist, H * 4, H * 4, scratch1);
  else P.i_hist = take(scratch1);
  P.o_dpart = P.i_hist; P.o_ext = P.i_hist + 4 * kTE * kDofRaw * D;
  const int dof_in0 = off;
  P.i_dof = slab_in(io.dof_state, D * 8, D * 8);
  P.i_act = slab_in(io.actions, D * 4, D * 4);
  P.i_tq = slab_in(io.torques, D * 4, D * 4);
  P.i_lact = slab_in(io.last_actions, D * 4, D * 4);
  P.i_ldv = slab_in(io.last_dof_vel, D * 4, D * 4);
  P.i_kp = slab_in(io.kp_scale, D * 4, D * 4);
  P.i_kd = slab_in(io.kd_scale, D * 4, D * 4);
  const int scratch2 = 4 * kTE * kBodyRaw * J;   // body partials live in the dead dof-input slabs
  if (off - dof_in0 < scratch2) take(scratch2 - (off - dof_in0));
  P.o_bpart = dof_in0;
  P.i_es = slab_in(io.episode_sums, c.n_episode_sums * 4, c.n_episode_sums * 4);
  P.i_envf = slab_in(io.env_f32, PBHC_ENVF_STRIDE * 4, PBHC_ENVF_STRIDE * 4);
  P.i_root = slab_in(io.root_state, 52, 52);
  P.i_org = slab_in(io.env_origins, 12, 12);
  P.i_ep = slab_in(io.episode_length_buf, 8, 8);
  P.i_mid = slab_in(io.motion_ids, 8, 8);
  P.i_com = slab_in(io.base_com_bias, 12, 12);
  P.i_fric = slab_in(io.friction, 4, 4);
  P.i_delay = slab_in(io.action_delay_idx, 8, 8);
  P.i_mass = L > 0 ? slab_in(io.link_mass_scale, L * 4, L * 4) : P.i_fric;
  P.o_rows = off;
  for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) {
    P.row_base[g] = (off - P.o_rows) >> 2;
    if (g < c.n_groups) slab_out(io.obs[g], c.group_dim[g] * 4);
  }
  P.q_rew = slab_out(io.rew_buf, c.use_vec_reward ? (c.n_rewards + 1) * 4 : 4);
  P.q_es = slab_out(io.episode_sums, c.n_episode_sums * 4);
  P.q_envf = slab_out(io.env_f32, PBHC_ENVF_STRIDE * 4);
  P.q_ep = slab_out(io.episode_length_buf, 8);
  P.q_reset = slab_out(io.reset_buf, 8);
  P.q_lact = slab_out(io.last_actions, D * 4);
  P.q_ldpos = slab_out(io.last_dof_pos, D * 4);
  P.q_ldvel = slab_out(io.last_dof_vel, D * 4);
  P.bulk_ok = ok ? 1 : 0;
  const size_t shmem = (size_t)off;
  if (shmem > 220 * 1024) { set_error("pbhc_track_step: shared memory %zu B too large", shmem); return PBHC_E_RANGE; }
  bool noise = false;
  for (int k = 0; k < kSrcT; ++k) noise |= c.obs_noise[k] != 0.0f;
  auto kern = noise ? track_tile_kernel<true> : track_tile_kernel<false>;
  static thread_local size_t configured[2] = {0, 0};
  static thread_local int per_sm_of[2] = {0, 0};
  const int variant = noise ? 1 : 0;
  if (shmem > configured[variant] || per_sm_of[variant] == 0) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_track_step(attr)");
    (void)cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured[variant] = shmem;
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kNT, shmem);
    per_sm_of[variant] = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
  }
  int blocks = P.n_tiles;
  const int cap = 148 * per_sm_of[variant];
  if (blocks > cap) blocks = cap;
  g_track_info[0] = 1; g_track_info[1] = blocks; g_track_info[2] = (int)shmem; g_track_info[3] = per_sm_of[variant];
  kern<<<(unsigned)blocks, kNT, shmem, stream>>>(P);
  *blocks_out = blocks;
  return PBHC_OK;
}

}  // namespace pbhc
