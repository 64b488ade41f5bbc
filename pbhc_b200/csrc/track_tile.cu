// track_tile.cu - K3 for LeggedRobotMotionTracking (config A/B): the fused env step as a TILE kernel.
//
// A block owns a tile of kTE = 8 consecutive envs per iteration of a grid-stride loop, one warp per env.  Every per-env
// array of the step ([N, row] tensors) is one contiguous slab for such a tile, so the whole input of a tile (rigid-body
// state, contact forces, dof state, actions, gains, history ring, episode sums, per-env scalars: ~3.9 KB / env) arrives
// in shared memory by ~20 TMA bulk copies (cp.async.bulk + mbarrier, issued by the lanes of one warp from a descriptor
// table) and the whole output (observation rows assembled in their final layout, rewards, episode sums, state:
// ~4.6 KB / env) leaves by ~10 bulk stores; the next tile's loads are issued as soon as the inputs are dead, before the
// stores.  No lane spends an instruction on moving simulator state or observation rows.
//
// Between the two block barriers of a tile (inputs arrived / outputs complete) the warps are independent: warp e runs
// env e start to finish - history ring -> rows, dofs (lane = dof), bodies (lane = body), feet, termination, rewards
// (lane = term), scalar features, state / reset - with per-env reductions through a small scratch row in shared memory
// (three lanes per sum, no 16-value shuffle butterflies).  The work that is scalar per env (frame pair / blend,
// root-frame velocities, projected gravity, heading quaternion) is split into pieces, piece w computed by warp w for
// all 8 envs of the NEXT tile (lane = env) from double-buffered copies of the small per-env inputs, so it is amortised
// over lanes and off the critical path.
//
// Arithmetic: this file is compiled with FMA contraction allowed.  The chains that feed bit-exact results - frame
// indices / blend, motion_end, the gravity and motion_far comparisons, the slerp branch decisions, simulator state
// written at resets - use __fmul_rn / __fadd_rn and reproduce the reference's unfused fp32 sequence; everything else is
// checked against the reference within 1e-5 rel / 1e-6 abs.  See DESIGN.md "arithmetic parity".
// Reference: legged_robot_base.py:297-338, 346-380, 408-489, 715-793, 944-1087; motion_tracking.py:330-358,
// 548-747, 1030-1328; helpers.py:128-152; history_handler.py:33-48 (paths under humanoidverse/envs|utils).
#ifndef __CUDACC_RTC__
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#endif
#include "track_common.cuh"

namespace pbhc {

constexpr int kTE = 8;             // envs per tile = warps per block
constexpr int kNT = kTE * 32;
constexpr int kBodyRaw = 10;       // per-body values of the body sums (already masked by the body's set memberships)
constexpr int kDofRaw = 9;         // per-dof values the dof sums are made of
constexpr int kRvT = 60;           // per-env scalar table: [0, 57) = enum Rv
constexpr int kSrcT = S_BASE_POS_Z + 1;  // observation sources of LeggedRobotMotionTracking
enum DofRaw { DR_DJA2, DR_DJV2, DR_TQ2, DR_DACT2, DR_DV2, DR_ACC2, DR_LPOS, DR_LVEL, DR_LTQ };
constexpr int kBodySums = 10;
constexpr int kPieces = 5;         // pieces of the per-env scalar pre-computation


// Launch constants of the hot loop.  The generic (ahead-of-time) kernel reads them from the launch parameters; the
// specialised kernel (jit.cu: the same source compiled by NVRTC for one env config) sees them as compile-time constants
// from a generated header, so shapes, shared-memory offsets and config flags fold into immediates and dead branches.
#define PBHC_HOTK_INT(X) \
  X(J) X(Jb) X(D) X(L) X(H) X(n_ext) X(n_feet) X(n_groups) X(R) X(n_sums) X(n_hd) X(hist_any_mul) X(has_air_time) X(n_feat) \
  X(contact_size) X(rec_stride) X(use_vec_reward) X(only_positive) X(has_termination) X(cur_torque_enable) X(term_gravity) \
  X(term_motion_far) X(term_motion_end) X(term_contact) X(term_low_height) X(term_dof_pos) X(term_dof_vel) X(term_torque) X(term_base) X(push_robots) X(dif_joint_used) X(ref_vel_used) X(ext_identity) X(ref_pos_out) X(ref_rot_out) X(default_dof_stride) \
  X(e_pos) X(e_vel) X(e_ang) X(e_dp) X(e_dv) X(e_c) \
  X(o_tab) X(o_stab) X(n_slots) X(o_hmap) X(o_hmul) X(o_rv) X(o_pre) X(o_rows) X(o_scr) X(scr_pitch) \
  X(i_body) X(i_cf) X(i_hist) X(i_dof) X(i_act) X(i_tq) X(i_lact) X(i_ldv) X(i_kp) X(i_kd) X(i_es0) X(i_es1) X(i_org0) X(i_org1) X(i_com0) X(i_com1) X(i_fric0) X(i_fric1) \
  X(i_delay0) X(i_delay1) X(i_mass0) X(i_mass1) X(i_root0) X(i_root1) X(i_ep0) X(i_ep1) X(i_envf0) X(i_envf1) X(i_mid0) X(i_mid1) \
    X(row_base0) X(row_base1) X(row_base2) X(row_base3) X(gdim0) X(gdim1) X(gdim2) X(gdim3) X(ghist0) X(ghist1) X(ghist2) X(ghist3)
#define PBHC_HOTK_FLT(X) X(dt) X(clip_obs) X(term_gravity_thr) X(term_min_height) X(term_dof_vel_scale) X(term_torque_scale) X(w_lower) X(w_upper) X(termination_scale) X(desired_feet_air_time) X(max_contact_force)
#define PBHC_HOTK_I64(X) X(max_episode_length) X(penalised_mask) X(term_contact_mask)
struct HotK {
#define X(n) int n;
  PBHC_HOTK_INT(X)
#undef X
#define X(n) float n;
  PBHC_HOTK_FLT(X)
#undef X
#define X(n) long long n;
  PBHC_HOTK_I64(X)
#undef X
};

// where each observation source goes - part of the specialised kernel's constants too (the emits of a known source fold
// into stores at immediate offsets); the generic kernel reads the same from a shared-memory table
struct SpecExtra { int n[20], base[20][4], stride[20][4], pushed[20], noisy[20]; float scale[20]; };

struct TileParams {
  PbhcTrackCfg c;
  PbhcTableDesc tab;
  PbhcTrackIO io;
  HotK k;
  SpecExtra x;
  int n, n_tiles;
  int hist_base[PBHC_MAX_HIST_KEYS];
};

#ifdef PBHC_SPEC
#include "track_spec.h"   // namespace spec { constexpr ... } generated for one env config (jit.cu)
#define KC(n) (spec::n)
#else
#define KC(n) (P.k.n)
#endif

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
// where one observation source goes: up to four obs groups {slab offset, env stride} and its history push slot
struct alignas(16) SrcDst { int base0, stride0, base1, stride1, base2, stride2, base3, stride3; int n, push, grp, pad; };
struct TileTables {
  float sigma[PBHC_NUM_SIGMA];
  float far_thr;
  uint32_t push_count;
  uint32_t noise_mask;
  float lim_lo[PBHC_MAX_DOF], lim_hi[PBHC_MAX_DOF], lim_vel[PBHC_MAX_DOF], lim_tq[PBHC_MAX_DOF];
  float t_vel[PBHC_MAX_DOF], t_tq[PBHC_MAX_DOF];  // termination limits: dof_vel_limits / torque_limits x termination scale (legged_robot_base.py:459, 470)
  int term_gates;                                 // bit 0 / 1 / 2: the dof_pos / dof_vel / torque termination applies in this step
  float r_scale[PBHC_MAX_REWARDS], r_pen[PBHC_MAX_REWARDS], r_sigma[PBHC_MAX_REWARDS];
  signed char r_kind[PBHC_MAX_REWARDS], r_arg[PBHC_MAX_REWARDS];
  signed char bs_slot[kBodySums + 2];
  unsigned char member[64];
  float bs_div[kBodySums];
  SrcDst sd[kSrcT];
  float obs_scale[kSrcT], obs_noise[kSrcT];
  unsigned char feat_src[64], feat_el[64];   // scalar feature i of an env: source and element
};

struct NoiseKeyT { uint64_t seed; int64_t env; uint32_t step; };
struct SmallSlot { const char* src; int row_bytes; int dst; };  // one 4-byte word of a late-read per-env array: global word of env 0, row size, shared-memory byte offset (row 0)
struct Rows { float* rows; float* ghist; int e; };   // observation slabs (shared), the env's global history row, tile-local env

__device__ __forceinline__ float clampf(float v, float c) { return fminf(fmaxf(v, -c), c); }


// (value + U(-1,1) * noise) * scale with an independent draw per destination (helpers.py:128-152)
__device__ __noinline__ void emit_noisy_t(const TileTables& T, const Rows& R, int k, int el, float raw, float clip, const NoiseKeyT& nk) {
  const Rng rng = {make_uint2((uint32_t)nk.seed, (uint32_t)(nk.seed >> 32)), (uint32_t)nk.env, (uint32_t)((uint64_t)nk.env >> 32), nk.step};
  const uint32_t st = 4096u + ((uint32_t)k * 256u + (uint32_t)el) * 2u;
  const uint4 a = rng.draw(st);
  const uint32_t b = rng.draw(st + 1u).x;
  const float ns = T.obs_noise[k], sc = T.obs_scale[k];
  const SrcDst& d = T.sd[k];
  const int base[4] = {d.base0, d.base1, d.base2, d.base3}, stride[4] = {d.stride0, d.stride1, d.stride2, d.stride3};
  for (int i = 0; i < d.n; ++i) {
    const int g = (d.grp >> (8 * i)) & 0xff;
    const uint32_t bits = g == 0 ? a.x : g == 1 ? a.y : g == 2 ? a.z : a.w;
    R.rows[base[i] + R.e * stride[i] + el] = clampf((raw + (u01(bits) * 2.0f - 1.0f) * ns) * sc, clip);
  }
  if (d.push >= 0) R.ghist[d.push + el] = (raw + (u01(b) * 2.0f - 1.0f) * ns) * sc;
}
// one element of observation source k to every row that shows it (clipped) and to the history ring (unclipped)
template <bool kNoise>
__device__ __forceinline__ void emit1(const TileTables& T, const Rows& R, int k, int el, float raw, float clip, const NoiseKeyT& nk) {
  if (kNoise && ((T.noise_mask >> k) & 1u)) { emit_noisy_t(T, R, k, el, raw, clip, nk); return; }
  const SrcDst& d = T.sd[k];
  const int4 a = *reinterpret_cast<const int4*>(&d.base0);
  const int2 np = *reinterpret_cast<const int2*>(&d.n);
  const float sv = raw * T.obs_scale[k], cv = clampf(sv, clip);
  float* r = R.rows + el;
  if (np.x > 0) r[a.x + R.e * a.y] = cv;
  if (np.x > 1) r[a.z + R.e * a.w] = cv;
  if (np.x > 2) { r[d.base2 + R.e * d.stride2] = cv; if (np.x > 3) r[d.base3 + R.e * d.stride3] = cv; }
  if (np.y >= 0) R.ghist[np.y + el] = sv;
}
template <bool kNoise>
__device__ __forceinline__ void emit3(const TileTables& T, const Rows& R, int k, int el, float x, float y, float z, float clip, const NoiseKeyT& nk) {
  if (kNoise && ((T.noise_mask >> k) & 1u)) {
    emit_noisy_t(T, R, k, el, x, clip, nk); emit_noisy_t(T, R, k, el + 1, y, clip, nk); emit_noisy_t(T, R, k, el + 2, z, clip, nk);
    return;
  }
  const SrcDst& d = T.sd[k];
  const int4 a = *reinterpret_cast<const int4*>(&d.base0);
  const int2 np = *reinterpret_cast<const int2*>(&d.n);
  const float sc = T.obs_scale[k];
  const float sx = x * sc, sy = y * sc, sz = z * sc;
  const float cx = clampf(sx, clip), cy = clampf(sy, clip), cz = clampf(sz, clip);
  float* r = R.rows + el;
  if (np.x > 0) { float* o = r + a.x + R.e * a.y; o[0] = cx; o[1] = cy; o[2] = cz; }
  if (np.x > 1) { float* o = r + a.z + R.e * a.w; o[0] = cx; o[1] = cy; o[2] = cz; }
  if (np.x > 2) {
    { float* o = r + d.base2 + R.e * d.stride2; o[0] = cx; o[1] = cy; o[2] = cz; }
    if (np.x > 3) { float* o = r + d.base3 + R.e * d.stride3; o[0] = cx; o[1] = cy; o[2] = cz; }
  }
  if (np.y >= 0) { float* h = R.ghist + np.y + el; h[0] = sx; h[1] = sy; h[2] = sz; }
}

// the same for a source known at compile time: in the specialised kernel the destinations are immediates
template <bool kNoise, int K>
__device__ __forceinline__ void emit1k(const TileTables& T, const Rows& R, int el, float raw, float clip, const NoiseKeyT& nk) {
#ifdef PBHC_SPEC
  if (kNoise && spec::sd_noisy[K]) { emit_noisy_t(T, R, K, el, raw, clip, nk); return; }
  const float sv = spec::sd_scale[K] == 1.0f ? raw : raw * spec::sd_scale[K], cv = clampf(sv, clip);
  float* r = R.rows + el;
  if (spec::sd_n[K] > 0) r[spec::sd_base[K][0] + R.e * spec::sd_stride[K][0]] = cv;
  if (spec::sd_n[K] > 1) r[spec::sd_base[K][1] + R.e * spec::sd_stride[K][1]] = cv;
  if (spec::sd_n[K] > 2) r[spec::sd_base[K][2] + R.e * spec::sd_stride[K][2]] = cv;
  if (spec::sd_n[K] > 3) r[spec::sd_base[K][3] + R.e * spec::sd_stride[K][3]] = cv;
  if (spec::sd_pushed[K]) R.ghist[T.sd[K].push + el] = sv;
#else
  emit1<kNoise>(T, R, K, el, raw, clip, nk);
#endif
}
template <bool kNoise, int K>
__device__ __forceinline__ void emit3k(const TileTables& T, const Rows& R, int el, float x, float y, float z, float clip, const NoiseKeyT& nk) {
#ifdef PBHC_SPEC
  if (kNoise && spec::sd_noisy[K]) { emit3<kNoise>(T, R, K, el, x, y, z, clip, nk); return; }
  if (spec::sd_n[K] == 0 && !spec::sd_pushed[K]) return;
  const float sc = spec::sd_scale[K];
  const float sx = x * sc, sy = y * sc, sz = z * sc;
  const float cx = clampf(sx, clip), cy = clampf(sy, clip), cz = clampf(sz, clip);
  float* r = R.rows + el;
  if (spec::sd_n[K] > 0) { float* o = r + spec::sd_base[K][0] + R.e * spec::sd_stride[K][0]; o[0] = cx; o[1] = cy; o[2] = cz; }
  if (spec::sd_n[K] > 1) { float* o = r + spec::sd_base[K][1] + R.e * spec::sd_stride[K][1]; o[0] = cx; o[1] = cy; o[2] = cz; }
  if (spec::sd_n[K] > 2) { float* o = r + spec::sd_base[K][2] + R.e * spec::sd_stride[K][2]; o[0] = cx; o[1] = cy; o[2] = cz; }
  if (spec::sd_n[K] > 3) { float* o = r + spec::sd_base[K][3] + R.e * spec::sd_stride[K][3]; o[0] = cx; o[1] = cy; o[2] = cz; }
  if (spec::sd_pushed[K]) { float* h = R.ghist + T.sd[K].push + el; h[0] = sx; h[1] = sy; h[2] = sz; }
#else
  emit3<kNoise>(T, R, K, el, x, y, z, clip, nk);
#endif
}

// ---- cp.async (LDGSTS) ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory"); }
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src) { asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory"); }
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) { asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---- TMA / mbarrier ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, int bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
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

__device__ void build_tile_tables(const TileParams& P, TileTables& T, short* hmap, float* hmul, const double* scal_g, double* scal) {
  const PbhcTrackCfg& c = P.c;
  const int tid = threadIdx.x, nt = blockDim.x;
  auto row_base = [&](int g) { return g == 0 ? KC(row_base0) : g == 1 ? KC(row_base1) : g == 2 ? KC(row_base2) : KC(row_base3); };
  auto ghist = [&](int g) { return g == 0 ? KC(ghist0) : g == 1 ? KC(ghist1) : g == 2 ? KC(ghist2) : KC(ghist3); };
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
    T.t_vel[tid] = xm(c.vel_limit[tid], c.term_dof_vel_scale);
    T.t_tq[tid] = xm(c.torque_limit[tid], c.term_torque_scale);
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
        default: break;
      }
      if (kind != K_RAW) sig = (float)scal[PBHC_S_SIGMA + arg];
    }
    T.r_kind[tid] = (signed char)kind; T.r_arg[tid] = (signed char)arg; T.r_sigma[tid] = sig;
    T.r_scale[tid] = tid < c.n_rewards ? c.reward_scale[tid] : 0.0f;
    T.r_pen[tid] = (tid < c.n_rewards && c.reward_is_penalty[tid]) ? (float)scal[PBHC_S_PENALTY] : 1.0f;
  }
  if (tid < kBodySums) {  // body sum k: divisor and rv slot (motion_tracking.py:646-676, 397-418)
    int slot = 0; float div = 1.0f;
    switch (tid) {
      case 0: div = (float)c.n_upper; slot = V_SIGMA_ERR + G_UPPER; break;
      case 1: div = (float)c.n_lower; slot = V_SIGMA_ERR + G_LOWER; break;
      case 2: div = (float)c.n_vr3; slot = V_SIGMA_ERR + G_VR3; break;
      case 3: div = (float)c.n_feet; slot = V_SIGMA_ERR + G_FEET; break;
      case 4: div = (float)c.J; slot = V_SIGMA_ERR + G_ROT; break;
      case 5: div = (float)c.J; slot = V_SIGMA_ERR + G_VEL; break;
      case 6: div = (float)c.J; slot = V_SIGMA_ERR + G_ANG_VEL; break;
      case 7: slot = V_UP_N; break;
      case 8: slot = V_LO_N; break;
      default: slot = V_VR_N; break;
    }
    T.bs_slot[tid] = (signed char)slot; T.bs_div[tid] = div;
  }
  if (tid < 64) {  // set memberships of body tid: bit 0 upper, 1 lower, 2 vr3, 3 feet
    int mb = 0;
    if ((c.upper_mask >> tid) & 1ull) mb |= 1;
    if ((c.lower_mask >> tid) & 1ull) mb |= 2;
    if ((c.vr3_mask >> tid) & 1ull) mb |= 4;
    for (int f = 0; f < c.n_feet; ++f) if (c.feet_ids[f] == tid) mb |= 8;
    T.member[tid] = (unsigned char)mb;
  }
  if (tid < kSrcT) {
    int base[4] = {0, 0, 0, 0}, stride[4] = {0, 0, 0, 0};
    SrcDst d;
    d.n = 0; d.push = -1; d.grp = 0; d.pad = 0;
    for (int g = 0; g < c.n_groups; ++g)
      if (c.src_dst[tid][g] >= 0) { base[d.n] = row_base(g) + c.src_dst[tid][g]; stride[d.n] = c.group_dim[g]; d.grp |= g << (8 * d.n); ++d.n; }
    d.base0 = base[0]; d.stride0 = stride[0]; d.base1 = base[1]; d.stride1 = stride[1];
    d.base2 = base[2]; d.stride2 = stride[2]; d.base3 = base[3]; d.stride3 = stride[3];
    for (int h = 0; h < c.n_hist_keys; ++h)
      if (c.hist_src[h] == tid) {  // ring push: one slot per key (HistoryHandler.add, history_handler.py:40-44)
        const int len = c.hist_len[h];
        d.push = P.hist_base[h] + (int)((len - ((push_count + 1) % len)) % len) * c.hist_dim[h];
        break;
      }
    T.sd[tid] = d;
    T.obs_scale[tid] = c.obs_scale[tid];
    T.obs_noise[tid] = c.cur_noise.enable ? (float)((double)c.obs_noise[tid] * scal[PBHC_S_NOISE_CUR]) : c.obs_noise[tid];
  }
  if (tid == 0) {
    uint32_t m = 0;
    for (int k = 0; k < kSrcT; ++k) if (c.obs_noise[k] != 0.0f) m |= 1u << k;
    T.noise_mask = m;
    // scalar features of an env (legged_robot_base.py:1114-1215, motion_tracking.py:944-1015)
    int n = 0;
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_BASE_LIN_VEL; T.feat_el[n++] = (unsigned char)i; }
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_BASE_ANG_VEL; T.feat_el[n++] = (unsigned char)i; }
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_PROJ_GRAV; T.feat_el[n++] = (unsigned char)i; }
    for (int i = 0; i < 3; ++i) { T.feat_src[n] = S_DR_COM; T.feat_el[n++] = (unsigned char)i; }
    T.feat_src[n] = S_PHASE; T.feat_el[n++] = 0;
    T.feat_src[n] = S_DR_FRICTION; T.feat_el[n++] = 0;
    T.feat_src[n] = S_DR_DELAY; T.feat_el[n++] = 0;
    T.feat_src[n] = S_BASE_POS_Z; T.feat_el[n++] = 0;
    for (int i = 0; i < c.L && n < 64; ++i) { T.feat_src[n] = S_DR_MASS; T.feat_el[n++] = (unsigned char)i; }
  }
  // history ring: physical slot of logical slot s of key k is (head_k + s) % len_k, head_k = -push_count mod len_k
  // (HistoryHandler keeps newest-first order, history_handler.py:40-48).  One plane per obs group that shows history:
  // hmap[plane][i] = float offset (inside the rows region, env 0) where physical ring element i goes in that group, -1 = none
  // (16-bit: the rows region of a tile is far below 32768 floats - checked by tile_plan).
  for (int i = tid; i < c.hist_row; i += nt) {
    int k = 0;
    while (k + 1 < c.n_hist_keys && i >= P.hist_base[k + 1]) ++k;
    const int len = c.hist_len[k], dim = c.hist_dim[k];
    const int r = i - P.hist_base[k], ps = r / dim, cc = r - ps * dim;
    const int head = (int)((len - (push_count % len)) % len);
    const int s = (ps - head + len) % len;  // logical slot
    int plane = 0;
    for (int g = 0; g < c.n_groups; ++g) {
      if (!ghist(g)) continue;
      int m = -1; float mul = 1.0f;
      for (int e = 0; e < c.n_hist_dst; ++e) {
        const PbhcHistDst& h = c.hist_dst[e];
        if (h.key == k && h.group == g && s < h.len) { m = row_base(g) + h.offset + s * dim + cc; mul = h.mul; }
      }
      hmap[plane * c.hist_row + i] = (short)m;
      if (KC(hist_any_mul)) hmul[plane * c.hist_row + i] = mul;
      ++plane;
    }
  }
  __syncthreads();
}

template <bool kNoise>
__device__ __forceinline__ void track_tile_body(const TileParams& P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const PbhcTrackCfg& c = P.c;
  const PbhcTrackIO& io = P.io;
  const PbhcTableDesc& tab = P.tab;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  TileTables& T = *reinterpret_cast<TileTables*>(smem + KC(o_tab));
  short* hmap = reinterpret_cast<short*>(smem + KC(o_hmap));
  float* hmul = reinterpret_cast<float*>(smem + KC(o_hmul));
  float* rv = reinterpret_cast<float*>(smem + KC(o_rv));            // [kTE][kRvT]
  float* rows = reinterpret_cast<float*>(smem + KC(o_rows));
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);

  for (int i = tid; i < kTE * kRvT; i += kNT) rv[i] = 0.0f;  // slots nobody writes are accumulated too
  {  // slot table of the late-read arrays (both buffers)
    constexpr int kArr = 10;
    const void* const ag[kArr] = {io.episode_sums, io.env_origins, io.base_com_bias, io.friction, io.action_delay_idx, io.link_mass_scale,
                                  io.root_state, io.episode_length_buf, io.env_f32, io.motion_ids};
    const int ab[kArr] = {KC(n_sums) * 4, 12, 12, 4, 8, KC(L) * 4, 52, 8, PBHC_ENVF_STRIDE * 4, 8};
    const int a0[kArr] = {KC(i_es0), KC(i_org0), KC(i_com0), KC(i_fric0), KC(i_delay0), KC(i_mass0), KC(i_root0), KC(i_ep0), KC(i_envf0), KC(i_mid0)};
    const int a1[kArr] = {KC(i_es1), KC(i_org1), KC(i_com1), KC(i_fric1), KC(i_delay1), KC(i_mass1), KC(i_root1), KC(i_ep1), KC(i_envf1), KC(i_mid1)};
    SmallSlot* st = reinterpret_cast<SmallSlot*>(smem + KC(o_stab));
    for (int i = tid; i < 2 * KC(n_slots); i += kNT) {
      const int buf = i >= KC(n_slots) ? 1 : 0;
      int w = i - buf * KC(n_slots), a = 0;
      while (a < kArr - 1 && w >= (ab[a] >> 2)) { w -= ab[a] >> 2; ++a; }
      st[i] = {reinterpret_cast<const char*>(ag[a]) + 4 * w, ab[a], (buf ? a1[a] : a0[a]) + 4 * w};
    }
  }
  build_tile_tables(P, T, hmap, hmul, io.scalars, reinterpret_cast<double*>(smem + KC(o_rows)));  // (the rows region is idle: scratch for the scalars)

  const int J = KC(J), Jb = KC(Jb), D = KC(D), H = KC(H);
  const int R = KC(R);
  const float clipv = KC(clip_obs), dt = KC(dt);
  const int e_pos = KC(e_pos), e_vel = KC(e_vel), e_ang = KC(e_ang), e_dp = KC(e_dp), e_dv = KC(e_dv), e_c = KC(e_c);
  const float far_thr = T.far_thr;
  const uint32_t push_count = T.push_count;
  double acc_reg = 0.0;  // slot = lane (< kAccLo), this warp's envs
  double acc_hi = 0.0;   // slot = 32 + lane: the base-class termination counts (PBHC_A_TERM_CONTACT ..)

  // Every warp runs its own pipeline over its own envs (row `e` = warp of every slab): no block barrier inside the loop.
  // Inputs arrive by cp.async (LDGSTS: 16 / 8 / 4 bytes per lane, whatever the row size allows - rows of 23 floats are
  // not 16-byte multiples): the rows read only by sections a-d (simulator state, history ring, dof-wise arrays) are
  // single-buffered and re-requested for the warp's next env as soon as section d is done, the ones read until the
  // end of an env are double-buffered; both arrive while sections e-i and the row write-out of the current env run.
  const int e = warp;
  const int64_t n_envs = P.n, step = (int64_t)gridDim.x * kTE;
  const SmallSlot* stab = reinterpret_cast<const SmallSlot*>(smem + KC(o_stab));
  auto row_load = [&](int soff, int bytes, const void* g, int64_t m, int64_t g_row_bytes) {
    const uint32_t dst = sbase + soff + e * bytes;
    const char* src = reinterpret_cast<const char*>(g) + m * g_row_bytes;
    if (((bytes | (int)g_row_bytes) & 15) == 0) {
      const uint32_t d = dst + lane * 16; const char* p = src + lane * 16;
      for (int o = 0; o < bytes; o += 512) if (o + lane * 16 < bytes) cp_async16(d + o, p + o);
    } else if (((bytes | (int)g_row_bytes) & 7) == 0) {
      const uint32_t d = dst + lane * 8; const char* p = src + lane * 8;
      for (int o = 0; o < bytes; o += 256) if (o + lane * 8 < bytes) cp_async8(d + o, p + o);
    } else {
      const uint32_t d = dst + lane * 4; const char* p = src + lane * 4;
      for (int o = 0; o < bytes; o += 128) if (o + lane * 4 < bytes) cp_async4(d + o, p + o);
    }
  };
  auto request = [&](int64_t m, int buf) {  // env m's inputs -> this warp's rows (late-read ones into buffer `buf`)
    row_load(KC(i_body), Jb * 52, io.body_state, m, io.body_env_stride * 4);
    row_load(KC(i_cf), Jb * 12, io.contact_forces, m, Jb * 12);
    if (H > 0) row_load(KC(i_hist), H * 4, io.hist, m, H * 4);
    row_load(KC(i_dof), D * 8, io.dof_state, m, D * 8);
    {  // the six dof-wise arrays: equal rows, equally sized consecutive slabs (tile_plan) - one offset serves all of them
      const int64_t o = m * (D * 4) + lane * 4;
      const uint32_t d = sbase + KC(i_act) + e * (D * 4) + lane * 4;
      constexpr int kDofWise = 6;
      const float* const g[kDofWise] = {io.actions, io.torques, io.last_actions, io.last_dof_vel, io.kp_scale, io.kd_scale};
      if (lane < D) {
#pragma unroll
        for (int a = 0; a < kDofWise; ++a) cp_async4(d + a * (kTE * D * 4), reinterpret_cast<const char*>(g[a]) + o);
      }
    }
    // the late-read arrays, a 4-byte word per lane and round through the slot table (rewards' episode sums, origins, domain-
    // randomisation scalars, root state, episode length, env scalars, motion id)
    for (int sl = lane; sl < KC(n_slots); sl += 32) {
      const SmallSlot t = stab[buf * KC(n_slots) + sl];
      cp_async4(sbase + t.dst + e * t.row_bytes, t.src + m * t.row_bytes);
    }
    cp_async_commit();
  };
  // the per-env scalars of the env in buffer `buf`, lane = piece (legged_robot_base.py:346-353, motion_tracking.py:548-558, 678-739)
  auto pre_pieces = [&](int buf) {
    if (lane >= kPieces) return;
    float* pe = reinterpret_cast<float*>(smem + KC(o_pre)) + e * kPre;
    const float* root = reinterpret_cast<const float*>(smem + (buf ? KC(i_root1) : KC(i_root0))) + e * 13;
    const Q4 rq = {root[3], root[4], root[5], root[6]};
    if (lane < 3) {  // base linear velocity, base angular velocity, projected gravity: one code, three vectors (the
      // reference's own operation sequence - the gravity feeds the termination check)
      const bool gv = lane == 2;
      const V3 v = {gv ? 0.0f : root[7 + 3 * lane], gv ? 0.0f : root[8 + 3 * lane], gv ? -1.0f : root[9 + 3 * lane]};
      const V3 o = quat_rotate_x(rq, v, true);
      float* d = pe + P_BLV + 3 * lane;
      d[0] = o.x; d[1] = o.y; d[2] = o.z;
      if (gv) { pe[P_GRAVN] = norm2(o.x, o.y); pe[P_ORIENT] = o.x * o.x + o.y * o.y; }
      if (lane == 0) { pe[P_RPX] = root[0]; pe[P_RPY] = root[1]; pe[P_RPZ] = root[2]; }
    } else if (lane == 3) {
      const Q4 hinv = heading_quat_inv_fast(rq);
      pe[P_HZ] = hinv.z; pe[P_HW] = hinv.w;
    } else {
      const int64_t ep = reinterpret_cast<const long long*>(smem + (buf ? KC(i_ep1) : KC(i_ep0)))[e] + 1;  // legged_robot_base.py:299
      const float start = reinterpret_cast<const float*>(smem + (buf ? KC(i_envf1) : KC(i_envf0)))[e * PBHC_ENVF_STRIDE];
      const int clip = resolve_clip(tab, io.clip_of_motion, reinterpret_cast<const long long*>(smem + (buf ? KC(i_mid1) : KC(i_mid0)))[e]);
      const float t = xa(xm((float)(ep + 1), dt), start);  // kick_motion_res, motion_tracking.py:554
      const FrameBlend fb = frame_blend_x(tab, clip, t);
      pe[P_F0] = __int_as_float((int)fb.f0); pe[P_F1] = __int_as_float((int)fb.f1); pe[P_BLEND] = fb.blend;
      pe[P_PHASE] = __fdiv_rn(t, tab.clip_length[clip]);  // motion_tracking.py:738-739
      pe[P_CLIP] = __int_as_float(clip);
    }
  };

  int64_t n = (int64_t)blockIdx.x * kTE + warp;
  if (n < n_envs) request(n, 0);
  for (int sb = 0; n < n_envs; n += step, sb ^= 1) {
    cp_async_wait_all();
    __syncwarp();
    pre_pieces(sb);
    __syncwarp();
    // ================================ warp = env =====================================================================
    {
      const float* pe = reinterpret_cast<const float*>(smem + KC(o_pre)) + e * kPre;
      const float* s_envf = reinterpret_cast<const float*>(smem + (sb ? KC(i_envf1) : KC(i_envf0))) + e * PBHC_ENVF_STRIDE;
      const int64_t ep_in = reinterpret_cast<const long long*>(smem + (sb ? KC(i_ep1) : KC(i_ep0)))[e];
      const float* s_body = reinterpret_cast<const float*>(smem + KC(i_body)) + e * (Jb * 13);
      const float* s_cf = reinterpret_cast<const float*>(smem + KC(i_cf)) + e * (Jb * 3);
      const float* s_org = reinterpret_cast<const float*>(smem + (sb ? KC(i_org1) : KC(i_org0))) + e * 3;
      float* scr = reinterpret_cast<float*>(smem + KC(o_scr) + e * KC(scr_pitch));
      float* rve = rv + e * kRvT;
      const Rows Rw = {rows, io.hist + n * H, e};
      const NoiseKeyT nk = {c.seed, n + io.env_offset, push_count};
      const float b = pe[P_BLEND], omb = 1.0f - b;
      const float* r0 = tab.records + (int64_t)__float_as_int(pe[P_F0]) * KC(rec_stride);
      const float* r1 = tab.records + (int64_t)__float_as_int(pe[P_F1]) * KC(rec_stride);

      // ---- a. history ring -> its places in the rows (HistoryHandler.query, history_handler.py:46-48) ------------------------
      {
        const float* sh = reinterpret_cast<const float*>(smem + KC(i_hist)) + e * H;
        if ((H & 3) == 0 && !KC(hist_any_mul)) {  // four ring elements per lane and pass
          auto plane4 = [&](int has, int plane, int gd, int i, const float4& v) {
            if (!has) return;
            const short4 m = reinterpret_cast<const short4*>(hmap)[plane * (H >> 2) + i];
            float* r = rows + e * gd;
            if (m.x >= 0) r[m.x] = v.x;
            if (m.y >= 0) r[m.y] = v.y;
            if (m.z >= 0) r[m.z] = v.z;
            if (m.w >= 0) r[m.w] = v.w;
          };
          for (int i = lane; i < (H >> 2); i += 32) {
            float4 v = reinterpret_cast<const float4*>(sh)[i];
            v.x = clampf(v.x, clipv); v.y = clampf(v.y, clipv); v.z = clampf(v.z, clipv); v.w = clampf(v.w, clipv);
            plane4(KC(ghist0), 0, KC(gdim0), i, v);
            plane4(KC(ghist1), KC(ghist0), KC(gdim1), i, v);
            plane4(KC(ghist2), KC(ghist0) + KC(ghist1), KC(gdim2), i, v);
            plane4(KC(ghist3), KC(ghist0) + KC(ghist1) + KC(ghist2), KC(gdim3), i, v);
          }
        } else {
          auto plane1 = [&](int has, int plane, int gd, int i, float v) {
            if (!has) return;
            const int m = hmap[plane * H + i];
            if (m >= 0) rows[m + e * gd] = clampf(KC(hist_any_mul) ? v * hmul[plane * H + i] : v, clipv);
          };
          for (int i = lane; i < H; i += 32) {
            const float v = sh[i];
            plane1(KC(ghist0), 0, KC(gdim0), i, v);
            plane1(KC(ghist1), KC(ghist0), KC(gdim1), i, v);
            plane1(KC(ghist2), KC(ghist0) + KC(ghist1), KC(gdim2), i, v);
            plane1(KC(ghist3), KC(ghist0) + KC(ghist1) + KC(ghist2), KC(gdim3), i, v);
          }
        }
      }
      __syncwarp();  // (the scratch row may alias this env's history row)

      // ---- b. dofs (lane = dof) -------------------------------------------------------------------------------------------
      float adja = 0.0f;  // |dif_joint_angles| of this lane's dof
      int dof_viol = 0;   // bit 0 / 1 / 2: this dof is beyond its position / velocity / torque termination limit
      if (lane < D) {
        const int d = lane, x = e * D + d;
        const float2 ds = reinterpret_cast<const float2*>(smem + KC(i_dof))[x];
        const float dpos = ds.x, dvel = ds.y;
        const float act = reinterpret_cast<const float*>(smem + KC(i_act))[x], tq = reinterpret_cast<const float*>(smem + KC(i_tq))[x];
        const float lact = reinterpret_cast<const float*>(smem + KC(i_lact))[x], ldv = reinterpret_cast<const float*>(smem + KC(i_ldv))[x];
        const float ddef = io.default_dof_pos[n * KC(default_dof_stride) + d];
        const float dja = (omb * __ldg(r0 + e_dp + d) + b * __ldg(r1 + e_dp + d)) - dpos;   // motion_tracking.py:661-665
        adja = fabsf(dja);
        const float djv = (omb * __ldg(r0 + e_dv + d) + b * __ldg(r1 + e_dv + d)) - dvel;
        const float dact = lact - act, dacc = (ldv - dvel) * (1.0f / dt);  // (tolerance-checked penalty term; the division's slow path is taken for a zero difference)
        const float lim_t = T.lim_tq[d];
        if (KC(term_base)) {  // legged_robot_base.py:449-476: sum over dofs of the clipped excess > 0  <=>  some dof exceeds
          if (KC(term_dof_pos) && (dpos < c.term_pos_lo[d] || dpos > c.term_pos_hi[d])) dof_viol |= 1;
          if (KC(term_dof_vel) && fabsf(dvel) > T.t_vel[d]) dof_viol |= 2;
          if (KC(term_torque) && fabsf(tq) > T.t_tq[d]) dof_viol |= 4;
        }
        float* dp_ = scr + d;
        dp_[DR_DJA2 * D] = dja * dja; dp_[DR_DJV2 * D] = djv * djv; dp_[DR_TQ2 * D] = tq * tq; dp_[DR_DACT2 * D] = dact * dact;
        dp_[DR_DV2 * D] = dvel * dvel; dp_[DR_ACC2 * D] = dacc * dacc;
        dp_[DR_LPOS * D] = -fminf(dpos - T.lim_lo[d], 0.0f) + fmaxf(dpos - T.lim_hi[d], 0.0f);  // legged_robot_base.py:974-1002
        dp_[DR_LVEL * D] = fminf(fmaxf(fabsf(dvel) - T.lim_vel[d], 0.0f), 1.0f);
        dp_[DR_LTQ * D] = KC(cur_torque_enable) ? fminf(fmaxf(fabsf(tq) - lim_t, 0.0f), 1.0f) : fmaxf(fabsf(tq) - lim_t, 0.0f);
        emit1k<kNoise, S_DOF_POS>(T, Rw, d, dpos - ddef, clipv, nk);
        emit1k<kNoise, S_DOF_VEL>(T, Rw, d, dvel, clipv, nk);
        emit1k<kNoise, S_ACTIONS>(T, Rw, d, act, clipv, nk);
        emit1k<kNoise, S_DR_KP>(T, Rw, d, reinterpret_cast<const float*>(smem + KC(i_kp))[x], clipv, nk);
        emit1k<kNoise, S_DR_KD>(T, Rw, d, reinterpret_cast<const float*>(smem + KC(i_kd))[x], clipv, nk);
        if (KC(dif_joint_used)) {
          emit1k<kNoise, S_DIF_JOINT>(T, Rw, d, dja, clipv, nk);
          emit1k<kNoise, S_DIF_JOINT_VEL>(T, Rw, d, djv, clipv, nk);
        }
        // _post_compute_observations_callback, legged_robot_base.py:398-405
        io.last_actions[n * D + d] = act; io.last_dof_pos[n * D + d] = dpos; io.last_dof_vel[n * D + d] = dvel;  // (small per-env results go straight to global: a bulk copy per array costs its issuer more than these stores)
      }
      {
        const float mx = warp_max(adja);
        if (lane == 0) rve[V_SIGMA_ERR + G_MAX_JOINT] = mx;
      }
      __syncwarp();
      {  // dof sums: three lanes per value (dofs part, part + 3, ...), combined in lane order - a fixed summation order
        const int k = lane / 3, part = lane - 3 * k;
        float v = 0.0f;
        if (k < kDofRaw) {
          const float* src = scr + k * D;
          for (int d = part; d < D; d += 3) v += src[d];
        }
        const float v1 = __shfl_down_sync(0xffffffffu, v, 1), v2 = __shfl_down_sync(0xffffffffu, v, 2);
        const float tot = (v + v1) + v2;
        if (part == 0 && k < kDofRaw) {
          // slots of DR_DJA2 .. DR_LTQ, one byte each
          const unsigned long long slots = (unsigned long long)(V_SIGMA_ERR + G_JOINT_POS) | ((unsigned long long)(V_SIGMA_ERR + G_JOINT_VEL) << 8) |
              ((unsigned long long)V_TQ2 << 16) | ((unsigned long long)V_DACT2 << 24) | ((unsigned long long)V_DV2 << 32) | ((unsigned long long)V_ACC2 << 40) |
              ((unsigned long long)V_LIM_POS << 48) | ((unsigned long long)V_LIM_VEL << 56);
          const int slot = k == DR_LTQ ? (int)V_LIM_TQ : (int)((slots >> (8 * k)) & 0xffull);
          rve[slot] = k <= DR_DJV2 ? tot * (1.0f / (float)D) : tot;  // (a mean: within 1 ulp of the division, whose slow path a zero sum would take)
          if (k == DR_DJA2) rve[V_JOINT_N] = sqrtf(tot);
        }
      }
      __syncwarp();

      // ---- c. bodies (lane = body): reference pose (lerp / slerp of the frame pair), extended bodies, differences, heading-frame features
      int far = 0, n_collide = 0, hit = 0;
      const float hz = pe[P_HZ], hw = pe[P_HW];
      for (int j0 = 0; j0 < J; j0 += 32) {
        const int j = j0 + lane;
        bool collide = false;
        if (j < J) {
          V3 p, v, w; Q4 q;
          if (j < Jb) {
            const float* bs = s_body + j * 13;
            p = {bs[0], bs[1], bs[2]}; q = {bs[3], bs[4], bs[5], bs[6]}; v = {bs[7], bs[8], bs[9]}; w = {bs[10], bs[11], bs[12]};
          } else {  // motion_tracking.py:619-643; the position chain is unfused (it feeds the motion_far comparison)
            const int x = j - Jb;
            const float* bs = s_body + c.ext_parent[x] * 13;
            const Q4 qp = {bs[3], bs[4], bs[5], bs[6]};
            const V3 off = {c.ext_pos[x][0], c.ext_pos[x][1], c.ext_pos[x][2]};
            const Q4 qe = {c.ext_rot_xyzw[x][0], c.ext_rot_xyzw[x][1], c.ext_rot_xyzw[x][2], c.ext_rot_xyzw[x][3]};
            V3 r = quat_rotate_x(qp, off, false);
            if (!KC(ext_identity)) r = quat_rotate_x(qe, r, false);  // (an identity rotation returns its argument bit for bit)
            p = {xa(r.x, bs[0]), xa(r.y, bs[1]), xa(r.z, bs[2])};
            q = KC(ext_identity) ? qp : quat_mul(qp, qe);           // (quat_mul by the identity: equal within 1 ulp, feeds tolerance-checked values only)
            w = {bs[10], bs[11], bs[12]};
            v = {bs[7] + (w.y * off.z - w.z * off.y), bs[8] + (w.z * off.x - w.x * off.z), bs[9] + (w.x * off.y - w.y * off.x)};
          }
          const float4 qa = __ldg(reinterpret_cast<const float4*>(r0) + j);
          const float4 qb = __ldg(reinterpret_cast<const float4*>(r1) + j);
          float rpz[3], rvv[3], rw[3];
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            rpz[k] = xa(xa(xm(omb, __ldg(r0 + e_pos + 3 * j + k)), xm(b, __ldg(r1 + e_pos + 3 * j + k))), s_org[k]);
            rvv[k] = omb * __ldg(r0 + e_vel + 3 * j + k) + b * __ldg(r1 + e_vel + 3 * j + k);
            rw[k] = omb * __ldg(r0 + e_ang + 3 * j + k) + b * __ldg(r1 + e_ang + 3 * j + k);
          }
          const Q4 rr = slerp_t({qa.x, qa.y, qa.z, qa.w}, {qb.x, qb.y, qb.z, qb.w}, b);
          if (KC(ref_pos_out)) {
            float* o = io.ref_body_pos_extend + n * (3 * J) + 3 * j;
            o[0] = rpz[0]; o[1] = rpz[1]; o[2] = rpz[2];
          }
          if (KC(ref_rot_out)) reinterpret_cast<float4*>(io.ref_body_rot_extend)[n * J + j] = make_float4(rr.x, rr.y, rr.z, rr.w);
          const V3 dp = {xs(rpz[0], p.x), xs(rpz[1], p.y), xs(rpz[2], p.z)};
          const float dq0 = rr.x - q.x, dq1 = rr.y - q.y, dq2 = rr.z - q.z, dq3 = rr.w - q.w;
          const V3 dv = {rvv[0] - v.x, rvv[1] - v.y, rvv[2] - v.z};
          const V3 dw = {rw[0] - w.x, rw[1] - w.y, rw[2] - w.z};
          const float nrm = norm3(dp.x, dp.y, dp.z);
          {
            const float m_pos = div3((dp.x * dp.x + dp.y * dp.y) + dp.z * dp.z);
            const int mb = T.member[j];
            float* bp = scr + j;
            bp[0 * J] = (mb & 1) ? m_pos : 0.0f; bp[1 * J] = (mb & 2) ? m_pos : 0.0f; bp[2 * J] = (mb & 4) ? m_pos : 0.0f; bp[3 * J] = (mb & 8) ? m_pos : 0.0f;
            bp[4 * J] = (((dq0 * dq0 + dq1 * dq1) + dq2 * dq2) + dq3 * dq3) * 0.25f;
            bp[5 * J] = div3((dv.x * dv.x + dv.y * dv.y) + dv.z * dv.z);
            bp[6 * J] = div3((dw.x * dw.x + dw.y * dw.y) + dw.z * dw.z);
            bp[7 * J] = (mb & 1) ? nrm : 0.0f; bp[8 * J] = (mb & 2) ? nrm : 0.0f; bp[9 * J] = (mb & 4) ? nrm : 0.0f;
          }
          if (nrm > far_thr) far = 1;
          if (KC(term_contact) && j < Jb && (((unsigned long long)KC(term_contact_mask) >> j) & 1ull)) {  // |f| > 1 (legged_robot_base.py:435); 0x1.000002p0 is the largest fp32 whose root is <= 1
            const float* fc = s_cf + 3 * j;
            if (__fmaf_rn(fc[2], fc[2], __fmaf_rn(fc[1], fc[1], fc[0] * fc[0])) > 0x1.000002p0f) hit = 1;
          }
          if (j < Jb && (((unsigned long long)KC(penalised_mask) >> j) & 1ull)) {
            const float* fc = s_cf + 3 * j;
            // |f| > 0.1 without the square root: sqrtf is monotonic and correctly rounded, 0x1.47ae16p-7 is the largest fp32 whose root is <= 0.1f
            collide = __fmaf_rn(fc[2], fc[2], __fmaf_rn(fc[1], fc[1], fc[0] * fc[0])) > 0x1.47ae16p-7f;
          }
          {
            const V3 dl = quat_rotate_z(hz, hw, dp);  // motion_tracking.py:694-697
            emit3k<kNoise, S_DIF_LOCAL>(T, Rw, 3 * j, dl.x, dl.y, dl.z, clipv, nk);
            const V3 lr = quat_rotate_z(hz, hw, {rpz[0] - pe[P_RPX], rpz[1] - pe[P_RPY], rpz[2] - pe[P_RPZ]});  // :705-707
            emit3k<kNoise, S_LOCAL_REF>(T, Rw, 3 * j, lr.x, lr.y, lr.z, clipv, nk);
          }
          if (KC(ref_vel_used)) {
            const V3 lv = quat_rotate_z(hz, hw, {rvv[0], rvv[1], rvv[2]});
            emit3k<kNoise, S_LOCAL_REF_VEL>(T, Rw, 3 * j, lv.x, lv.y, lv.z, clipv, nk);
            emit3k<kNoise, S_GLOBAL_REF_VEL>(T, Rw, 3 * j, rvv[0], rvv[1], rvv[2], clipv, nk);
          }
        }
        n_collide += __popc(__ballot_sync(0xffffffffu, collide));
      }
      const bool any_far = __any_sync(0xffffffffu, far);
      __syncwarp();
      {  // body sums: three lanes per sum (bodies part, part + 3, ...), combined in lane order
        const int k = lane / 3, part = lane - 3 * k;
        float v = 0.0f;
        if (k < kBodySums) {
          const float* src = scr + k * J;
          for (int j = part; j < J; j += 3) v += src[j];
        }
        const float v1 = __shfl_down_sync(0xffffffffu, v, 1), v2 = __shfl_down_sync(0xffffffffu, v, 2);
        if (part == 0 && k < kBodySums) rve[T.bs_slot[k]] = ((v + v1) + v2) / T.bs_div[k];
      }

      // ---- d. feet (lane = foot): contacts, feet_air_time, contact-force / slip terms ---------------------------------------------
      float x_air = 0.f, x_force = 0.f, x_slip = 0.f, x_fslip = 0.f, x_cm = 0.f, contact_f = 0.f, air_f = 0.f;
      bool x_stumble = false;
      if (lane < KC(n_feet)) {
        const int fb = c.feet_ids[lane];
        const float* fc = s_cf + 3 * fb;
        const float* bs = s_body + 13 * fb;
        const float last_c = s_envf[4 + lane];
        air_f = s_envf[2 + lane];
        const float fn = norm3_z(fc[0], fc[1], fc[2]);
        contact_f = fn > 1.0f ? 1.0f : 0.0f;                                   // legged_robot_base.py:352-353
        const float filt = (contact_f != 0.0f || last_c != 0.0f) ? 1.0f : 0.0f;
        if (KC(contact_size)) x_cm = fabsf(filt - (omb * __ldg(r0 + e_c + lane) + b * __ldg(r1 + e_c + lane)));
        const bool fz = fc[2] > 1.0f || last_c != 0.0f;                        // feet_air_time, motion_tracking.py:1307-1319
        if (KC(has_air_time)) {
          const float first = (air_f > 0.0f && fz) ? 1.0f : 0.0f;
          air_f = xa(air_f, dt);
          x_air = (air_f - KC(desired_feet_air_time)) * first;
          air_f *= fz ? 0.0f : 1.0f;
        }
        x_force = fmaxf(fn - KC(max_contact_force), 0.0f);                        // :1321-1323
        x_stumble = norm2_z(fc[0], fc[1]) > 5.0f * fabsf(fc[2]);                  // :1325-1328
        x_slip = norm3_z(bs[7], bs[8], bs[9]) * contact_f;                        // legged_robot_base.py:1004-1010
        x_fslip = contact_f * norm2_z(bs[7], bs[8]);                              // :1012-1017
      }
      x_air += __shfl_xor_sync(0xffffffffu, x_air, 1); x_force += __shfl_xor_sync(0xffffffffu, x_force, 1);
      x_slip += __shfl_xor_sync(0xffffffffu, x_slip, 1); x_fslip += __shfl_xor_sync(0xffffffffu, x_fslip, 1);
      x_cm += __shfl_xor_sync(0xffffffffu, x_cm, 1);
      const bool stumble = __any_sync(0xffffffffu, x_stumble);
      float air1 = __shfl_sync(0xffffffffu, air_f, 1), contact1 = __shfl_sync(0xffffffffu, contact_f, 1);  // lane 0 packs the state

      // ---- the single-buffered input rows (simulator state, history ring / scratch, dof-wise arrays) are not read past this point:
      // the warp's next env is requested, so that its inputs arrive while sections e-i and the row write-out run --------------------
      __syncwarp();
      if (n + step < n_envs) request(n + step, sb ^ 1);

      // ---- e. termination (legged_robot_base.py:408-489, motion_tracking.py:330-358) -------------------------------------------
      const int64_t last_ep = ep_in + 1;
      int64_t ep = last_ep;
      float start = s_envf[0], mlen = s_envf[1], etr = s_envf[6];
      const bool t_grav = KC(term_gravity) && pe[P_GRAVN] > KC(term_gravity_thr);
      const bool t_far = KC(term_motion_far) && any_far;
      const bool t_out = ep > KC(max_episode_length);
      const bool t_end = KC(term_motion_end) && (xa(xm((float)ep, dt), start) > mlen);
      const bool time_out = t_out || t_end;
      bool t_contact = false, t_low = false, t_dpos = false, t_dvel = false, t_tq = false;
      if (KC(term_base)) {
        const int gates = T.term_gates;
        t_contact = KC(term_contact) && __any_sync(0xffffffffu, hit);
        t_low = KC(term_low_height) && pe[P_RPZ] < KC(term_min_height);                  // root z (legged_robot_base.py:445)
        t_dpos = KC(term_dof_pos) && (gates & 1) && __any_sync(0xffffffffu, dof_viol & 1);
        t_dvel = KC(term_dof_vel) && (gates & 2) && __any_sync(0xffffffffu, dof_viol & 2);
        t_tq = KC(term_torque) && (gates & 4) && __any_sync(0xffffffffu, dof_viol & 4);
      }
      const bool terminated = t_grav || t_far || time_out || t_contact || t_low || t_dpos || t_dvel || t_tq;
      const bool reset = terminated || (io.force_reset != nullptr && io.force_reset[n] != 0);  // forced: resample_motion (motion_tracking.py:385-389)
      if (lane == 0) {
        rve[V_T_GRAV] = t_grav; rve[V_T_FAR] = t_far; rve[V_T_OUT] = time_out; rve[V_T_END] = t_end;
        rve[V_RESET] = reset; rve[V_ONE] = 1.0f; rve[V_ALIVE] = 1.0f;
        rve[V_CONTACT] = 1.0f - x_cm / (float)KC(n_feet); rve[V_CONTACT_V2] = 0.5f - x_cm / (float)KC(n_feet);
        rve[V_ORIENT] = pe[P_ORIENT];
        rve[V_AIR] = x_air; rve[V_FORCE] = x_force; rve[V_STUMBLE] = stumble ? 1.0f : 0.0f; rve[V_SLIP] = x_slip; rve[V_FSLIP] = x_fslip;
        rve[V_COLLIDE] = (float)n_collide;
      }
      __syncwarp();

      // ---- f. rewards (lane = term; legged_robot_base.py:715-761) + episode sums ---------------------------------------------------
      {
        float rew = 0.0f;
        if (lane < R) {
          const int kind = T.r_kind[lane];
          float v = rve[T.r_arg[lane]];
          if (kind != K_RAW) {
            v = expf(-v / T.r_sigma[lane]);
            if (kind == K_EXP2)  // teleop_body_position_extend, motion_tracking.py:1177-1199
              v = v * KC(w_lower) + expf(-rve[V_SIGMA_ERR + G_UPPER] / T.sigma[G_UPPER]) * KC(w_upper);
          }
          rew = v * T.r_scale[lane] * T.r_pen[lane];
        }
        const float term_rew = KC(has_termination) ? ((terminated && !time_out) ? 1.0f : 0.0f) * KC(termination_scale) : 0.0f;
        if (KC(use_vec_reward)) {
          float v = rew;
          if (KC(only_positive)) v = fmaxf(v, 0.0f);
          if (lane == R - 1) v += term_rew;  // the reference's stale loop index, legged_robot_base.py:744
          if (lane < R) io.rew_buf[n * (R + 1) + lane] = v;
          if (lane == R) io.rew_buf[n * (R + 1) + R] = 0.0f;
        } else {  // rew_buf += rew term by term (legged_robot_base.py:733): a sequential fp32 sum in term order
          float tot = 0.0f;
          for (int r = 0; r < R; ++r) tot += __shfl_sync(0xffffffffu, rew, r);
          if (KC(only_positive)) tot = fmaxf(tot, 0.0f);
          if (lane == 0) io.rew_buf[n] = tot + term_rew;
        }
        if (lane < KC(n_sums)) {
          float ev = reinterpret_cast<const float*>(smem + (sb ? KC(i_es1) : KC(i_es0)))[e * KC(n_sums) + lane] + (lane < R ? rew : term_rew);
          if (reset) { io.last_episode_sums[n * KC(n_sums) + lane] = ev; ev = 0.0f; }  // extras["episode"], legged_robot_base.py:509-513
          io.episode_sums[n * KC(n_sums) + lane] = ev;
        }
      }

      // ---- g. scalar features into the rows (legged_robot_base.py:1114-1215; motion_tracking.py:944-1015) -------------------------
      int64_t delay = reinterpret_cast<const long long*>(smem + (sb ? KC(i_delay1) : KC(i_delay0)))[e];
      auto emit_scalar_features = [&](int64_t delay_) {
        for (int i = lane; i < KC(n_feat); i += 32) {
          const int k = T.feat_src[i], el = T.feat_el[i];
          float val;
          if (i < 9) val = pe[P_BLV + i];
          else if (k == S_DR_COM) val = reinterpret_cast<const float*>(smem + (sb ? KC(i_com1) : KC(i_com0)))[e * 3 + el];
          else if (k == S_PHASE) val = pe[P_PHASE];        // stale for envs that reset, like the reference
          else if (k == S_DR_FRICTION) val = reinterpret_cast<const float*>(smem + (sb ? KC(i_fric1) : KC(i_fric0)))[e];
          else if (k == S_DR_DELAY) val = (float)delay_;
          else if (k == S_BASE_POS_Z) val = pe[P_RPZ];
          else val = reinterpret_cast<const float*>(smem + (sb ? KC(i_mass1) : KC(i_mass0)))[e * KC(L) + el];
          emit1<kNoise>(T, Rw, k, el, val, clipv, nk);
        }
      };
      if (__builtin_expect(!reset, 1)) emit_scalar_features(delay);

      // ---- h. reset (legged_robot_base.py:491-516, 647-686, 587-635; motion_tracking.py:274-296, 369-384, 445-543) ---------------
      float air0 = air_f, contact0 = contact_f;  // meaningful in lane 0
      if (__builtin_expect(reset, 0)) {  // (rare: keep the block out of the hot instruction stream)
        const int clip = __float_as_int(pe[P_CLIP]);
        const bool ld = lane < D;
        const Rng rng = {make_uint2((uint32_t)c.seed, (uint32_t)(c.seed >> 32)), (uint32_t)(n + io.env_offset), (uint32_t)((uint64_t)(n + io.env_offset) >> 32), push_count};
        if (ld) io.actions[n * D + lane] = 0.0f;
        if (lane == 0 && io.last_episode_length_buf) io.last_episode_length_buf[n] = last_ep;  // extras["episode"]["end_epis_length"], legged_robot_base.py:515
        if (lane == 0 && io.refresh_buf) io.refresh_buf[n] = (int64_t)push_count + 1;             // need_to_refresh_envs, :493
        air0 = 0.0f; air1 = 0.0f; contact0 = 0.0f; contact1 = 0.0f;
        // HistoryHandler.reset: zero the ring and what the observation rows already took from it
        for (int i = lane; i < H; i += 32) {
          io.hist[n * H + i] = 0.0f;
          auto zero1 = [&](int has, int plane, int gd) {
            if (!has) return;
            const int m = hmap[plane * H + i];
            if (m >= 0) rows[m + e * gd] = 0.0f;
          };
          zero1(KC(ghist0), 0, KC(gdim0));
          zero1(KC(ghist1), KC(ghist0), KC(gdim1));
          zero1(KC(ghist2), KC(ghist0) + KC(ghist1), KC(gdim2));
          zero1(KC(ghist3), KC(ghist0) + KC(ghist1) + KC(ghist2), KC(gdim3));
        }
        __syncwarp();
        ep = 0;
        // Random draws (when the caller does not inject the reference's): Philox4x32-10 keyed by (seed, global env id, step);
        // every call yields four words, so one call serves the four DR scales of a dof, one a Box-Muller pair, one the
        // env-level scalars - a reset costs six calls per lane instead of sixteen (the other lanes of the tile wait for it).
        auto gauss2 = [](uint32_t a, uint32_t b_, float& n0, float& n1) {
          const float r = sqrtf(-2.0f * __logf(fmaxf(u01(a), 1.0f / 16777216.0f)));
          float sn, cs;
          __sincosf(6.283185307179586f * u01(b_), &sn, &cs);
          n0 = r * cs; n1 = r * sn;
        };
        const uint4 r_env = rng.draw(0);   // x: start time, y: control delay, z: root-rotation angle
        float kp2 = 1.f, kd2 = 1.f, dpos2 = 0.f, dvel2 = 0.f;
        if (ld) {  // _episodic_domain_randomization
          const uint4 r_dr = rng.draw(16 + lane);
          kp2 = io.kp_scale[n * D + lane]; kd2 = io.kd_scale[n * D + lane];
          if (c.randomize_pd_gain) {
            kp2 = io.inject_kp ? io.inject_kp[n * D + lane] : xa(xm(xs(c.kp_hi, c.kp_lo), u01(r_dr.x)), c.kp_lo);
            kd2 = io.inject_kd ? io.inject_kd[n * D + lane] : xa(xm(xs(c.kd_hi, c.kd_lo), u01(r_dr.y)), c.kd_lo);
            io.kp_scale[n * D + lane] = kp2; io.kd_scale[n * D + lane] = kd2;
          }
          if (c.randomize_rfi_lim)
            io.rfi_lim_scale[n * D + lane] = io.inject_rfi ? io.inject_rfi[n * D + lane] : xa(xm(xs(c.rfi_hi, c.rfi_lo), u01(r_dr.z)), c.rfi_lo);
          if (c.use_rao)
            io.rao_scale[n * D + lane] = io.inject_rao ? io.inject_rao[n * D + lane] : xa(xm(xs(c.rao_lim, -c.rao_lim), u01(r_dr.w)), -c.rao_lim);
        }
        if (c.randomize_ctrl_delay) {
          delay = io.inject_delay ? io.inject_delay[n] : c.delay_lo + (int64_t)(r_env.y % (uint32_t)(c.delay_hi - c.delay_lo + 1));
          if (lane == 0) io.action_delay_idx[n] = delay;
        }
        etr = __fdiv_rn(xa(xm((float)last_ep, dt), start), mlen);           // motion_tracking.py:281-282
        mlen = tab.clip_length[clip];                                       // _resample_motion_times
        start = io.inject_start_times ? io.inject_start_times[n] : (c.start_time_zero ? 0.0f : xm(u01(r_env.x), mlen));
        // _reset_dofs / _reset_root_states from the motion state at the new time (unfused: these are simulator state)
        const FrameBlend f2 = frame_blend_x(tab, clip, xa(xm((float)(ep + 1), dt), start));
        const float b2 = f2.blend, omb2 = 1.0f - b2;
        const float* q0 = tab.records + f2.f0 * KC(rec_stride);
        const float* q1 = tab.records + f2.f1 * KC(rec_stride);
        const bool any_noise = c.reset_noise[0] != 0.0f || c.reset_noise[1] != 0.0f || c.reset_noise[2] != 0.0f || c.reset_noise[3] != 0.0f ||
                               c.reset_noise[4] != 0.0f || c.reset_noise[5] != 0.0f;
        float g0 = 0.0f, g1 = 0.0f, g2 = 0.0f;   // this lane's standard normals: dof position, dof velocity, root-state component
        Q4 nq = {0.0f, 0.0f, 0.0f, 1.0f};  // small_random_quaternions(n, max_angle): the same value in every lane
        if (any_noise) {
          const uint4 r_n = rng.draw(144 + lane);
          gauss2(r_n.x, r_n.y, g0, g1);
          { float unused; gauss2(r_n.z, r_n.w, g2, unused); }
          if (c.reset_noise[3] != 0.0f) {
            const uint4 r_q = rng.draw(2);
            V3 ax; float unused;
            gauss2(r_q.x, r_q.y, ax.x, ax.y);
            gauss2(r_q.z, r_q.w, ax.z, unused);
            const float an = norm3(ax.x, ax.y, ax.z);
            const float ang = c.reset_noise[3] * u01(r_env.z);
            float sh, ch;
            __sincosf(ang * 0.5f, &sh, &ch);
            nq = {sh * (ax.x / an), sh * (ax.y / an), sh * (ax.z / an), ch};
          }
        }
        if (ld) {
          dpos2 = xa(xm(omb2, q0[e_dp + lane]), xm(b2, q1[e_dp + lane]));
          dvel2 = xa(xm(omb2, q0[e_dv + lane]), xm(b2, q1[e_dv + lane]));
          if (c.reset_noise[0] != 0.0f) dpos2 = xa(dpos2, xm(g0, c.reset_noise[0]));
          if (c.reset_noise[1] != 0.0f) dvel2 = xa(dvel2, xm(g1, c.reset_noise[1]));
          reinterpret_cast<float2*>(io.dof_state)[n * D + lane] = make_float2(dpos2, dvel2);
        }
        {
          const float4 a = reinterpret_cast<const float4*>(q0)[0], d4 = reinterpret_cast<const float4*>(q1)[0];
          const Q4 rr = slerp_t({a.x, a.y, a.z, a.w}, {d4.x, d4.y, d4.z, d4.w}, b2);
          const Q4 o = quat_mul(nq, rr);
          if (lane < 13) {
            float v;
            if (lane >= 3 && lane < 7) v = lane == 3 ? o.x : (lane == 4 ? o.y : (lane == 5 ? o.z : o.w));
            else {
              const int el = lane < 3 ? e_pos + lane : (lane < 10 ? e_vel + (lane - 7) : e_ang + (lane - 10));
              v = xa(xm(omb2, q0[el]), xm(b2, q1[el]));
              if (lane < 3) v = xa(v, s_org[lane]);
              const float ns = lane < 3 ? c.reset_noise[2] : (lane < 10 ? c.reset_noise[4] : c.reset_noise[5]);
              if (ns != 0.0f) v = xa(v, xm(g2, ns));
            }
            io.root_state[n * 13 + lane] = v;
          }
        }
        // the observations of an env that reset: dof-wise features from the new state; the scalar features go in (again)
        // after the ring was zeroed (ref_motion_phase & co. stay stale like the reference's)
        if (ld) {
          const float ddef = io.default_dof_pos[n * KC(default_dof_stride) + lane];
          emit1k<kNoise, S_DOF_POS>(T, Rw, lane, dpos2 - ddef, clipv, nk);
          emit1k<kNoise, S_DOF_VEL>(T, Rw, lane, dvel2, clipv, nk);
          emit1k<kNoise, S_ACTIONS>(T, Rw, lane, 0.0f, clipv, nk);
          emit1k<kNoise, S_DR_KP>(T, Rw, lane, kp2, clipv, nk);
          emit1k<kNoise, S_DR_KD>(T, Rw, lane, kd2, clipv, nk);
          const int x = e * D + lane;
          io.last_actions[n * D + lane] = 0.0f; io.last_dof_pos[n * D + lane] = dpos2; io.last_dof_vel[n * D + lane] = dvel2;
        }
        emit_scalar_features(delay);
      }

      // ---- i. state write-back, cross-env sums ----------------------------------------------------------------------------------------
      if (lane == 0) {
        float4* of = reinterpret_cast<float4*>(io.env_f32) + n * 2;
        const int pw = KC(push_robots) ? push_step(c, io, n, __float_as_int(s_envf[7]), reset, P.c.seed, push_count) : __float_as_int(s_envf[7]);
        of[0] = make_float4(start, mlen, air0, KC(n_feet) > 1 ? air1 : 0.0f);
        of[1] = make_float4(contact0, KC(n_feet) > 1 ? contact1 : 0.0f, etr, __int_as_float(pw));
        io.episode_length_buf[n] = ep;
        io.reset_buf[n] = reset ? 1 : 0;
        io.time_out_buf[n] = time_out ? 1 : 0;
      }
      if (lane < kAccLo) {  // per-lane fp64 accumulation of the cross-env sums (slot = lane, layout PBHC_A_*)
        double a = (double)rve[lane];
        if (lane == V_RESET_EPLEN) a = reset ? (double)last_ep : 0.0;
        if (lane == V_ETR) a = (double)etr;
        if (lane == V_ETR_SQ) a = (double)etr * (double)etr;
        acc_reg += a;
      }
      if (KC(term_base)) {
        const int slot = 32 + lane;
        acc_hi += slot == PBHC_A_TERM_CONTACT ? (double)t_contact : slot == PBHC_A_TERM_LOW_HEIGHT ? (double)t_low : slot == PBHC_A_TERM_DOF_POS ? (double)t_dpos :
                  slot == PBHC_A_TERM_DOF_VEL ? (double)t_dvel : slot == PBHC_A_TERM_TORQUE ? (double)t_tq : 0.0;
      }
      // ---- j. the env's observation rows leave (16 bytes per lane where the row size allows) -----------------------------------
      __syncwarp();
      auto row_out = [&](int has, int row_base, int gd, float* g) {
        if (!has) return;
        const float* src = rows + row_base + e * gd;
        float* dst = g + n * gd;
        const uintptr_t al = reinterpret_cast<uintptr_t>(g);  // (a storage slot the caller redirected the group to may start anywhere)
        if ((gd & 3) == 0 && (al & 15) == 0) { for (int i = lane; i < (gd >> 2); i += 32) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i]; }
        else if ((gd & 1) == 0 && (al & 7) == 0) { for (int i = lane; i < (gd >> 1); i += 32) reinterpret_cast<float2*>(dst)[i] = reinterpret_cast<const float2*>(src)[i]; }
        else { for (int i = lane; i < gd; i += 32) dst[i] = src[i]; }
      };
      row_out(KC(n_groups) > 0, KC(row_base0), KC(gdim0), io.obs[0]);
      row_out(KC(n_groups) > 1, KC(row_base1), KC(gdim1), io.obs[1]);
      row_out(KC(n_groups) > 2, KC(row_base2), KC(gdim2), io.obs[2]);
      row_out(KC(n_groups) > 3, KC(row_base3), KC(gdim3), io.obs[3]);
      __syncwarp();  // (the rows are written again by the next env)
    }
  }
  cp_async_wait_all();
  // block partials: 64 threads add the warps' sums up in warp order
  __syncthreads();
  double* wacc = reinterpret_cast<double*>(smem + KC(o_rows));
  if (lane < PBHC_ACC / 2) { wacc[warp * 32 + lane] = lane < kAccLo ? acc_reg : 0.0; wacc[(kTE + warp) * 32 + lane] = acc_hi; }
  __syncthreads();
  if (tid < PBHC_ACC) {
    double v = 0.0;
    for (int w = 0; w < kTE; ++w) v += wacc[((tid >> 5) * kTE + w) * 32 + (tid & 31)];
    io.partials[(int64_t)blockIdx.x * PBHC_ACC + tid] = v;
  }
}

#ifdef __CUDACC_RTC__
}  // namespace pbhc
// the specialised kernel: this file compiled by NVRTC with -DPBHC_SPEC and the generated "track_spec.h" (jit.cu)
extern "C" __global__ void __launch_bounds__(pbhc::kNT, 3) pbhc_track_tile_spec(const __grid_constant__ pbhc::TileParams P) {
  pbhc::track_tile_body<PBHC_NOISE != 0>(P);
}
#else

template <bool kNoise>
__global__ void __launch_bounds__(kNT, 3) track_tile_kernel(const __grid_constant__ TileParams P) { track_tile_body<kNoise>(P); }

extern int g_track_info[4];
int jit_get(const char* what, const std::string& spec_header, bool noise, const void** kernel);   // jit.cu: 0 = *kernel is the cudaKernel_t of the specialised kernel

std::string tile_spec_header(const HotK& k, const SpecExtra& x) {
  char buf[128];
  std::string h = "// generated by pbhc_b200 (track_tile.cu tile_spec_header): launch constants of one env config\nnamespace spec {\n";
#define X(n) snprintf(buf, sizeof(buf), "constexpr int %s = %d;\n", #n, k.n); h += buf;
  PBHC_HOTK_INT(X)
#undef X
#define X(n) snprintf(buf, sizeof(buf), "constexpr float %s = %af;\n", #n, (double)k.n); h += buf;
  PBHC_HOTK_FLT(X)
#undef X
#define X(n) snprintf(buf, sizeof(buf), "constexpr long long %s = (long long)0x%llxULL;\n", #n, (unsigned long long)k.n); h += buf;
  PBHC_HOTK_I64(X)
#undef X
  auto arr_i = [&](const char* name, const int* v, int n0, int n1) {
    h += std::string("constexpr int ") + name + "[" + std::to_string(n0) + "]" + (n1 ? "[" + std::to_string(n1) + "]" : std::string()) + " = {";
    for (int i = 0; i < n0; ++i) {
      if (n1) { h += "{"; for (int j = 0; j < n1; ++j) { h += std::to_string(v[i * n1 + j]); h += j + 1 < n1 ? "," : ""; } h += "}"; }
      else h += std::to_string(v[i]);
      h += i + 1 < n0 ? "," : "";
    }
    h += "};\n";
  };
  arr_i("sd_n", x.n, 20, 0); arr_i("sd_base", &x.base[0][0], 20, 4); arr_i("sd_stride", &x.stride[0][0], 20, 4);
  arr_i("sd_pushed", x.pushed, 20, 0); arr_i("sd_noisy", x.noisy, 20, 0);
  h += "constexpr float sd_scale[20] = {";
  for (int i = 0; i < 20; ++i) { snprintf(buf, sizeof(buf), "%af%s", (double)x.scale[i], i < 19 ? "," : ""); h += buf; }
  h += "};\n";
  h += "}  // namespace spec\n";
  return h;
}

// the kernel specialised for these constants (kernel == nullptr: compile into the on-disk cache only)
static int jit_track_tile(const HotK& k, const SpecExtra& x, bool noise, const void** kernel) {
  static thread_local HotK last_k;
  static thread_local SpecExtra last_x;
  static thread_local bool last_noise = false, have = false;
  static thread_local const void* last_fn = nullptr;
  static thread_local int last_rc = 0;
  if (kernel && have && last_noise == noise && !memcmp(&last_k, &k, sizeof(k)) && !memcmp(&last_x, &x, sizeof(x))) { *kernel = last_fn; return last_rc; }
  const void* fn = nullptr;
  const int rc = jit_get("track_tile", tile_spec_header(k, x), noise, kernel ? &fn : nullptr);
  if (kernel) { last_k = k; last_x = x; last_noise = noise; have = true; last_fn = fn; last_rc = rc; *kernel = fn; }
  return rc;
}

// shared-memory layout + hot constants + descriptor tables of one launch; returns the dynamic shared memory size (< 0: error)
static int tile_plan(const PbhcTrackCfg& c, const PbhcTableDesc& tab, const PbhcTrackIO& io, int64_t n, TileParams& P) {
  P.c = c; P.tab = tab; P.io = io;
  if (n > 0x7fffffffLL / 64) { set_error("pbhc_track_step: too many envs for one launch"); return PBHC_E_RANGE; }
  P.n = (int)n; P.n_tiles = (int)((n + kTE - 1) / kTE);
  HotK& k = P.k;
  memset(&k, 0, sizeof(k));
  const int D = c.D, J = c.J, Jb = c.Jb, H = c.hist_row, L = c.L;
  if (c.n_feet > 2) { set_error("pbhc_track_step: at most two feet"); return PBHC_E_RANGE; }
  if (c.n_rewards > 30) { set_error("pbhc_track_step: more than 30 reward terms"); return PBHC_E_RANGE; }
  if (L > 48) { set_error("pbhc_track_step: more than 48 randomised link masses"); return PBHC_E_RANGE; }
  k.J = J; k.Jb = Jb; k.D = D; k.L = L; k.H = H; k.n_ext = c.n_ext; k.n_feet = c.n_feet; k.n_groups = c.n_groups; k.R = c.n_rewards;
  k.n_sums = c.n_episode_sums; k.n_feat = 16 + L; k.contact_size = tab.contact_size; k.rec_stride = tab.rec_stride;
  k.use_vec_reward = c.use_vec_reward; k.only_positive = c.only_positive_rewards; k.has_termination = c.has_termination;
  k.term_contact = c.term_contact; k.term_low_height = c.term_low_height; k.term_dof_pos = c.term_dof_pos; k.term_dof_vel = c.term_dof_vel;
  k.push_robots = c.push_robots; k.term_torque = c.term_torque; k.term_base = (c.term_contact || c.term_low_height || c.term_dof_pos || c.term_dof_vel || c.term_torque) ? 1 : 0;
  k.term_min_height = c.term_min_height; k.term_dof_vel_scale = c.term_dof_vel_scale; k.term_torque_scale = c.term_torque_scale;
  k.term_contact_mask = (long long)c.term_contact_mask;
  k.cur_torque_enable = c.cur_torque.enable; k.term_gravity = c.term_gravity; k.term_motion_far = c.term_motion_far; k.term_motion_end = c.term_motion_end;
  k.dif_joint_used = (c.src_used_mask & ((1ull << S_DIF_JOINT) | (1ull << S_DIF_JOINT_VEL))) ? 1 : 0;
  k.ref_vel_used = (c.src_used_mask & ((1ull << S_LOCAL_REF_VEL) | (1ull << S_GLOBAL_REF_VEL))) ? 1 : 0;
  k.ext_identity = 1;
  for (int x = 0; x < c.n_ext; ++x) if (c.ext_rot_xyzw[x][0] != 0.0f || c.ext_rot_xyzw[x][1] != 0.0f || c.ext_rot_xyzw[x][2] != 0.0f || c.ext_rot_xyzw[x][3] != 1.0f) k.ext_identity = 0;
  k.ref_pos_out = io.ref_body_pos_extend ? 1 : 0; k.ref_rot_out = io.ref_body_rot_extend ? 1 : 0;
  k.default_dof_stride = (int)io.default_dof_stride;
  k.dt = c.dt; k.clip_obs = c.clip_obs; k.term_gravity_thr = c.term_gravity_thr; k.w_lower = c.w_lower; k.w_upper = c.w_upper;
  k.termination_scale = c.termination_scale; k.desired_feet_air_time = c.desired_feet_air_time; k.max_contact_force = c.max_contact_force;
  k.max_episode_length = c.max_episode_length; k.penalised_mask = (long long)c.penalised_mask;
  for (int r = 0; r < c.n_rewards; ++r) if (c.reward_id[r] == R_FEET_AIR_TIME) k.has_air_time = 1;
  int hrow = 0;
  for (int i = 0; i < PBHC_MAX_HIST_KEYS; ++i) { P.hist_base[i] = hrow; if (i < c.n_hist_keys) hrow += c.hist_len[i] * c.hist_dim[i]; }
  int ghist[PBHC_MAX_OBS_GROUPS] = {0, 0, 0, 0};
  for (int e = 0; e < c.n_hist_dst; ++e) {
    const PbhcHistDst& h = c.hist_dst[e];
    ghist[h.group] = 1;
    if (h.mul != 1.0f) k.hist_any_mul = 1;
  }
  k.ghist0 = ghist[0]; k.ghist1 = ghist[1]; k.ghist2 = ghist[2]; k.ghist3 = ghist[3];
  k.n_hd = ghist[0] + ghist[1] + ghist[2] + ghist[3];
  k.e_pos = 4 * J; k.e_vel = 7 * J; k.e_ang = 10 * J; k.e_dp = 13 * J; k.e_dv = k.e_dp + D; k.e_c = k.e_dv + D;
  // ---- shared-memory layout ----
  int off = 0;
  auto take = [&](int bytes) { const int o = off; off += (bytes + 15) & ~15; return o; };
  k.o_tab = take((int)sizeof(TileTables));
  k.n_slots = c.n_episode_sums + 3 + 3 + 1 + 2 + L + 13 + 2 + PBHC_ENVF_STRIDE + 2;  // 4-byte words of the late-read arrays
  k.o_stab = take(2 * k.n_slots * (int)sizeof(SmallSlot));
  k.o_hmap = take(2 * (k.n_hd > 0 ? k.n_hd : 1) * (H > 0 ? H : 1));
  k.o_hmul = k.hist_any_mul ? take(4 * k.n_hd * H) : k.o_hmap;
  k.o_rv = take(4 * kTE * kRvT);
  k.o_pre = take(4 * kTE * kPre);
  bool ok = true;  // every per-env array starts on a 16-byte boundary (what the vector widths of the loads / stores assume)
  auto slab_in = [&](const void* g, int row_bytes, int64_t g_stride) {
    if (((uintptr_t)g & 15) || (g_stride & 3)) ok = false;
    return take(kTE * row_bytes);
  };
  auto slab_small = [&](const void* g, int row_bytes, int* o0, int* o1) {
    *o0 = take(kTE * row_bytes); *o1 = take(kTE * row_bytes);
    if ((uintptr_t)g & 15) ok = false;
  };
  auto slab_out = [&](void* g, int row_bytes) {
    if ((uintptr_t)g & 3) ok = false;
    return take(kTE * row_bytes);
  };
  k.i_body = slab_in(io.body_state, Jb * 13 * 4, io.body_env_stride * 4);
  k.i_cf = slab_in(io.contact_forces, Jb * 12, Jb * 12);
  // per-env scratch row (dof partials, then body partials): the env's history row once it is scattered, if that is large enough
  const int scr_floats = kDofRaw * D > kBodyRaw * J ? kDofRaw * D : kBodyRaw * J;  // (10 x 27 = 270 <= 316 for G1 23-DoF obs=main)
  if (H > 0) k.i_hist = slab_in(io.hist, H * 4, H * 4);
  else k.i_hist = off;
  if (H >= scr_floats) { k.o_scr = k.i_hist; k.scr_pitch = H * 4; }
  else { k.scr_pitch = ((scr_floats * 4) + 15) & ~15; k.o_scr = take(kTE * k.scr_pitch); }
  k.i_dof = slab_in(io.dof_state, D * 8, D * 8);
  k.i_act = slab_in(io.actions, D * 4, D * 4);
  k.i_tq = slab_in(io.torques, D * 4, D * 4);
  k.i_lact = slab_in(io.last_actions, D * 4, D * 4);
  k.i_ldv = slab_in(io.last_dof_vel, D * 4, D * 4);
  k.i_kp = slab_in(io.kp_scale, D * 4, D * 4);
  k.i_kd = slab_in(io.kd_scale, D * 4, D * 4);
  if (k.i_kd != k.i_act + 5 * kTE * D * 4) { set_error("pbhc_track_step: internal layout error (dof-wise slabs)"); return PBHC_E_BADARG; }
  // read late in a tile (rewards, scalar features, reset): double-buffered like the per-env scalars, so that the big slabs above
  // can be re-filled for the next tile while this one still runs
  slab_small(io.episode_sums, c.n_episode_sums * 4, &k.i_es0, &k.i_es1);
  slab_small(io.env_origins, 12, &k.i_org0, &k.i_org1);
  slab_small(io.base_com_bias, 12, &k.i_com0, &k.i_com1);
  slab_small(io.friction, 4, &k.i_fric0, &k.i_fric1);
  slab_small(io.action_delay_idx, 8, &k.i_delay0, &k.i_delay1);
  if (L > 0) slab_small(io.link_mass_scale, L * 4, &k.i_mass0, &k.i_mass1);
  else { k.i_mass0 = k.i_fric0; k.i_mass1 = k.i_fric1; }
  slab_small(io.root_state, 52, &k.i_root0, &k.i_root1);
  slab_small(io.episode_length_buf, 8, &k.i_ep0, &k.i_ep1);
  slab_small(io.env_f32, PBHC_ENVF_STRIDE * 4, &k.i_envf0, &k.i_envf1);
  slab_small(io.motion_ids, 8, &k.i_mid0, &k.i_mid1);
  k.o_rows = off;
  int row_base[PBHC_MAX_OBS_GROUPS], gdim[PBHC_MAX_OBS_GROUPS];
  for (int g = 0; g < PBHC_MAX_OBS_GROUPS; ++g) {
    row_base[g] = (off - k.o_rows) >> 2;
    gdim[g] = g < c.n_groups ? c.group_dim[g] : 0;
    if (g < c.n_groups) slab_out(io.obs[g], c.group_dim[g] * 4);
  }
  k.row_base0 = row_base[0]; k.row_base1 = row_base[1]; k.row_base2 = row_base[2]; k.row_base3 = row_base[3];
  k.gdim0 = gdim[0]; k.gdim1 = gdim[1]; k.gdim2 = gdim[2]; k.gdim3 = gdim[3];
  memset(&P.x, 0, sizeof(P.x));
  static_assert(kSrcT <= 20, "SpecExtra is sized for 20 sources");
  for (int s = 0; s < kSrcT; ++s) {
    for (int g = 0; g < c.n_groups; ++g)
      if (c.src_dst[s][g] >= 0) { P.x.base[s][P.x.n[s]] = row_base[g] + c.src_dst[s][g]; P.x.stride[s][P.x.n[s]] = c.group_dim[g]; ++P.x.n[s]; }
    for (int h = 0; h < c.n_hist_keys; ++h) if (c.hist_src[h] == s) P.x.pushed[s] = 1;
    P.x.noisy[s] = c.obs_noise[s] != 0.0f ? 1 : 0;
    P.x.scale[s] = c.obs_scale[s];
  }
  if (off - k.o_rows < 16 * 32 * kTE) take(16 * 32 * kTE - (off - k.o_rows));  // (the rows region also parks the scalars / the warps' sums)
  const int rows_end = off;
  for (const void* g : {(const void*)io.rew_buf, (const void*)io.env_f32, (const void*)io.last_dof_pos}) if ((uintptr_t)g & 15) ok = false;
  if (!ok) return 0;  // (0: not for this kernel - the caller falls back to the warp-per-env kernel)
  if (off > 220 * 1024) { set_error("pbhc_track_step: shared memory %d B too large", off); return PBHC_E_RANGE; }
  if (((rows_end - k.o_rows) >> 2) > 32767) { set_error("pbhc_track_step: observation rows too wide for the 16-bit history map"); return PBHC_E_RANGE; }
  return off;
}

struct KernelState { size_t configured = 0; int per_sm = 0; };

// host side: plan + kernel choice + launch (called by pbhc_track_step for general == 0).  The kernel specialised for this env
// config is used when it can be had (compiled by NVRTC once per config and cached on disk, PBHC_JIT=0 turns it off);
// otherwise the generic kernel built with the library runs the same source with the constants read from the parameters.
int track_tile_launch(const PbhcTrackCfg& c, const PbhcTableDesc& tab, const PbhcTrackIO& io, int64_t n, cudaStream_t stream, int* blocks_out) {
  static thread_local TileParams P;
  const int shmem = tile_plan(c, tab, io, n, P);
  if (shmem < 0) return shmem;
  if (shmem == 0) return 1;  // unaligned tensors: not an error, the warp-per-env kernel takes the step
  bool noise = false;
  for (int s = 0; s < kSrcT; ++s) noise |= c.obs_noise[s] != 0.0f;
  const void* kern = noise ? (const void*)track_tile_kernel<true> : (const void*)track_tile_kernel<false>;
  static thread_local KernelState generic[2];
  static thread_local std::unordered_map<const void*, KernelState> jitted;
  KernelState* st = &generic[noise ? 1 : 0];
  int which = 1;
  const void* jk = nullptr;
  if (jit_track_tile(P.k, P.x, noise, &jk) == 0 && jk) { kern = jk; st = &jitted[jk]; which = 2; }
  if ((size_t)shmem > st->configured || st->per_sm == 0) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, shmem) != cudaSuccess) return check_launch("pbhc_track_step(attr)");
    (void)cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    st->configured = (size_t)shmem;
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kNT, (size_t)shmem);
    st->per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
  }
  int blocks = P.n_tiles;
  const int cap = 148 * st->per_sm;
  if (blocks > cap) blocks = cap;
  g_track_info[0] = which; g_track_info[1] = blocks; g_track_info[2] = shmem; g_track_info[3] = st->per_sm;
  void* args[] = {(void*)&P};
  if (cudaLaunchKernel(kern, dim3((unsigned)blocks), dim3(kNT), args, (size_t)shmem, stream) != cudaSuccess) return check_launch("pbhc_track_step(tile)");
  *blocks_out = blocks;
  return PBHC_OK;
}

// the specialised kernel of a config without a launch (build-time warm-up of the on-disk cache; needs no GPU)
int track_tile_prepare(const PbhcTrackCfg& c, const PbhcTableDesc& tab, const PbhcTrackIO& io) {
  static thread_local TileParams P;
  const int shmem = tile_plan(c, tab, io, 8, P);
  if (shmem <= 0) return shmem < 0 ? shmem : PBHC_E_ALIGN;
  g_track_info[0] = 2; g_track_info[1] = 0; g_track_info[2] = shmem; g_track_info[3] = 0;  // (pbhc_track_launch_info after a warm-up: the planned shared memory)
  bool noise = false;
  for (int s = 0; s < kSrcT; ++s) noise |= c.obs_noise[s] != 0.0f;
  return jit_track_tile(P.k, P.x, noise, nullptr) == 0 ? PBHC_OK : PBHC_E_CUDA;
}

}  // namespace pbhc
#endif  // __CUDACC_RTC__
