// motion_prep.cu - load-time clip preparation on the device: from FK output to
// the packed per-frame records of the motion table.
//   pass 1: raw body velocity (np.gradient / dt) and angular velocity (q[t+1] * conj(q[t]) ->
//           angle * axis / dt)                                   torch_humanoid_batch.py:272-290
//   pass 2: Gaussian filter sigma=2 (scipy gaussian_filter1d, mode="nearest", computed in double
//           with correlate1d's symmetric summation order), dof_pos = pose_aa.sum(-1)[1:Jb],
//           dof_vel forward difference padded with its row T-3 (:216-224), pack the record.
#include "common.cuh"

namespace pbhc {

constexpr int kMaxRadius = 16;
struct GaussW { int radius; double w[2 * kMaxRadius + 1]; };
// the frames of several clips back to back: per-frame clip id and per-clip (first frame, frames, dt); NULL = one clip
struct ClipMap { const int32_t* clip_of_frame; const int64_t* clip_start; const int32_t* clip_frames; const float* clip_dt; int T; float dt; };
struct ClipAt { int64_t base; int t, T; float dt; };
__device__ __forceinline__ ClipAt clip_at(const ClipMap& M, int64_t f) {
  if (!M.clip_of_frame) return {0, (int)f, M.T, M.dt};
  const int c = M.clip_of_frame[f];
  const int64_t b = M.clip_start[c];
  return {b, (int)(f - b), M.clip_frames[c], M.clip_dt[c]};
}

__global__ void clip_raw_vel_kernel(const float* __restrict__ pos_all, const float* __restrict__ quat_all, int64_t F, int J, const ClipMap M,
                                    float* __restrict__ scratch) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t tot = F * J;
  if (i >= tot) return;
  const int64_t f = i / J;
  const int j = (int)(i - f * J);
  const ClipAt ca = clip_at(M, f);
  const int t = ca.t, T = ca.T;
  const float dt = ca.dt;
  const float* pos = pos_all + ca.base * J * 3;
  const float* quat = quat_all + ca.base * J * 4;
  float* v = scratch + i * 6;
  // np.gradient: one-sided at the ends, central elsewhere (float32), then / dt
  const int tm = t > 0 ? t - 1 : t, tp = t < T - 1 ? t + 1 : t;
  const float* pm = pos + ((int64_t)tm * J + j) * 3;
  const float* pp = pos + ((int64_t)tp * J + j) * 3;
  for (int c = 0; c < 3; ++c) {
    float g = (T == 1) ? 0.0f : ((tp - tm == 2) ? (pp[c] - pm[c]) / 2.0f : (pp[c] - pm[c]));
    v[c] = g / dt;
  }
  // angular velocity
  Q4 dq = {0.f, 0.f, 0.f, 1.f};
  if (t < T - 1) {
    const float* a = quat + ((int64_t)(t + 1) * J + j) * 4;
    const float* b = quat + ((int64_t)t * J + j) * 4;
    Q4 m = quat_mul({a[0], a[1], a[2], a[3]}, quat_conj({b[0], b[1], b[2], b[3]}));
    float n = norm4(m.x, m.y, m.z, m.w);
    n = fmaxf(n, 1e-9f);
    dq = {m.x / n, m.y / n, m.z / n, m.w / n};
  }
  float s = 2.0f * (dq.w * dq.w) - 1.0f;
  s = fminf(fmaxf(s, -1.0f), 1.0f);
  const float ang = acosf(s);
  const float an = fmaxf(norm3(dq.x, dq.y, dq.z), 1e-9f);
  v[3] = (dq.x / an) * ang / dt;
  v[4] = (dq.y / an) * ang / dt;
  v[5] = (dq.z / an) * ang / dt;
}

__global__ void clip_pack_kernel(const float* __restrict__ pos_all, const float* __restrict__ quat_all, const float* __restrict__ pose_all,
                                 const float* __restrict__ contact_all, const float* __restrict__ scratch_all, const ClipMap M, int J, int Jb, int D,
                                 int C, const __grid_constant__ GaussW G, float* __restrict__ records, int rec_stride) {
  const int64_t f = blockIdx.x;
  const ClipAt ca = clip_at(M, f);
  const int t = ca.t, T = ca.T;
  const float dt = ca.dt;
  const float* pos = pos_all + ca.base * J * 3;
  const float* quat = quat_all + ca.base * J * 4;
  const float* pose_aa = pose_all + ca.base * J * 3;
  const float* contact = contact_all ? contact_all + ca.base * C : nullptr;
  const float* scratch = scratch_all + ca.base * J * 6;
  float* rec = records + f * rec_stride;
  const int e_pos = 4 * J, e_vel = 7 * J, e_ang = 10 * J, e_dp = 13 * J, e_dv = e_dp + D, e_c = e_dv + D, e_end = e_c + C;
  for (int e = threadIdx.x; e < rec_stride; e += blockDim.x) {
    float out = 0.0f;
    if (e < e_pos) out = quat[(int64_t)t * 4 * J + e];
    else if (e < e_vel) out = pos[(int64_t)t * 3 * J + (e - e_pos)];
    else if (e < e_dp) {
      const bool is_ang = e >= e_ang;
      const int idx = is_ang ? e - e_ang : e - e_vel;
      const int j = idx / 3, c = idx - j * 3 + (is_ang ? 3 : 0);
      auto at = [&](int tt) { tt = tt < 0 ? 0 : (tt > T - 1 ? T - 1 : tt); return (double)scratch[((int64_t)tt * J + j) * 6 + c]; };
      const int r = G.radius;
      double acc = at(t) * G.w[r];
      for (int l = -r; l < 0; ++l) acc += (at(t + l) + at(t - l)) * G.w[l + r];
      out = (float)acc;
    } else if (e < e_c) {
      const bool vel = e >= e_dv;
      const int d = vel ? e - e_dv : e - e_dp;
      auto dof = [&](int tt) { const float* a = pose_aa + ((int64_t)tt * J + (d + 1)) * 3; return (a[0] + a[1]) + a[2]; };
      if (!vel) out = dof(t);
      else {
        int tt = t < T - 1 ? t : T - 3;  // cat([dv, dv[-2:-1]]): the pad row is dv[T-3]
        out = (T >= 3 || t < T - 1) ? (dof(tt + 1) - dof(tt)) / dt : 0.0f;
      }
    } else if (e < e_end) out = contact[(int64_t)t * C + (e - e_c)];
    rec[e] = out;
  }
}

}  // namespace pbhc

static int clips_prepare(const float* pos, const float* quat, const float* pose_aa, const float* contact, int64_t F, const pbhc::ClipMap& M,
                         int32_t J, int32_t Jb, int32_t D, int32_t C, const double* gauss_weights_host, int32_t radius, float* scratch,
                         float* records, int32_t rec_stride, void* stream, const char* what) {
  using namespace pbhc;
  if (!pos || !quat || !pose_aa || !scratch || !records || !gauss_weights_host || (C > 0 && !contact)) { set_error("%s: null argument", what); return PBHC_E_BADARG; }
  if (F <= 0 || F > 0x7fffffffLL || J <= 0 || Jb > J || D + 1 > J || radius < 0 || radius > kMaxRadius || rec_stride < 13 * J + 2 * D + C) {
    set_error("%s: bad shape", what); return PBHC_E_BADARG;
  }
  GaussW G;
  G.radius = radius;
  for (int i = 0; i < 2 * radius + 1; ++i) G.w[i] = gauss_weights_host[i];
  const int64_t tot = F * J;
  clip_raw_vel_kernel<<<(unsigned)((tot + 127) / 128), 128, 0, (cudaStream_t)stream>>>(pos, quat, F, J, M, scratch);
  clip_pack_kernel<<<(unsigned)F, 128, 0, (cudaStream_t)stream>>>(pos, quat, pose_aa, contact, scratch, M, J, Jb, D, C, G, records, rec_stride);
  return check_launch(what);
}

extern "C" int pbhc_clip_prepare(
    const float* pos, const float* quat, const float* pose_aa, const float* contact, int32_t T, int32_t J, int32_t Jb,
    int32_t D, int32_t C, float dt, const double* gauss_weights_host, int32_t radius, float* scratch, float* records,
    int32_t rec_stride, void* stream) {
  const pbhc::ClipMap M = {nullptr, nullptr, nullptr, nullptr, T, dt};
  return clips_prepare(pos, quat, pose_aa, contact, T, M, J, Jb, D, C, gauss_weights_host, radius, scratch, records, rec_stride, stream, "pbhc_clip_prepare");
}

extern "C" int pbhc_clips_prepare(
    const float* pos, const float* quat, const float* pose_aa, const float* contact, int64_t total_frames, const int32_t* clip_of_frame,
    const int64_t* clip_start, const int32_t* clip_frames, const float* clip_dt, int32_t J, int32_t Jb, int32_t D, int32_t C,
    const double* gauss_weights_host, int32_t radius, float* scratch, float* records, int32_t rec_stride, void* stream) {
  if (!clip_of_frame || !clip_start || !clip_frames || !clip_dt) { pbhc::set_error("pbhc_clips_prepare: null clip map"); return PBHC_E_BADARG; }
  const pbhc::ClipMap M = {clip_of_frame, clip_start, clip_frames, clip_dt, 0, 0.0f};
  return clips_prepare(pos, quat, pose_aa, contact, total_frames, M, J, Jb, D, C, gauss_weights_host, radius, scratch, records, rec_stride, stream, "pbhc_clips_prepare");
}
