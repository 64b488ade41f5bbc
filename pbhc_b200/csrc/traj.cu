// traj.cu - trajectory evaluation: resampling of a reference motion at the recorded policy times and
// the accuracy / smoothness metrics of a recorded trajectory against it.
// Reference: humanoidverse/measure_traj.py:67-133 (blend_motion), :146-225 (eval_accuracy), :227-290 (eval_smoothness).
#include "common.cuh"

namespace pbhc {

// measure_traj.py:68-84 (_calc_frame_blend of blend_motion; `len` = num_frames / fps, not the motion-lib length)
struct Blend { int64_t f0, f1; float b; };
__device__ __forceinline__ Blend traj_frame_blend(float time, float len, int64_t num_frames, float dt) {
  float phase = time / len;
  phase = fminf(fmaxf(phase, 0.0f), 1.0f);
  if (time < 0.0f) time = 0.0f;
  Blend r;
  r.f0 = (int64_t)(phase * (float)(num_frames - 1));
  r.f1 = r.f0 + 1 < num_frames - 1 ? r.f0 + 1 : num_frames - 1;
  r.b = clip01((time - (float)r.f0 * dt) / dt);
  return r;
}

// out[t, c] for one key: lerp `b * x1 + (1 - b) * x0` (measure_traj.py:128) or slerp over groups of 4 (xyzw, :121)
__global__ void __launch_bounds__(256) blend_frames_kernel(const float* __restrict__ src, int64_t num_frames, int row, int quats,
                                                           const float* __restrict__ times, int64_t n_out, float len, float dt,
                                                           float* __restrict__ out) {
  const int units = quats ? row / 4 : row;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_out * units) return;
  const int64_t t = i / units;
  const int u = (int)(i - t * units);
  const Blend fb = traj_frame_blend(times[t], len, num_frames, dt);
  if (quats) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src + fb.f0 * row) + u);
    const float4 b = __ldg(reinterpret_cast<const float4*>(src + fb.f1 * row) + u);
    const Q4 q = slerp({a.x, a.y, a.z, a.w}, {b.x, b.y, b.z, b.w}, fb.b);
    reinterpret_cast<float4*>(out + t * row)[u] = make_float4(q.x, q.y, q.z, q.w);
  } else {
    const float x0 = __ldg(src + fb.f0 * row + u), x1 = __ldg(src + fb.f1 * row + u);
    out[t * row + u] = fb.b * x1 + (1.0f - fb.b) * x0;
  }
}

// ---- metrics -------------------------------------------------------------------------------------------------------
// One warp per frame t.  Finite differences are taken like the reference, one order at a time in fp32
// (`(x[1:] - x[:-1]) * delta`), so every per-element value is the reference's; the means are accumulated in fp64.
constexpr int kTrajWarps = 8;

struct TrajParams {
  const float *pol_pos, *ref_pos, *pol_dof, *ref_dof, *pol_contact, *ref_contact;
  int64_t T;
  int J, D, C;
  float delta;
};

// first three difference orders of x at frame t from x[t..t+3] (entries beyond T are unused by the caller)
struct Diffs { float v, a, j; };
__device__ __forceinline__ Diffs diffs(float x0, float x1, float x2, float x3, float d) {
  const float v0 = (x1 - x0) * d, v1 = (x2 - x1) * d, v2 = (x3 - x2) * d;
  const float a0 = (v1 - v0) * d, a1 = (v2 - v1) * d;
  return {v0, a0, (a1 - a0) * d};
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(kTrajWarps * 32) traj_metrics_kernel(const TrajParams P, double* __restrict__ partials) {
  __shared__ double sm[kTrajWarps][PBHC_TRAJ_METRICS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double acc[PBHC_TRAJ_METRICS];
#pragma unroll
  for (int i = 0; i < PBHC_TRAJ_METRICS; ++i) acc[i] = 0.0;
  const float d = P.delta;
  for (int64_t t = (int64_t)blockIdx.x * kTrajWarps + warp; t < P.T; t += (int64_t)gridDim.x * kTrajWarps) {
    const int avail = (int)(P.T - 1 - t < 3 ? P.T - 1 - t : 3);  // difference orders that exist at t
    // body positions: |.| over xyz, mean over (t, body)
    const int64_t fs = (int64_t)P.J * 3;
    float rp[3], rr[3];  // roots of frame t (root-relative error)
#pragma unroll
    for (int c = 0; c < 3; ++c) { rp[c] = P.pol_pos[t * fs + c]; rr[c] = P.ref_pos[t * fs + c]; }
    for (int j = lane; j < P.J; j += 32) {
      float e0[3], ev[3], ea[3], rel[3], pv[3], pa[3], pj[3], rv[3], ra[3], rj[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float p[4], r[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t o = (t + (k <= avail ? k : avail)) * fs + j * 3 + c;
          p[k] = P.pol_pos[o]; r[k] = P.ref_pos[o];
        }
        const Diffs dp = diffs(p[0], p[1], p[2], p[3], d), dr = diffs(r[0], r[1], r[2], r[3], d);
        e0[c] = p[0] - r[0];
        rel[c] = (p[0] - rp[c]) - (r[0] - rr[c]);
        ev[c] = dp.v - dr.v; ea[c] = dp.a - dr.a;
        pv[c] = dp.v; pa[c] = dp.a; pj[c] = dp.j; rv[c] = dr.v; ra[c] = dr.a; rj[c] = dr.j;
      }
      acc[PBHC_TM_GMPBPE] += norm3(e0[0], e0[1], e0[2]);
      acc[PBHC_TM_MPBPE] += norm3(rel[0], rel[1], rel[2]);
      if (avail >= 1) {
        const float n = norm3(ev[0], ev[1], ev[2]);
        acc[PBHC_TM_PBVE] += n;
        if (j == 0) acc[PBHC_TM_ROOT_VEL] += n;
        acc[PBHC_TM_L2_VEL] += norm3(pv[0], pv[1], pv[2]);
        acc[PBHC_TM_L2_REF_VEL] += norm3(rv[0], rv[1], rv[2]);
      }
      if (avail >= 2) {
        const float n = norm3(ea[0], ea[1], ea[2]);
        acc[PBHC_TM_PBAE] += n;
        if (j == 0) acc[PBHC_TM_ROOT_ACC] += n;
        acc[PBHC_TM_L2_ACC] += norm3(pa[0], pa[1], pa[2]);
        acc[PBHC_TM_L2_REF_ACC] += norm3(ra[0], ra[1], ra[2]);
      }
      if (avail >= 3) {
        acc[PBHC_TM_L2_JERK] += norm3(pj[0], pj[1], pj[2]);
        acc[PBHC_TM_L2_REF_JERK] += norm3(rj[0], rj[1], rj[2]);
      }
    }
    // dof vectors: |.| over all D dofs of a frame, mean over t
    float s[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = lane; k < P.D; k += 32) {
      float p[4], r[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t o = (t + (q <= avail ? q : avail)) * P.D + k;
        p[q] = P.pol_dof[o]; r[q] = P.ref_dof[o];
      }
      const Diffs dp = diffs(p[0], p[1], p[2], p[3], d), dr = diffs(r[0], r[1], r[2], r[3], d);
      const float e = p[0] - r[0], ev = dp.v - dr.v, ea = dp.a - dr.a;
      s[0] += e * e; s[1] += ev * ev; s[2] += ea * ea;
      s[3] += dp.v * dp.v; s[4] += dp.a * dp.a; s[5] += dp.j * dp.j;
      s[6] += dr.v * dr.v; s[7] += dr.a * dr.a; s[8] += dr.j * dr.j;
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) s[i] = warp_sum(s[i]);
    if (lane == 0) {
      acc[PBHC_TM_MPJPE] += sqrtf(s[0]);
      if (avail >= 1) { acc[PBHC_TM_MPJVE] += sqrtf(s[1]); acc[PBHC_TM_L2_DOF_VEL] += sqrtf(s[3]); acc[PBHC_TM_L2_REF_DOF_VEL] += sqrtf(s[6]); }
      if (avail >= 2) { acc[PBHC_TM_MPJAE] += sqrtf(s[2]); acc[PBHC_TM_L2_DOF_ACC] += sqrtf(s[4]); acc[PBHC_TM_L2_REF_DOF_ACC] += sqrtf(s[7]); }
      if (avail >= 3) { acc[PBHC_TM_L2_DOF_JERK] += sqrtf(s[5]); acc[PBHC_TM_L2_REF_DOF_JERK] += sqrtf(s[8]); }
    }
    if (P.pol_contact)
      for (int k = lane; k < P.C; k += 32) acc[PBHC_TM_CONTACT] += fabsf(P.pol_contact[t * P.C + k] - P.ref_contact[t * P.C + k]);
  }
  // lanes -> warp -> block, fixed order
#pragma unroll
  for (int i = 0; i < PBHC_TRAJ_METRICS; ++i) {
    double v = acc[i];
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) sm[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < PBHC_TRAJ_METRICS) {
    double v = 0.0;
    for (int w = 0; w < kTrajWarps; ++w) v += sm[w][threadIdx.x];
    partials[(int64_t)blockIdx.x * PBHC_TRAJ_METRICS + threadIdx.x] = v;
  }
}

__global__ void traj_metrics_finish_kernel(const double* __restrict__ partials, int blocks, TrajParams P, double* __restrict__ out) {
  const int i = threadIdx.x;
  if (i >= PBHC_TRAJ_METRICS) return;
  double v = 0.0;
  for (int b = 0; b < blocks; ++b) v += partials[(int64_t)b * PBHC_TRAJ_METRICS + i];
  // number of terms of each mean (an empty mean is NaN, like torch)
  int order = 0, per_frame = P.J;
  switch (i) {
    case PBHC_TM_GMPBPE: case PBHC_TM_MPBPE: break;
    case PBHC_TM_PBVE: case PBHC_TM_L2_VEL: case PBHC_TM_L2_REF_VEL: order = 1; break;
    case PBHC_TM_PBAE: case PBHC_TM_L2_ACC: case PBHC_TM_L2_REF_ACC: order = 2; break;
    case PBHC_TM_L2_JERK: case PBHC_TM_L2_REF_JERK: order = 3; break;
    case PBHC_TM_ROOT_VEL: order = 1; per_frame = 1; break;
    case PBHC_TM_ROOT_ACC: order = 2; per_frame = 1; break;
    case PBHC_TM_MPJPE: per_frame = 1; break;
    case PBHC_TM_MPJVE: case PBHC_TM_L2_DOF_VEL: case PBHC_TM_L2_REF_DOF_VEL: order = 1; per_frame = 1; break;
    case PBHC_TM_MPJAE: case PBHC_TM_L2_DOF_ACC: case PBHC_TM_L2_REF_DOF_ACC: order = 2; per_frame = 1; break;
    case PBHC_TM_L2_DOF_JERK: case PBHC_TM_L2_REF_DOF_JERK: order = 3; per_frame = 1; break;
    case PBHC_TM_CONTACT: per_frame = P.pol_contact ? P.C : 0; break;
  }
  const double n = (double)(P.T - order > 0 ? P.T - order : 0) * (double)per_frame;
  out[i] = v / n;
}

}  // namespace pbhc

extern "C" int pbhc_blend_frames(const float* src, int64_t num_frames, int32_t row_floats, int32_t quats, const float* times,
                                 int64_t n_out, float fps, float* out, void* stream) {
  using namespace pbhc;
  if (!src || !times || !out || num_frames <= 0 || row_floats <= 0 || n_out < 0 || !(fps > 0.0f)) {
    set_error("pbhc_blend_frames: null / non-positive argument"); return PBHC_E_BADARG;
  }
  if (quats && ((row_floats & 3) || ((uintptr_t)src & 15) || ((uintptr_t)out & 15))) {
    set_error("pbhc_blend_frames: quaternion rows need a multiple of 4 floats and 16-byte aligned buffers"); return PBHC_E_ALIGN;
  }
  if (n_out == 0) return PBHC_OK;
  // measure_traj.py:87-96: preblend_motion_length = num_frames / fps, dt = 1 / fps (python floats, then fp32 tensor math)
  const float len = (float)((double)num_frames / (double)fps), dt = (float)(1.0 / (double)fps);
  const int64_t n = n_out * (quats ? row_floats / 4 : row_floats);
  blend_frames_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src, num_frames, row_floats, quats, times, n_out, len, dt, out);
  return check_launch("pbhc_blend_frames");
}

extern "C" int pbhc_traj_metrics(const float* pol_pos, const float* ref_pos, const float* pol_dof, const float* ref_dof,
                                 const float* pol_contact, const float* ref_contact, int64_t T, int32_t J, int32_t D, int32_t C,
                                 float delta, double* workspace, int64_t workspace_doubles, double* out, void* stream) {
  using namespace pbhc;
  if (!pol_pos || !ref_pos || !pol_dof || !ref_dof || !workspace || !out || T < 0 || J <= 0 || D <= 0 || (!pol_contact) != (!ref_contact) ||
      (pol_contact && C <= 0)) {
    set_error("pbhc_traj_metrics: null / non-positive argument"); return PBHC_E_BADARG;
  }
  int64_t blocks = (T + kTrajWarps - 1) / kTrajWarps;
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (blocks < 1) blocks = 1;
  if (workspace_doubles < blocks * PBHC_TRAJ_METRICS) {
    set_error("pbhc_traj_metrics: workspace of %lld doubles, need %lld", (long long)workspace_doubles, (long long)(blocks * PBHC_TRAJ_METRICS));
    return PBHC_E_RANGE;
  }
  TrajParams P{pol_pos, ref_pos, pol_dof, ref_dof, pol_contact, ref_contact, T, J, D, C, delta};
  traj_metrics_kernel<<<(unsigned)blocks, kTrajWarps * 32, 0, (cudaStream_t)stream>>>(P, workspace);
  int rc = check_launch("pbhc_traj_metrics");
  if (rc) return rc;
  traj_metrics_finish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(workspace, (int)blocks, P, out);
  return check_launch("pbhc_traj_metrics(finish)");
}
