// lookup.cu - K1: motion-state lookup.  One warp per query: frame index + blend (bit-exact
// restatement of _calc_frame_blend), then the two adjacent packed records are streamed with
// float4 loads (quaternions -> slerp) and coalesced scalar loads (everything else -> lerp) and
// scattered into the reference's output dict layout.
// Reference: utils/motion_lib/motion_lib_base.py:123-259, 503-513; isaac_utils/rotations.py:210-232.
#include "common.cuh"
#include "table.cuh"

namespace pbhc {

struct LookupParams {
  PbhcTableDesc tab;
  PbhcLookupOut out;
  const int32_t* clip_of_motion;
  const int64_t* motion_ids;
  const float* motion_times;
  const float* offset;
  int64_t n;
};

__device__ __forceinline__ void emit3(float* ext, float* base, float* root, int64_t q, int J, int Jb, int idx, float v) {
  if (ext) ext[q * (3 * J) + idx] = v;
  if (base && idx < 3 * Jb) base[q * (3 * Jb) + idx] = v;
  if (root && idx < 3) root[q * 3 + idx] = v;
}

constexpr int kWarps = 8;

__global__ void __launch_bounds__(kWarps * 32) lookup_kernel(const LookupParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t q = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (q >= p.n) return;
  const PbhcTableDesc& tab = p.tab;
  const PbhcLookupOut& o = p.out;
  const int J = tab.num_joints, Jb = tab.num_bodies, D = tab.num_dof, C = tab.contact_size;

  const int clip = resolve_clip(tab, p.clip_of_motion, p.motion_ids[q]);
  const FrameBlend fb = frame_blend(tab, clip, p.motion_times[q]);
  const float b = fb.blend, omb = 1.0f - b;
  const float* r0 = tab.records + fb.f0 * tab.rec_stride;
  const float* r1 = tab.records + fb.f1 * tab.rec_stride;
  if (lane == 0) {
    if (o.f0l) o.f0l[q] = fb.f0;
    if (o.f1l) o.f1l[q] = fb.f1;
    if (o.blend) o.blend[q] = b;
  }
  float off[3] = {0.f, 0.f, 0.f};
  if (p.offset) { off[0] = p.offset[q * 3]; off[1] = p.offset[q * 3 + 1]; off[2] = p.offset[q * 3 + 2]; }

  // rotations: one quaternion per lane
  for (int j = lane; j < J; j += 32) {
    float4 a = __ldg(reinterpret_cast<const float4*>(r0) + j);
    float4 c = __ldg(reinterpret_cast<const float4*>(r1) + j);
    Q4 s = slerp({a.x, a.y, a.z, a.w}, {c.x, c.y, c.z, c.w}, b);
    float4 v = make_float4(s.x, s.y, s.z, s.w);
    if (o.rg_rot_t) reinterpret_cast<float4*>(o.rg_rot_t)[q * J + j] = v;
    if (o.rb_rot && j < Jb) reinterpret_cast<float4*>(o.rb_rot)[q * Jb + j] = v;
    if (o.root_rot && j == 0) reinterpret_cast<float4*>(o.root_rot)[q] = v;
  }
  // everything else: lerp, element e of the record
  const int e_pos = 4 * J, e_vel = 7 * J, e_ang = 10 * J, e_dp = 13 * J, e_dv = e_dp + D, e_c = e_dv + D, e_end = e_c + C;
#pragma unroll 4
  for (int e = e_pos + lane; e < e_end; e += 32) {
    float v = omb * __ldg(r0 + e) + b * __ldg(r1 + e);
    if (e < e_vel) {
      int idx = e - e_pos;
      emit3(o.rg_pos_t, o.rg_pos, o.root_pos, q, J, Jb, idx, v + off[idx % 3]);
    } else if (e < e_ang) {
      emit3(o.body_vel_t, o.body_vel, o.root_vel, q, J, Jb, e - e_vel, v);
    } else if (e < e_dp) {
      emit3(o.body_ang_vel_t, o.body_ang_vel, o.root_ang_vel, q, J, Jb, e - e_ang, v);
    } else if (e < e_dv) {
      if (o.dof_pos) o.dof_pos[q * D + (e - e_dp)] = v;
    } else if (e < e_c) {
      if (o.dof_vel) o.dof_vel[q * D + (e - e_dv)] = v;
    } else {
      if (o.contact_mask) o.contact_mask[q * C + (e - e_c)] = v;
    }
  }
  if (o.motion_aa) for (int i = lane; i < 3 * Jb; i += 32) o.motion_aa[q * (3 * Jb) + i] = 0.0f;
  if (o.motion_bodies && lane < 17) o.motion_bodies[q * 17 + lane] = 0.0f;
}

}  // namespace pbhc

extern "C" int pbhc_lookup(const PbhcTableDesc* tab, const int32_t* clip_of_motion, const int64_t* motion_ids,
                           const float* motion_times, const float* offset, int64_t n, const PbhcLookupOut* out,
                           void* stream) {
  using namespace pbhc;
  if (!tab || !out || n < 0) { set_error("pbhc_lookup: null argument"); return PBHC_E_BADARG; }
  if (n > 0 && (!motion_ids || !motion_times)) { set_error("pbhc_lookup: null motion_ids / motion_times"); return PBHC_E_BADARG; }
  if (!tab->records || tab->num_clips <= 0 || tab->num_joints <= 0 || tab->num_bodies > tab->num_joints) {
    set_error("pbhc_lookup: bad table descriptor"); return PBHC_E_BADARG;
  }
  const int need = 13 * tab->num_joints + 2 * tab->num_dof + tab->contact_size;
  if (tab->rec_stride < need || (tab->rec_stride & 3)) { set_error("pbhc_lookup: rec_stride %d (need >= %d, multiple of 4)", tab->rec_stride, need); return PBHC_E_BADARG; }
  const uintptr_t al = (uintptr_t)tab->records | (uintptr_t)out->rg_rot_t | (uintptr_t)out->rb_rot | (uintptr_t)out->root_rot;
  if (al & 15) { set_error("pbhc_lookup: records / quaternion outputs must be 16-byte aligned"); return PBHC_E_ALIGN; }
  if (n == 0) return PBHC_OK;
  LookupParams p{*tab, *out, clip_of_motion, motion_ids, motion_times, offset, n};
  const int64_t blocks = (n + kWarps - 1) / kWarps;
  if (blocks > 0x7fffffffLL) { set_error("pbhc_lookup: too many queries"); return PBHC_E_RANGE; }
  lookup_kernel<<<(unsigned)blocks, kWarps * 32, 0, (cudaStream_t)stream>>>(p);
  return check_launch("pbhc_lookup");
}
