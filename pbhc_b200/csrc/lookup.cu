// lookup.cu - K1: motion-state lookup.  One warp per query, eight consecutive queries per block: frame index +
// blend (bit-exact restatement of _calc_frame_blend), then the two adjacent packed records are streamed with
// float4 loads (quaternions -> slerp) and coalesced scalar loads (everything else -> lerp).  The results are NOT
// scattered from registers: each warp assembles its query's rows in shared memory in the reference's output
// layout, one staging region per output tensor holding the block's eight rows back to back - which is exactly
// the tensor's global layout for those eight queries - and each region then leaves with ONE TMA bulk store
// (cp.async.bulk shared -> global; 8 rows are always a multiple of 32 bytes), issued by one thread per tensor.
// No 324-byte rows written in 4-byte pieces, no 64-bit index arithmetic per element.  Ragged last blocks and
// tensors whose base is not 16-byte aligned are flushed by a plain cooperative copy.
// Reference: utils/motion_lib/motion_lib_base.py:123-259, 503-513; isaac_utils/rotations.py:210-232.
#include "common.cuh"
#include "table.cuh"

namespace pbhc {

constexpr int kQ = 8;  // queries (= warps) per block
// staging regions, in PbhcLookupOut's field order (the struct is 20 pointers)
enum LookupArr { A_ROOT_POS, A_ROOT_ROT, A_DOF_POS, A_ROOT_VEL, A_ROOT_ANG, A_DOF_VEL, A_AA, A_BODIES, A_RG_POS, A_RB_ROT,
                 A_BODY_VEL, A_BODY_ANG, A_RG_POS_T, A_RG_ROT_T, A_BODY_VEL_T, A_BODY_ANG_T, A_CONTACT, A_F0, A_F1, A_BLEND, kArr };
static_assert(sizeof(PbhcLookupOut) == kArr * sizeof(void*), "PbhcLookupOut is not 20 pointers");

struct LookupParams {
  PbhcTableDesc tab;
  void* dst[kArr];     // output tensors (NULL: not requested)
  int row[kArr];       // row length in 32-bit words (the int64 frame indices: 2)
  int off[kArr];       // word offset of the tensor's staging region (kQ rows)
  uint32_t bulk_mask;  // tensors whose base pointer is 16-byte aligned
  const int32_t* clip_of_motion;
  const int64_t* motion_ids;
  const float* motion_times;
  const float* offset;
  int64_t n;
};

#ifndef K1_MINB
#define K1_MINB 6  // 40 registers: six blocks (48 warps) per SM hide the ids -> clip -> frame -> record load chain (64 registers / 4 blocks: 0.094 ms, 40 / 6: 0.082 ms)
#endif
__global__ void __launch_bounds__(kQ * 32, K1_MINB) lookup_kernel(const __grid_constant__ LookupParams p) {
  extern __shared__ __align__(16) float sm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t q0 = (int64_t)blockIdx.x * kQ, q = q0 + w;
  const int nq = (int)min((int64_t)kQ, p.n - q0);
  const PbhcTableDesc& tab = p.tab;
  const int J = tab.num_joints, Jb = tab.num_bodies, D = tab.num_dof, C = tab.contact_size;
  auto O = [&](int a) { return p.off[a] + w * p.row[a]; };  // word offset of this warp's row of tensor a

  // the all-zero tensors (motion_aa, motion_bodies): the block clears both staging regions with 128-bit stores
  {
    float4* z = reinterpret_cast<float4*>(sm + p.off[A_AA]);  // A_AA and A_BODIES are adjacent regions, kQ rows each (multiples of 32 B)
    const int nz = (kQ * (p.row[A_AA] + p.row[A_BODIES])) >> 2;
    for (int i = threadIdx.x; i < nz; i += kQ * 32) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (q < p.n) {
    const int clip = resolve_clip(tab, p.clip_of_motion, p.motion_ids[q]);
    const FrameBlend fb = frame_blend(tab, clip, p.motion_times[q]);
    const float b = fb.blend, omb = 1.0f - b;
    const float* r0 = tab.records + fb.f0 * tab.rec_stride;
    const float* r1 = tab.records + fb.f1 * tab.rec_stride;
    if (lane == 0) {
      *reinterpret_cast<long long*>(sm + O(A_F0)) = fb.f0;
      *reinterpret_cast<long long*>(sm + O(A_F1)) = fb.f1;
      sm[O(A_BLEND)] = b;
    }
    // offset component of element i = lane + 32 k of a [J,3] row: (lane + 2 k) % 3 - three values that rotate with k
    float oa = 0.f, ob = 0.f, oc = 0.f;
    if (p.offset) {
      const float ox = p.offset[q * 3], oy = p.offset[q * 3 + 1], oz = p.offset[q * 3 + 2];
      const int m = lane % 3;
      oa = m == 0 ? ox : m == 1 ? oy : oz;  // k = 0: component m
      ob = m == 0 ? oz : m == 1 ? ox : oy;  // k = 1: component (m + 2) % 3
      oc = m == 0 ? oy : m == 1 ? oz : ox;  // k = 2: component (m + 1) % 3
    }

    // rotations: one quaternion per lane
    {
      float4* ext = reinterpret_cast<float4*>(sm + O(A_RG_ROT_T));
      float4* base = reinterpret_cast<float4*>(sm + O(A_RB_ROT));
      for (int j = lane; j < J; j += 32) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(r0) + j);
        const float4 c = __ldg(reinterpret_cast<const float4*>(r1) + j);
        const Q4 s = slerp({a.x, a.y, a.z, a.w}, {c.x, c.y, c.z, c.w}, b);
        const float4 v = make_float4(s.x, s.y, s.z, s.w);
        ext[j] = v;
        if (j < Jb) base[j] = v;
        if (j == 0) *reinterpret_cast<float4*>(sm + O(A_ROOT_ROT)) = v;
      }
    }
    // positions / velocities / angular velocities: element i of the three [J,3] record segments per pass - six loads in
    // flight, one set of predicates for the extended, base-body and root rows.  (Requesting every record element up
    // front - 30 loads in flight per lane, 64-80 registers - was measured SLOWER, 0.088 vs 0.082 ms: resident warps
    // matter more than per-warp memory parallelism here.)
    {
      const int J3 = 3 * J, Jb3 = 3 * Jb;
      const float* a0 = r0 + 4 * J + lane;
      const float* a1 = r1 + 4 * J + lane;
      float* e_pos = sm + O(A_RG_POS_T) + lane; float* e_vel = sm + O(A_BODY_VEL_T) + lane; float* e_ang = sm + O(A_BODY_ANG_T) + lane;
      float* b_pos = sm + O(A_RG_POS) + lane;   float* b_vel = sm + O(A_BODY_VEL) + lane;   float* b_ang = sm + O(A_BODY_ANG) + lane;
      for (int i = lane, k = 0; i < J3; i += 32, k += 32) {
        const float p0 = __ldg(a0 + k), p1 = __ldg(a1 + k);
        const float v0 = __ldg(a0 + J3 + k), v1 = __ldg(a1 + J3 + k);
        const float w0 = __ldg(a0 + 2 * J3 + k), w1 = __ldg(a1 + 2 * J3 + k);
        const float vp = (omb * p0 + b * p1) + oa;
        const float vv = omb * v0 + b * v1;
        const float vw = omb * w0 + b * w1;
        const float t = oa; oa = ob; ob = oc; oc = t;
        e_pos[k] = vp; e_vel[k] = vv; e_ang[k] = vw;
        if (i < Jb3) { b_pos[k] = vp; b_vel[k] = vv; b_ang[k] = vw; }
        if (i < 3) { sm[O(A_ROOT_POS) + i] = vp; sm[O(A_ROOT_VEL) + i] = vv; sm[O(A_ROOT_ANG) + i] = vw; }
      }
    }
    // dof_pos | dof_vel | contact are adjacent in the record
    {
      const int e0 = 13 * J;
      const int o_dp = O(A_DOF_POS), o_dv = O(A_DOF_VEL) - D, o_cm = O(A_CONTACT) - 2 * D;
      for (int i = lane; i < 2 * D + C; i += 32) {
        const float v = omb * __ldg(r0 + e0 + i) + b * __ldg(r1 + e0 + i);
        sm[(i < D ? o_dp : i < 2 * D ? o_dv : o_cm) + i] = v;
      }
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the bulk stores
  __syncthreads();
  // flush: thread a owns tensor a
  const int a = threadIdx.x;
  if (a < kArr && p.dst[a] != nullptr && nq == kQ && ((p.bulk_mask >> a) & 1u)) {
    const uint32_t bytes = (uint32_t)(kQ * p.row[a]) * 4u;
    char* g = reinterpret_cast<char*>(p.dst[a]) + q0 * (int64_t)p.row[a] * 4;
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(sm + p.off[a]);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(s), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  // ragged last block / unaligned tensors: cooperative word copy (block-uniform conditions)
  if (nq < kQ || p.bulk_mask != (1u << kArr) - 1u) {
    for (int t = 0; t < kArr; ++t) {
      if (p.dst[t] == nullptr || (nq == kQ && ((p.bulk_mask >> t) & 1u))) continue;
      const uint32_t* s = reinterpret_cast<const uint32_t*>(sm + p.off[t]);
      uint32_t* g = reinterpret_cast<uint32_t*>(p.dst[t]) + q0 * (int64_t)p.row[t];
      for (int i = threadIdx.x; i < nq * p.row[t]; i += kQ * 32) g[i] = s[i];
    }
  }
  if (a < kArr) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the staging rows must outlive the copies' reads
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
  LookupParams p;
  p.tab = *tab; p.clip_of_motion = clip_of_motion; p.motion_ids = motion_ids; p.motion_times = motion_times; p.offset = offset; p.n = n;
  const int J = tab->num_joints, Jb = tab->num_bodies, D = tab->num_dof, C = tab->contact_size;
  const int rows[kArr] = {3, 4, D, 3, 3, D, 3 * Jb, 17, 3 * Jb, 4 * Jb, 3 * Jb, 3 * Jb, 3 * J, 4 * J, 3 * J, 3 * J, C, 2, 2, 1};
  void* const* dst = reinterpret_cast<void* const*>(out);
  int words = 0;
  p.bulk_mask = 0;
  for (int a = 0; a < kArr; ++a) {
    p.dst[a] = rows[a] > 0 ? dst[a] : nullptr;
    p.row[a] = rows[a];
    p.off[a] = words;                         // kQ rows = a multiple of 8 words: every region stays 32-byte aligned
    words += kQ * rows[a];
    if (p.dst[a] == nullptr || ((uintptr_t)p.dst[a] & 15) == 0) p.bulk_mask |= 1u << a;
  }
  const size_t shmem = (size_t)words * 4;
  static thread_local size_t configured = 48 * 1024;
  if (shmem > configured) {
    if (cudaFuncSetAttribute(lookup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_lookup(attr)");
    configured = shmem;
  }
  const int64_t blocks = (n + kQ - 1) / kQ;
  if (blocks > 0x7fffffffLL) { set_error("pbhc_lookup: too many queries"); return PBHC_E_RANGE; }
  lookup_kernel<<<(unsigned)blocks, kQ * 32, shmem, (cudaStream_t)stream>>>(p);
  return check_launch("pbhc_lookup");
}
