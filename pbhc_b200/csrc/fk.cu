// fk.cu - K2: G1 forward kinematics.  A warp owns a batch of 32 poses (thread per pose walks the
// kinematic tree in DFS order with the parent transform in registers; branch-point transforms
// are parked in per-warp shared memory).  Global I/O is staged through shared memory in chunks of
// kChunk joints so every global access is a run of contiguous floats per pose.
// Reference: utils/motion_lib/torch_humanoid_batch.py:168-269; isaac_utils/rotations.py:519-647.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace pbhc {

constexpr int kChunk = 3;         // joints per staging chunk
constexpr int kFkWarps = 1;         // one warp per block: 15 blocks (= pose tiles) fit an SM, 65536 poses are a single wave
constexpr int kMaxSlots = 8;
// output staging: one region per output array, [pose][columns of the chunk] with odd strides (conflict-free per-thread
// rows); a region mirrors the global layout of the chunk, so the flush is a linear walk without index arithmetic
constexpr int kPosRow = kChunk * 3, kRotRow = kChunk * 9, kQuatRow = kChunk * 4 + 1;
constexpr int kStageRow = kPosRow + kRotRow + kQuatRow;
static_assert((kPosRow & 1) && (kRotRow & 1) && (kQuatRow & 1), "odd shared-memory strides");
constexpr int kInRow = kChunk * 3 + 1;

struct FkParams {
  int J, n_slots;
  int8_t parent[PBHC_MAX_JOINTS];
  int8_t save_slot[PBHC_MAX_JOINTS];  // slot that keeps joint i's world transform, or -1
  int8_t load_slot[PBHC_MAX_JOINTS];  // slot holding the parent's transform, or -1 (= previous joint)
  float off[PBHC_MAX_JOINTS * 3];
  float lrot[PBHC_MAX_JOINTS * 9];
  // the tree is cut into n_parts joint ranges [part_begin, part_end) walked by different warps of the same 32 poses (a thread
  // per pose leaves a B200 with 14 warps per SM at 65536 poses: latency-bound); a part first recomputes, without output,
  // the ancestors of its range that lie outside it (`silent`, root first)
  int n_parts, part_begin[2], part_end[2], n_silent[2];
  int8_t silent[2][8];
};

struct M3 { float m[9]; };

__device__ __forceinline__ M3 matmul3(const M3& a, const M3& b) {
  M3 c;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) c.m[i * 3 + j] = __fmaf_rn(a.m[i * 3 + 2], b.m[6 + j], __fmaf_rn(a.m[i * 3 + 1], b.m[3 + j], a.m[i * 3] * b.m[j]));
  return c;
}

// axis_angle_to_quaternion (rotations.py:554-578) -> quaternion_to_matrix (rotations.py:519-550)
__device__ __forceinline__ M3 aa_to_mat(float ax, float ay, float az) {
  const float ang = norm3(ax, ay, az);
  const float half = ang * 0.5f;
  const float k = (fabsf(ang) < 1e-6f) ? (0.5f - (ang * ang) / 48.0f) : (sinf(half) / ang);
  const float r = cosf(half), i = ax * k, j = ay * k, kk = az * k;
  const float two_s = 2.0f / (((r * r + i * i) + j * j) + kk * kk);
  M3 o;
  o.m[0] = 1.0f - two_s * (j * j + kk * kk); o.m[1] = two_s * (i * j - kk * r); o.m[2] = two_s * (i * kk + j * r);
  o.m[3] = two_s * (i * j + kk * r); o.m[4] = 1.0f - two_s * (i * i + kk * kk); o.m[5] = two_s * (j * kk - i * r);
  o.m[6] = two_s * (i * kk - j * r); o.m[7] = two_s * (j * kk + i * r); o.m[8] = 1.0f - two_s * (i * i + j * j);
  return o;
}

// The same two conversions for the per-step fk_batch (kFast): within the 1e-5 rel / 1e-6 abs bar of the reference but not
// operation-for-operation - one sincosf, reciprocal square root / approximate division (2 ulp) instead of IEEE sqrt and
// division, FMA everywhere; matrix_to_quaternion takes ONE square root (the candidate with the largest 1 +- m00 +- m11 +- m22
// is the one with the largest sqrt).  Load-time FK (load_motions -> the motion table, whose angular velocities amplify
// last-ulp differences) keeps the exact sequence below.
__device__ __forceinline__ M3 aa_to_mat_fast(float ax, float ay, float az) {
  const float s2 = __fmaf_rn(az, az, __fmaf_rn(ay, ay, ax * ax));
  const float rinv = rsqrtf(fmaxf(s2, 1e-30f)), ang = s2 * rinv;
  const float half = ang * 0.5f;
  float sh, r;
  sincosf(half, &sh, &r);
  const float k = (ang < 1e-6f) ? __fmaf_rn(s2, -1.0f / 48.0f, 0.5f) : sh * rinv;
  const float i = ax * k, j = ay * k, kk = az * k;
  const float two_s = __fdividef(2.0f, __fmaf_rn(kk, kk, __fmaf_rn(j, j, __fmaf_rn(i, i, r * r))));
  const float ii = two_s * i, jj = two_s * j, k2 = two_s * kk;
  M3 o;
  o.m[0] = __fmaf_rn(-jj, j, __fmaf_rn(-k2, kk, 1.0f)); o.m[1] = __fmaf_rn(ii, j, -k2 * r); o.m[2] = __fmaf_rn(ii, kk, jj * r);
  o.m[3] = __fmaf_rn(ii, j, k2 * r); o.m[4] = __fmaf_rn(-ii, i, __fmaf_rn(-k2, kk, 1.0f)); o.m[5] = __fmaf_rn(jj, kk, -ii * r);
  o.m[6] = __fmaf_rn(ii, kk, -jj * r); o.m[7] = __fmaf_rn(jj, kk, ii * r); o.m[8] = __fmaf_rn(-ii, i, __fmaf_rn(-jj, j, 1.0f));
  return o;
}
__device__ __forceinline__ void mat_to_quat_xyzw_fast(const M3& R, float* q) {
  const float m00 = R.m[0], m01 = R.m[1], m02 = R.m[2], m10 = R.m[3], m11 = R.m[4], m12 = R.m[5], m20 = R.m[6], m21 = R.m[7], m22 = R.m[8];
  const float p0 = 1.0f + m00 + m11 + m22, p1 = 1.0f + m00 - m11 - m22, p2 = 1.0f - m00 + m11 - m22, p3 = 1.0f - m00 - m11 + m22;
  int best = 0;
  float pb = fmaxf(p0, 0.0f);
  if (p1 > pb) { best = 1; pb = p1; }
  if (p2 > pb) { best = 2; pb = p2; }
  if (p3 > pb) { best = 3; pb = p3; }
  float c0, c1, c2, c3;  // wxyz candidate of `best`
  if (best == 0) { c0 = pb; c1 = m21 - m12; c2 = m02 - m20; c3 = m10 - m01; }
  else if (best == 1) { c0 = m21 - m12; c1 = pb; c2 = m10 + m01; c3 = m02 + m20; }
  else if (best == 2) { c0 = m02 - m20; c1 = m10 + m01; c2 = pb; c3 = m12 + m21; }
  else { c0 = m10 - m01; c1 = m20 + m02; c2 = m21 + m12; c3 = pb; }
  const float qa = pb * rsqrtf(fmaxf(pb, 1e-30f));
  const float inv = __fdividef(0.5f, fmaxf(qa, 0.1f));
  q[0] = c1 * inv; q[1] = c2 * inv; q[2] = c3 * inv; q[3] = c0 * inv;
}

// matrix_to_quaternion (rotations.py:589-647) then wxyz_to_xyzw (torch_humanoid_batch.py:188)
__device__ __forceinline__ void mat_to_quat_xyzw(const M3& R, float* q) {
  const float m00 = R.m[0], m01 = R.m[1], m02 = R.m[2], m10 = R.m[3], m11 = R.m[4], m12 = R.m[5], m20 = R.m[6], m21 = R.m[7], m22 = R.m[8];
  float pre[4] = {1.0f + m00 + m11 + m22, 1.0f + m00 - m11 - m22, 1.0f - m00 + m11 - m22, 1.0f - m00 - m11 + m22};
  float qa[4];
  int best = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) { qa[i] = pre[i] > 0.0f ? sqrtf(pre[i]) : 0.0f; }
#pragma unroll
  for (int i = 1; i < 4; ++i) if (qa[i] > qa[best]) best = i;
  float c[4];
  if (best == 0) { c[0] = qa[0] * qa[0]; c[1] = m21 - m12; c[2] = m02 - m20; c[3] = m10 - m01; }
  else if (best == 1) { c[0] = m21 - m12; c[1] = qa[1] * qa[1]; c[2] = m10 + m01; c[3] = m02 + m20; }
  else if (best == 2) { c[0] = m02 - m20; c[1] = m10 + m01; c[2] = qa[2] * qa[2]; c[3] = m12 + m21; }
  else { c[0] = m10 - m01; c[1] = m20 + m02; c[2] = m21 + m12; c[3] = qa[3] * qa[3]; }
  const float d = 2.0f * fmaxf(qa[best], 0.1f);
  q[0] = c[1] / d; q[1] = c[2] / d; q[2] = c[3] / d; q[3] = c[0] / d;
}

// Row staging between global memory ([pose][row_len], poses of a warp are contiguous) and shared memory
// ([lane = pose][odd stride], conflict-free for the per-thread compute).  The copies run pose-major: for each pose the
// lanes cover the chunk's columns, so the per-lane shared / global offsets are loop invariants and an iteration is
// one LDS/LDG + one STG/STS (no integer division in the loop).
__device__ __forceinline__ void fill_rows(float* sm, int stride, const float* g, int row_len, int col0, int ncol, int npose, int lane) {
  if (lane < ncol) {
    const float* gp = g + col0 + lane;
    float* sp = sm + lane;
#pragma unroll 4
    for (int t = 0; t < npose; ++t) sp[t * stride] = __ldg(gp + t * row_len);
  }
}
template <int NCOMP>
__device__ __forceinline__ void flush_rows(const float* sm, int stride, int sm_col0, int sm_jstride, float* g, int row_len, int col0,
                                           int nj, int npose, int lane) {
  const int ncol = nj * NCOMP;
  for (int c = lane; c < ncol; c += 32) {
    const int jj = c / NCOMP, cc = c - jj * NCOMP;
    const float* sp = sm + jj * sm_jstride + sm_col0 + cc;
    float* gp = g + col0 + c;
#pragma unroll 4
    for (int t = 0; t < npose; ++t) gp[t * row_len] = sp[t * stride];
  }
}

// Full tiles (32 poses, kChunk joints): every lane moves one element per step - element idx = lane + 32 k of the
// [pose][column] block, so a step stores 32 consecutive elements (runs of NC contiguous floats per pose) and no lane
// idles; NC is a compile-time constant, the divisions are multiplications.
template <int NC>
__device__ __forceinline__ void fill_full(float* sm, const float* g, int row_len, int col0, int lane) {
#pragma unroll 3
  for (int k = 0; k < NC; ++k) {
    const int idx = lane + 32 * k, t = idx / NC, c = idx - t * NC;
    sm[t * kInRow + c] = __ldg(g + t * row_len + col0 + c);
  }
}
// the same copy queued with cp.async (LDGSTS): the next chunk's pose rows arrive while the current chunk is computed
template <int NC>
__device__ __forceinline__ void fill_full_async(float* sm, const float* g, int row_len, int col0, int lane) {
  const uint32_t sd = (uint32_t)__cvta_generic_to_shared(sm);
#pragma unroll 3
  for (int k = 0; k < NC; ++k) {
    const int idx = lane + 32 * k, t = idx / NC, c = idx - t * NC;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sd + 4 * (t * kInRow + c)), "l"(g + t * row_len + col0 + c) : "memory");
  }
}
// full tile (32 poses): element idx = lane + 32 k of the [pose][NC columns] block, walked with running pointers
template <int NC, int S>
__device__ __forceinline__ void flush_lin(const float* sm, float* g, int row_len, int lane) {
  constexpr int DT = 32 / NC, DC = 32 % NC;
  int t = lane / NC, c = lane - t * NC;
  const float* sp = sm + t * S + c;
  float* gp = g + (int64_t)t * row_len + c;
  const int g_step = DT * row_len + DC, g_wrap = row_len - NC;
#pragma unroll 9
  for (int k = 0; k < NC; ++k) {  // 32 NC elements / 32 lanes
    *gp = *sp;
    c += DC; sp += DT * S + DC; gp += g_step;
    if (c >= NC) { c -= NC; sp += S - NC; gp += g_wrap; }
  }
}

template <bool kFast>
__global__ void __launch_bounds__(kFkWarps * 32, 16) fk_kernel(const __grid_constant__ FkParams P, const float* __restrict__ pose_aa,
                                                            const float* __restrict__ trans, int64_t n, float* __restrict__ pos,
                                                            float* __restrict__ rotmat, float* __restrict__ quat, int tile) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int J = P.J;
  const int per_warp = 2 * 32 * kInRow + 32 * kStageRow + P.n_slots * 12 * 32;
  float* sm_in0 = smem + warp * per_warp;  // two input buffers (full tiles prefetch the next chunk)
  float* sm_out = sm_in0 + 2 * 32 * kInRow;
  float* sm_pos = sm_out, *sm_rot = sm_pos + 32 * kPosRow, *sm_quat = sm_rot + 32 * kRotRow;
  float* sm_slot = sm_out + 32 * kStageRow;
  const int part = (int)(blockIdx.x % (unsigned)P.n_parts);
  const int64_t p0 = ((int64_t)(blockIdx.x / (unsigned)P.n_parts) * kFkWarps + warp) * tile;  // tile = poses per warp (32, or fewer for small batches)
  const int jb = P.part_begin[part], je = P.part_end[part];
  if (p0 >= n) return;
  const int npose = (int)min((int64_t)tile, n - p0);
  const bool live = lane < npose;
  const float* g_in = pose_aa + p0 * (3 * J);
  M3 Rc;  // world rotation of the previous joint
  float pc[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 9; ++i) Rc.m[i] = 0.f;

  const bool tile32 = npose == 32;
  if (tile32 && je - jb >= kChunk) fill_full_async<3 * kChunk>(sm_in0, g_in, 3 * J, 3 * jb, lane);
  if (live) {  // ancestors of this part's range outside it: world transform only (their outputs belong to another part)
    for (int s = 0; s < P.n_silent[part]; ++s) {
      const int i = P.silent[part][s];
      const float* a = g_in + (int64_t)lane * (3 * J) + 3 * i;
      const M3 Mj = kFast ? aa_to_mat_fast(__ldg(a), __ldg(a + 1), __ldg(a + 2)) : aa_to_mat(__ldg(a), __ldg(a + 1), __ldg(a + 2));
      if (P.parent[i] < 0) {
        Rc = Mj;
        const float* tr = trans + (p0 + lane) * 3;
        pc[0] = tr[0]; pc[1] = tr[1]; pc[2] = tr[2];
      } else {  // (the silent chain is root first: the previous one is the parent)
        M3 L;
#pragma unroll
        for (int c = 0; c < 9; ++c) L.m[c] = P.lrot[i * 9 + c];
        const M3 LMj = matmul3(L, Mj);
        const float ox = P.off[i * 3], oy = P.off[i * 3 + 1], oz = P.off[i * 3 + 2];
        float np_[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) np_[c] = __fmaf_rn(Rc.m[c * 3 + 2], oz, __fmaf_rn(Rc.m[c * 3 + 1], oy, Rc.m[c * 3] * ox)) + pc[c];
        pc[0] = np_[0]; pc[1] = np_[1]; pc[2] = np_[2];
        Rc = matmul3(Rc, LMj);
      }
      const int ss = P.save_slot[i];
      if (ss >= 0) {
        float* sl = sm_slot + ss * 12 * 32 + lane;
#pragma unroll
        for (int c = 0; c < 9; ++c) sl[c * 32] = Rc.m[c];
#pragma unroll
        for (int c = 0; c < 3; ++c) sl[(9 + c) * 32] = pc[c];
      }
    }
  }
  int buf = 0;
  for (int j0 = jb; j0 < je; j0 += kChunk, buf ^= 1) {
    const int nj = min(kChunk, je - j0);
    float* sm_in = sm_in0 + buf * 32 * kInRow;
    __syncwarp();
    const bool full = tile32 && nj == kChunk;
    if (full) {
      asm volatile("cp.async.wait_all;" ::: "memory");
      if (j0 + 2 * kChunk <= je) fill_full_async<3 * kChunk>(sm_in0 + (buf ^ 1) * 32 * kInRow, g_in, 3 * J, 3 * (j0 + kChunk), lane);
    } else {
      fill_rows(sm_in, kInRow, g_in, 3 * J, 3 * j0, 3 * nj, npose, lane);
    }
    __syncwarp();
    if (live) {
      // Three unrolled passes per chunk so that the work that does not depend on the parent (axis-angle -> matrix,
      // local_rot x joint matrix, matrix -> quaternion: three quarters of the arithmetic) is independent
      // straight-line code the scheduler can interleave; only `parent x local` walks the chain.
      M3 LM[kChunk], Rw[kChunk];
#pragma unroll
      for (int jj = 0; jj < kChunk; ++jj) {
        if (jj < nj) {
          const int i = j0 + jj;
          const float* a = sm_in + lane * kInRow + 3 * jj;
          const M3 Mj = kFast ? aa_to_mat_fast(a[0], a[1], a[2]) : aa_to_mat(a[0], a[1], a[2]);
          M3 L;
#pragma unroll
          for (int c = 0; c < 9; ++c) L.m[c] = P.lrot[i * 9 + c];
          LM[jj] = P.parent[i] < 0 ? Mj : matmul3(L, Mj);
        }
      }
#pragma unroll
      for (int jj = 0; jj < kChunk; ++jj) {
        if (jj < nj) {
          const int i = j0 + jj;
          if (P.parent[i] < 0) {
            Rc = LM[jj];
            const float* tr = trans + (p0 + lane) * 3;
            pc[0] = tr[0]; pc[1] = tr[1]; pc[2] = tr[2];
          } else {
            M3 Rp = Rc;
            float pp[3] = {pc[0], pc[1], pc[2]};
            const int ls = P.load_slot[i];
            if (ls >= 0) {
              const float* s = sm_slot + ls * 12 * 32 + lane;
#pragma unroll
              for (int c = 0; c < 9; ++c) Rp.m[c] = s[c * 32];
#pragma unroll
              for (int c = 0; c < 3; ++c) pp[c] = s[(9 + c) * 32];
            }
            const float ox = P.off[i * 3], oy = P.off[i * 3 + 1], oz = P.off[i * 3 + 2];
#pragma unroll
            for (int c = 0; c < 3; ++c) pc[c] = __fmaf_rn(Rp.m[c * 3 + 2], oz, __fmaf_rn(Rp.m[c * 3 + 1], oy, Rp.m[c * 3] * ox)) + pp[c];
            Rc = matmul3(Rp, LM[jj]);
          }
          const int ss = P.save_slot[i];
          if (ss >= 0) {
            float* s = sm_slot + ss * 12 * 32 + lane;
#pragma unroll
            for (int c = 0; c < 9; ++c) s[c * 32] = Rc.m[c];
#pragma unroll
            for (int c = 0; c < 3; ++c) s[(9 + c) * 32] = pc[c];
          }
          float* op = sm_pos + lane * kPosRow + jj * 3;
          op[0] = pc[0]; op[1] = pc[1]; op[2] = pc[2];
          float* orot = sm_rot + lane * kRotRow + jj * 9;
#pragma unroll
          for (int c = 0; c < 9; ++c) orot[c] = Rc.m[c];
          Rw[jj] = Rc;
        }
      }
      if (quat) {
#pragma unroll
        for (int jj = 0; jj < kChunk; ++jj)
          if (jj < nj) { if (kFast) mat_to_quat_xyzw_fast(Rw[jj], sm_quat + lane * kQuatRow + jj * 4); else mat_to_quat_xyzw(Rw[jj], sm_quat + lane * kQuatRow + jj * 4); }
      }
    }
    __syncwarp();
    if (full) {
      flush_lin<3 * kChunk, kPosRow>(sm_pos, pos + p0 * (3 * J) + 3 * j0, 3 * J, lane);
      if (rotmat) flush_lin<9 * kChunk, kRotRow>(sm_rot, rotmat + p0 * (9 * J) + 9 * j0, 9 * J, lane);
      if (quat) flush_lin<4 * kChunk, kQuatRow>(sm_quat, quat + p0 * (4 * J) + 4 * j0, 4 * J, lane);
    } else {
      flush_rows<3>(sm_pos, kPosRow, 0, 3, pos + p0 * (3 * J), 3 * J, 3 * j0, nj, npose, lane);
      if (rotmat) flush_rows<9>(sm_rot, kRotRow, 0, 9, rotmat + p0 * (9 * J), 9 * J, 9 * j0, nj, npose, lane);
      if (quat) flush_rows<4>(sm_quat, kQuatRow, 0, 4, quat + p0 * (4 * J), 4 * J, 4 * j0, nj, npose, lane);
    }
  }
}


// ---- the limb kernel: lane = (pose, limb) ------------------------------------------------------------------------------------------
// A warp owns 8 consecutive poses; the tree is cut into up to 4 joint ranges ("limbs") and the 4 lanes of a pose walk one
// each - first, without quaternion output, the ancestors of its range that lie outside it, then the range.  Every world
// transform a lane computes goes to the warp's staging area, which has the layout of the global outputs of the 8 poses
// ([pose][joint][..], contiguous and 16-byte aligned in global memory), so a parent that is not the previous step is read back from
// there (the same lane wrote it), ancestors computed by several limbs are written with identical values, and the tile
// leaves as three contiguous copies of 16 bytes per lane.  Per pose ~2.3x fewer warp-instructions than a thread per pose
// (whose 27-joint chain is serial and whose staged chunks leave float by float), and a batch of 4096 poses still fills the GPU.
#ifndef F_FK_WARPS
#define F_FK_WARPS 2
#endif
#ifndef F_FK_INBUFS
#define F_FK_INBUFS 1   // 1: one input buffer, refilled as soon as the tile's joints are computed (13 KB of staging per warp, 16 warps per SM);
                        // 2: the next tile's axis-angles arrive while this one is computed (15.8 KB, 14 warps per SM).
                        // B200, 65536 poses of the 23-DoF tree: 0.0489 ms vs 0.0522 ms (and 0.0567 ms with the quaternions staged too)
#endif
constexpr int kLimbs = 4, kLimbPoses = 8, kLimbWarps = F_FK_WARPS, kLimbInBufs = F_FK_INBUFS;
// floats of a warp's staging: input buffer(s) + world positions and rotations of the 8 poses in the layout of the global outputs
// (the quaternions leave straight from registers, 16 bytes per joint: nothing reads them back)
__host__ __device__ inline int fk_limb_in_floats(int J) { return (kLimbPoses * 3 * J + kLimbPoses * 3 + 3) & ~3; }
__host__ __device__ inline int fk_limb_warp_floats(int J) { return kLimbInBufs * fk_limb_in_floats(J) + kLimbPoses * 12 * J; }
struct FkLimbParams {
  int J, n_limbs, max_steps;
  int8_t parent[PBHC_MAX_JOINTS];
  int8_t step_joint[kLimbs][PBHC_MAX_JOINTS];  // joint of step s of limb k (its outside ancestors root first, then its range), -1 = no step
  int8_t n_silent[kLimbs];
  float off[PBHC_MAX_JOINTS * 3];
  float lrot[PBHC_MAX_JOINTS * 9];
};

template <bool kFast>
__global__ void __launch_bounds__(kLimbWarps * 32) fk_limb_kernel(const __grid_constant__ FkLimbParams P, const float* __restrict__ pose_aa,
                                                                 const float* __restrict__ trans, int64_t n, float* __restrict__ pos,
                                                                 float* __restrict__ rotmat, float* __restrict__ quat, int aligned) {
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int J = P.J;
  // block constants: [J][12] = local rotation (9) + offset (3), then parent / step tables as ints
  float* s_tree = smem;
  int* s_parent = reinterpret_cast<int*>(smem + 12 * J);
  int* s_step = s_parent + J;                       // [kLimbs][max_steps]
  const int const_floats = (12 * J + J + kLimbs * P.max_steps + 3) & ~3;
  for (int i = threadIdx.x; i < J; i += blockDim.x) {
    for (int c = 0; c < 9; ++c) s_tree[i * 12 + c] = P.lrot[i * 9 + c];
    for (int c = 0; c < 3; ++c) s_tree[i * 12 + 9 + c] = P.off[i * 3 + c];
    s_parent[i] = P.parent[i];
  }
  for (int i = threadIdx.x; i < kLimbs * P.max_steps; i += blockDim.x) s_step[i] = P.step_joint[i / P.max_steps][i % P.max_steps];
  __syncthreads();
  // per warp: input buffer(s) [8][3J] + [8][3], staging pos [8][3J] | rot [8][9J]
  const int in_floats = fk_limb_in_floats(J);
  float* w_base = smem + const_floats + warp * fk_limb_warp_floats(J);
  float* st_pos = w_base + kLimbInBufs * in_floats, *st_rot = st_pos + kLimbPoses * 3 * J;
  const int p = lane >> 2, k = lane & 3;
  const int n_silent = P.n_silent[k];
  const int64_t n_tiles = (n + kLimbPoses - 1) / kLimbPoses, tstep = (int64_t)gridDim.x * kLimbWarps;

  auto request = [&](int64_t t, int buf) {  // the tile's axis-angles and root translations -> input buffer `buf`
    float* dst = w_base + buf * in_floats;
    const int64_t p0 = t * kLimbPoses;
    const int np = (int)min((int64_t)kLimbPoses, n - p0);
    const float* g = pose_aa + p0 * (3 * J);
    const float* gt = trans + p0 * 3;
    if (aligned && np == kLimbPoses) {
      const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
      for (int i = lane; i < (kLimbPoses * 3 * J) / 4; i += 32) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + 16 * i), "l"(g + 4 * i) : "memory");
      if (lane < kLimbPoses * 3 / 4) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + 4 * (kLimbPoses * 3 * J) + 16 * lane), "l"(gt + 4 * lane) : "memory");
    } else {
      const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
      for (int i = lane; i < np * 3 * J; i += 32) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 4 * i), "l"(g + i) : "memory");
      if (lane < np * 3) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 4 * (kLimbPoses * 3 * J + lane)), "l"(gt + lane) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  int64_t t = (int64_t)blockIdx.x * kLimbWarps + warp;
  if (t < n_tiles) request(t, 0);
  for (int buf = 0; t < n_tiles; t += tstep, buf ^= (kLimbInBufs - 1)) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    if (kLimbInBufs == 2 && t + tstep < n_tiles) request(t + tstep, buf ^ 1);
    const int64_t p0 = t * kLimbPoses;
    const int np = (int)min((int64_t)kLimbPoses, n - p0);
    const float* in = w_base + buf * in_floats;
    if (p < np && k < P.n_limbs) {
      const float* a_row = in + p * (3 * J);
      M3 Rc;
      float pc[3] = {0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 9; ++c) Rc.m[c] = 0.f;
      int prev = -2;
      for (int s = 0; s < P.max_steps; ++s) {
        const int i = s_step[k * P.max_steps + s];
        if (i < 0) break;
        const float* a = a_row + 3 * i;
        const M3 Mj = kFast ? aa_to_mat_fast(a[0], a[1], a[2]) : aa_to_mat(a[0], a[1], a[2]);
        const int par = s_parent[i];
        if (par < 0) {
          Rc = Mj;
          const float* tr = in + kLimbPoses * 3 * J + p * 3;
          pc[0] = tr[0]; pc[1] = tr[1]; pc[2] = tr[2];
        } else {
          const float* tc = s_tree + i * 12;
          M3 L;
#pragma unroll
          for (int c = 0; c < 9; ++c) L.m[c] = tc[c];
          const M3 LMj = matmul3(L, Mj);
          if (par != prev) {  // a branch point further up: this lane staged its transform earlier
            const float* sr = st_rot + (p * J + par) * 9;
            const float* sp = st_pos + (p * J + par) * 3;
#pragma unroll
            for (int c = 0; c < 9; ++c) Rc.m[c] = sr[c];
            pc[0] = sp[0]; pc[1] = sp[1]; pc[2] = sp[2];
          }
          const float ox = tc[9], oy = tc[10], oz = tc[11];
          float np_[3];
#pragma unroll
          for (int c = 0; c < 3; ++c) np_[c] = __fmaf_rn(Rc.m[c * 3 + 2], oz, __fmaf_rn(Rc.m[c * 3 + 1], oy, Rc.m[c * 3] * ox)) + pc[c];
          pc[0] = np_[0]; pc[1] = np_[1]; pc[2] = np_[2];
          Rc = matmul3(Rc, LMj);
        }
        prev = i;
        float* op = st_pos + (p * J + i) * 3;
        op[0] = pc[0]; op[1] = pc[1]; op[2] = pc[2];
        float* orot = st_rot + (p * J + i) * 9;
#pragma unroll
        for (int c = 0; c < 9; ++c) orot[c] = Rc.m[c];
        if (quat && s >= n_silent) {
          float q4[4];
          if (kFast) mat_to_quat_xyzw_fast(Rc, q4); else mat_to_quat_xyzw(Rc, q4);
          float* gq = quat + ((p0 + p) * J + i) * 4;
          if (aligned) *reinterpret_cast<float4*>(gq) = make_float4(q4[0], q4[1], q4[2], q4[3]);
          else { gq[0] = q4[0]; gq[1] = q4[1]; gq[2] = q4[2]; gq[3] = q4[3]; }
        }
      }
    }
    __syncwarp();
    if (kLimbInBufs == 1 && t + tstep < n_tiles) request(t + tstep, 0);  // (the input buffer is free again)
    // the tile leaves: three contiguous runs of np poses
    auto flush = [&](const float* st, float* g, int row) {
      float* dst = g + p0 * row;
      const int nf = np * row;
      if (aligned && np == kLimbPoses) { for (int i = lane; i < nf / 4; i += 32) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(st)[i]; }
      else { for (int i = lane; i < nf; i += 32) dst[i] = st[i]; }
    };
    flush(st_pos, pos, 3 * J);
    if (rotmat) flush(st_rot, rotmat, 9 * J);
    __syncwarp();
  }
}

}  // namespace pbhc

static int fk_launch(const PbhcTree* tree, const float* pose_aa, const float* trans, int64_t n, float* pos,
                     float* rotmat, float* quat_xyzw, bool fast, void* stream) {
  using namespace pbhc;
  if (!tree || !tree->parents || !tree->offsets || !tree->local_rot_mat || !pose_aa || !trans || !pos || n < 0) {
    set_error("pbhc_fk: null argument"); return PBHC_E_BADARG;
  }
  const int J = tree->num_joints;
  if (J <= 0 || J > PBHC_MAX_JOINTS) { set_error("pbhc_fk: num_joints %d outside 1..%d", J, PBHC_MAX_JOINTS); return PBHC_E_RANGE; }
  FkParams P;
  P.J = J;
  P.n_slots = 0;
  for (int i = 0; i < J; ++i) {
    const int p = tree->parents[i];
    if (p >= i || p < -1 || (i == 0 && p != -1)) { set_error("pbhc_fk: parents[%d]=%d must be < %d (DFS order)", i, p, i); return PBHC_E_BADARG; }
    P.parent[i] = (int8_t)p; P.save_slot[i] = -1; P.load_slot[i] = -1;
    for (int c = 0; c < 3; ++c) P.off[i * 3 + c] = tree->offsets[i * 3 + c];
    for (int c = 0; c < 9; ++c) P.lrot[i * 9 + c] = tree->local_rot_mat[i * 9 + c];
  }
  for (int i = 1; i < J; ++i) {
    const int p = P.parent[i];
    if (p < 0 || p == i - 1) continue;
    if (P.save_slot[p] < 0) {
      if (P.n_slots == kMaxSlots) { set_error("pbhc_fk: more than %d branch points", kMaxSlots); return PBHC_E_RANGE; }
      P.save_slot[p] = (int8_t)P.n_slots++;
    }
    P.load_slot[i] = P.save_slot[p];
  }
  if (n == 0) return PBHC_OK;
  {  // the limb kernel when the tree cuts into 2..4 ranges (any pre-order numbered tree with a few branches does)
    static thread_local FkLimbParams LP;
    static thread_local int lp_valid = 0;           // 0: not built, 1: usable, -1: this tree does not split
    static thread_local int8_t lp_parents[PBHC_MAX_JOINTS];
    static thread_local int lp_J = 0;
    bool same = lp_valid != 0 && lp_J == J;
    for (int i = 0; same && i < J; ++i) same = lp_parents[i] == P.parent[i];
    if (!same) {
      lp_J = J;
      for (int i = 0; i < J; ++i) lp_parents[i] = P.parent[i];
      int depth[PBHC_MAX_JOINTS];
      for (int i = 0; i < J; ++i) depth[i] = P.parent[i] < 0 ? 0 : depth[P.parent[i]] + 1;
      // pre-order walk (children in index order): in this order every parent of a contiguous range that lies outside the
      // range is an ancestor of the range's first joint.  (The joint numbering itself need not be pre-order: the extended
      // bodies of torch_humanoid_batch.py:80-102 are appended at the end, whatever their parent.)
      int ord[PBHC_MAX_JOINTS], n_ord = 0;
      {
        int stack[PBHC_MAX_JOINTS], sp = 0;
        stack[sp++] = 0;
        while (sp > 0) {
          const int v = stack[--sp];
          ord[n_ord++] = v;
          for (int c2 = J - 1; c2 > v; --c2) if (P.parent[c2] == v) stack[sp++] = c2;  // (pushed in reverse: popped in index order)
        }
      }
      auto walkable = [&](int b, int e) {  // positions [b, e) of ord
        bool anc[PBHC_MAX_JOINTS] = {false}, in[PBHC_MAX_JOINTS] = {false};
        for (int a = P.parent[ord[b]]; a >= 0; a = P.parent[a]) anc[a] = true;
        for (int i = b; i < e; ++i) {
          const int q = P.parent[ord[i]];
          if (!((q >= 0 && (in[q] || anc[q])) || (q < 0 && i == 0))) return false;
          in[ord[i]] = true;
        }
        return true;
      };
      int best_cost = 1 << 30, best_cut[kLimbs + 1] = {0, J, J, J, J}, best_n = 1;
      // cuts 0 = c0 < c1 <= c2 <= c3 <= c4 = J (equal cuts = fewer limbs); cost = the longest walk (ancestors + range)
      for (int c1 = 1; c1 <= J && n_ord == J; ++c1)
        for (int c2 = c1; c2 <= J; ++c2)
          for (int c3 = c2; c3 <= J; ++c3) {
            const int cut[kLimbs + 1] = {0, c1, c2, c3, J};
            int cost = 0, nl = 0;
            bool ok = true;
            for (int k = 0; k < kLimbs && ok; ++k) {
              const int bb = cut[k], ee = cut[k + 1];
              if (ee == bb) continue;
              if (nl != k) { ok = false; break; }  // empty limbs only at the end
              ++nl;
              const int w = depth[ord[bb]] + (ee - bb);
              if (w > cost) cost = w;
              if (w >= best_cost || !walkable(bb, ee)) ok = false;
            }
            if (ok && cost < best_cost) { best_cost = cost; best_n = nl; for (int k = 0; k <= kLimbs; ++k) best_cut[k] = cut[k]; }
          }
      if (best_n < 2) lp_valid = -1;
      else {
        LP.J = J; LP.n_limbs = best_n; LP.max_steps = best_cost;
        for (int i = 0; i < J; ++i) LP.parent[i] = P.parent[i];
        for (int i = 0; i < 3 * J; ++i) LP.off[i] = 0.f;
        for (int k = 0; k < kLimbs; ++k) {
          for (int q = 0; q < PBHC_MAX_JOINTS; ++q) LP.step_joint[k][q] = -1;
          LP.n_silent[k] = 0;
          if (k >= best_n) continue;
          const int bb = best_cut[k], ee = best_cut[k + 1];
          int chain[PBHC_MAX_JOINTS], d = 0;
          for (int a = P.parent[ord[bb]]; a >= 0; a = P.parent[a]) chain[d++] = a;
          LP.n_silent[k] = (int8_t)d;
          int q = 0;
          for (int i = d - 1; i >= 0; --i) LP.step_joint[k][q++] = (int8_t)chain[i];
          for (int i = bb; i < ee; ++i) LP.step_joint[k][q++] = (int8_t)ord[i];
        }
        lp_valid = 1;
        if (const char* v = getenv("PBHC_FK_VERBOSE")) if (v[0] == '1')
          fprintf(stderr, "[pbhc fk] %d joints -> %d limbs, cuts %d %d %d, longest walk %d steps\n", J, best_n, best_cut[1], best_cut[2], best_cut[3], best_cost);
      }
    }
    if (lp_valid == 1) {
      for (int i = 0; i < 3 * J; ++i) LP.off[i] = P.off[i];     // (the constants may change between calls with the same topology)
      for (int i = 0; i < 9 * J; ++i) LP.lrot[i] = P.lrot[i];
      const int const_floats = (12 * J + J + kLimbs * LP.max_steps + 3) & ~3;
      const size_t shmem = (size_t)(const_floats + kLimbWarps * fk_limb_warp_floats(J)) * sizeof(float);
      auto kern = fast ? fk_limb_kernel<true> : fk_limb_kernel<false>;
      static thread_local size_t configured[2] = {0, 0};
      static thread_local int per_sm[2] = {0, 0};
      if (shmem > configured[fast]) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_fk(attr)");
        (void)cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured[fast] = shmem;
        int m = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&m, kern, kLimbWarps * 32, shmem);
        per_sm[fast] = m < 1 ? 1 : m;
      }
      const int64_t n_tiles = (n + kLimbPoses - 1) / kLimbPoses;
      int64_t blocks = (n_tiles + kLimbWarps - 1) / kLimbWarps;
      const int64_t cap = 148LL * per_sm[fast];
      if (blocks > cap) blocks = cap;
      const int aligned = !(((uintptr_t)pose_aa | (uintptr_t)trans | (uintptr_t)pos | (uintptr_t)rotmat | (uintptr_t)quat_xyzw) & 15);
      kern<<<(unsigned)blocks, kLimbWarps * 32, shmem, (cudaStream_t)stream>>>(LP, pose_aa, trans, n, pos, rotmat, quat_xyzw, aligned);
      return check_launch("pbhc_fk");
    }
  }
  // (trees that do not split - a single chain - take the thread-per-pose kernel below)
  // two parts when the batch alone cannot fill the machine (<= 64 warps per SM) and the tree splits: the cut s is the joint
  // nearest J/2 such that every joint >= s has its parent either >= s or among the ancestors of s (a chain of at most 8)
  P.n_parts = 1; P.part_begin[0] = 0; P.part_end[0] = J; P.n_silent[0] = 0; P.n_silent[1] = 0;
  if (J >= 8 && n <= 148LL * 32 * 32) {
    int best = -1;
    for (int s = 2; s < J - 1; ++s) {
      bool anc[PBHC_MAX_JOINTS] = {false};
      int depth = 0;
      for (int a = P.parent[s]; a >= 0; a = P.parent[a]) { anc[a] = true; ++depth; }
      bool ok = depth <= 8;
      for (int i = s; i < J && ok; ++i) ok = P.parent[i] >= s || (P.parent[i] >= 0 && anc[P.parent[i]]);
      if (ok && (best < 0 || abs(2 * s - J) < abs(2 * best - J))) best = s;
    }
    if (best > 0) {
      P.n_parts = 2; P.part_end[0] = best; P.part_begin[1] = best; P.part_end[1] = J;
      int chain[8], d = 0;
      for (int a = P.parent[best]; a >= 0; a = P.parent[a]) chain[d++] = a;
      P.n_silent[1] = d;
      for (int i = 0; i < d; ++i) P.silent[1][i] = (int8_t)chain[d - 1 - i];  // root first
    }
  }
  const size_t shmem = (size_t)kFkWarps * (2 * 32 * kInRow + 32 * kStageRow + P.n_slots * 12 * 32) * sizeof(float);
  auto kern = fast ? fk_kernel<true> : fk_kernel<false>;
  static thread_local size_t configured[2] = {0, 0};
  static thread_local bool carved[2] = {false, false};
  if (!carved[fast]) {  // shared memory bounds the resident warps: ask for the largest carve-out
    (void)cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carved[fast] = true;
  }
  if (shmem > 48 * 1024 && shmem > configured[fast]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_fk(attr)");
    configured[fast] = shmem;
  }
  // small batches: with 32 poses per warp fewer warps than SMs would exist and the kernel is pure latency; narrower
  // tiles (idle lanes) keep every SM busy and shorten the per-pose staging loops
  int tile = 32;
  while (tile > 8 && (n + tile - 1) / tile < 2 * 148) tile >>= 1;
  const int64_t per_block = (int64_t)kFkWarps * tile;
  const int64_t blocks = (n + per_block - 1) / per_block * P.n_parts;
  if (blocks > 0x7fffffffLL) { set_error("pbhc_fk: too many poses"); return PBHC_E_RANGE; }
  kern<<<(unsigned)blocks, kFkWarps * 32, shmem, (cudaStream_t)stream>>>(P, pose_aa, trans, n, pos, rotmat, quat_xyzw, tile);
  return check_launch("pbhc_fk");
}

extern "C" int pbhc_fk(const PbhcTree* tree, const float* pose_aa, const float* trans, int64_t n, float* pos,
                       float* rotmat, float* quat_xyzw, void* stream) {
  return fk_launch(tree, pose_aa, trans, n, pos, rotmat, quat_xyzw, true, stream);
}
extern "C" int pbhc_fk_exact(const PbhcTree* tree, const float* pose_aa, const float* trans, int64_t n, float* pos,
                             float* rotmat, float* quat_xyzw, void* stream) {
  return fk_launch(tree, pose_aa, trans, n, pos, rotmat, quat_xyzw, false, stream);
}
