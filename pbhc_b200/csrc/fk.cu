// fk.cu - K2: G1 forward kinematics.  A warp owns a batch of 32 poses (thread per pose walks the
// kinematic tree in DFS order with the parent transform in registers; branch-point transforms
// are parked in per-warp shared memory).  Global I/O is staged through shared memory in chunks of
// kChunk joints so every global access is a run of contiguous floats per pose.
// Reference: utils/motion_lib/torch_humanoid_batch.py:168-269; isaac_utils/rotations.py:519-647.
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
  const int64_t p0 = ((int64_t)blockIdx.x * kFkWarps + warp) * tile;  // tile = poses per warp (32, or fewer for small batches)
  if (p0 >= n) return;
  const int npose = (int)min((int64_t)tile, n - p0);
  const bool live = lane < npose;
  const float* g_in = pose_aa + p0 * (3 * J);
  M3 Rc;  // world rotation of the previous joint
  float pc[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 9; ++i) Rc.m[i] = 0.f;

  const bool tile32 = npose == 32;
  if (tile32 && J >= kChunk) fill_full_async<3 * kChunk>(sm_in0, g_in, 3 * J, 0, lane);
  int buf = 0;
  for (int j0 = 0; j0 < J; j0 += kChunk, buf ^= 1) {
    const int nj = min(kChunk, J - j0);
    float* sm_in = sm_in0 + buf * 32 * kInRow;
    __syncwarp();
    const bool full = tile32 && nj == kChunk;
    if (full) {
      asm volatile("cp.async.wait_all;" ::: "memory");
      if (j0 + 2 * kChunk <= J) fill_full_async<3 * kChunk>(sm_in0 + (buf ^ 1) * 32 * kInRow, g_in, 3 * J, 3 * (j0 + kChunk), lane);
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
          const M3 Mj = aa_to_mat(a[0], a[1], a[2]);
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
          if (jj < nj) mat_to_quat_xyzw(Rw[jj], sm_quat + lane * kQuatRow + jj * 4);
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

}  // namespace pbhc

extern "C" int pbhc_fk(const PbhcTree* tree, const float* pose_aa, const float* trans, int64_t n, float* pos,
                       float* rotmat, float* quat_xyzw, void* stream) {
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
  const size_t shmem = (size_t)kFkWarps * (2 * 32 * kInRow + 32 * kStageRow + P.n_slots * 12 * 32) * sizeof(float);
  static thread_local size_t configured = 0;
  static thread_local bool carved = false;
  if (!carved) {  // shared memory bounds the resident warps: ask for the largest carve-out
    (void)cudaFuncSetAttribute(fk_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carved = true;
  }
  if (shmem > 48 * 1024 && shmem > configured) {
    if (cudaFuncSetAttribute(fk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return check_launch("pbhc_fk(attr)");
    configured = shmem;
  }
  // small batches: with 32 poses per warp fewer warps than SMs would exist and the kernel is pure latency; narrower
  // tiles (idle lanes) keep every SM busy and shorten the per-pose staging loops
  int tile = 32;
  while (tile > 8 && (n + tile - 1) / tile < 2 * 148) tile >>= 1;
  const int64_t per_block = (int64_t)kFkWarps * tile;
  const int64_t blocks = (n + per_block - 1) / per_block;
  if (blocks > 0x7fffffffLL) { set_error("pbhc_fk: too many poses"); return PBHC_E_RANGE; }
  fk_kernel<<<(unsigned)blocks, kFkWarps * 32, shmem, (cudaStream_t)stream>>>(P, pose_aa, trans, n, pos, rotmat, quat_xyzw, tile);
  return check_launch("pbhc_fk");
}
