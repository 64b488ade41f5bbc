// common.cuh - shared device helpers.  Compiled with -fmad=false: every a*b+c below is two
// roundings exactly like the reference's unfused torch ops; explicit __fmaf_rn marks the places
// where ATen itself fuses (vector norm).
#pragma once
#ifdef __CUDACC_RTC__  // NVRTC (csrc/jit.cu): no host headers
typedef signed char int8_t; typedef unsigned char uint8_t; typedef int int32_t; typedef unsigned int uint32_t;
typedef long long int64_t; typedef unsigned long long uint64_t; typedef unsigned long long uintptr_t;
#else
#include <cuda_runtime.h>
#include <stdint.h>
#endif
#include "../../include/pbhc_b200.h"

namespace pbhc {

#ifndef __CUDACC_RTC__
void set_error(const char* fmt, ...);
int check_launch(const char* what);
#endif

struct Q4 { float x, y, z, w; };  // xyzw
struct V3 { float x, y, z; };

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ float clip01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }

// torch.linalg.vector_norm over 3 contiguous floats on CPU = sqrt(fma(z,z,fma(y,y,x*x)))
// (probed against torch 2.11; see DESIGN.md "arithmetic parity").
__device__ __forceinline__ float norm3(float x, float y, float z) {
  return sqrtf(__fmaf_rn(z, z, __fmaf_rn(y, y, x * x)));
}
// ... but over 4 contiguous floats ATen's vectorised path gives the plain sequential sum (probed).
__device__ __forceinline__ float norm4(float x, float y, float z, float w) { return sqrtf(((x * x + y * y) + z * z) + w * w); }
__device__ __forceinline__ float norm2(float x, float y) { return sqrtf(__fmaf_rn(y, y, x * x)); }

// Philox4x32-10, counter-based RNG of the device-side domain randomisation / noise (keyed by seed; counter = ids + stream)
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u; key.y += 0xBB67AE85u;
  }
  return ctr;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x & 0xFFFFFFu) * (1.0f / 16777216.0f); }  // [0, 1), 24 bits

// slerp, isaac_utils/rotations.py:210-232 (not renormalised)
__device__ __forceinline__ Q4 slerp(Q4 a, Q4 b, float t) {
  float c = ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w;
  if (c < 0.0f) { b.x = -b.x; b.y = -b.y; b.z = -b.z; b.w = -b.w; }
  c = fabsf(c);
  float h = acosf(c);
  float s = sqrtf(1.0f - c * c);
  float ra = sinf((1.0f - t) * h) / s;
  float rb = sinf(t * h) / s;
  Q4 q = {ra * a.x + rb * b.x, ra * a.y + rb * b.y, ra * a.z + rb * b.z, ra * a.w + rb * b.w};
  if (fabsf(s) < 0.001f) q = {0.5f * a.x + 0.5f * b.x, 0.5f * a.y + 0.5f * b.y, 0.5f * a.z + 0.5f * b.z, 0.5f * a.w + 0.5f * b.w};
  if (fabsf(c) >= 1.0f) q = a;
  return q;
}

// quat_mul (w_last), rotations.py:414-441 - the 8-multiplication form, association order kept
__device__ __forceinline__ Q4 quat_mul(Q4 a, Q4 b) {
  float ww = (a.z + a.x) * (b.x + b.y);
  float yy = (a.w - a.y) * (b.w + b.z);
  float zz = (a.w + a.y) * (b.w - b.z);
  float xx = ww + yy + zz;
  float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
  Q4 r;
  r.w = qq - ww + (a.z - a.y) * (b.y - b.z);
  r.x = qq - xx + (a.x + a.w) * (b.x + b.w);
  r.y = qq - yy + (a.w - a.x) * (b.y + b.z);
  r.z = qq - zz + (a.z + a.y) * (b.w - b.x);
  return r;
}
__device__ __forceinline__ Q4 quat_conj(Q4 a) { return {-a.x, -a.y, -a.z, a.w}; }

// my_quat_rotate / quat_rotate (rotations.py:83-99, 244-253): a + b + c ; inverse: a - b + c
__device__ __forceinline__ V3 quat_rotate_impl(Q4 q, V3 v, bool inverse) {
  float s = 2.0f * (q.w * q.w) - 1.0f;
  V3 a = {v.x * s, v.y * s, v.z * s};
  V3 cr = {q.y * v.z - q.z * v.y, q.z * v.x - q.x * v.z, q.x * v.y - q.y * v.x};
  V3 b = {cr.x * q.w * 2.0f, cr.y * q.w * 2.0f, cr.z * q.w * 2.0f};
  float d = (q.x * v.x + q.y * v.y) + q.z * v.z;  // bmm [1,3]x[3,1]
  V3 c = {q.x * d * 2.0f, q.y * d * 2.0f, q.z * d * 2.0f};
  if (inverse) return {a.x - b.x + c.x, a.y - b.y + c.y, a.z - b.z + c.z};
  return {a.x + b.x + c.x, a.y + b.y + c.y, a.z + b.z + c.z};
}
__device__ __forceinline__ V3 quat_rotate(Q4 q, V3 v) { return quat_rotate_impl(q, v, false); }
__device__ __forceinline__ V3 quat_rotate_inverse(Q4 q, V3 v) { return quat_rotate_impl(q, v, true); }

// ---- helpers of LeggedRobotGeneralTracking (utils/torch_utils.py, isaac_utils/rotations.py) ----
// quat_apply (torch_utils.py:51-58): b + w*t + xyz x t with t = 2 (xyz x b)
__device__ __forceinline__ V3 quat_apply(Q4 a, V3 b) {
  const V3 t = {(a.y * b.z - a.z * b.y) * 2.0f, (a.z * b.x - a.x * b.z) * 2.0f, (a.x * b.y - a.y * b.x) * 2.0f};
  const V3 u = {a.y * t.z - a.z * t.y, a.z * t.x - a.x * t.z, a.x * t.y - a.y * t.x};
  return {(b.x + a.w * t.x) + u.x, (b.y + a.w * t.y) + u.y, (b.z + a.w * t.z) + u.z};
}
// yaw_quat (torch_utils.py:239-270, w_last): (0, 0, sin(yaw/2), cos(yaw/2)) renormalised like maths.normalize
__device__ __forceinline__ Q4 yaw_quat(Q4 q) {
  const float yaw = atan2f(2.0f * (q.w * q.z + q.x * q.y), 1.0f - 2.0f * (q.y * q.y + q.z * q.z));
  const float s = sinf(yaw / 2.0f), c = cosf(yaw / 2.0f);
  const float n = fmaxf(norm4(0.0f, 0.0f, s, c), 1e-9f);
  return {0.0f / n, 0.0f / n, s / n, c / n};
}
// sqrtf / norms whose argument is very often an exact zero (contact forces of bodies in the air, velocities of standing feet):
// sqrtf takes an out-of-line slow path for zero and denormal arguments, and one such lane sends the whole warp through it.
// The zero lanes compute sqrtf(1) instead and the result is put back: identical values, no slow path.
__device__ __forceinline__ float sqrt_z(float x) { const bool z = x == 0.0f; const float r = sqrtf(z ? 1.0f : x); return z ? 0.0f : r; }
__device__ __forceinline__ float norm3_z(float x, float y, float z) { return sqrt_z(__fmaf_rn(z, z, __fmaf_rn(y, y, x * x))); }
__device__ __forceinline__ float norm2_z(float x, float y) { return sqrt_z(__fmaf_rn(y, y, x * x)); }
// first two columns of matrix_from_quat (torch_utils.py:274-297, w_last), row-major: m00 m01 m10 m11 m20 m21
__device__ __forceinline__ void mat_cols2(Q4 q, float (&o)[6]) {
  const float i = q.x, j = q.y, k = q.z, r = q.w;
  const float two_s = 2.0f / (((i * i + j * j) + k * k) + r * r);
  o[0] = 1.0f - two_s * (j * j + k * k); o[1] = two_s * (i * j - k * r);
  o[2] = two_s * (i * j + k * r);        o[3] = 1.0f - two_s * (i * i + k * k);
  o[4] = two_s * (i * k - j * r);        o[5] = two_s * (j * k + i * r);
}
// angle of quat_to_angle_axis (rotations.py:185-207): normalize_angle(2 acos(w)), 0 when sqrt(1 - w^2) <= 1e-5 (or NaN).
// normalize_angle is atan2(sin a, cos a) there; for a in [0, 2 pi] that is a, or a - 2 pi above pi, which is what is
// computed here (agrees to ~1e-7 relative; every caller squares the angle).
__device__ __forceinline__ float quat_angle(Q4 q) {
  const float st = sqrtf(1.0f - q.w * q.w);
  const float a = 2.0f * acosf(q.w);
  const float an = a > 3.14159265358979f ? a - 6.28318530717959f : a;
  return fabsf(st) > 1e-5f ? an : 0.0f;
}
// roll and pitch of get_euler_xyz_in_tensor (rotations.py:368-387; copysign of maths.py:16-19)
__device__ __forceinline__ void euler_roll_pitch(Q4 q, float& roll, float& pitch) {
  const float sinr = 2.0f * (q.w * q.x + q.y * q.z);
  const float cosr = ((q.w * q.w - q.x * q.x) - q.y * q.y) + q.z * q.z;
  roll = atan2f(sinr, cosr);
  const float sinp = 2.0f * (q.w * q.y - q.z * q.x);
  const float sgn = sinp > 0.0f ? 1.0f : (sinp < 0.0f ? -1.0f : 0.0f);
  pitch = fabsf(sinp) >= 1.0f ? fabsf(1.5707963267948966f) * sgn : asinf(sinp);
}

// yaw of get_euler_xyz_in_tensor (rotations.py:381-385)
__device__ __forceinline__ float euler_yaw(Q4 q) {
  const float siny = 2.0f * (q.w * q.z + q.x * q.y);
  const float cosy = ((q.w * q.w + q.x * q.x) - q.y * q.y) - q.z * q.z;
  return atan2f(siny, cosy);
}
// wrap_to_pi (rotations.py:49-53) as the reference RUNS it: the function is @torch.jit.script and TorchScript compiles
// `angles %= 2*np.pi` to aten::fmod_ (C fmod, sign of the dividend) - negative angles stay negative - then subtracts
// fp32(2 pi) where the value exceeds fp32(pi)
__device__ __forceinline__ float wrap_to_pi(float a) {
  const float two_pi = 6.283185307179586f;
  const float r = fmodf(a, two_pi);
  return r - (r > 3.141592653589793f ? two_pi : 0.0f);
}

}  // namespace pbhc
