// table.cuh - frame index / blend arithmetic shared by the stand-alone lookup (K1) and the fused step (K3).
#pragma once
#include "common.cuh"

namespace pbhc {

struct FrameBlend { int64_t f0, f1; float blend; };

// motion_lib_base.py:503-513 with torch's promotions: (num_frames-1) and idx0 are int64 -> fp32
__device__ __forceinline__ FrameBlend frame_blend(const PbhcTableDesc& tab, int clip, float time) {
  int nf = tab.clip_num_frames[clip];
  float len = tab.clip_length[clip];
  float dt = tab.clip_dt[clip];
  float phase = clip01(time / len);
  if (time < 0.0f) time = 0.0f;
  int64_t idx0 = (int64_t)(phase * (float)(nf - 1));
  int64_t idx1 = min(idx0 + 1, (int64_t)(nf - 1));
  float blend = clip01((time - (float)idx0 * dt) / dt);
  int64_t start = tab.clip_frame_start[clip];
  return {idx0 + start, idx1 + start, blend};
}

__device__ __forceinline__ int resolve_clip(const PbhcTableDesc& tab, const int32_t* clip_of_motion, int64_t mid) {
  int64_t c = clip_of_motion ? (int64_t)clip_of_motion[mid] : mid;
  if (c < 0) c = 0;
  if (c >= tab.num_clips) c = tab.num_clips - 1;
  return (int)c;
}

}  // namespace pbhc
