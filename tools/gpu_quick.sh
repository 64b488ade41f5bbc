#!/bin/bash
# quick GPU check of the fused step: parity tests of the tracking kernel, then the bench line (kernel ms / roofline frac)
# usage (under gpurun): bash tools/gpu_quick.sh [extra bench args]
timeout 900 python -m pytest tests/test_gpu_track.py -x -q -m gpu 2>&1 | tail -4
for j in 1 0; do
  PBHC_JIT=$j timeout 300 python bench.py --steps 50 --warmup 5 --no-e2e --no-cpu-baseline "$@" 2>gpurun_out/quick.err | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('JIT=$j', d['config'].get('track_launch'), 'step_ms %.4f track_ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac']), d['stages_ms'])"
done
tail -2 gpurun_out/quick.err
