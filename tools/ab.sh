#!/bin/bash
# development aid: time the track kernel with each pbhc_b200/variant_*.so (built by hand with -D feature flags)
for v in pbhc_b200/variant_*.so; do
  PBHC_B200_LIB=$PWD/$v python bench.py --no-cpu-baseline --no-e2e --no-fk --steps 100 > gpurun_out/ab_$(basename $v .so).log 2>&1
  python -c "
import json,sys
d=json.loads(open('gpurun_out/ab_$(basename $v .so).log').read().strip().splitlines()[-1]); print('$v', round(d['roofline']['kernel_ms'],4), round(d['roofline']['frac'],4))"
done
