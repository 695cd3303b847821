#!/bin/bash
# A/B of the 2-D state marches without the v reconstruction against the general ones (AWFL_GENERAL_2D=1), resident lines only
cd "$(dirname "$0")/.."
F="--no-cpu-baseline --no-e2e --no-strong-c3 --no-gpu-reference --steps 20 --warmup 5"
for g in 0 1; do
  for w in "C2" "C4 --ord 3" "C4" "C4 --ord 9" "C5"; do
    AWFL_GENERAL_2D=$g python bench.py --workload $w $F 2> gpurun_out/ab_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('general_2d=$g', '$w', '%.1f M/s %.3f ms'%(d['value']/1e6, d['ms_per_step']))
"
  done
done
