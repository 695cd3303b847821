#!/bin/bash
# ncu evidence of one round: launch list of a short default bench run + --set full capture of a stage-2 pair of state marches
tag=${1:-r02}
cd "$(dirname "$0")/.."
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-gpu-reference --no-strong-c3"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 30 -c 400 --csv --log-file gpurun_out/${tag}_launches_c2.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_march_state -s 14 -c 2 -f -o gpurun_out/${tag}_state_prof $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
ls -la gpurun_out/${tag}_*
