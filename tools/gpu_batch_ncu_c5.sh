#!/bin/bash
# ncu --set full of the C5 (10 tracers) tracer kernels: fused FCT + update, tracer x / z marches
tag=${1:-r02}
cd "$(dirname "$0")/.."
CMD="python bench.py --workload C5 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-gpu-reference --no-strong-c3"
$CMD > gpurun_out/${tag}_c5_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_fct_update_2d|k_march_tracer" -s 9 -c 3 -f -o gpurun_out/${tag}_c5_tracer_prof $CMD > gpurun_out/${tag}_c5_ncu.log 2>&1
ls -la gpurun_out/${tag}_c5_*
