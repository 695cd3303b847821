#!/bin/bash
# A/B of the first tracer carried by the 2-D state marches (AWFL_FUSED_TRACER: bit 0 = x march, bit 1 = z march) against the
# tracer march for every tracer (0).   usage: tools/ab_fused_tracer.sh ["C4 --ord 7" ...]
cd "$(dirname "$0")/.."
F="--no-cpu-baseline --no-e2e --no-strong-c3 --no-gpu-reference --steps 20 --warmup 5"
if [ $# -gt 0 ]; then W=("$@"); else W=("C2" "C4 --ord 3" "C4 --ord 9" "C5"); fi
for w in "${W[@]}"; do
  for g in 3 1 0; do
    AWFL_FUSED_TRACER=$g python bench.py --workload $w $F 2> gpurun_out/ab_err.log > /tmp/ab.json
    echo "fused_tracer=$g $w: $(python tools/bench_summary.py /tmp/ab.json | cut -d: -f2-)"
  done
done
