#!/bin/bash
# A/B of the tracer march with two tracers per warp (default) against two lines of one tracer (AWFL_TRACER_PAIRS=0)
cd "$(dirname "$0")/.."
F="--no-cpu-baseline --no-e2e --no-strong-c3 --no-gpu-reference --steps 20 --warmup 5"
for p in 1 0; do
  AWFL_TRACER_PAIRS=$p python bench.py --workload C5 $F 2> gpurun_out/ab_err.log > /tmp/ab.json
  echo "tracer_pairs=$p C5: $(python tools/bench_summary.py /tmp/ab.json | cut -d: -f2-)"
done
