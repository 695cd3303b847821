#!/bin/bash
# Resident numbers of libawfl_b200.so and of every variant library under pam_b200/variants/ (tools/build_variant.sh).
#   tools/variant_sweep.sh ["--workload C4 --ord 3"]      (default: C2)
cd "$(dirname "$0")/.."
for lib in pam_b200/libawfl_b200.so pam_b200/variants/libawfl_*.so; do
  [ -f "$lib" ] || continue
  AWFL_B200_LIB=$PWD/$lib python bench.py $1 --steps 10 --no-e2e --no-cpu-baseline --no-gpu-reference --no-strong-c3 2>/dev/null > /tmp/v.json
  echo "$(basename $lib): $(python tools/bench_summary.py /tmp/v.json | cut -d: -f2-)"
done
