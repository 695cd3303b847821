#!/bin/bash
# Sweep of the host-buffer entry point's slab count / compute streams (C2, one B200): e2e M cell-updates/s per setting
cd "$(dirname "$0")/.."
for cs in 2 3 4; do for sl in 6 8 10 13 16 20 26; do
  v=$(AWFL_HOST_CSTREAMS=$cs python bench.py --steps 10 --host-slabs $sl --no-cpu-baseline --no-gpu-reference --no-strong-c3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.1f %.2f' % (d['e2e']['value']/1e6, d['e2e']['ms_per_step']))")
  echo "cstreams=$cs slabs=$sl e2e=$v"
done; done
