#!/bin/bash
# Driver-style bench lines of every BASELINE.json configuration on one B200 (both arms), into gpurun_out/<tag>_*.json
tag=${1:-r02}
cd "$(dirname "$0")/.."
run() { name=$1; shift; python bench.py "$@" > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err || echo "FAILED $name rc=$?"; }
run bench_c2_default --steps 20 --warmup 5
run bench_c3 --workload C3 --steps 20 --warmup 5
run bench_c4_ord3 --workload C4 --ord 3 --steps 20 --warmup 5
run bench_c4_ord5 --workload C4 --steps 20 --warmup 5
run bench_c4_ord9 --workload C4 --ord 9 --steps 20 --warmup 5
run bench_c5 --workload C5 --steps 20 --warmup 5
run bench_c1 --workload C1 --steps 20 --warmup 5 --no-gpu-reference
run ref_c1 --impl reference --workload C1 --steps 20 --warmup 2
run ref_c2 --impl reference --workload C2 --steps 10 --warmup 1
run ref_c3 --impl reference --workload C3 --steps 5 --warmup 1
run ref_c4_ord3 --impl reference --workload C4 --ord 3 --steps 5 --warmup 1
run ref_c4_ord5 --impl reference --workload C4 --steps 5 --warmup 1
run ref_c4_ord9 --impl reference --workload C4 --ord 9 --steps 5 --warmup 1
run ref_c5 --impl reference --workload C5 --steps 5 --warmup 1
python tools/bench_summary.py gpurun_out/${tag}_bench_*.json
for f in gpurun_out/${tag}_ref_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], "%.3f M/s"%(d["value"]/1e6), d["cpu_baseline"]["cores"], "cores")
PY
done
