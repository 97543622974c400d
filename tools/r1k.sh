#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
B="python bench.py --warmup 2 --no-cpu-baseline --steps 3"
run() { name=$1; shift; env "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run k_default $B
run k_c3 $B --workload c3_irr_n10000
run k_c1b $B --workload c1b_reg36_n1000
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_k_default.json').read().strip().splitlines()[-1])
print(d['roofline']['profile_ms'])
PY
