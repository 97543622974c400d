#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
B="python bench.py --warmup 2 --no-cpu-baseline --steps 3"
run() { name=$1; shift; env "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run j_default $B
run j_unfused LUTLDPC_VN_UNFUSED=1 $B
