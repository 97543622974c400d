#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_e.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_e.log
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
run() { name=$1; shift; env "$@" $B > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run e_default X=1
run e_chunk4096 LUTLDPC_HOST_CHUNK=4096
run e_chunk1024 LUTLDPC_HOST_CHUNK=1024
run e_chunk8192 LUTLDPC_HOST_CHUNK=8192
