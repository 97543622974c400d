#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "stagger" > gpurun_out/pytest_d.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_d.log
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
run() { name=$1; shift; env "$@" $B > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run p_cw2 LUTLDPC_STAGGER_CN_WARPS=2
run p_cw1 LUTLDPC_STAGGER_CN_WARPS=1
run p_cw3 LUTLDPC_STAGGER_CN_WARPS=3
run p_cw4 LUTLDPC_STAGGER_CN_WARPS=4
run p_cw2_ops64 LUTLDPC_STAGGER_CN_WARPS=2 LUTLDPC_STAGGER_VN_OPS=64
run p_cw2_r8 LUTLDPC_STAGGER_CN_WARPS=2 LUTLDPC_STAGGER_CN_ROWS=8
