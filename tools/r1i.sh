#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --warmup 2 --no-cpu-baseline --steps 2"
run() { name=$1; shift; env "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
for t in 34 64 128 256 512 1024; do run i_tile$t LUTLDPC_VN_TILE_OPS=$t $B; done
run i_unfused LUTLDPC_VN_UNFUSED=1 $B
run i_cnvec4 LUTLDPC_CN_VEC=4 $B
