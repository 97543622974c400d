#!/bin/bash
mkdir -p gpurun_out
python tools/pcie_probe.py
B="python bench.py --warmup 3 --no-cpu-baseline --steps 3"
run() { name=$1; shift; env "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run g_c2048 LUTLDPC_HOST_CHUNK=2048 $B
run g_c4096 LUTLDPC_HOST_CHUNK=4096 $B
run g_c8192 LUTLDPC_HOST_CHUNK=8192 $B
run g_16k_c4096 LUTLDPC_HOST_CHUNK=4096 $B --frames 16384
run g_16k_c8192 LUTLDPC_HOST_CHUNK=8192 $B --frames 16384
