#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --warmup 2 --no-cpu-baseline --steps 2"
run() { name=$1; shift; env "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run l_32k $B --frames 32768
run l_64k $B --frames 65536
run l_16k_snr16 $B --snr 1.6
run l_16k_snr20 $B --snr 2.0
