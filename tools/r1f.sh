#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_f.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_f.log
B="python bench.py --warmup 3 --no-cpu-baseline"
run() { name=$1; shift; env "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; tail -2 gpurun_out/bench_$name.err; }
run f_default $B --steps 3
run f_16k $B --steps 2 --frames 16384
run f_snr16 $B --steps 3 --snr 1.6
run f_c3 $B --steps 3 --workload c3_irr_n10000
