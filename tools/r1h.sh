#!/bin/bash
mkdir -p gpurun_out
nproc
( time python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>&1 | grep real
cat gpurun_out/bench_default.json | cut -c1-3000
tail -3 gpurun_out/bench_default.err
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2>&1 | grep real
cat gpurun_out/bench_ref.json
tail -3 gpurun_out/bench_ref.err
