#!/bin/bash
# ncu evidence of the default bench command: launch list (per-launch gpu__time_duration) and one --set full capture
# of the dominant kernels, each after the same command exited 0 without ncu.
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/launch_ncu.log 2>&1
echo launches rc=$?
$B > gpurun_out/ncu_plain2.json 2> gpurun_out/ncu_plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:"k_vn_fused|k_cn_exact" -s 120 -c 4 -o gpurun_out/prof_top $B > gpurun_out/full_ncu.log 2>&1
echo full rc=$?
tail -2 gpurun_out/full_ncu.log
