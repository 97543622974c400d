#!/bin/bash
# ncu --set full of the two dominant kernels in the middle of a decode (all frames still live)
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/ncu_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_cn_exact|k_vn_fused' --launch-skip 64 --launch-count 4 -o gpurun_out/prof_top -f $B > gpurun_out/ncu_top.log 2>&1
echo rc=$?
tail -3 gpurun_out/ncu_top.log
