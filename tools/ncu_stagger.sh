#!/bin/bash
mkdir -p gpurun_out
export LUTLDPC_STAGGER_CN_WARPS=${CW:-4}
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/ncu_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stagger --launch-skip 30 --launch-count 2 -o gpurun_out/prof_stagger -f $B > gpurun_out/ncu_stagger.log 2>&1
echo rc=$?
tail -3 gpurun_out/ncu_stagger.log
