#!/bin/bash
# ncu evidence of the default bench command (run under gpurun, one GPU): launch list (per-launch gpu__time_duration)
# and --set full captures of the dominant kernels, each after the same command exited 0 without ncu.
#   tools/ncu_round.sh            -> gpurun_out/launches.csv, prof_top.ncu-rep, prof_channel.ncu-rep, prof_onchip.ncu-rep
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/launch_ncu.log 2>&1
echo launches rc=$?
$B > gpurun_out/ncu_plain2.json 2> gpurun_out/ncu_plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:"k_vn_fused|k_cn_exact" -s 120 -c 4 -f -o gpurun_out/prof_top $B > gpurun_out/full_ncu.log 2>&1
echo full rc=$?; tail -1 gpurun_out/full_ncu.log
ncu --set full --clock-control none --import-source on -k regex:"k_channel|k_encode_bits|k_unpack_bits|k_pack" -c 6 -f -o gpurun_out/prof_channel $B > gpurun_out/full_ncu_channel.log 2>&1
echo channel rc=$?; tail -1 gpurun_out/full_ncu_channel.log
C="python bench.py --workload c1b_reg36_n1000 --steps 1 --warmup 1 --no-cpu-baseline"
$C > gpurun_out/ncu_plain_c1b.json 2> gpurun_out/ncu_plain_c1b.err && \
ncu --set full --clock-control none --import-source on -k regex:k_onchip -s 2 -c 1 -f -o gpurun_out/prof_onchip $C > gpurun_out/full_ncu_onchip.log 2>&1
echo onchip rc=$?; tail -1 gpurun_out/full_ncu_onchip.log
for r in prof_top prof_channel prof_onchip; do ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/${r}_raw.csv 2>/dev/null; done
