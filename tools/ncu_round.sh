#!/bin/bash
# ncu evidence of the default bench command (run under gpurun, one GPU): launch list (per-launch gpu__time_duration)
# and --set full captures of the dominant kernels, each after the same command exited 0 without ncu.  Only CSV / text
# summaries are left in gpurun_out/ (the .ncu-rep files are tens of MB each; gpurun brings back at most 64 MiB).
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/launch_ncu.log 2>&1
echo launches rc=$?
cap() {  # name, kernel regex, skip, count, command...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c $cnt -f -o /tmp/$name "$@" > gpurun_out/full_ncu_$name.log 2>&1
  echo "$name rc=$?"; tail -1 gpurun_out/full_ncu_$name.log
  ncu -i /tmp/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  python tools/ncu_srcagg.py /tmp/$name.ncu-rep > gpurun_out/${name}_source.txt 2>/dev/null
  rm -f /tmp/$name.ncu-rep
}
cap prof_vn "k_vn_fused" 60 1 $B
cap prof_cn "k_cn_exact" 90 3 $B
cap prof_channel "k_channel" 0 1 $B
cap prof_encode "k_encode_bits" 0 1 $B
cap prof_io "k_pack|k_unpack" 0 4 $B
C="python bench.py --workload c1b_reg36_n1000 --steps 1 --warmup 1 --no-cpu-baseline"
$C > gpurun_out/ncu_plain_c1b.json 2> gpurun_out/ncu_plain_c1b.err && cap prof_onchip "k_onchip" 2 1 $C
ls -la gpurun_out | head -40; du -sh gpurun_out
