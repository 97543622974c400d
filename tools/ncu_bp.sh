#!/bin/bash
# ncu evidence for the BP comparison decoder (run under gpurun, one GPU): launch list of one Monte-Carlo call and
# --set full captures of its check and variable passes; only CSV / text summaries stay in gpurun_out/.
mkdir -p gpurun_out
B="python tools/bp_probe.py c5_n64800 1.3 8192 50 1"
$B > gpurun_out/bp_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/bp_launches.csv $B > gpurun_out/bp_launch_ncu.log 2>&1
echo launches rc=$?
cap() {
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c $cnt -f -o /tmp/$name "$@" > gpurun_out/full_ncu_$name.log 2>&1
  echo "$name rc=$?"; tail -1 gpurun_out/full_ncu_$name.log
  ncu -i /tmp/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  python tools/ncu_srcagg.py /tmp/$name.ncu-rep > gpurun_out/${name}_source.txt 2>/dev/null
  rm -f /tmp/$name.ncu-rep
}
cap bp_cn "k_bp_cn" 3 1 $B
cap bp_vn "k_bp_vn" 3 1 $B
du -sh gpurun_out
