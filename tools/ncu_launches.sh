#!/bin/bash
# launch list of the default bench command (per-launch gpu__time_duration), after a plain run exited 0
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/launch_plain.json 2> gpurun_out/launch_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/launch_ncu.log 2>&1
echo rc=$?
python tools/benchline.py plain < gpurun_out/launch_plain.json
