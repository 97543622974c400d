B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/ncu_plain3.json 2> gpurun_out/ncu_plain3.err && \
ncu --set full --clock-control none --import-source on -k regex:k_cn_exact -s 90 -c 3 -o gpurun_out/prof_cn $B > gpurun_out/full_ncu_cn.log 2>&1
echo rc=$?; tail -2 gpurun_out/full_ncu_cn.log
