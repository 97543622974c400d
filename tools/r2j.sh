#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2j_pytest.log
bash tools/ncu_round.sh
for wl in c1b_reg36_n1000 c2_irr_n500 c3_irr_n10000 c4_reg632_high c1a_reg632_wide; do
  python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2j_bench_$wl.json 2> gpurun_out/r2j_bench_$wl.err; echo "bench $wl rc=$?"
  python tools/benchline.py $wl < gpurun_out/r2j_bench_$wl.json | cut -c1-200
done
python bench.py --steps 5 --warmup 3 > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err; echo "bench rc=$?"; python tools/benchline.py c5 < gpurun_out/r2j_bench.json | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2j_bench_ref.json 2> gpurun_out/r2j_bench_ref.err; echo "bench ref rc=$?"; cat gpurun_out/r2j_bench_ref.json | cut -c1-600
