#!/bin/bash
# round 2, call a: GPU tests, no-alloc test alone, frame search for the deeper C5/C3 fixtures, sim probe, bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2a_pytest.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2a_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 gpurun_out/r2a_noalloc_alone.log
python tools/find_frames.py c5_n64800 1.15,1.3 65536 10 > gpurun_out/r2a_find_c5.log 2>&1; echo "find c5 rc=$?"; grep "returns seen" gpurun_out/r2a_find_c5.log
python tools/find_frames.py c3_irr_n10000 1.2,1.4 65536 8 > gpurun_out/r2a_find_c3.log 2>&1; echo "find c3 rc=$?"; grep "returns seen" gpurun_out/r2a_find_c3.log
python tools/sim_probe.py c5_n64800 1.3 32768 8192 > gpurun_out/r2a_sim_8192.log 2>&1; tail -2 gpurun_out/r2a_sim_8192.log
python tools/sim_probe.py c5_n64800 1.3 32768 32768 > gpurun_out/r2a_sim_32768.log 2>&1; tail -2 gpurun_out/r2a_sim_32768.log
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"; cat gpurun_out/r2a_bench.json
