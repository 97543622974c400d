#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_generator.py -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2h_pytest.log
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2h_bench.err
python tools/benchline.py c5 < gpurun_out/r2h_bench.json
python -c "
import json; d=json.loads(open('gpurun_out/r2h_bench.json').read().strip().splitlines()[-1]); print('random_cw', d['random_codeword'])"
