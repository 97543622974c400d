#!/bin/bash
# round-1 measurement script (c): staggered schedule parity + bench variants
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_c.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_c.log
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
run() { name=$1; shift; env "$@" $B > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "$name rc=$?"; python tools/benchline.py $name < gpurun_out/bench_$name.json; }
run stg_v2 LUTLDPC_STAGGER_VEC=2
run stg_v4 LUTLDPC_STAGGER_VEC=4
run nostg LUTLDPC_NO_STAGGER=1
run stg_v2_ops128 LUTLDPC_STAGGER_VEC=2 LUTLDPC_STAGGER_VN_OPS=128
run stg_v2_ops32_r1 LUTLDPC_STAGGER_VEC=2 LUTLDPC_STAGGER_VN_OPS=32 LUTLDPC_STAGGER_CN_ROWS=1
run stg_v2_r4 LUTLDPC_STAGGER_VEC=2 LUTLDPC_STAGGER_CN_ROWS=4
python - <<'PY'
import sys, os
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "tests")]
import _golden, lut_ldpc_b200 as L
d = L.Decoder(_golden.bundle_path("c1b_reg36_n1000")); d.set_exit_conditions(d.max_iters, 1, 1)
for seed in (5, 6):
    c = d.simulate_snr(2.0, 400000, seed=seed, zero_codeword=True)
    print("c1b 2.0dB seed", seed, c.as_dict(), "FER", c.frame_errors / c.frames)
PY
