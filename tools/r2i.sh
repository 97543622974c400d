#!/bin/bash
# round 2, 8-GPU call: multi-GPU harness test, weak-scaling bench at 1/2/4/8 GPUs, BER curves on the reference grids
mkdir -p gpurun_out/ber
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi -L | head -8
python -m pytest tests/test_gpu_ber.py -m gpu -x -q -k multi_gpu > gpurun_out/r2i_multi.log 2>&1; echo "multi-gpu harness test rc=$?"; tail -2 gpurun_out/r2i_multi.log
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_scale_1.json 2> gpurun_out/r2i_scale_1.err; echo "bench 1 rc=$?"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/r2i_scale_$n.json 2> gpurun_out/r2i_scale_$n.err; echo "bench $n rc=$?"
done
for n in 1 2 4 8; do python tools/benchline.py N$n < gpurun_out/r2i_scale_$n.json | cut -c1-150; done
# BER curves (zero codeword, psc = pisc = 1)
$TR --nproc-per-node 8 --master-port 29701 tools/ber_curves.py c2_irr_n500 0:.5:4 4e6 1e12 0 0 65536 gpurun_out/ber/c2_irr_n500_8gpu.json 2> gpurun_out/ber/c2.err | tail -1 | cut -c1-400
$TR --nproc-per-node 8 --master-port 29702 tools/ber_curves.py c1b_reg36_n1000 1:.5:4 2e6 1e12 0 0 65536 gpurun_out/ber/c1b_reg36_n1000_8gpu.json 2> gpurun_out/ber/c1b.err | tail -1 | cut -c1-400
for n in 1 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29710+n)) tools/ber_curves.py c3_irr_n10000 1.1:.1:1.7 1e6 1e12 0 0 32768 gpurun_out/ber/c3_irr_n10000_${n}gpu.json 2> gpurun_out/ber/c3_$n.err | tail -1 | cut -c1-400
done
$TR --nproc-per-node 8 --master-port 29720 tools/ber_curves.py c5_n64800 1.1:.05:1.8 6e7 100 1e-7 0 32768 gpurun_out/ber/c5_n64800_8gpu.json 2> gpurun_out/ber/c5.err | tail -1 | cut -c1-600
