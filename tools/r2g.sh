#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_generator.py -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2g_pytest.log
python - <<'PY' > gpurun_out/r2g_c5_random.log 2>&1
import sys, time, json
sys.path[:0]=['.','tests']
import _golden, lut_ldpc_b200 as L
src=L.Decoder(_golden.bundle_path("c5_n64800"))
t0=time.time(); sysd,perm=src.systematic(); t1=time.time()
print("systematic construction (rank + generator on device) s:", round(t1-t0,2), "rank", sysd.nchk_lin_indep)
for d,zc in ((src,True),(sysd,False)):
    d.set_exit_conditions(50,True,True)
    d.simulate_snr(1.3,32768,seed=1,zero_codeword=zc,batch_frames=32768)
    d.set_profiling(True)
    c=d.simulate_snr(1.3,32768,seed=2,zero_codeword=zc,batch_frames=32768)
    st=d.last_call_stats()
    print(json.dumps({"zero_cw":zc,"device_ms":st["device_ms"],"gbit_s":32768*d.ninfo/st["device_ms"]/1e6,"fer":c.frame_errors/c.frames,"prof":{k:round(v[0],2) for k,v in d.get_profile().items() if v[0]>1}}))
PY
cat gpurun_out/r2g_c5_random.log | tail -5
python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/benchline.py c5
