"""Per-launch timeline of one decode step (LUTLDPC_PROFILE_DUMP=1): kind, ms.  Usage: timeline.py [frames] [snr]"""
import os, sys
os.environ["LUTLDPC_PROFILE_DUMP"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch, _golden, lut_ldpc_b200 as L
import bench
F = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
snr = float(sys.argv[2]) if len(sys.argv) > 2 else 1.3
dec = L.Decoder(_golden.bundle_path("c5_n64800")); dec.set_exit_conditions(dec.max_iters, True, True)
if os.environ.get("LUTLDPC_POLL"):
    dec.set_poll_interval(int(os.environ["LUTLDPC_POLL"]))
st = torch.cuda.Stream(); dec.set_stream(st.cuda_stream)
with torch.cuda.stream(st):
    cha, msg = bench.make_inputs(torch, dec, F, snr, 1000)
out = torch.empty((F, dec.nvar), dtype=torch.uint8, device="cuda"); its = torch.empty((F,), dtype=torch.int32, device="cuda")
for _ in range(2):
    dec.decode_quantized_device(F, cha.data_ptr(), msg.data_ptr(), out.data_ptr(), its.data_ptr())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(3):
    dec.decode_quantized_device(F, cha.data_ptr(), msg.data_ptr(), out.data_ptr(), its.data_ptr())
e1.record(st); torch.cuda.synchronize()
print("unprofiled_step_ms", e0.elapsed_time(e1) / 3, file=sys.stderr)
dec.set_profiling(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
dec.decode_quantized_device(F, cha.data_ptr(), msg.data_ptr(), out.data_ptr(), its.data_ptr())
e1.record(st); torch.cuda.synchronize()
print("step_ms", e0.elapsed_time(e1), file=sys.stderr)
dec.get_profile()
