"""Host-buffer decode (lutldpc_decode_quantized, pinned memory) timed from the host, with per-chunk times
(LUTLDPC_HOST_TRACE=1) and the raw PCIe copy rates.  Usage: e2e_trace.py [frames]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch, numpy as np, _golden, lut_ldpc_b200 as L, bench
F = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
dec = L.Decoder(_golden.bundle_path("c5_n64800")); dec.set_exit_conditions(dec.max_iters, True, True)
st = torch.cuda.Stream(); dec.set_stream(st.cuda_stream)
with torch.cuda.stream(st):
    cha, msg = bench.make_inputs(torch, dec, F, 1.3, 1000)
N = dec.nvar
cha_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True); msg_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True)
cha_h.copy_(cha); msg_h.copy_(msg); torch.cuda.synchronize(); del cha, msg
out_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True); its_h = torch.empty((F,), dtype=torch.int32, pin_memory=True)
a, b, o, i = cha_h.numpy(), msg_h.numpy(), out_h.numpy(), its_h.numpy()
dec.decode_quantized(a, b, out=o, iters=i)
for rep in range(2):
    t0 = time.perf_counter(); dec.decode_quantized(a, b, out=o, iters=i); dt = time.perf_counter() - t0
    print("e2e F %d: %.2f ms -> %.3f Gbit/s" % (F, dt * 1e3, F * dec.ninfo / dt / 1e9), file=sys.stderr)
t0 = time.perf_counter(); d = cha_h.cuda(non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("H2D %.1f GB/s" % (F * N / dt / 1e9), file=sys.stderr)
t0 = time.perf_counter(); out_h.copy_(d, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("D2H %.1f GB/s" % (F * N / dt / 1e9), file=sys.stderr)
