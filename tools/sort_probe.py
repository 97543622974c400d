"""Does a cheap per-frame channel metric predict the iteration count?  (frame ordering experiment)"""
import os, sys
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [R, os.path.join(R, "tests")]
import _golden, lut_ldpc_b200 as L
d = L.Decoder(_golden.bundle_path("c5_n64800")); d.set_exit_conditions(d.max_iters, 1, 1)
F, N = 8192, d.nvar
snr = float(sys.argv[1]) if len(sys.argv) > 1 else 1.3
sigma = np.sqrt(0.5 / (d.rate * 10 ** (snr / 10)))
rng = np.random.default_rng(5)
cha = np.empty((F, N), np.uint8); msg = np.empty((F, N), np.uint8)
m1 = np.empty(F); m2 = np.empty(F); m3 = np.empty(F)
for f0 in range(0, F, 256):
    llr = (2.0 / sigma ** 2) * (1.0 + sigma * rng.standard_normal((256, N)))
    c, m = d.quantize(llr)
    cha[f0:f0 + 256], msg[f0:f0 + 256] = c, m
    rel = np.abs(c.astype(np.float64) - 7.5)
    m1[f0:f0 + 256] = rel.mean(axis=1)               # mean label reliability
    m2[f0:f0 + 256] = (rel < 2.0).mean(axis=1)       # fraction of weak labels
    m3[f0:f0 + 256] = np.abs(llr).mean(axis=1)       # mean |LLR|
out, it = d.decode_quantized(cha, msg)
ai = np.abs(it).astype(np.int64)
print("mean iters", ai.mean(), "fails", int((it < 0).sum()))
def cost(order, g):
    a = ai[order]
    return sum(a[i:i + g].max() for i in range(0, F, g)) / (F // g)
for name, key in (("random", rng.random(F)), ("mean_rel", -m1), ("weak_frac", m2), ("mean_abs_llr", -m3), ("oracle", ai.astype(float))):
    o = np.argsort(key, kind="stable")
    print("%-13s corr %.3f  mean over groups of max iters: g=8 %.2f  g=256 %.2f  g=1024 %.2f" % (
        name, np.corrcoef(key, ai)[0, 1], cost(o, 8), cost(o, 256), cost(o, 1024)))
