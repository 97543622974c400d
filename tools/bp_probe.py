#!/usr/bin/env python3
"""Throughput of the BP comparison decoder's Monte-Carlo step: tools/bp_probe.py [cfg] [snr_db] [frames] [max_iter] [repack]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import _alist  # noqa: E402
from lut_ldpc_b200 import capi  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "c5_n64800"
snr = float(sys.argv[2]) if len(sys.argv) > 2 else 1.3
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
max_iter = int(sys.argv[4]) if len(sys.argv) > 4 else 50
repack = int(sys.argv[5]) if len(sys.argv) > 5 else 1
g = _alist.bundle_graph(cfg)
if g["nvar"] >= 50000:
    g["nchk_lin_indep"] = g["nchk"]
dec = capi.BpDecoder(g, max_iters=max_iter)
dec.set_repack(repack)
dec.simulate_snr(snr, min(frames, 1024), seed=1)
for rep in range(2):
    t0 = time.perf_counter()
    c = dec.simulate_snr(snr, frames, seed=2 + rep, batch_frames=frames)
    wall = time.perf_counter() - t0
    st = dec.last_call_stats()
    E = g["num_edges"]
    # algorithmic bytes: every frame-iteration reads and writes each edge message twice (check pass, variable pass)
    # and reads the channel value once: (4 E + N) * 4 bytes
    alg = st["frame_iterations"] * (4 * E + g["nvar"]) * 4
    print(json.dumps(dict(cfg=cfg, snr_db=snr, frames=frames, max_iter=max_iter, repack=repack, device_ms=round(st["device_ms"], 2),
                          wall_ms=round(wall * 1e3, 2), launches=st["kernel_launches"],
                          info_gbit_s=round(frames * dec.ninfo / st["device_ms"] / 1e6, 4),
                          mean_iters=round(c.sum_iters / c.frames, 2), fer=c.frame_errors / c.frames,
                          ber=c.data_bit_errors / c.data_bits, alg_GBps=round(alg / st["device_ms"] / 1e6, 1))), flush=True)
