#!/usr/bin/env python3
"""Times lutldpc_simulate_snr (the fused Monte-Carlo step) with the per-kind profile.
  python tools/sim_probe.py <cfg> <snr> <frames> [batch_frames]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _golden  # noqa: E402
import lut_ldpc_b200 as L  # noqa: E402

cfg, snr, frames = sys.argv[1], float(sys.argv[2]), int(sys.argv[3])
batch = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dec = L.Decoder(_golden.bundle_path(cfg))
dec.set_exit_conditions(dec.max_iters, True, True)
dec.simulate_snr(snr, frames, seed=1, batch_frames=batch)
for prof in (False, True):
    dec.set_profiling(prof)
    c = dec.simulate_snr(snr, frames, seed=2, batch_frames=batch)
    st = dec.last_call_stats()
    line = {"cfg": cfg, "snr": snr, "frames": frames, "batch": batch, "device_ms": round(st["device_ms"], 3),
            "info_gbit_s": round(frames * dec.ninfo / st["device_ms"] / 1e6, 4), "fer": c.frame_errors / c.frames,
            "mean_iters": c.sum_iters / c.frames, "launches": st["kernel_launches"]}
    if prof:
        line["profile_ms"] = {k: round(v[0], 3) for k, v in dec.get_profile().items()}
    print(json.dumps(line), flush=True)
