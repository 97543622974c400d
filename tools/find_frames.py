#!/usr/bin/env python3
"""GPU-assisted search for rare decoder outcomes (run under gpurun): decodes seeded AWGN frames of a golden config on
the B200 and writes the channel LLRs (float64) of a selection -- frames that return -max_iters, +max_iters (valid only
after the decision pass), early, late -- to gpurun_out/<cfg>_sel.llr.f64 + .json.  The selection is then pushed through
the UNMODIFIED reference (`lutldpc_ref fixture --llr-from`, tests/golden/make_golden.py), whose outputs become the
committed fixture; nothing the GPU computed here is stored as expected output.
  python tools/find_frames.py <cfg> <snr,snr,..> <frames per snr> <n_select per snr>"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _golden  # noqa: E402
import lut_ldpc_b200 as L  # noqa: E402

cfg, snrs, total, nsel = sys.argv[1], [float(x) for x in sys.argv[2].split(",")], int(sys.argv[3]), int(sys.argv[4])
dec = L.Decoder(_golden.bundle_path(cfg))
dec.set_exit_conditions(dec.max_iters, True, True)
N, M = dec.nvar, dec.max_iters
chunk = max(256, min(4096, (1 << 31) // (8 * N) // 256 * 256))
sel_llr, meta = [], []
for snr in snrs:
    N0 = 10.0 ** (-snr / 10.0) / dec.rate
    sigma = (N0 / 2.0) ** 0.5
    g = torch.Generator(device="cuda")
    g.manual_seed(int(snr * 1000))
    picked = {}
    hist = np.zeros(2 * M + 1, np.int64)
    for f0 in range(0, total, chunk):
        llr = (1.0 + sigma * torch.randn((chunk, N), dtype=torch.float64, device="cuda", generator=g)) * (4.0 / N0)
        out = torch.empty((chunk, N), dtype=torch.uint8, device="cuda")
        its = torch.empty((chunk,), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        dec.decode_llr_device(chunk, llr.data_ptr(), out.data_ptr(), its.data_ptr())
        torch.cuda.synchronize()
        it = its.cpu().numpy()
        np.add.at(hist, it + M, 1)
        want = {"fail": np.flatnonzero(it == -M), "at_max": np.flatnonzero(it == M), "late": np.flatnonzero((it > 0.8 * M) & (it < M)),
                "early": np.flatnonzero((it > 0) & (it <= np.percentile(it[it > 0], 5) if (it > 0).any() else 0)),
                "typical": np.flatnonzero((it > 0) & (it < 0.8 * M))}
        quota = {"fail": max(2, nsel // 4), "at_max": 2, "late": 2, "early": 2, "typical": nsel}
        for k, idx in want.items():
            have = picked.setdefault(k, [])
            for i in idx[: max(0, quota[k] - len(have))]:
                have.append((llr[i].cpu().numpy().copy(), int(it[i]), f0 + int(i)))
        if all(len(picked.get(k, [])) >= quota[k] for k in ("fail", "at_max", "late", "early")):
            break
    order = []
    for k in ("fail", "at_max", "late", "early", "typical"):
        for x in picked.get(k, []):
            if len(order) < nsel and not any(x[2] == y[2] for y in order):
                order.append(x + (k,))
    for x in order:
        sel_llr.append(x[0])
        meta.append({"snr": snr, "gpu_ret": x[1], "frame": x[2], "kind": x[3]})
    print("snr", snr, "returns seen:", {int(i - M): int(c) for i, c in enumerate(hist) if c}, flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.stack(sel_llr).astype("<f8").tofile(os.path.join(ROOT, "gpurun_out", cfg + "_sel.llr.f64"))
json.dump(meta, open(os.path.join(ROOT, "gpurun_out", cfg + "_sel.json"), "w"), indent=1)
print(json.dumps(meta))
