#!/usr/bin/env python3
"""BER/FER points of the reference's BP comparison decoder (oracle/_ref/lutldpc_ref bpsim: the frame loop of
LDPC_BER_Sim::sim_snr_point around IT++'s LDPC_Code::decode, [BP] defaults max_iter 30, QLLR 12/300/7/28) ->
tests/golden/bp_ber_reference.json.  Needs /root/reference and oracle/_ref; run once, output committed."""
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import _alist  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "lutldpc_ref")
POINTS = {  # cfg: (snr list, frames, extra options)
    "c1b_reg36_n1000": ("1.0,1.5,2.0,2.5", 6000, ""),
    "c2_irr_n500": ("1.0,2.0,3.0", 6000, ""),
    "c2_irr_n500:minsum": ("2.0,3.0", 6000, "--d2 0"),
    "c3_irr_n10000": ("0.9,1.1", 400, "--max-iter 50"),
}

out = {}
with tempfile.TemporaryDirectory() as tmp:
    for key, (snrs, frames, extra) in POINTS.items():
        cfg = key.split(":")[0]
        alist = os.path.join(tmp, cfg + ".alist")
        _alist.write_alist(cfg, alist)
        txt = subprocess.check_output([REF, "bpsim", alist, "--snr", snrs, "--frames", str(frames), "--seed", "5"] + extra.split(),
                                      text=True)
        out[key] = {"_args": extra}
        for line in txt.splitlines():
            if line.startswith("{"):
                r = json.loads(line)
                out[key]["%.2f" % r["snr_db"]] = {k: r[k] for k in ("frames", "data_bits", "data_bit_errors", "uncoded_bits",
                                                                     "uncoded_bit_errors", "frame_errors")}
                print(key, line)
json.dump(out, open(os.path.join(HERE, "bp_ber_reference.json"), "w"), indent=1, sort_keys=True)
