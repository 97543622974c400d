#!/usr/bin/env python3
"""Turns the output of the GPU frame search (tools/find_frames.py under gpurun: gpurun_out/<cfg>_sel.llr.f64 + .json)
into the committed INPUT of the `sel` fixtures: tests/golden/<cfg>.sel.llr_i16.xz (+ .json).

The channel LLRs are snapped to the centres of the 0.01-wide cells the reference's quantiser thresholds are aligned to
(src/LDPC_DE.cpp:562,579: thresholds are multiples of 0.01): llr' = (k + 0.5) * 0.01, stored as int16 k.  A centre is
0.005 away from every threshold, so quant_nonlin gives llr' the label of its cell no matter how the double compare
rounds, and the file is a quarter of the float64 size.  Frame order: one failing and one converging frame first (they are
the ones `lutldpc_ref fixture --trace 2` dumps per iteration), then the rest as found.

usage: python tests/golden/make_sel_inputs.py <cfg>"""
import json
import lzma
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def main():
    cfg = sys.argv[1]
    meta = json.load(open(os.path.join(ROOT, "gpurun_out", cfg + "_sel.json")))
    llr = np.fromfile(os.path.join(ROOT, "gpurun_out", cfg + "_sel.llr.f64"), "<f8").reshape(len(meta), -1)
    first_fail = next(i for i, m in enumerate(meta) if m["kind"] == "fail")
    first_conv = next(i for i, m in enumerate(meta) if m["kind"] in ("typical", "early"))
    order = [first_fail, first_conv] + [i for i in range(len(meta)) if i not in (first_fail, first_conv)]
    k = np.floor(llr[order] / 0.01)
    assert np.abs(k).max() < 32000
    k = k.astype("<i2")
    with lzma.open(os.path.join(HERE, cfg + ".sel.llr_i16.xz"), "wb", preset=9) as f:
        f.write(k.tobytes())
    info = {"what": "channel LLR cell index k, llr = (k + 0.5) * 0.01; int16 little endian [frames][nvar]; all-zero codeword",
            "found_by": "tools/find_frames.py (GPU search for rare lut_decode returns; the GPU's own results are NOT stored "
                        "as expected outputs -- the fixture is what the reference makes of these inputs)",
            "frames": [{"snr_db": meta[i]["snr"], "picked_as": meta[i]["kind"]} for i in order]}
    json.dump(info, open(os.path.join(HERE, cfg + ".sel.json"), "w"), indent=1)
    print(cfg, k.shape, os.path.getsize(os.path.join(HERE, cfg + ".sel.llr_i16.xz")))


if __name__ == "__main__":
    main()
