#!/usr/bin/env python3
"""BER/FER points of the REFERENCE ber_sim frame loop (oracle/_ref/lutldpc_ref sim = the unmodified
reference decoder + IT++ BPSK/AWGN/BERC/BLERC, src/LDPC_BER_Sim.cpp:246-311), written to
tests/golden/ber_reference.json.  The GPU tests compare the fused Philox pipeline against these
within the 95 % Monte-Carlo confidence interval.  Runs only in the build container."""
import json
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

POINTS = [  # (cfg, snr list, frames per seed, seeds); points already in ber_reference.json are kept, not recomputed
    ("c1b_reg36_n1000", [1.5, 2.0, 2.5], 4000, 8),
    ("c2_irr_n500", [1.0, 2.0, 2.5], 4000, 8),
    ("c3_irr_n10000", [1.2, 1.3], 60, 8),
    # round 2: the SNR grids of the shipped params files (params/ber.ini.irregular.example:10 -> 0:.5:4;
    # the (3,6) decoder on 1:.5:4), at the frame counts this container's 8 cores afford
    ("c2_irr_n500", [0.0, 0.5, 1.5, 3.0, 3.5, 4.0], 1000, 8),
    ("c1b_reg36_n1000", [1.0, 3.0, 3.5, 4.0], 1000, 8),
    ("c3_irr_n10000", [1.1, 1.4], 60, 8),
    ("c5_n64800", [1.15, 1.3], 20, 8),
    # QCHA initial messages, file trees with random codewords (params/ber.ini.regular.example), the 6-input root LUT
    ("x_qcha_irr_n1000", [1.5, 2.0, 2.5], 1000, 8),
    ("c1a_reg632_wide", [3.5, 4.0, 4.5], 1000, 8),
    ("c4_reg632_high", [4.0, 4.5], 1000, 8),
]


def run_one(cfg, snr, frames, seed):
    alist, dargs, _ = mg.CONFIGS[cfg]
    apath = os.path.join(mg.CODES, alist + ".alist")
    if cfg in mg.CODEC_FOR and "--gen" not in dargs:
        # the committed codec file the reference wrote (save_code): skips the design step (N=64800: 9 minutes of rank)
        sys.path.insert(0, os.path.dirname(HERE))
        import _golden
        src = [_golden.codec_path(cfg)]
    else:
        src = [apath] + dargs.split()
    cmd = [mg.REF_BIN, "sim"] + src + ["--snr", str(snr), "--frames", str(frames), "--seed", str(seed), "--count-iters"]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def main():
    out_path = os.path.join(HERE, "ber_reference.json")
    res = json.load(open(out_path)) if os.path.exists(out_path) else {}
    jobs = []
    with ThreadPoolExecutor(max_workers=int(os.environ.get("BER_REF_JOBS", "8"))) as ex:
        for cfg, snrs, frames, seeds in POINTS:
            for snr in snrs:
                if "%.2f" % snr in res.get(cfg, {}):
                    continue
                for s in range(seeds):
                    jobs.append((cfg, snr, frames, seeds, ex.submit(run_one, cfg, snr, frames, s)))
        for cfg, snr, frames, seeds, fut in jobs:
            r = fut.result()
            acc = res.setdefault(cfg, {}).setdefault("%.2f" % snr, {k: 0 for k in (
                "frames", "frame_errors", "data_bits", "data_bit_errors", "uncoded_bits", "uncoded_bit_errors",
                "sum_iters", "decode_failures")})
            for k in acc:
                acc[k] += int(r[k])
            res[cfg].setdefault("_how", "lutldpc_ref sim <alist or codec.it> %s --snr S --frames <frames/8> --seed 0..7 "
                                        "--count-iters" % mg.CONFIGS[cfg][1])
            res[cfg]["_zero_codeword"] = "--gen" not in mg.CONFIGS[cfg][1]
            res[cfg]["_init_mode"] = "qcha" if "--init-mode qcha" in mg.CONFIGS[cfg][1] else "cont"
    for cfg in res:
        res[cfg].setdefault("_init_mode", "qcha" if "--init-mode qcha" in mg.CONFIGS[cfg][1] else "cont")
    json.dump(res, open(out_path, "w"), indent=1, sort_keys=True)
    print("wrote ber_reference.json")


if __name__ == "__main__":
    main()
