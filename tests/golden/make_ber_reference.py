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

POINTS = {  # cfg -> (snr list, frames per seed, seeds)
    "c1b_reg36_n1000": ([1.5, 2.0, 2.5], 4000, 8),
    "c2_irr_n500": ([1.0, 2.0, 2.5], 4000, 8),
    "c3_irr_n10000": ([1.2, 1.3], 60, 8),
}


def run_one(cfg, snr, frames, seed):
    alist, dargs, _ = mg.CONFIGS[cfg]
    apath = os.path.join(mg.CODES, alist + ".alist")
    cmd = [mg.REF_BIN, "sim", apath] + dargs.split() + ["--snr", str(snr), "--frames", str(frames), "--seed", str(seed),
                                                       "--count-iters"]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def main():
    res = {}
    jobs = []
    with ThreadPoolExecutor(max_workers=8) as ex:
        for cfg, (snrs, frames, seeds) in POINTS.items():
            for snr in snrs:
                for s in range(seeds):
                    jobs.append((cfg, snr, ex.submit(run_one, cfg, snr, frames, s)))
        for cfg, snr, fut in jobs:
            r = fut.result()
            acc = res.setdefault(cfg, {}).setdefault("%.2f" % snr, {k: 0 for k in (
                "frames", "frame_errors", "data_bits", "data_bit_errors", "uncoded_bits", "uncoded_bit_errors",
                "sum_iters", "decode_failures")})
            for k in acc:
                acc[k] += int(r[k])
    for cfg in res:
        res[cfg]["_how"] = "lutldpc_ref sim <alist> %s --snr S --frames %d --seed 0..%d --count-iters" % (
            mg.CONFIGS[cfg][1], POINTS[cfg][1], POINTS[cfg][2] - 1)
        res[cfg]["_zero_codeword"] = "--gen" not in mg.CONFIGS[cfg][1]
    json.dump(res, open(os.path.join(HERE, "ber_reference.json"), "w"), indent=1, sort_keys=True)
    print("wrote ber_reference.json")


if __name__ == "__main__":
    main()
