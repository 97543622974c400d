#!/usr/bin/env python3
"""Golden vectors of the BP comparison decoder, made by the UNMODIFIED reference (oracle/_ref/lutldpc_ref bpfixture:
IT++'s LDPC_Code::bp_decode with the LLR_calc_unit the [BP] section names).  Needs /root/reference and oracle/_ref;
run once in the build container, the outputs (tests/golden/bp/*.fix.xz + bp_manifest.json) are committed.

Per frame the fixture holds the channel LLRs (small codes), to_qllr of them, bp_decode's soft output and return value
and the bits LDPC_Code::decode hands to ber_sim.  Graphs: the codes of the LUT fixtures (alist rebuilt from the golden
bundle, tests/_alist.py, so that the tests need no /root/reference) and two synthetic matrices with degree mixes
no shipped alist has (check degrees 2..9 and 40 with the hard-coded degree-6 update; (3,48)-regular)."""
import hashlib
import json
import lzma
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import _alist  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "lutldpc_ref")
OUT = os.path.join(HERE, "bp")

# name: (graph, reference options).  graph = golden bundle cfg or synth:<kind>
CASES = {
    "c1b_default": ("c1b_reg36_n1000", "--snr 1.5 --frames 12 --max-iter 30 --seed 1"),
    "c1b_minsum_q6": ("c1b_reg36_n1000", "--snr 2.0 --frames 10 --max-iter 30 --d1 2 --d2 0 --d4 6 --seed 2 --with-llr 0"),
    "c1b_saturating": ("c1b_reg36_n1000", "--snr 2.0 --frames 8 --max-iter 30 --d4 14 --llr-scale 3 --seed 3"),
    "c1b_nopsc": ("c1b_reg36_n1000", "--snr 3.0 --frames 6 --max-iter 8 --psc 0 --pisc 0 --seed 4 --with-llr 0"),
    "c1b_highsnr_pisc": ("c1b_reg36_n1000", "--snr 9.0 --frames 8 --max-iter 30 --seed 5 --with-llr 0"),
    "c2_default": ("c2_irr_n500", "--snr 1.0 --frames 12 --max-iter 30 --seed 6 --with-llr 0"),
    "c2_d4_16_nopsc": ("c2_irr_n500", "--snr 3.0 --frames 8 --max-iter 30 --d4 16 --llr-scale 4 --psc 0 --seed 7 --with-llr 0"),
    "c2_coarse_table": ("c2_irr_n500", "--snr 1.5 --frames 8 --max-iter 40 --d1 8 --d2 64 --d3 4 --d4 12 --seed 8"),
    "c4_dc32": ("c4_reg632_high", "--snr 4.0 --frames 6 --max-iter 30 --seed 9 --with-llr 0"),
    "c3_n10000": ("c3_irr_n10000", "--snr 1.2 --frames 3 --max-iter 50 --with-llr 0 --seed 10"),
    "c3_n10000_lowsnr": ("c3_irr_n10000", "--snr 0.85 --frames 3 --max-iter 50 --with-llr 0 --seed 16"),
    "c5_n64800": ("c5_n64800", "--snr 1.2 --frames 1 --max-iter 50 --with-llr 0 --seed 11"),
    "synth_mixed": ("synth:mixed", "--snr 2.0 --frames 10 --max-iter 20 --seed 12"),
    "synth_mixed_minsum": ("synth:mixed", "--snr 3.0 --frames 8 --max-iter 20 --d2 0 --seed 13 --with-llr 0"),
    "synth_wide_dc48": ("synth:wide", "--snr 5.0 --frames 8 --max-iter 20 --seed 14 --with-llr 0"),
}


def main():
    os.makedirs(OUT, exist_ok=True)
    manifest = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, (graph, opts) in CASES.items():
            alist = os.path.join(tmp, name + ".alist")
            if graph.startswith("synth:"):
                N, rows = _alist.synth_rows(graph[6:])
                _alist.write_alist_rows(N, rows, alist)
            else:
                _alist.write_alist(graph, alist)
            fix = os.path.join(tmp, name + ".fix")
            subprocess.check_call([REF, "bpfixture", alist, fix] + opts.split())
            raw = open(fix, "rb").read()
            with lzma.open(os.path.join(OUT, name + ".fix.xz"), "wb", preset=9 | lzma.PRESET_EXTREME) as f:
                f.write(raw)
            manifest[name] = dict(graph=graph, args=opts, sha256=hashlib.sha256(raw).hexdigest(), bytes=len(raw))
            print(name, len(raw), os.path.getsize(os.path.join(OUT, name + ".fix.xz")))
        # bp_codec.it as the reference writes it (LDPC_Code::save_code): hashes for the writer test, and one file with a
        # generator (+ three (info, codeword) pairs of its encoder) for the reader test
        codec = {}
        for cfg in ("c1b_reg36_n1000", "c2_irr_n500", "c4_reg632_high", "c3_irr_n10000"):
            alist = os.path.join(tmp, cfg + ".alist")
            _alist.write_alist(cfg, alist)
            out = os.path.join(tmp, cfg + ".bp_codec.it")
            subprocess.check_call([REF, "bpcodec", alist, out], stdout=subprocess.DEVNULL)
            codec[cfg] = hashlib.sha256(open(out, "rb").read()).hexdigest()
        alist = os.path.join(tmp, "c2_irr_n500.alist")
        out, pairs = os.path.join(tmp, "g.it"), os.path.join(OUT, "c2_gen.pairs.bin")
        subprocess.check_call([REF, "bpcodec", alist, out, "--gen", "1", "--pairs", pairs, "--n", "3", "--seed", "21"],
                              stdout=subprocess.DEVNULL)
        raw = open(out, "rb").read()
        with lzma.open(os.path.join(OUT, "c2_gen.bp_codec.it.xz"), "wb", preset=9 | lzma.PRESET_EXTREME) as f:
            f.write(raw)
        codec["c2_gen"] = hashlib.sha256(raw).hexdigest()
        json.dump(codec, open(os.path.join(HERE, "bp_codec_sha256.json"), "w"), indent=1, sort_keys=True)
        # frames through the decoder the reference builds from that file: after the generator's column permutation the
        # lists of H are no longer sorted, and the order of a node's inputs is the order of its operations
        name, opts = "c2_gen_stored_order", "--snr 1.2 --frames 10 --max-iter 30 --with-llr 0 --seed 15"
        fix = os.path.join(tmp, name + ".fix")
        subprocess.check_call([REF, "bpfixture", out, fix] + opts.split())
        raw = open(fix, "rb").read()
        with lzma.open(os.path.join(OUT, name + ".fix.xz"), "wb", preset=9 | lzma.PRESET_EXTREME) as f:
            f.write(raw)
        manifest[name] = dict(graph="file:c2_gen.bp_codec.it", args=opts, sha256=hashlib.sha256(raw).hexdigest(), bytes=len(raw))
    json.dump(manifest, open(os.path.join(HERE, "bp_manifest.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
