#!/usr/bin/env python3
"""Generate the golden decoder bundles and fixtures under tests/golden/ by RUNNING THE REFERENCE.

Runs only in the build container (needs /root/reference and oracle/_ref/lutldpc_ref, the unmodified
reference decoder compiled by oracle/Makefile).  Everything it writes is an OUTPUT of the reference:
  <cfg>.codec.it.xz    (bench configs only) the codec file written by the reference's save_code
  <cfg>.bundle.xz      flat decoder bundle (graph arrays, quantiser thresholds, the reference's own
                       text serialisation of its designed LUT trees)         -> oracle/ref_driver.cpp
  <cfg>.<tag>.fix.xz   frames pushed through the reference's lut_decode: quantised inputs, hard
                       decisions, return values, per-iteration message dumps (output_verbosity=3)
  manifest.json        the exact command lines, sizes and sha256 of the uncompressed files

usage: python tests/golden/make_golden.py [cfg ...]     (default: all; c5 takes ~10 min)
"""
import hashlib
import json
import lzma
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "lutldpc_ref")
REF = os.environ.get("LUTLDPC_REFERENCE", "/root/reference")
CODES = os.path.join(REF, "codes")
TREES = os.path.join(REF, "trees")

IRR = "rate0.50_dv02-17_dc08-09_lut_q4"
# name -> (alist, design args, [(tag, fixture args)])
CONFIGS = {
    # BASELINE.json configs[0] as described: regular (3,6) N=1000, q=4, 50 iterations
    "c1b_reg36_n1000": ("rate0.50_dv03_dc06_N1000", "--design-thr 0.88 --max-iter 50 --qcha 4 --qmsg 4", [
        ("snr1.5", "--frames 8 --trace 2 --snr 1.5 --seed 1"),
        ("snr3.0", "--frames 16 --trace 1 --snr 3.0 --seed 2"),
        ("nopsc", "--frames 4 --trace 1 --snr 2.0 --seed 3 --psc 0 --pisc 0"),
        ("snr9", "--frames 8 --trace 0 --snr 9.0 --seed 4"),          # initial syndrome check fires
    ]),
    # params/ber.ini.regular.example as shipped: (6,32) N=2048, file trees 6_32_wide, q 4/3, QCHA, 8 its
    "c1a_reg632_wide": ("rate0.84_reg_v6c32_N2048",
                        "--design-snr 3.9 --max-iter 8 --qcha 4 --qmsg 3 --init-mode qcha --gen "
                        "--tree filename=%s/6_32_wide.ini" % TREES, [
        ("snr4.0", "--frames 8 --trace 2 --snr 4.0 --seed 1"),
        ("snr5.5", "--frames 8 --trace 1 --snr 5.5 --seed 2"),
    ]),
    # params/ber.ini.irregular.example: irregular N=500, q 4/4, random codewords
    "c2_irr_n500": (IRR + "_N500", "--design-thr 0.88 --max-iter 50 --qcha 4 --qmsg 4 --gen", [
        ("snr1.0", "--frames 8 --trace 2 --snr 1.0 --seed 1"),
        ("snr2.5", "--frames 16 --trace 1 --snr 2.5 --seed 2"),
    ]),
    # configs[2]: irregular N=10000
    "c3_irr_n10000": (IRR + "_N10000", "--design-thr 0.88 --max-iter 50 --qcha 4 --qmsg 4", [
        ("snr1.3", "--frames 4 --trace 0 --snr 1.3 --seed 1 --with-llr 0"),
        # 16 frames picked on the GPU for their outcome (2 x -50, 2 x +50, late, early at two SNRs), decoded by the
        # reference from the committed LLR input; per-iteration dumps of one failing and one converging frame
        ("sel", "--trace 2 --with-llr 0 --llr-from c3_irr_n10000.sel.llr_i16"),
    ]),
    # configs[3]: (6,32) N=2048 with trees/6_32_high.ini (single 6-input root LUT); q_msg 2 is the
    # largest resolution the reference's own design step can build
    "c4_reg632_high": ("rate0.84_reg_v6c32_N2048",
                       "--design-snr 3.9 --max-iter 8 --qcha 4 --qmsg 2 --tree filename=%s/6_32_high.ini" % TREES, [
        ("snr4.5", "--frames 8 --trace 2 --snr 4.5 --seed 1"),
    ]),
    # configs[4]: LUT-optimised N=64800 code
    "c5_n64800": ("rate0.50_dv02-08_dc07-08_lut_q4_N64800", "--design-thr 0.88 --max-iter 50 --qcha 4 --qmsg 4", [
        ("snr1.3", "--frames 2 --trace 0 --snr 1.3 --seed 1 --with-llr 0"),
        # 20 frames picked on the GPU for their outcome (4 x -50, 4 x +50, late, early, typical at 1.15 and 1.3 dB),
        # decoded by the reference from the committed LLR input; the per-iteration dumps of one failing and one
        # converging frame (100 + ~70 records of 239023 messages) are stored as SHA-256 digests (LUTFIX02)
        ("sel", "--trace 2 --with-llr 0 --llr-from c5_n64800.sel.llr_i16 --hash-traces"),
    ]),
    # per-iteration message resolution (LUT.qbits_messages) + LUT reuse schedule (LUT.reuse_lut)
    "x_varq_reuse": ("rate0.50_dv03_dc06_N1000",
                     "--design-thr 0.80 --max-iter 10 --qcha 4 --qmsg 3 --qmsg-list 4,4,4,3,3,3,3,2,2,2 "
                     "--reuse 0,1,0,0,1,1,0,0,1,0", [
        ("snr2.5", "--frames 8 --trace 2 --snr 2.5 --seed 1"),
    ]),
    # LUT check nodes (LUT.min_lut=false)
    "x_lutcn_irr_n500": (IRR + "_N500", "--design-thr 0.85 --max-iter 12 --qcha 4 --qmsg 3 --min-lut 0", [
        ("snr2.0", "--frames 8 --trace 2 --snr 2.0 --seed 1"),
    ]),
    # 3-bit channel / QCHA initial messages with Nq_Cha != Nq_Msg on an irregular code
    "x_qcha_irr_n1000": (IRR + "_N1000", "--design-thr 0.85 --max-iter 20 --qcha 4 --qmsg 3 --init-mode qcha", [
        ("snr2.0", "--frames 8 --trace 1 --snr 2.0 --seed 1"),
    ]),
}


CODEC_FOR = ("c5_n64800", "c3_irr_n10000", "c1b_reg36_n1000")


def sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def xz(src, dst):
    with open(src, "rb") as f, lzma.open(dst, "wb", preset=6) as g:
        g.write(f.read())


def expand_llr_input(name, dst):
    """tests/golden/<name>.xz (int16 cell indices, see make_sel_inputs.py) -> float64 LLR file for --llr-from."""
    import numpy as np
    k = np.frombuffer(lzma.open(os.path.join(HERE, name + ".xz"), "rb").read(), "<i2")
    ((k.astype(np.float64) + 0.5) * 0.01).astype("<f8").tofile(dst)


def hash_traces(path):
    """LUTFIX01 -> LUTFIX02: every trace record's message array is replaced by its SHA-256 digest (32 bytes)."""
    import struct
    b = open(path, "rb").read()
    assert b[:8] == b"LUTFIX01"
    n, e, nf, nt, with_llr, zero_cw = struct.unpack("<6i", b[8:32])
    off = 40 + nf * ((8 * n if with_llr else 0) + 4 * n + 4)
    out = [b"LUTFIX02", b[8:off]]
    for _ in range(nt):
        (nr,) = struct.unpack_from("<i", b, off)
        out.append(b[off:off + 4])
        off += 4
        for _ in range(nr):
            out.append(b[off:off + 8])
            out.append(hashlib.sha256(b[off + 8:off + 8 + e]).digest())
            off += 8 + e
    assert off == len(b)
    open(path, "wb").write(b"".join(out))


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    return subprocess.run(cmd, check=True, capture_output=True, text=True).stdout


def main():
    names = sys.argv[1:] or list(CONFIGS)
    mpath = os.path.join(HERE, "manifest.json")
    manifest = json.load(open(mpath)) if os.path.exists(mpath) else {}
    for name in names:
        alist, dargs, fixtures = CONFIGS[name]
        apath = os.path.join(CODES, alist + ".alist")
        qcha = "--init-mode qcha" in dargs
        with tempfile.TemporaryDirectory() as tmp:
            pre = os.path.join(tmp, name)
            # LUTLDPC_PREDESIGNED=<prefix>: reuse <prefix>.it/.bundle/.log written earlier by the same
            # `design` command (the N=64800 design spends 8 min in the reference's dense rank)
            predesigned = os.environ.get("LUTLDPC_PREDESIGNED")
            only = os.environ.get("LUTLDPC_FIXTURES_ONLY")
            if only and name in CODEC_FOR and name in manifest:
                # new fixtures for a config whose committed codec file stays: LUTLDPC_FIXTURES_ONLY=tag[,tag]
                codec = os.path.join(tmp, name + ".it")
                with open(codec, "wb") as f:
                    f.write(lzma.open(os.path.join(HERE, name + ".codec.it.xz"), "rb").read())
                assert sha(codec) == manifest[name]["codec_sha256"]
                for tag, fargs in fixtures:
                    if tag not in only.split(","):
                        continue
                    fx = "%s.%s.fix" % (pre, tag)
                    fl = fargs.split()
                    hashed = "--hash-traces" in fl
                    if hashed:
                        fl.remove("--hash-traces")
                    if "--llr-from" in fl:
                        i = fl.index("--llr-from")
                        tmp_llr = os.path.join(tmp, "llr_in.f64")
                        expand_llr_input(fl[i + 1], tmp_llr)
                        fl[i + 1] = tmp_llr
                    run([REF_BIN, "fixture", codec, fx] + fl)
                    if hashed:
                        hash_traces(fx)
                    manifest[name]["fixtures"][tag] = {"args": fargs, "sha256": sha(fx), "bytes": os.path.getsize(fx)}
                    xz(fx, os.path.join(HERE, "%s.%s.fix.xz" % (name, tag)))
                json.dump(manifest, open(mpath, "w"), indent=1, sort_keys=True)
                continue
            if predesigned and len(names) == 1:
                pre = predesigned
                info = open(pre + ".log").read().split("\n")[0]
            else:
                info = run([REF_BIN, "design", apath, pre] + dargs.split())
            entry = {"alist": alist, "design": "lutldpc_ref design <codes>/%s.alist <out> %s" % (alist, dargs),
                     "design_report": json.loads(info.strip().splitlines()[-1]),
                     "bundle_sha256": sha(pre + ".bundle"), "bundle_bytes": os.path.getsize(pre + ".bundle"),
                     "init_mode": "qcha" if qcha else "cont", "fixtures": {}}
            xz(pre + ".bundle", os.path.join(HERE, name + ".bundle.xz"))
            if name in CODEC_FOR:  # the reference's own codec file (save_code), for the timed CPU baseline
                xz(pre + ".it", os.path.join(HERE, name + ".codec.it.xz"))
                entry["codec_sha256"] = sha(pre + ".it")
            # with a generator the reference cannot re-load its own codec file (ref_driver.cpp note),
            # so those fixtures rebuild the decoder from the alist with the same design arguments
            src = [apath] + dargs.split() if "--gen" in dargs else [pre + ".it"]
            for tag, fargs in fixtures:
                fx = "%s.%s.fix" % (pre, tag)
                fl = fargs.split()
                hashed = "--hash-traces" in fl
                if hashed:
                    fl.remove("--hash-traces")
                if "--llr-from" in fl:
                    i = fl.index("--llr-from")
                    tmp_llr = os.path.join(tmp, "llr_in.f64")
                    expand_llr_input(fl[i + 1], tmp_llr)
                    fl[i + 1] = tmp_llr
                run([REF_BIN, "fixture", src[0], fx] + src[1:] + fl +
                    (["--init-mode", "qcha"] if qcha and "--gen" not in dargs else []))
                if hashed:
                    hash_traces(fx)
                entry["fixtures"][tag] = {"args": fargs, "sha256": sha(fx), "bytes": os.path.getsize(fx)}
                xz(fx, os.path.join(HERE, "%s.%s.fix.xz" % (name, tag)))
            manifest[name] = entry
        json.dump(manifest, open(mpath, "w"), indent=1, sort_keys=True)
    print("wrote", mpath)


if __name__ == "__main__":
    main()
