"""Fused Monte-Carlo pipeline (Philox AWGN -> LLR -> quantise -> decode -> count) against the reference's
own ber_sim frame loop (tests/golden/ber_reference.json, made by tests/golden/make_ber_reference.py from the compiled
reference).  What is asserted per code, over every SNR point of the reference grid: each FER point lies inside the
99.9 % two-sample interval (hard bound), at most ONE point per code lies outside the 95 % interval (one in twenty does
by chance), the BER difference is inside a 99.9 % bound built from the burst statistics, and the uncoded BER matches
Q(sqrt(2 R Eb/N0)).  Plus internal consistency of the counters, the multi-GPU harness and the reference-shaped
command line."""
import json
import math
import os
import subprocess

import numpy as np
import pytest

import _golden

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = json.load(open(os.path.join(_golden.GOLD, "ber_reference.json")))


@pytest.fixture(scope="module")
def L():
    import lut_ldpc_b200
    return lut_ldpc_b200


def _agree(k1, n1, k2, n2, z=1.96):
    """Two binomial proportions agree at the 95 % level (pooled two-sample z test; Wilson-style slack when
    both counts are tiny)."""
    if n1 == 0 or n2 == 0:
        return True
    p1, p2 = k1 / n1, k2 / n2
    pool = (k1 + k2) / (n1 + n2)
    se = math.sqrt(max(pool * (1 - pool), 1e-300) * (1 / n1 + 1 / n2))
    return abs(p1 - p2) <= z * se + 3.0 / min(n1, n2)


N_GPU = {"c1b_reg36_n1000": 40000, "c2_irr_n500": 40000, "c3_irr_n10000": 8192, "c5_n64800": 4096,
         "x_qcha_irr_n1000": 40000, "c1a_reg632_wide": 40000, "c4_reg632_high": 40000}


@pytest.mark.parametrize("cfg", sorted(REF))
def test_ber_fer_within_confidence_interval(L, cfg):
    """Every SNR point of the reference grids (c2: 0:.5:4 of params/ber.ini.irregular.example; c1b: 1:.5:4; QCHA
    initial messages; file trees with random codewords as params/ber.ini.regular.example; N=10000 and N=64800)."""
    dec = L.Decoder(_golden.bundle_path(cfg))
    dec.set_exit_conditions(dec.max_iters, True, True)
    dec.set_initial_message_mode(L.INIT_QCHA if REF[cfg]["_init_mode"] == "qcha" else L.INIT_CONT)
    zero_cw = REF[cfg]["_zero_codeword"]
    n_gpu = N_GPU[cfg]
    inside95 = []
    for snr_s, r in sorted(REF[cfg].items()):
        if snr_s.startswith("_"):
            continue
        c = dec.simulate_snr(float(snr_s), n_gpu, seed=77, zero_codeword=zero_cw)
        assert c.frames == n_gpu and c.data_bits == n_gpu * dec.ninfo and c.uncoded_bits == n_gpu * dec.nvar
        # frame errors: binomial over frames.  Every point must sit inside the 99.9 % interval (a hard bound
        # that a correct implementation violates once in a thousand points), and the 95 % interval is
        # tallied below (one point in twenty falls outside it by chance).
        assert _agree(c.frame_errors, c.frames, r["frame_errors"], r["frames"], z=3.29), (cfg, snr_s, c.as_dict(), r)
        inside95.append(_agree(c.frame_errors, c.frames, r["frame_errors"], r["frames"]))
        # bit errors come in bursts (one frame error = many bit errors, with a heavy-tailed count per frame):
        # per-frame bit errors X = B * [frame error] with B roughly exponential around the mean burst b, so
        # Var(X) ~ p (2 b^2) - (p b)^2; the BER difference is tested against that dispersion at the 99.9 % level
        ber_g, ber_r = c.data_bit_errors / c.data_bits, r["data_bit_errors"] / r["data_bits"]
        fe = max(1, c.frame_errors + r["frame_errors"])
        burst = (c.data_bit_errors + r["data_bit_errors"]) / fe / dec.ninfo
        pool = (c.frame_errors + r["frame_errors"]) / (c.frames + r["frames"])
        se = burst * math.sqrt(max(2 * pool - pool * pool, 1e-12) * (1 / c.frames + 1 / r["frames"]))
        assert abs(ber_g - ber_r) <= 3.29 * se + 3.0 * burst / min(c.frames, r["frames"]) + 1e-12, (cfg, snr_s, ber_g, ber_r)
        # uncoded BER = Q(sqrt(2 R Eb/N0)): a direct check of sigma and the slicer
        ub = c.uncoded_bit_errors / c.uncoded_bits
        q = 0.5 * math.erfc(math.sqrt(dec.rate * 10 ** (float(snr_s) / 10)))
        assert abs(ub - q) < 4 * math.sqrt(q * (1 - q) / c.uncoded_bits) + 1e-9
        # mean iterations track the reference too
        assert abs(c.sum_iters / c.frames - r["sum_iters"] / r["frames"]) < 0.1 * dec.max_iters
    assert sum(inside95) >= len(inside95) - 1, inside95


def test_zero_and_random_codewords_agree(L):
    """Linear code + symmetric channel + symmetric decoder: error rates do not depend on the codeword."""
    dec = L.Decoder(_golden.bundle_path("c2_irr_n500"))
    dec.set_exit_conditions(dec.max_iters, True, True)
    a = dec.simulate_snr(2.0, 30000, seed=5, zero_codeword=True)
    b = dec.simulate_snr(2.0, 30000, seed=6, zero_codeword=False)
    assert _agree(a.frame_errors, a.frames, b.frame_errors, b.frames)
    assert a.frame_errors > 100


def test_counters_are_reproducible_and_split_invariant(L):
    """Frames are Philox streams keyed by the global frame index: any split into calls gives the same sums."""
    dec = L.Decoder(_golden.bundle_path("c2_irr_n500"))
    dec.set_exit_conditions(dec.max_iters, True, True)
    whole = dec.simulate_snr(1.5, 3000, seed=9, first_frame=100, zero_codeword=False).as_array()
    acc = L.SimCounters()
    dec.simulate_snr(1.5, 1000, seed=9, first_frame=100, zero_codeword=False, acc=acc, batch_frames=256)
    dec.simulate_snr(1.5, 2000, seed=9, first_frame=1100, zero_codeword=False, acc=acc)
    assert np.array_equal(whole, acc.as_array())
    assert int(np.sum(whole[8:])) == 3000  # iteration histogram covers every frame


def test_simulate_matches_decode_of_same_frames(L):
    """Counter sanity of the fused path.  BLERC counts blocks of the K data bits (src/LDPC_BER_Sim.cpp:259,
    :291), so a decode failure (negative lut_decode return) whose wrong bits are all parity bits is not a
    frame error: the two counters are not ordered."""
    dec = L.Decoder(_golden.bundle_path("c1b_reg36_n1000"))
    dec.set_exit_conditions(dec.max_iters, True, True)
    c = dec.simulate_snr(2.0, 5000, seed=3, zero_codeword=True)
    assert 0 < c.decode_failures <= c.frames and 0 < c.frame_errors <= c.frames
    assert c.data_bit_errors >= c.frame_errors
    assert c.frames <= c.sum_iters <= c.frames * dec.max_iters


def test_ber_sim_command_line(L, tmp_path):
    exe = os.path.join(ROOT, "lut_ldpc_b200", "ber_sim")
    if not os.path.exists(exe):
        pytest.skip("ber_sim binary not built")
    ini = tmp_path / "ber.ini"
    ini.write_text("[Sim]\nSNRdB = 1:1:4\nNframes = 4000\nNfers = 100\ncodec_filename = %s\n[LDPC]\nzero_codeword = false\n"
                   "[LUT]\nmax_iter = 50\n" % _golden.bundle_path("c2_irr_n500"))
    out = subprocess.run([exe, "-p", str(ini), "-b", str(tmp_path), "-s", "3"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    lines = [l for l in out.stdout.splitlines() if l.startswith("SNR = ")]
    assert lines and "Data BER:" in lines[0] and "FER:" in lines[0]
    res = [os.path.join(dp, f) for dp, _, fs in os.walk(tmp_path / "results") for f in fs if f.endswith(".json")]
    assert len(res) == 1 and "_rseed0003" in res[0]
    r = json.load(open(res[0]))
    assert r["sim_SNRdB"] == [1, 2, 3, 4] and r["ldpc_nvar"] == 500
    assert r["sim_frame_errors"][0] > 100 and r["sim_Nframes"][0] < 4000 + 8192  # Nfers stop
    # error rates fall with SNR; once FER < fer_min the sweep stops and later points stay empty
    assert r["sim_frame_errors"][1] < r["sim_frame_errors"][0]
    # the IT++ results file (LDPC_BER_Sim_Results::write_itfile) holds the same numbers: read back with IT++'s own
    # it_ifile through the compiled reference when it is there, and with a plain parser of the v3 layout always
    itf = res[0][:-5] + ".it"
    assert os.path.exists(itf)
    got = _read_itfile(itf)
    for k in ("sim_SNRdB", "sim_Nframes", "sim_Ndatabits", "sim_frame_errors", "sim_data_bit_errors", "sim_uncoded_bit_errors"):
        assert got[k] == [float(x) for x in r[k]], k
    assert got["ldpc_nvar"] == [500.0] and got["ldpc_nchk"] == [250.0] and isinstance(got["gitversion"], str)
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "lutldpc_ref")
    if os.path.exists(ref_bin):
        o = subprocess.run([ref_bin, "results", itf], capture_output=True, text=True, timeout=60)
        assert o.returncode == 0, o.stderr
        viaitpp = json.loads(o.stdout.strip().splitlines()[-1])
        for k in ("sim_SNRdB", "sim_Nframes", "sim_frame_errors", "sim_data_bit_errors", "ldpc_code_rate"):
            assert viaitpp[k] == got[k], k


def _read_itfile(path):
    """IT++ it_file v3: 'IT++' 3, then blocks u64 hdr, data, block; name, type, desc (NUL terminated); data."""
    import struct
    b = open(path, "rb").read()
    assert b[:5] == b"IT++\x03"
    off, out = 5, {}
    while off < len(b):
        hdr, data, block = struct.unpack("<QQQ", b[off:off + 24])
        name, typ, _ = b[off + 24:off + hdr].split(b"\0")[:3]
        d = b[off + hdr:off + hdr + data]
        if typ == b"dvec":
            (n,) = struct.unpack("<Q", d[:8])
            out[name.decode()] = list(struct.unpack("<%dd" % n, d[8:8 + 8 * n]))
        elif typ == b"float64":
            out[name.decode()] = struct.unpack("<d", d)[0]
        elif typ == b"string":
            (n,) = struct.unpack("<Q", d[:8])
            out[name.decode()] = d[8:8 + n].decode()
        off += block
    return out


def test_python_harness_single_gpu(L):
    from lut_ldpc_b200 import ber_sim as bs
    p = bs.SimParams(None, SNRdB=[1.0, 2.0], Nframes=6000, Nfers=10 ** 9, codec_filename=_golden.bundle_path("c2_irr_n500"),
                     zero_codeword=False, max_iter=50)
    sim = bs.LDPC_BER_Sim_LUT(p, rand_seed=1, chunk_frames=2048).load()
    r = sim.run(verbose=False)
    assert r["sim_Nframes"] == [6000, 6000]
    ref = REF["c2_irr_n500"]
    for i, s in enumerate(("1.00", "2.00")):
        # 99.9 % bound: a fixed-seed check of the harness plumbing, not a second statistical study (that is
        # test_ber_fer_within_confidence_interval)
        assert _agree(r["sim_frame_errors"][i], 6000, ref[s]["frame_errors"], ref[s]["frames"], z=3.29)


def test_multi_gpu_harness_is_split_invariant(L):
    """The real multi-GPU Monte-Carlo: ber_sim.py under torchrun with one rank per GPU (NCCL all-reduce of the counters
    after every chunk, global stop rules) gives exactly the counts of the single-process run -- frames are Philox
    streams keyed by their global index.  Needs two visible devices."""
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    worker = os.path.join(ROOT, "tests", "ber_multi_worker.py")
    args = ["c2_irr_n500", "6000", "1000000000", "1024", "1.0,2.0"]
    one = subprocess.run([sys.executable, worker] + args, capture_output=True, text=True, timeout=600)
    assert one.returncode == 0, one.stderr[-2000:]
    two = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29531", worker] + args,
                         capture_output=True, text=True, timeout=900)
    assert two.returncode == 0, two.stderr[-2000:]
    r1 = json.loads(one.stdout.strip().splitlines()[-1])
    r2 = json.loads([l for l in two.stdout.splitlines() if l.startswith("{")][-1])
    assert r2["world"] == 2 and r2["allreduces"] >= 2 * 3  # 6000 frames in chunks of 2 x 1024 per SNR point
    for k in ("sim_Nframes", "sim_frame_errors", "sim_data_bit_errors", "sim_uncoded_bit_errors", "sim_sum_iterations"):
        assert r1[k] == r2[k], k
    assert r1["sim_frame_errors"][0] > 100


def test_ber_sim_designs_its_decoder_like_the_shipped_examples(L, tmp_path):
    """ber_sim without a codec file, as the reference ships its examples: params/ber.ini.irregular.example (alist +
    design_thr, auto trees, random codewords) and params/ber.ini.regular.example ((6,32) code, tree template file, 4/3
    bits, QCHA initial messages, design_SNRdB).  The alist import, the LUT design (host) and the systematic generator
    (device) all run inside the binary; FER must agree with the reference's own runs of the same configurations."""
    from _alist import write_alist
    exe = os.path.join(ROOT, "lut_ldpc_b200", "ber_sim")
    if not os.path.exists(exe):
        pytest.skip("ber_sim binary not built")
    (tmp_path / "codes").mkdir()
    (tmp_path / "trees").mkdir()
    write_alist("c2_irr_n500", str(tmp_path / "codes" / "rate0.50_dv02-17_dc08-09_lut_q4_N500.alist"))
    write_alist("c4_reg632_high", str(tmp_path / "codes" / "rate0.84_reg_v6c32_N2048.alist"))
    (tmp_path / "trees" / "6_32_wide.ini").write_text(
        "[var_iter_000]\n   var_deg_006 = riiim/m//im/m///m//c/\n\n[DT]\n   var_deg_006 = rim/m/m//im/m/m//c/\n")
    irr = tmp_path / "ber.ini.irregular"
    irr.write_text("[Sim]\nSNRdB = 1:1:2\nNframes = 20000\nNfers = 1000000\nfer_min = 0\nber_min = 0\n"
                   "[LDPC]\nparity_filename = rate0.50_dv02-17_dc08-09_lut_q4_N500\nzero_codeword = false\n"
                   "[LUT]\nmax_iter = 50\ndesign_thr = 0.88\nqbits_channel = 4\nqbits_message_uniform = 4\n")
    reg = tmp_path / "ber.ini.regular"
    reg.write_text("[Sim]\nSNRdB = 4:.5:4.5\nNframes = 20000\nNfers = 1000000\nfer_min = 0\nber_min = 0\n"
                   "[LDPC]\nparity_filename = rate0.84_reg_v6c32_N2048\nzero_codeword = false\n"
                   "[LUT]\nmax_iter = 8\ndesign_SNRdB = 3.9\nqbits_channel = 4\nqbits_message_uniform = 3\n"
                   "tree_mode = file\ntrees_filename = 6_32_wide.ini\n"
                   "initial_message_mode = from_quantized_channel_llrs\n")
    for ini, cfg, snrs in ((irr, "c2_irr_n500", ("1.00", "2.00")), (reg, "c1a_reg632_wide", ("4.00", "4.50"))):
        out = subprocess.run([exe, "-p", str(ini), "-b", str(tmp_path), "-s", "1"], capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, (out.stdout[-1000:], out.stderr[-1000:])
        assert "Designed LUT decoder" in out.stdout
        lines = [l for l in out.stdout.splitlines() if l.startswith("SNR = ")]
        assert len(lines) == 2
        for line, s in zip(lines, snrs):
            frames = int(line.split("Simulated ")[1].split(" frames")[0])
            fer = float(line.split("FER: ")[1])
            ref = REF[cfg][s]
            assert frames == 20000
            assert _agree(round(fer * frames), frames, ref["frame_errors"], ref["frames"], z=3.29), (cfg, s, fer, ref)


def test_stimuli_pairs_of_the_fused_step(L, tmp_path):
    """output_verbosity > 0: the reference prints, per frame, the quantised channel labels and the decoded word
    ("Stimuli Pair", src/LDPC_Code_LUT.cpp:228-238, the golden vectors of its VHDL flow).  The fused step hands them back
    (lutldpc_simulate_snr_stimuli): decoding the returned labels again gives the returned words, the counters are those
    of the same frames, and ber_sim prints the pairs in the reference's format."""
    dec = L.Decoder(_golden.bundle_path("c2_irr_n500"))
    dec.set_exit_conditions(dec.max_iters, True, True)
    acc, cha, out = dec.simulate_snr_stimuli(1.5, 700, seed=12, first_frame=40, zero_codeword=True)
    plain = dec.simulate_snr(1.5, 700, seed=12, first_frame=40, zero_codeword=True)
    assert np.array_equal(acc.as_array(), plain.as_array())
    assert cha.max() < 16 and set(np.unique(out)) <= {0, 1}
    # this decoder's channel and message quantisers coincide (Nq_Cha = Nq_Msg(0) = 16): llr_in_msg = llr_in_cha
    out2, it2 = dec.decode_quantized(cha, cha)
    assert np.array_equal(out2, out)
    assert int((out[:, :dec.ninfo] != 0).any(axis=1).sum()) == acc.frame_errors
    assert int(out[:, :dec.ninfo].sum()) == acc.data_bit_errors and int(np.abs(it2).sum()) == acc.sum_iters
    exe = os.path.join(ROOT, "lut_ldpc_b200", "ber_sim")
    if not os.path.exists(exe):
        pytest.skip("ber_sim binary not built")
    ini = tmp_path / "v.ini"
    ini.write_text("[Sim]\nSNRdB = 2\nNframes = 20\nNfers = 20\nfer_min = 0\nber_min = 0\ncodec_filename = %s\n"
                   "[LDPC]\nzero_codeword = true\n[LUT]\nmax_iter = 50\noutput_verbosity = 1\n" % _golden.bundle_path("c2_irr_n500"))
    o = subprocess.run([exe, "-p", str(ini), "-b", str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert o.returncode == 0, o.stderr[-1000:]
    lines = o.stdout.splitlines()
    heads = [i for i, l in enumerate(lines) if l.startswith("Stimuli Pair (Quantized channel LLR decoder inputs in hex format")]
    assert len(heads) == 20
    labels = lines[heads[0] + 1].split()
    bits = lines[heads[0] + 2].split()
    assert len(labels) == 500 and len(bits) == 500 and all(len(x) == 8 and int(x, 16) < 16 for x in labels)
    assert set(bits) <= {"0", "1"}
    o1, _ = dec.decode_quantized(np.array([[int(x, 16) for x in labels]], np.uint8), np.array([[int(x, 16) for x in labels]], np.uint8))
    assert [str(b) for b in o1[0]] == bits


def test_designed_decoder_is_saved_as_the_reference_codec_file(L, tmp_path):
    """Sim.save_codec (src/LDPC_BER_Sim.cpp:59,511-520): the harness that designs its decoder from an alist leaves
    results/<name>/lut_codec.it when the seed matches; the file is the reference's format (a handle created from it
    decodes the reference's fixtures bit-exactly, and it equals the conversion of the golden bundle)."""
    import pyoracle
    from _alist import write_alist
    from lut_ldpc_b200 import ber_sim as bs
    (tmp_path / "codes").mkdir()
    write_alist("c1b_reg36_n1000", str(tmp_path / "codes" / "reg36.alist"))
    ini = tmp_path / "p.ini"
    ini.write_text("[Sim]\nSNRdB = 3\nNframes = 2048\nsave_codec = 7\n[LDPC]\nparity_filename = reg36\n"
                   "[LUT]\nmax_iter = %d\ndesign_thr = %s\nqbits_channel = %d\nqbits_message_uniform = %d\n"
                   % _design_keys("c1b_reg36_n1000"))
    sim = bs.LDPC_BER_Sim_LUT(bs.SimParams(str(ini)), rand_seed=7, chunk_frames=2048).load(str(tmp_path))
    sim.run(verbose=False)
    saved = tmp_path / "results" / sim.gen_filename() / "lut_codec.it"
    assert saved.exists()
    dec = L.Decoder(str(saved))
    fx = pyoracle.read_fixture(_golden.fixture_path("c1b_reg36_n1000", "snr1.5"))
    dec.set_exit_conditions(dec.max_iters, True, True)
    bits, iters = dec.decode_quantized(fx["cha"], fx["msg"])
    assert np.array_equal(iters, fx["ret"]) and np.array_equal(bits, fx["out"])
    assert saved.read_bytes() == open(_golden.codec_path("c1b_reg36_n1000"), "rb").read()


def _design_keys(cfg):
    """(max_iter, design_thr, qbits_channel, qbits_message) of the reference design run recorded in the manifest"""
    a = _golden.MANIFEST[cfg]["design"].split()
    kv = {a[i]: a[i + 1] for i in range(len(a) - 1) if a[i].startswith("--")}
    return int(kv["--max-iter"]), kv["--design-thr"], int(kv["--qcha"]), int(kv["--qmsg"])
