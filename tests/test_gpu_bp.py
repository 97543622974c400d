"""BP comparison decoder (the [BP] section of ber_sim: IT++'s LDPC_Code::bp_decode, batched on the GPU) through the C
ABI: bit-exact against the reference's golden vectors (tests/golden/bp/) and against the C oracle on random batches
with frames that return at different iterations; the Monte-Carlo step against the reference's BER/FER points."""
import json
import math
import os

import numpy as np
import pytest

import _alist
import _bpgolden
import _golden
import pyoracle

pytestmark = pytest.mark.gpu
BER = json.load(open(os.path.join(_golden.GOLD, "bp_ber_reference.json")))


@pytest.fixture(scope="module")
def capi():
    from lut_ldpc_b200 import capi
    return capi


def _dec(capi, g, par, **kw):
    if "codec_file" in g:  # a stored bp_codec.it: the graph in the order the file holds it
        import tempfile
        with tempfile.NamedTemporaryFile(suffix=".it") as f:
            f.write(_bpgolden.stored_codec(g["codec_file"]))
            f.flush()
            return capi.BpDecoder.from_file(f.name, max_iters=par.max_iter, psc=par.psc, pisc=par.pisc, qllr_scale_res=par.d1,
                                            qllr_table_size=par.d2, qllr_spacing_res=par.d3, qllr_total_res=par.d4)
    return capi.BpDecoder(g, max_iters=par.max_iter, psc=par.psc, pisc=par.pisc, qllr_scale_res=par.d1,
                          qllr_table_size=par.d2, qllr_spacing_res=par.d3, qllr_total_res=par.d4, **kw)


@pytest.mark.parametrize("name", _bpgolden.CASES)
def test_bp_decode_matches_reference(capi, name):
    """bp_decode: soft output and return value of every golden frame; decode(): the bits ber_sim counts"""
    fx = _bpgolden.fixture(name)
    g = _bpgolden.graph(name)
    dec = _dec(capi, g, fx["par"])
    assert dec.nvar == fx["nvar"] and dec.ninfo == fx["ninfo"]
    out, ret = dec.decode_qllr(fx["qin"])
    assert np.array_equal(ret, fx["ret"])
    assert np.array_equal(out, fx["qout"])
    if fx["llr"] is not None:
        bits, ret2 = dec.decode_llr(fx["llr"])
        assert np.array_equal(ret2, fx["ret"])
        assert np.array_equal(bits[:, :fx["ninfo"]], fx["syst"])
        assert np.array_equal(bits, (fx["qout"] < 0).astype(np.uint8))


def _noisy_llr(g_nvar, n, snr_db, rate, seed):
    rng = np.random.default_rng(seed)
    N0 = 10 ** (-snr_db / 10) / rate
    x = 1.0 + math.sqrt(N0 / 2) * rng.standard_normal((n, g_nvar))
    return 4 * x / N0


@pytest.mark.parametrize("name,snr,kw", [
    ("c2_default", 1.6, {}),
    ("c2_default", 2.2, dict(qllr_table_size=0)),
    ("c1b_default", 1.8, {}),
    ("synth_mixed", 2.5, {}),
    ("synth_wide_dc48", 5.0, dict(qllr_scale_res=4, qllr_spacing_res=2, qllr_total_res=12)),
])
def test_bp_batch_vs_oracle_with_and_without_repacking(capi, name, snr, kw):
    """700 frames that return between iteration 0 and never: every frame equals the oracle's, whether the batch is
    narrowed while it runs or not (soft output: fixed slots; bits: repacked batches)"""
    g = _bpgolden.graph(name)
    par = pyoracle.BpParams(max_iter=25, d1=kw.get("qllr_scale_res", 12), d2=kw.get("qllr_table_size", 300),
                            d3=kw.get("qllr_spacing_res", 7), d4=kw.get("qllr_total_res", 28))
    dec = _dec(capi, g, par)
    llr = _noisy_llr(g["nvar"], 700, snr, dec.rate, seed=11)
    llr[5] = np.abs(llr[5])  # a frame whose initial syndrome check passes
    qin = pyoracle.bp_to_qllr(llr, par.d1, par.d4)
    want_out, want_ret = pyoracle.bp_decode(g["dv_vec"], g["dc_vec"], g["cn_msg_idx"], qin, par)
    assert want_ret[5] == 0 and (want_ret > 0).sum() > 100 and len(set(want_ret.tolist())) > 5
    out, ret = dec.decode_qllr(qin)
    assert np.array_equal(ret, want_ret) and np.array_equal(out, want_out)
    for on in (True, False):
        dec.set_repack(on)
        bits, ret2 = dec.decode_llr(llr)
        assert np.array_equal(ret2, want_ret)
        assert np.array_equal(bits, (want_out < 0).astype(np.uint8))


def test_bp_alist_reader_and_create_errors(capi, tmp_path):
    """the graph through lutldpc_alist_graph equals the one built in Python; degrees the reference refuses are refused"""
    N, rows = _alist.synth_rows("mixed")
    p = str(tmp_path / "mixed.alist")
    _alist.write_alist_rows(N, rows, p)
    ga, gp = capi.alist_graph(p), _alist.graph_from_rows(N, rows)
    for k in ("dv_vec", "dc_vec", "cn_msg_idx"):
        assert np.array_equal(ga[k], gp[k])
    bad = _alist.graph_from_rows(6, [[0, 1, 2], [3], [2, 4, 5]])
    with pytest.raises(capi.LutLdpcError, match=r"sumX2\(j\)=1"):
        capi.BpDecoder(bad)
    with pytest.raises(capi.LutLdpcError, match="Resolution too large"):
        capi.BpDecoder(gp, qllr_total_res=29)


def _agree(k1, n1, k2, n2, z=3.29):
    pool = (k1 + k2) / (n1 + n2)
    se = math.sqrt(max(pool * (1 - pool), 1e-300) * (1 / n1 + 1 / n2))
    return abs(k1 / n1 - k2 / n2) <= z * se + 3.0 / min(n1, n2)


@pytest.mark.parametrize("key", sorted(BER))
def test_bp_ber_fer_against_reference_points(capi, key):
    """FER inside the 99.9 % two-sample interval at every reference point; BER within a factor set by the burst
    statistics (errors come in frames); the slicer's error rate equals Q(sqrt(2 R Eb/N0))"""
    cfg = key.split(":")[0]
    g = _alist.bundle_graph(cfg)
    extra = BER[key]["_args"].split()
    kv = {extra[i][2:]: int(extra[i + 1]) for i in range(0, len(extra) - 1, 2)}
    dec = capi.BpDecoder(g, max_iters=kv.get("max-iter", 30), qllr_table_size=kv.get("d2", 300))
    n = 20000 if g["nvar"] <= 1000 else 2048
    for snr_s, r in sorted(BER[key].items()):
        if snr_s.startswith("_"):
            continue
        c = dec.simulate_snr(float(snr_s), n, seed=9)
        assert c.frames == n and c.data_bits == n * dec.ninfo and c.uncoded_bits == n * dec.nvar
        assert _agree(c.frame_errors, c.frames, r["frame_errors"], r["frames"]), (key, snr_s, c.frame_errors, r)
        if r["frame_errors"] >= 30:
            per_frame_ref = r["data_bit_errors"] / r["frame_errors"]
            per_frame = c.data_bit_errors / max(1, c.frame_errors)
            assert 0.6 < per_frame / per_frame_ref < 1.6, (key, snr_s, per_frame, per_frame_ref)
        p_unc = 0.5 * math.erfc(math.sqrt(dec.rate * 10 ** (float(snr_s) / 10)))
        assert abs(c.uncoded_bit_errors / c.uncoded_bits - p_unc) < 4 * math.sqrt(p_unc / c.uncoded_bits)
        assert c.sum_iters > 0 and sum(c.iter_hist) == n


def test_bp_simulate_invariances(capi):
    """the counters do not depend on repacking, on the batch size or on how the frame range is split (what the
    multi-GPU harness relies on), and the noise is the LUT decoder's for the same seed"""
    import lut_ldpc_b200 as L
    g = _alist.bundle_graph("c2_irr_n500")
    dec = capi.BpDecoder(g)
    a = dec.simulate_snr(2.0, 3000, seed=4).as_array()
    dec.set_repack(False)
    assert np.array_equal(dec.simulate_snr(2.0, 3000, seed=4).as_array(), a)
    dec.set_repack(True)
    assert np.array_equal(dec.simulate_snr(2.0, 3000, seed=4, batch_frames=512).as_array(), a)
    acc = dec.simulate_snr(2.0, 1000, seed=4)
    acc = dec.simulate_snr(2.0, 2000, seed=4, first_frame=1000, acc=acc)
    assert np.array_equal(acc.as_array(), a)
    lut = L.Decoder(_golden.bundle_path("c2_irr_n500"))
    assert lut.simulate_snr(2.0, 3000, seed=4).uncoded_bit_errors == a[5]


def test_bp_generator_and_random_codewords(capi):
    """LDPC_Generator_Systematic + LDPC_Code(H, G): codewords satisfy the permuted H, decode to themselves, and the
    Monte-Carlo step with random codewords sees the error rates of the all-zero one"""
    g = _alist.bundle_graph("c1b_reg36_n1000")
    dec = capi.BpDecoder(g, with_generator=True)
    assert dec.has_gen and dec.ninfo + dec.nchk_lin_indep == dec.nvar
    assert sorted(dec.col_perm.tolist()) == list(range(dec.nvar))
    rng = np.random.default_rng(3)
    info = rng.integers(0, 2, (300, dec.ninfo), dtype=np.uint8)
    cw = dec.encode(info)
    assert np.array_equal(cw[:, :dec.ninfo], info)
    # H of the permuted code: column v of the original is column col_perm[v]
    orc_rows = _alist.synth_rows  # noqa: F841 (rows of the bundle graph below)
    dv, idx = g["dv_vec"], g["cn_msg_idx"]
    edge_vn = np.repeat(np.arange(g["nvar"]), dv)
    cs = np.concatenate([[0], np.cumsum(g["dc_vec"])])
    for c in range(0, g["nchk"], 7):
        cols = dec.col_perm[edge_vn[idx[cs[c]:cs[c + 1]]]]
        assert not (cw[:, cols].sum(axis=1) & 1).any()
    llr = (1.0 - 2.0 * cw) * 6.0
    bits, ret = dec.decode_llr(llr)
    assert np.array_equal(bits, cw) and (ret == 0).all()
    z = dec.simulate_snr(1.8, 20000, seed=1, zero_codeword=True)
    r = dec.simulate_snr(1.8, 20000, seed=2, zero_codeword=False)
    assert _agree(z.frame_errors, z.frames, r.frame_errors, r.frames)
    assert abs(z.uncoded_bit_errors - r.uncoded_bit_errors) < 6 * math.sqrt(z.uncoded_bit_errors)


def test_ber_sim_with_a_bp_section(capi, tmp_path):
    """a params file with a [BP] section (and no [LUT]) through both harnesses -- the C++ ber_sim binary and
    ber_sim.py -- gives the reference's FER at its own points; random codewords (generator built inside) included"""
    import subprocess
    from lut_ldpc_b200 import ber_sim as bs
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    (tmp_path / "codes").mkdir()
    _alist.write_alist("c2_irr_n500", str(tmp_path / "codes" / "irr500.alist"))
    ini = tmp_path / "bp.ini"
    ini.write_text("[Sim]\nSNRdB = 1:1:2\nNframes = 20000\nNfers = 1000000\nfer_min = 0\nber_min = 0\n"
                   "[LDPC]\nparity_filename = irr500\nzero_codeword = false\n[BP]\nmax_iter = 30\n")
    ref = BER["c2_irr_n500"]
    # Python harness
    p = bs.SimParams(str(ini))
    sim = bs.make_sim(p, rand_seed=3, chunk_frames=8192)
    assert type(sim) is bs.LDPC_BER_Sim
    res = sim.load(str(tmp_path)).run(verbose=False)
    for k, s in enumerate(("1.00", "2.00")):
        assert res["sim_Nframes"][k] == 20000
        assert _agree(res["sim_frame_errors"][k], 20000, ref[s]["frame_errors"], ref[s]["frames"]), (s, res)
    # C++ harness
    exe = os.path.join(root, "lut_ldpc_b200", "ber_sim")
    if not os.path.exists(exe):
        pytest.skip("ber_sim binary not built")
    out = subprocess.run([exe, "-p", str(ini), "-b", str(tmp_path), "-s", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, (out.stdout[-1000:], out.stderr[-1000:])
    lines = [l for l in out.stdout.splitlines() if l.startswith("SNR = ")]
    assert len(lines) == 2
    for line, s in zip(lines, ("1.00", "2.00")):
        frames = int(line.split("Simulated ")[1].split(" frames")[0])
        fer = float(line.split("FER: ")[1])
        assert frames == 20000 and _agree(round(fer * frames), frames, ref[s]["frame_errors"], ref[s]["frames"]), (s, line)
    assert any(d.startswith("RES_N500_R0.5_maxIter30_zcw0_frames20000") for d in os.listdir(tmp_path / "results"))


def test_cpp_facade_one_frame_per_call(tmp_path):
    """host/ldpc_code_bp.hpp: Channel_Code::decode(vec), bp_decode, encode -- one frame per call like the reference's
    callers -- against a reference fixture"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "facade_check_bp")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-o", exe, os.path.join(root, "tests", "facade_check_bp.cpp"),
                           "-I" + root, "-L" + os.path.join(root, "lut_ldpc_b200"), "-llutldpc_b200",
                           "-Wl,-rpath," + os.path.join(root, "lut_ldpc_b200")])
    alist = str(tmp_path / "reg36.alist")
    _alist.write_alist("c1b_reg36_n1000", alist)
    _bpgolden.fixture("c1b_default")  # unpacks into the cache
    fix = os.path.join(_golden.CACHE, "bp_c1b_default.fix")
    p = subprocess.run([exe, alist, fix], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-2000:])
    r = json.loads(p.stdout.strip().splitlines()[-1])
    assert r["frames"] == 12 and r["mismatches"] == 0 and r["checked"] >= 3 * 12


def _random_rows(rs, N, M, dcs):
    """a random parity-check matrix: row c has dcs[c] distinct columns; every column is used at least once"""
    rows = [sorted(rs.choice(N, size=d, replace=False).tolist()) for d in dcs]
    used = set(v for r in rows for v in r)
    for v in range(N):
        if v not in used:
            c = int(rs.randint(M))
            rows[c] = sorted(set(rows[c]) | {v})
    return rows


@pytest.mark.parametrize("seed", range(12))
def test_bp_random_graphs_and_llr_units_vs_oracle(capi, seed):
    """random matrices (check degrees 2..40, variable degrees from 1) with random LLR_calc_unit parameters -- table
    sizes from 0 (min-sum) up, coarse and fine spacings, total resolutions from 5 bits (saturation everywhere, the lower
    clamp of Boxplus live) to 28 -- and random exit conditions: soft outputs and return values equal the oracle's"""
    rs = np.random.RandomState(1000 + seed)
    N, M = int(rs.randint(60, 400)), int(rs.randint(20, 120))
    dcs = rs.choice([2, 3, 4, 5, 6, 7, 8, 9, 12, 17, 33, 40], size=M, p=[.1, .1, .1, .1, .15, .1, .1, .05, .05, .05, .05, .05])
    dcs = np.minimum(dcs, N - 1)
    g = _alist.graph_from_rows(N, _random_rows(rs, N, M, dcs))
    d1 = int(rs.choice([2, 4, 8, 12]))
    d4 = int(rs.choice([5, 8, 12, 16, 28]))
    d4 = max(d4, d1 + 2) if rs.rand() < 0.7 else d4
    par = pyoracle.BpParams(max_iter=int(rs.randint(1, 20)), psc=bool(rs.randint(2)), pisc=bool(rs.randint(2)),
                            d1=d1, d2=int(rs.choice([0, 8, 64, 300])), d3=int(rs.randint(0, d1 + 1)), d4=d4)
    dec = _dec(capi, g, par)
    llr = rs.standard_normal((300, N)) * rs.choice([1.0, 4.0, 20.0]) + rs.choice([0.0, 2.0, 6.0])
    qin = pyoracle.bp_to_qllr(llr, par.d1, par.d4)
    want_out, want_ret = pyoracle.bp_decode(g["dv_vec"], g["dc_vec"], g["cn_msg_idx"], qin, par)
    out, ret = dec.decode_qllr(qin)
    assert np.array_equal(ret, want_ret), (seed, vars(par))
    assert np.array_equal(out, want_out), (seed, vars(par))
    bits, ret2 = dec.decode_llr(llr)
    assert np.array_equal(ret2, want_ret) and np.array_equal(bits, (want_out < 0).astype(np.uint8))


@pytest.mark.parametrize("d1,d2,d3,d4", [(12, 300, 7, 12), (12, 300, 7, 8), (12, 64, 9, 5), (6, 300, 0, 6)])
def test_bp_resolutions_below_the_table_values(capi, d1, d2, d3, d4):
    """qllr_max below logexp_table[0]: Boxplus saturates on both sides, and with Dint4 < Dint1 + 1 the table entries
    themselves are to_qllr's overflow value QLLR_MAX (itpp/itpp/comm/llr.h:266-272).  Still the oracle's numbers."""
    g = _bpgolden.graph("synth_mixed")
    par = pyoracle.BpParams(max_iter=6, psc=True, pisc=False, d1=d1, d2=d2, d3=d3, d4=d4)
    tab = pyoracle.bp_logexp_table(d1, d2, d3, d4)
    assert tab[0] > (1 << (d4 - 1)) - 1
    dec = _dec(capi, g, par)
    llr = np.random.RandomState(d4).standard_normal((200, g["nvar"])) * 3 + 1
    qin = pyoracle.bp_to_qllr(llr, d1, d4)
    want_out, want_ret = pyoracle.bp_decode(g["dv_vec"], g["dc_vec"], g["cn_msg_idx"], qin, par)
    out, ret = dec.decode_qllr(qin)
    assert np.array_equal(ret, want_ret) and np.array_equal(out, want_out)


def _unxz(name, tmp_path):
    import lzma
    dst = str(tmp_path / name)
    open(dst, "wb").write(lzma.open(os.path.join(_bpgolden.BP_DIR, name + ".xz")).read())
    return dst


def _pairs(path):
    raw = open(path, "rb").read()
    n, K, N = np.frombuffer(raw, "<i4", 3)
    a = np.frombuffer(raw, np.uint8, offset=12).reshape(n, K + N)
    return a[:, :K].copy(), a[:, K:].copy()


def test_bp_codec_file_with_generator_round_trip(capi, tmp_path):
    """a bp_codec.it the reference wrote WITH its generator (random column order): the handle created from it encodes
    like the reference's LDPC_Generator_Systematic::encode, decodes its codewords, and saves the same bytes again"""
    src = _unxz("c2_gen.bp_codec.it", tmp_path)
    dec = capi.BpDecoder.from_file(src, max_iters=30)
    assert dec.has_gen and dec.nvar == 500 and dec.ninfo == 250
    info, cw = _pairs(os.path.join(_bpgolden.BP_DIR, "c2_gen.pairs.bin"))
    assert np.array_equal(dec.encode(info), cw)
    bits, ret = dec.decode_llr((1.0 - 2.0 * cw) * 5.0)
    assert np.array_equal(bits, cw) and (ret == 0).all()
    out = str(tmp_path / "again.it")
    dec.save_code(out)
    assert open(out, "rb").read() == open(src, "rb").read()
    c = dec.simulate_snr(2.0, 4096, seed=3, zero_codeword=False)
    assert c.frames == 4096 and 0 < c.frame_errors < 400


def test_bp_generator_built_here_is_accepted_by_the_reference(capi, tmp_path):
    """alist -> generator on the device -> save_code: the unmodified reference loads the file (LDPC_Code(file, G)),
    writes the same bytes back, and its encoder gives the codewords lutldpc_bp_encode gives"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = os.path.join(root, "oracle", "_ref", "lutldpc_ref")
    if not os.path.exists(ref):
        pytest.skip("compiled reference not present")
    dec = capi.BpDecoder(_alist.bundle_graph("c1b_reg36_n1000"), with_generator=True)
    mine, back, pairs = str(tmp_path / "bp_codec.it"), str(tmp_path / "back.it"), str(tmp_path / "pairs.bin")
    dec.save_code(mine)
    subprocess.run([ref, "bpreload", mine, back, "--pairs", pairs, "--n", "4", "--seed", "5"], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert open(back, "rb").read() == open(mine, "rb").read()
    info, cw = _pairs(pairs)
    assert np.array_equal(dec.encode(info), cw)
    again = capi.BpDecoder.from_file(mine)
    assert again.has_gen and np.array_equal(again.encode(info), cw)


def test_bp_harness_saves_and_reloads_its_codec(capi, tmp_path):
    """Sim.save_codec / Sim.codec_filename + codec_type = BP (src/LDPC_BER_Sim.cpp:207-239): a run leaves
    results/<name>/bp_codec.it, a second run that loads it (no alist, no [BP] section needed to pick the decoder)
    reproduces the first run's counters; the C++ binary does the same"""
    import shutil
    import subprocess
    from lut_ldpc_b200 import ber_sim as bs
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    (tmp_path / "codes").mkdir()
    _alist.write_alist("c2_irr_n500", str(tmp_path / "codes" / "irr500.alist"))
    a = tmp_path / "a.ini"
    a.write_text("[Sim]\nSNRdB = 2\nNframes = 8192\nNfers = 1000000\nfer_min = 0\nber_min = 0\nsave_codec = 3\n"
                 "[LDPC]\nparity_filename = irr500\nzero_codeword = false\n[BP]\nmax_iter = 30\n")
    sim = bs.make_sim(bs.SimParams(str(a)), rand_seed=3, chunk_frames=8192).load(str(tmp_path))
    first = sim.run(verbose=False)
    saved = tmp_path / "results" / sim.gen_filename() / "bp_codec.it"
    assert saved.exists()
    shutil.copy(saved, tmp_path / "codes" / "stored_bp.it")
    b = tmp_path / "b.ini"
    b.write_text("[Sim]\nSNRdB = 2\nNframes = 8192\nNfers = 1000000\nfer_min = 0\nber_min = 0\nsave_codec = -1\n"
                 "codec_filename = stored_bp.it\ncodec_type = BP\n[LDPC]\nzero_codeword = false\n")
    p = bs.SimParams(str(b))
    assert p.has_bp_section and not p.has_lut_section
    again = bs.make_sim(p, rand_seed=3, chunk_frames=8192).load(str(tmp_path)).run(verbose=False)
    assert again == first and first["sim_frame_errors"][0] > 0
    exe = os.path.join(root, "lut_ldpc_b200", "ber_sim")
    if os.path.exists(exe):
        out = subprocess.run([exe, "-p", str(b), "-b", str(tmp_path), "-s", "3"], capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, (out.stdout[-1000:], out.stderr[-1000:])
        line = [l for l in out.stdout.splitlines() if l.startswith("SNR = ")][0]
        assert int(line.split("Obtained ")[1].split(" data bit errors")[0]) == first["sim_data_bit_errors"][0]
