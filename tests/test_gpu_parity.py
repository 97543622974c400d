"""GPU parity tests: the CUDA path, called through the C ABI, against the reference's own fixtures
(bit-exact: hard decisions, lut_decode return values, every per-iteration message dump) and against
the oracle restatement on seeded random inputs."""
import numpy as np
import pytest

import _golden
import pyoracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    import lut_ldpc_b200
    return lut_ldpc_b200


_decoders = {}


def _dec(L, cfg):
    if cfg not in _decoders:
        _decoders[cfg] = L.Decoder(_golden.bundle_path(cfg))
    return _decoders[cfg]


@pytest.mark.parametrize("cfg,tag", _golden.all_fixtures())
def test_fixture_bit_exact(L, cfg, tag):
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
    a = _golden.fixture_args(cfg, tag)
    dec = _dec(L, cfg)
    dec.set_exit_conditions(dec.max_iters, a["psc"], a["pisc"])
    out, it = dec.decode_quantized(fx["cha"], fx["msg"])
    assert np.array_equal(it, fx["ret"])
    assert np.array_equal(out, fx["out"])
    for f, recs in enumerate(fx["traces"]):
        o, i2, mine = dec.decode_quantized_traced(fx["cha"], fx["msg"], f, len(recs) + 4)
        assert np.array_equal(i2, fx["ret"]) and np.array_equal(o, fx["out"])
        assert len(mine) == len(recs), (len(mine), len(recs))
        for (k0, i0, m0), (k1, i1, m1) in zip(recs, mine):
            assert (k0, i0) == (k1, i1)
            assert pyoracle.same_msgs(m0, m1), (cfg, tag, f, k0, i0)


@pytest.mark.parametrize("cfg,tag", [(c, t) for c, t in _golden.all_fixtures() if "with-llr 0" not in
                                     _golden.MANIFEST[c]["fixtures"][t]["args"]])
def test_llr_front_end(L, cfg, tag):
    """decode(const vec&): quant_nonlin front end + lut_decode from the reference's double LLRs."""
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
    a = _golden.fixture_args(cfg, tag)
    dec = _dec(L, cfg)
    dec.set_exit_conditions(dec.max_iters, a["psc"], a["pisc"])
    dec.set_initial_message_mode(L.INIT_QCHA if a["init_mode"] == "qcha" else L.INIT_CONT)
    cha, msg = dec.quantize(fx["llr"])
    assert np.array_equal(cha, fx["cha"]) and np.array_equal(msg, fx["msg"])
    out, it = dec.decode_llr(fx["llr"])
    assert np.array_equal(it, fx["ret"]) and np.array_equal(out, fx["out"])


@pytest.mark.parametrize("cfg", ["c3_irr_n10000", "c5_n64800"])
def test_llr_front_end_on_selected_frames(L, cfg):
    """decode(const vec&) on the large codes: the committed LLR inputs of the `sel` fixtures (frames that end in -50,
    +50, late, early; the reference quantised and decoded exactly these doubles) through quant_nonlin + lut_decode."""
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, "sel"))
    llr = _golden.sel_llr(cfg)
    dec = _dec(L, cfg)
    dec.set_exit_conditions(dec.max_iters, 1, 1)
    dec.set_initial_message_mode(L.INIT_CONT)
    cha, msg = dec.quantize(llr)
    assert np.array_equal(cha, fx["cha"]) and np.array_equal(msg, fx["msg"])
    out, it = dec.decode_llr(llr)
    assert np.array_equal(it, fx["ret"]) and np.array_equal(out, fx["out"])
    assert (it == -dec.max_iters).sum() >= 2 and (it == dec.max_iters).sum() >= 1  # decision pass, both outcomes


@pytest.mark.parametrize("cfg,n_frames", [(c, n) for c in ("c1b_reg36_n1000", "c2_irr_n500", "c1a_reg632_wide",
                                                            "c4_reg632_high", "x_varq_reuse", "x_lutcn_irr_n500",
                                                            "x_qcha_irr_n1000") for n in (1, 8, 333)] +
                         [("c3_irr_n10000", 12), ("c5_n64800", 6)])
def test_random_inputs_vs_oracle(L, cfg, n_frames):
    """Seeded noisy labels (ragged frame counts, several 256-frame column blocks) against the oracle."""
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    dec = _dec(L, cfg)
    rng = np.random.default_rng(1234 + n_frames)
    nq_cha, nq_msg = orc.nq_cha, int(orc.nq_msg[0])
    # all-zero codeword through a coarse "channel": mostly positive labels with noise
    base = rng.normal(1.0, 0.9, size=(n_frames, orc.nvar))
    cha = np.clip(np.floor((base + 2.0) / 4.0 * nq_cha), 0, nq_cha - 1).astype(np.uint8)
    msg = np.clip(np.floor((base + 2.0) / 4.0 * nq_msg), 0, nq_msg - 1).astype(np.uint8)
    for psc, pisc in ((1, 1), (0, 0), (1, 0)):
        orc.set_exit_conditions(orc.max_iters, psc, pisc)
        dec.set_exit_conditions(dec.max_iters, psc, pisc)
        want_out, want_it = orc.lut_decode_batch(cha, msg)
        out, it = dec.decode_quantized(cha, msg)
        assert np.array_equal(it, want_it), (psc, pisc)
        assert np.array_equal(out, want_out), (psc, pisc)



def test_fresh_handle_has_the_reference_constructor_defaults(L):
    """LDPC_Code_LUT is constructed with psc(true), pisc(false) (src/LDPC_Code_LUT.cpp:46-47): a handle that never
    saw lutldpc_set_exit_conditions must not take the initial syndrome exit.  The snr9 frames are codewords at the
    channel output: pisc would return 0, the reference default runs iteration 0 and returns >= 1."""
    cfg = "c1b_reg36_n1000"
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, "snr9"))
    assert (fx["ret"] == 0).any()  # the fixture itself was made with pisc=1
    dec = L.Decoder(_golden.bundle_path(cfg))
    out, it = dec.decode_quantized(fx["cha"], fx["msg"])
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    orc.set_exit_conditions(orc.max_iters, 1, 0)
    want_out, want_it = orc.lut_decode_batch(fx["cha"], fx["msg"])
    assert (want_it != 0).all() and (want_it[fx["ret"] == 0] >= 1).all()
    assert np.array_equal(it, want_it) and np.array_equal(out, want_out)


def test_empty_batch(L):
    dec = _dec(L, "c1b_reg36_n1000")
    out, it = dec.decode_quantized(np.zeros((0, dec.nvar), np.uint8), np.zeros((0, dec.nvar), np.uint8))
    assert out.shape == (0, dec.nvar) and it.shape == (0,)


def test_encode_and_syndrome(L):
    for cfg in ("c2_irr_n500", "c1a_reg632_wide"):
        orc = pyoracle.Oracle(_golden.bundle_path(cfg))
        dec = _dec(L, cfg)
        rng = np.random.default_rng(7)
        info = rng.integers(0, 2, size=(37, orc.ninfo), dtype=np.uint8)
        cw = dec.encode(info)
        want = np.stack([orc.encode(u) for u in info])
        assert np.array_equal(cw, want)
        ok = dec.syndrome_check(cw)
        assert ok.all()
        bad = cw.copy()
        bad[5, 3] ^= 1
        bad[20, orc.nvar - 1] ^= 1
        ok = dec.syndrome_check(bad)
        assert [i for i in range(37) if not ok[i]] == [5, 20]


def test_large_code_failures_and_decision_pass_across_column_blocks(L):
    """N=64800: the 20 selected reference frames (4 x -50, 4 x +50 -- the decision pass with both final-syndrome
    outcomes --, late, early, typical), shuffled over a ragged batch of several 256-frame column blocks, so that
    failing and converging frames share words and the batch is repacked while the failing ones go on."""
    cfg = "c5_n64800"
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, "sel"))
    assert (fx["ret"] == -50).sum() >= 2 and (fx["ret"] == 50).sum() >= 1
    dec = _dec(L, cfg)
    dec.set_exit_conditions(dec.max_iters, 1, 1)
    idx = np.random.default_rng(64800).integers(0, len(fx["ret"]), 601)
    before = dec.repack_count
    out, it = dec.decode_quantized(fx["cha"][idx], fx["msg"][idx])
    assert np.array_equal(it, fx["ret"][idx])
    assert np.array_equal(out, fx["out"][idx])
    assert dec.repack_count > before


def test_fast_vn_path_is_used(L):
    """The auto_bin_balanced 4-bit decoders must run on the specialised VN kernel, file trees on the interpreter."""
    for cfg, fast in (("c5_n64800", True), ("c3_irr_n10000", True), ("c2_irr_n500", True), ("c1b_reg36_n1000", True),
                      ("c4_reg632_high", False), ("c1a_reg632_wide", False)):
        d = _dec(L, cfg)
        assert d.n_var_programs > 0
        assert (d.n_fast_var_programs == d.n_var_programs) == fast, (cfg, d.n_fast_var_programs, d.n_var_programs)


def test_generic_interpreter_still_bit_exact(L):
    """LUTLDPC_GENERIC_VN=1 forces the node-program interpreter for the balanced trees too."""
    import os
    fx = pyoracle.read_fixture(_golden.fixture_path("c2_irr_n500", "snr1.0"))
    os.environ["LUTLDPC_GENERIC_VN"] = "1"
    try:
        d = L.Decoder(_golden.bundle_path("c2_irr_n500"))
    finally:
        del os.environ["LUTLDPC_GENERIC_VN"]
    assert d.n_fast_var_programs == 0
    d.set_exit_conditions(d.max_iters, 1, 1)
    out, it = d.decode_quantized(fx["cha"], fx["msg"])
    assert np.array_equal(it, fx["ret"]) and np.array_equal(out, fx["out"])


@pytest.mark.parametrize("cfg,snr,n", [("c2_irr_n500", 1.5, 3000), ("c1b_reg36_n1000", 1.8, 2100),
                                       ("c3_irr_n10000", 1.1, 1300), ("c5_n64800", 1.25, 1100)])
def test_frame_repacking_is_not_observable(L, cfg, snr, n):
    """Frames that return early are dropped from the frame-interleaved batch by gathering the unfinished
    ones into a narrower layout (several times per batch at an aggressive threshold): decisions and
    lut_decode returns equal those of the fixed layout and, on a sample, the oracle; ragged frame count,
    frames that finish at the initial check, early, late and never."""
    dec, ref = L.Decoder(_golden.bundle_path(cfg)), L.Decoder(_golden.bundle_path(cfg))
    for d in (dec, ref):  # repacking belongs to the HBM regime; the small codes would take the on-chip one
        d.set_onchip(False)
    ref.set_repack(False)
    ref_count0 = ref.repack_count
    rng = np.random.default_rng(5)
    sigma = np.sqrt(0.5 / (dec.rate * 10 ** (snr / 10)))
    llr = (2.0 / sigma ** 2) * (1.0 + sigma * rng.standard_normal((n, dec.nvar)))
    llr[::7] = np.abs(llr[::7]) + 1.0  # clean frames: return 0 at the initial syndrome check (pisc)
    cha, msg = dec.quantize(llr)
    for alpha in (0.97, 0.8, 0.5):
        dec.set_repack(True, alpha)
        for psc, pisc in ((1, 1), (1, 0), (0, 1)):
            dec.set_exit_conditions(dec.max_iters, psc, pisc)
            ref.set_exit_conditions(ref.max_iters, psc, pisc)
            before = dec.repack_count
            o1, i1 = dec.decode_quantized(cha, msg)
            o2, i2 = ref.decode_quantized(cha, msg)
            assert np.array_equal(i1, i2) and np.array_equal(o1, o2), (alpha, psc, pisc)
            if psc and alpha > 0.9:
                assert dec.repack_count > before, "no repack happened"
            if psc and pisc:
                assert len(np.unique(i1)) > 3
    assert ref.repack_count == ref_count0
    ref.set_repack(True)
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    orc.set_exit_conditions(dec.max_iters, 1, 1)
    dec.set_exit_conditions(dec.max_iters, 1, 1)
    dec.set_repack(True, 0.9)
    out, it = dec.decode_quantized(cha, msg)
    k = {"c2_irr_n500": 60, "c1b_reg36_n1000": 60, "c3_irr_n10000": 12, "c5_n64800": 8}[cfg]  # oracle: seconds per frame
    sel = np.r_[0:k, n - k:n]
    want_out, want_it = orc.lut_decode_batch(cha[sel], msg[sel])
    assert np.array_equal(want_it, it[sel]) and np.array_equal(want_out, out[sel])


def test_host_path_is_chunked_and_pipelined_bit_exact(L, monkeypatch):
    """lutldpc_decode_quantized / _llr cut a call into chunks that overlap their copies with the decoder
    (double-buffered staging, three streams).  Forced to 256-frame chunks: 5 chunks incl. a ragged last one,
    pageable host memory, results identical to the oracle frame by frame and to the one-chunk call."""
    cfg = "c2_irr_n500"
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    dec = L.Decoder(_golden.bundle_path(cfg))
    dec.set_onchip(False)  # the chunk pipeline is regime-independent; the HBM regime has more launches to interleave
    dec.set_exit_conditions(dec.max_iters, 1, 1)
    orc.set_exit_conditions(orc.max_iters, 1, 1)
    n = 1100
    rng = np.random.default_rng(21)
    sigma = np.sqrt(0.5 / (dec.rate * 10 ** (1.5 / 10)))
    llr = (2.0 / sigma ** 2) * (1.0 + sigma * rng.standard_normal((n, dec.nvar)))
    cha, msg = dec.quantize(llr)
    whole_out, whole_it = dec.decode_quantized(cha, msg)
    monkeypatch.setenv("LUTLDPC_HOST_CHUNK", "256")
    for _ in range(2):  # second call reuses the staging buffers and events
        out, it = dec.decode_quantized(cha, msg)
        assert np.array_equal(it, whole_it) and np.array_equal(out, whole_out)
    out2, it2 = dec.decode_llr(llr)
    assert np.array_equal(it2, whole_it) and np.array_equal(out2, whole_out)
    monkeypatch.delenv("LUTLDPC_HOST_CHUNK")
    # long calls ramp their chunk sizes (small first and last chunk, large ones in between): nominal 1024
    # -> chunks 512, 1024, [middle], 1024, 512 over 6000 frames (ragged), same results as chunk by chunk
    big = np.tile(np.arange(n), 6)[:6000 - 7]
    monkeypatch.setenv("LUTLDPC_HOST_RAMP_NOMINAL", "1024")
    out3, it3 = dec.decode_quantized(cha[big], msg[big])
    monkeypatch.delenv("LUTLDPC_HOST_RAMP_NOMINAL")
    assert np.array_equal(it3, whole_it[big]) and np.array_equal(out3, whole_out[big])
    want_out, want_it = orc.lut_decode_batch(cha, msg)
    assert np.array_equal(want_it, whole_it) and np.array_equal(want_out, whole_out)
    assert len(np.unique(whole_it)) > 3


@pytest.mark.parametrize("cfg", ["c1b_reg36_n1000", "c3_irr_n10000", "c5_n64800"])
def test_decoder_from_reference_codec_file(L, cfg):
    """lutldpc_create_from_bundle on the reference's own lut_codec.it (LDPC_Code_LUT::save_code): same
    decoder as from the bundle, bit-exact on the reference's fixture frames."""
    d = L.Decoder(_golden.codec_path(cfg))
    b = _dec(L, cfg)
    assert (d.nvar, d.nchk, d.num_edges, d.max_iters, d.lookups_per_frame_iter) == \
           (b.nvar, b.nchk, b.num_edges, b.max_iters, b.lookups_per_frame_iter)
    for tag in sorted(_golden.MANIFEST[cfg]["fixtures"]):
        fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
        a = _golden.fixture_args(cfg, tag)
        d.set_exit_conditions(d.max_iters, a["psc"], a["pisc"])
        out, it = d.decode_quantized(fx["cha"], fx["msg"])
        assert np.array_equal(it, fx["ret"]) and np.array_equal(out, fx["out"])


# decoder qualifies for the on-chip regime (min-sum checks, state of a few columns fits shared memory) / takes it by default
ONCHIP = {"c1b_reg36_n1000": (True, True), "c2_irr_n500": (True, False), "c1a_reg632_wide": (True, False),
          "c4_reg632_high": (True, False), "x_varq_reuse": (True, False), "x_qcha_irr_n1000": (True, False),
          "x_lutcn_irr_n500": (False, False), "c3_irr_n10000": (False, False), "c5_n64800": (False, False)}


def test_onchip_regime_qualification(L):
    """min-sum decoders whose per-column state fits shared memory can decode on chip; by default only those for which
    it was measured faster than the HBM regime do (compile-time variable programs, degree <= 8).  LUT check nodes and
    the large codes stream through HBM; asking for the regime on a decoder that does not qualify is an error."""
    for cfg, (fits, default) in ONCHIP.items():
        d = L.Decoder(_golden.bundle_path(cfg))
        assert bool(d.onchip) == default, (cfg, d.onchip, d.onchip_smem_bytes)
        assert (d.onchip_smem_bytes > 0) == fits
        if fits:
            d.set_onchip(True)
            d.set_onchip(False)
        else:
            with pytest.raises(L.LutLdpcError):
                d.set_onchip(True)


@pytest.mark.parametrize("cfg", [c for c, v in ONCHIP.items() if v[0]])
def test_onchip_regime_equals_hbm_regime_and_oracle(L, cfg):
    """Same seeded noisy frames through the persistent shared-memory kernel and through the per-pass HBM kernels: same
    decisions and lut_decode returns (returns at the initial check, early, late, never; ragged batch over several
    columns per CTA), and both equal the oracle on a sample; all exit configurations."""
    oc, hbm = L.Decoder(_golden.bundle_path(cfg)), L.Decoder(_golden.bundle_path(cfg))
    oc.set_onchip(True)
    hbm.set_onchip(False)
    assert oc.onchip and not hbm.onchip
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    qcha = _golden.MANIFEST[cfg]["init_mode"] == "qcha"
    for d in (oc, hbm):
        d.set_initial_message_mode(L.INIT_QCHA if qcha else L.INIT_CONT)
    snr = {"c1b_reg36_n1000": 1.8, "c2_irr_n500": 1.5, "c1a_reg632_wide": 3.8, "c4_reg632_high": 4.0,
           "x_varq_reuse": 2.5, "x_qcha_irr_n1000": 1.8}[cfg]
    n = 2500 + 37
    rng = np.random.default_rng(99)
    sigma = np.sqrt(0.5 / (oc.rate * 10 ** (snr / 10)))
    llr = (2.0 / sigma ** 2) * (1.0 + sigma * rng.standard_normal((n, oc.nvar)))
    llr[::9] = np.abs(llr[::9]) + 1.0  # clean frames: return 0 at the initial syndrome check (pisc)
    cha, msg = oc.quantize(llr)
    for psc, pisc in ((1, 1), (1, 0), (0, 0)):
        for d in (oc, hbm):
            d.set_exit_conditions(d.max_iters, psc, pisc)
        o1, i1 = oc.decode_quantized(cha, msg)
        o2, i2 = hbm.decode_quantized(cha, msg)
        assert np.array_equal(i1, i2), (psc, pisc, np.flatnonzero(i1 != i2)[:10], i1[i1 != i2][:10], i2[i1 != i2][:10])
        assert np.array_equal(o1, o2), (psc, pisc)
        if psc and pisc:
            assert len(np.unique(i1)) > 3
        sel = np.r_[0:24, n - 24:n]
        orc.set_exit_conditions(orc.max_iters, psc, pisc)
        want_out, want_it = orc.lut_decode_batch(cha[sel], msg[sel])
        assert np.array_equal(want_it, i1[sel]) and np.array_equal(want_out, o1[sel])
    # the LLR front end and the fused Monte-Carlo step take the regime too: same counters either way
    oc.set_exit_conditions(oc.max_iters, 1, 1)
    hbm.set_exit_conditions(hbm.max_iters, 1, 1)
    o3, i3 = oc.decode_llr(llr)
    assert np.array_equal(i3, i1 if False else oc.decode_quantized(cha, msg)[1])
    a = oc.simulate_snr(snr, 3000, seed=4, zero_codeword=True).as_array()
    b = hbm.simulate_snr(snr, 3000, seed=4, zero_codeword=True).as_array()
    assert np.array_equal(a, b)


@pytest.mark.parametrize("cfg,tag", [("c2_irr_n500", "snr1.0"), ("c1a_reg632_wide", "snr4.0"), ("c5_n64800", "sel")])
def test_packed_entry_points_equal_the_byte_form(L, cfg, tag):
    """lutldpc_decode_packed[_device]: both labels of a symbol in one byte in, one BIT per decision out (73 KB instead of
    194 KB per N=64800 frame across PCIe).  Same decoder, same results: the reference's fixture frames, shuffled over a
    ragged batch of several column blocks (nvar = 500 also has a ragged last byte)."""
    import torch
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
    a = _golden.fixture_args(cfg, tag)
    dec = _dec(L, cfg)
    dec.set_exit_conditions(dec.max_iters, a["psc"], a["pisc"])
    idx = np.random.default_rng(3).integers(0, len(fx["ret"]), 777)
    both = (fx["cha"][idx] | (fx["msg"][idx] << 4)).astype(np.uint8)
    packed, it = dec.decode_packed(both)
    nb = (dec.nvar + 7) // 8
    assert packed.shape == (777, nb)
    bits = np.unpackbits(packed, axis=1, bitorder="little")[:, :dec.nvar]
    assert np.array_equal(it, fx["ret"][idx])
    assert np.array_equal(bits, fx["out"][idx])
    assert not np.unpackbits(packed, axis=1, bitorder="little")[:, dec.nvar:].any()  # padding bits of the last byte
    d_in = torch.from_numpy(both).cuda()
    d_out = torch.zeros((777, nb), dtype=torch.uint8, device="cuda")
    d_it = torch.zeros(777, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    dec.decode_packed_device(777, d_in.data_ptr(), d_out.data_ptr(), d_it.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy(), packed) and np.array_equal(d_it.cpu().numpy(), it)


@pytest.mark.parametrize("cfg", ["c1b_reg36_n1000", "x_varq_reuse", "x_lutcn_irr_n500", "c4_reg632_high", "c3_irr_n10000"])
def test_decoder_built_from_alist_and_own_design_is_bit_exact(L, cfg, tmp_path):
    """The whole set-up chain without any reference-written decoder file: alist import + graph flattening
    (lutldpc_alist_graph), LUT design by density evolution on the host (lutldpc_design_luts), lutldpc_create from the
    resulting arrays -- and then the reference's fixture frames (decisions, returns, every per-iteration message dump)
    must come out exactly as the reference decoded them with ITS OWN design of the same code."""
    import test_design_host as T
    from _alist import write_alist
    from lut_ldpc_b200 import capi
    alist = str(tmp_path / (cfg + ".alist"))
    write_alist(cfg, alist)
    kv = T._args(cfg)
    graph = capi.alist_graph(alist, rank_mode=1)
    max_iter = int(kv["max-iter"])
    nq = np.full(max_iter, 1 << int(kv["qmsg"]), np.int32)
    for i, q in enumerate(kv.get("qmsg-list", "").split(",")[:max_iter] if "qmsg-list" in kv else []):
        nq[i] = 1 << int(q)
    reuse = np.zeros(max_iter, np.int32)
    for i, r in enumerate(kv.get("reuse", "").split(",")[:max_iter] if "reuse" in kv else []):
        reuse[i] = int(r) != 0
    rate = 1.0 - graph["nchk_lin_indep"] / graph["nvar"]
    sigma2 = float(kv["design-thr"]) ** 2 if "design-thr" in kv else 10.0 ** (-float(kv["design-snr"]) / 10) / (2 * rate)
    method = "auto_bin_balanced"
    if "tree" in kv:
        name = kv["tree"].split("/")[-1]
        (tmp_path / name).write_text(T.TEMPLATES[name])
        method = "filename=" + str(tmp_path / name)
    min_lut = kv.get("min-lut", "1") != "0"
    design = capi.design_luts(graph["dv_vec"], graph["dc_vec"], max_iter, 1 << int(kv["qcha"]), nq, sigma2, reuse_vec=reuse,
                              min_lut=min_lut, tree_method=method)
    dec = L.Decoder.from_design(graph, design, max_iter, 1 << int(kv["qcha"]), nq, reuse_vec=reuse, min_lut=min_lut)
    ref = _dec(L, cfg)
    assert (dec.nvar, dec.nchk_lin_indep, dec.num_edges, dec.lookups_per_frame_iter) == \
           (ref.nvar, ref.nchk_lin_indep, ref.num_edges, ref.lookups_per_frame_iter)
    for tag in sorted(_golden.MANIFEST[cfg]["fixtures"]):
        fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
        a = _golden.fixture_args(cfg, tag)
        dec.set_exit_conditions(dec.max_iters, a["psc"], a["pisc"])
        out, it = dec.decode_quantized(fx["cha"], fx["msg"])
        assert np.array_equal(it, fx["ret"]) and np.array_equal(out, fx["out"])
        for f, recs in enumerate(fx["traces"][:1]):
            o, i2, mine = dec.decode_quantized_traced(fx["cha"], fx["msg"], f, len(recs) + 4)
            assert len(mine) == len(recs)
            for (k0, i0, m0), (k1, i1, m1) in zip(recs, mine):
                assert (k0, i0) == (k1, i1) and pyoracle.same_msgs(m0, m1), (cfg, tag, f, k0, i0)
