"""Pins the plain-C restatement (oracle/lutldpc_oracle.c) against outputs of the reference itself:
quantised inputs, hard decisions, lut_decode return values and every per-iteration message dump the
reference printed at output_verbosity=3 (src/LDPC_Code_LUT.cpp:291-337)."""
import numpy as np
import pytest

import _golden
import pyoracle


@pytest.mark.parametrize("cfg,tag", _golden.all_fixtures())
def test_restatement_matches_reference(cfg, tag):
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    a = _golden.fixture_args(cfg, tag)
    orc.set_exit_conditions(orc.max_iters, a["psc"], a["pisc"])
    assert orc.nvar == fx["nvar"] and orc.num_edges == fx["num_edges"]
    if fx["llr"] is not None:  # quantiser front end, src/LDPC_Code_LUT.cpp:207-221
        mode = pyoracle.INIT_QCHA if a["init_mode"] == "qcha" else pyoracle.INIT_CONT
        cha, msg = orc.quantise(fx["llr"], mode)
        assert np.array_equal(cha, fx["cha"]) and np.array_equal(msg, fx["msg"])
    out, ret = orc.lut_decode_batch(fx["cha"], fx["msg"])
    assert np.array_equal(ret, fx["ret"])
    assert np.array_equal(out, fx["out"])
    for f, recs in enumerate(fx["traces"]):
        o, r, mine = orc.lut_decode(fx["cha"][f], fx["msg"][f], max_trace=len(recs) + 4)
        assert r == fx["ret"][f]
        assert len(mine) == len(recs)
        for (k0, i0, m0), (k1, i1, m1) in zip(recs, mine):
            assert (k0, i0) == (k1, i1)
            assert pyoracle.same_msgs(m0, m1), (cfg, tag, f, k0, i0)


def test_encoder_matches_reference_codewords():
    """Random-codeword fixtures: [u, P*u] of the bundle reproduces the reference's encode()."""
    for cfg in ("c2_irr_n500", "c1a_reg632_wide"):
        orc = pyoracle.Oracle(_golden.bundle_path(cfg))
        assert orc.has_gen
        tag = sorted(_golden.MANIFEST[cfg]["fixtures"])[0]
        fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
        assert not fx["zero_cw"] and fx["cw"].any()
        for cw in fx["cw"]:
            assert np.array_equal(orc.encode(cw[:orc.ninfo]), cw)
            assert orc.syndrome_ok(cw)
