"""The C restatement of the BP comparison decoder (oracle/bp_oracle.c: LLR_calc_unit + LDPC_Code::bp_decode) against
the golden vectors the unmodified reference produced: to_qllr, the logexp table through its effects, bp_decode's soft
output and return value, and the bits LDPC_Code::decode returns -- all bit-exact."""
import numpy as np
import pytest

import _bpgolden
import pyoracle


@pytest.mark.parametrize("name", _bpgolden.CASES)
def test_bp_oracle_reproduces_reference(name):
    fx = _bpgolden.fixture(name)
    g = _bpgolden.graph(name)
    par = fx["par"]
    assert g["nvar"] == fx["nvar"]
    n = len(fx["ret"]) if g["nvar"] < 20000 else 1
    if fx["llr"] is not None:
        assert np.array_equal(pyoracle.bp_to_qllr(fx["llr"], par.d1, par.d4), fx["qin"])
    out, ret = pyoracle.bp_decode(g["dv_vec"], g["dc_vec"], g["cn_msg_idx"], fx["qin"][:n], par)
    assert np.array_equal(ret, fx["ret"][:n])
    assert np.array_equal(out, fx["qout"][:n])
    assert np.array_equal((out[:, :fx["ninfo"]] < 0).astype(np.uint8), fx["syst"][:n])


def test_fixtures_cover_the_branches():
    """what the golden set exercises: every return kind, saturation at qllr_max, the QLLR_MAX overflow quirk of
    to_qllr, min-sum (empty table), the degree-6 check update, degrees 1 and 2 variables, a degree above 32"""
    rets = np.concatenate([_bpgolden.fixture(c)["ret"] for c in _bpgolden.SMALL])
    assert (rets == 0).any() and (rets > 0).any() and (rets < 0).any()
    fx = _bpgolden.fixture("c1b_saturating")
    qmax = (1 << (fx["par"].d4 - 1)) - 1
    assert (fx["qin"] == (2**31 - 1) >> 4).any() and (fx["qin"] == -qmax).any() and (np.abs(fx["qout"]) == qmax).any()
    assert _bpgolden.fixture("c1b_minsum_q6")["par"].d2 == 0
    g = _bpgolden.graph("synth_mixed")
    assert {1, 2, 3}.issubset(set(g["dv_vec"].tolist())) and {2, 3, 4, 5, 6, 7, 40}.issubset(set(g["dc_vec"].tolist()))
    assert _bpgolden.graph("synth_wide_dc48")["dc_vec"].max() == 48


def test_logexp_table_matches_its_definition():
    """construct_logexp_table (itpp/itpp/comm/llr.cpp:82-92): to_qllr(log(1 + exp(-x))) at x = 2^(Dint3-Dint1) i"""
    t = pyoracle.bp_logexp_table(12, 300, 7, 28)
    x = 2.0 ** (7 - 12) * np.arange(300)
    assert np.array_equal(t, np.floor(0.5 + 4096 * np.log(1 + np.exp(-x))).astype(np.int32))
    assert t[0] == 2839 and t[-1] == 0
