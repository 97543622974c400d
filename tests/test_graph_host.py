"""Alist import, graph flattening and GF(2) rank on the host: lutldpc_alist_graph must reproduce the
arrays the reference's decoder_parameterization wrote into the golden bundles -- dv_vec, dc_vec, cn_msg_idx element for
element -- and its rank must equal the reference's nchk_lin_indep (dense GF2mat::row_rank)."""
import numpy as np
import pytest

import _golden
from _alist import write_alist

CFGS = ["c1b_reg36_n1000", "c2_irr_n500", "c3_irr_n10000", "c4_reg632_high", "x_qcha_irr_n1000", "c5_n64800"]


@pytest.mark.parametrize("cfg", CFGS)
def test_alist_graph_equals_reference_flattening(tmp_path, cfg):
    import lut_ldpc_b200 as L
    from lut_ldpc_b200 import capi
    path = str(tmp_path / (cfg + ".alist"))
    orc = write_alist(cfg, path)
    small = orc.nvar <= 10000
    g = capi.alist_graph(path, rank_mode=1 if small else 0)
    assert (g["nvar"], g["nchk"], g["num_edges"]) == (orc.nvar, orc.nchk, orc.num_edges)
    assert np.array_equal(g["dv_vec"], orc.dv) and np.array_equal(g["dc_vec"], orc.dc)
    assert np.array_equal(g["cn_msg_idx"], orc.cn_msg_idx)
    if small:  # (6,32) N=2048: 384 checks of rank 325 (README.md:239)
        assert g["nchk_lin_indep"] == orc.nvar - orc.ninfo
    else:
        assert g["nchk_lin_indep"] == -1


def test_alist_zero_padding_and_errors(tmp_path):
    from lut_ldpc_b200 import capi
    import lut_ldpc_b200 as L
    p0, p1 = str(tmp_path / "a.alist"), str(tmp_path / "b.alist")
    orc = write_alist("c2_irr_n500", p0, pad=0)   # index lists padded with -1 ...
    write_alist("c2_irr_n500", p1, pad=1)         # ... or with 0, as most published alists are
    a, b = capi.alist_graph(p0), capi.alist_graph(p1)
    assert np.array_equal(a["cn_msg_idx"], b["cn_msg_idx"]) and np.array_equal(a["cn_msg_idx"], orc.cn_msg_idx)
    bad = tmp_path / "bad.alist"
    lines = open(p1).read().split("\n")
    first = lines[4].split()                      # first column list: point its first entry at another check,
    first[0] = str(1 + (int(first[0]) % orc.nchk))  # so that it no longer matches the row lists
    lines[4] = " ".join(first)
    bad.write_text("\n".join(lines))
    with pytest.raises(L.LutLdpcError) as e:
        capi.alist_graph(str(bad))
    assert "disagree" in str(e.value)
    with pytest.raises(L.LutLdpcError):
        capi.alist_graph(str(tmp_path / "missing.alist"))
