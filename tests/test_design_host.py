"""LUT design on the host: lutldpc_design_luts -- tree templates, discrete density evolution with LUT
export, the information-optimal quantiser -- must reproduce what the reference's design_luts wrote into the golden
bundles BYTE FOR BYTE: the text serialisation of every LUT tree, the quantiser thresholds and Nq_Cha_2_Nq_Msg_map.
The design arguments are the ones tests/golden/manifest.json records for each bundle."""
import os

import numpy as np
import pytest

import _golden

# tree templates of the two file-tree configurations (the grammar of trees/README.md: pre-order over r i m c, "/" closes
# a node), written to a template ini by the test
TEMPLATES = {
    "6_32_wide.ini": "[var_iter_000]\n   var_deg_006 = riiim/m//im/m///m//c/\n\n[DT]\n   var_deg_006 = rim/m/m//im/m/m//c/\n",
    "6_32_high.ini": "[var_iter_000]\n   var_deg_006 = rm/m/m/m/m/c//\n\n\n[DT]\n   var_deg_006 = rm/m/m/m/m/m/c//   \n",
}


def _args(cfg):
    a = _golden.MANIFEST[cfg]["design"].split()
    kv = {}
    for i, t in enumerate(a):
        if t.startswith("--"):
            kv[t[2:]] = a[i + 1] if i + 1 < len(a) and not a[i + 1].startswith("--") else "1"
    return kv


@pytest.mark.parametrize("cfg", ["c1b_reg36_n1000", "c2_irr_n500", "c3_irr_n10000", "c5_n64800", "x_varq_reuse",
                                 "x_qcha_irr_n1000", "x_lutcn_irr_n500", "c1a_reg632_wide", "c4_reg632_high"])
def test_designed_tables_equal_the_reference(cfg, tmp_path):
    from lut_ldpc_b200 import capi
    b = _golden.read_bundle(_golden.bundle_path(cfg))
    kv = _args(cfg)
    max_iter = int(kv["max-iter"])
    nq = np.full(max_iter, 1 << int(kv["qmsg"]), np.int32)
    if "qmsg-list" in kv:
        for i, q in enumerate(kv["qmsg-list"].split(",")[:max_iter]):
            nq[i] = 1 << int(q)
    reuse = np.zeros(max_iter, np.int32)
    if "reuse" in kv:
        for i, r in enumerate(kv["reuse"].split(",")[:max_iter]):
            reuse[i] = int(r) != 0
    assert np.array_equal(nq, b["Nq_Msg"]) and np.array_equal(reuse, b["reuse_vec"])
    if "design-thr" in kv:
        sigma2 = float(kv["design-thr"]) ** 2
    else:  # src/LDPC_BER_Sim.cpp:481-483 with rate = 1 - nchk_lin_indep / nvar (src/LDPC_Code_LUT.hpp:246-248)
        rate = 1.0 - float(b["nchk_lin_indep"]) / b["nvar"]
        sigma2 = 10.0 ** (-float(kv["design-snr"]) / 10) / (2 * rate)
    method = "auto_bin_balanced"
    if "tree" in kv and kv["tree"].startswith("filename="):
        name = os.path.basename(kv["tree"][len("filename="):])
        (tmp_path / name).write_text(TEMPLATES[name])
        method = "filename=" + str(tmp_path / name)
    r = capi.design_luts(b["dv_vec"], b["dc_vec"], max_iter, 1 << int(kv["qcha"]), nq, sigma2, reuse_vec=reuse,
                         min_lut=kv.get("min-lut", "1") != "0", tree_method=method)
    assert np.array_equal(r["qb_Cha"], b["qb_Cha"]), (r["qb_Cha"], b["qb_Cha"])
    assert np.array_equal(r["qb_Msg"], b["qb_Msg"])
    assert np.array_equal(r["Nq_Cha_2_Nq_Msg_map"], b["Nq_Cha_2_Nq_Msg_map"])
    if r["var_tree_string"] != b["var_tree_string"]:
        x, y = r["var_tree_string"].split(b"\n"), b["var_tree_string"].split(b"\n")
        first = next((i for i, (p, q) in enumerate(zip(x, y)) if p != q), min(len(x), len(y)))
        raise AssertionError("var trees differ from line %d of %d/%d: %r vs %r" % (first, len(x), len(y), x[first][:80], y[first][:80]))
    assert r["chk_tree_string"] == b["chk_tree_string"]
