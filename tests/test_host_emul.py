"""CPU check of the packed-nibble arithmetic (core.cuh) and the compiled node programs (model.cpp):
the same per-thread code the kernels run, driven on the host, against the reference's fixtures."""
import numpy as np
import pytest

import _emul
import _golden
import pyoracle


@pytest.mark.parametrize("cfg,tag", _golden.all_fixtures())
def test_emulated_kernels_match_reference(cfg, tag):
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
    a = _golden.fixture_args(cfg, tag)
    em = _emul.Emul(_golden.bundle_path(cfg))
    out, it, _ = em.decode(fx["cha"], fx["msg"], a["psc"], a["pisc"])
    assert np.array_equal(it, fx["ret"])
    assert np.array_equal(out, fx["out"])
    for f, recs in enumerate(fx["traces"]):
        o, i2, mine = em.decode(fx["cha"], fx["msg"], a["psc"], a["pisc"], trace_frame=f,
                                num_edges=fx["num_edges"], max_rec=len(recs) + 4)
        assert len(mine) == len(recs), (len(mine), len(recs))
        for (k0, i0, m0), (k1, i1, m1) in zip(recs, mine):
            assert (k0, i0) == (k1, i1)
            assert pyoracle.same_msgs(m0, m1), (cfg, tag, f, k0, i0)


def test_program_sharing_reduces_lookups():
    em = _emul.Emul(_golden.bundle_path("c5_n64800"))
    ops, naive = em.stats()
    assert naive == 952522          # sum dv*(dv-1) for the N=64800 code
    assert ops < naive


def test_guarded_minsum_equals_plain():
    """core.cuh: the guard-bit min-sum used by the vectorised CN kernel == the plain SWAR min-sum."""
    assert _emul.lib().emul_cng_selftest(200000, 12345) == 0


@pytest.mark.parametrize("cfg", ["c1b_reg36_n1000", "c3_irr_n10000", "c5_n64800"])
def test_reference_codec_file_loads_like_the_bundle(cfg):
    """The reference's own codec file (IT++ it_file written by LDPC_Code_LUT::save_code,
    src/LDPC_Code_LUT.cpp:643-697) is read directly: same compiled programs and the same decisions on the
    reference's fixture frames as the bundle dumped from the very same reference object."""
    a = _emul.Emul(_golden.codec_path(cfg))
    b = _emul.Emul(_golden.bundle_path(cfg))
    assert a.stats() == b.stats()
    tag = sorted(_golden.MANIFEST[cfg]["fixtures"])[0]
    fx = pyoracle.read_fixture(_golden.fixture_path(cfg, tag))
    fa = _golden.fixture_args(cfg, tag)
    n = 1 if cfg == "c5_n64800" else 4
    out, it, _ = a.decode(fx["cha"][:n], fx["msg"][:n], fa["psc"], fa["pisc"])
    assert np.array_equal(it, fx["ret"][:n]) and np.array_equal(out, fx["out"][:n])


def test_codec_file_errors_are_reported(tmp_path):
    p = tmp_path / "bad.it"
    p.write_bytes(b"IT++\x03" + b"\x00" * 40)
    with pytest.raises(RuntimeError):
        _emul.Emul(str(p))
    good = open(_golden.codec_path("c1b_reg36_n1000"), "rb").read()
    p.write_bytes(good[:len(good) // 2])  # truncated in the middle of the tree string
    with pytest.raises(RuntimeError):
        _emul.Emul(str(p))
