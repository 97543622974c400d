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
            assert np.array_equal(m0, m1), (cfg, tag, f, k0, i0, np.flatnonzero(m0 != m1)[:8])


def test_program_sharing_reduces_lookups():
    em = _emul.Emul(_golden.bundle_path("c5_n64800"))
    ops, naive = em.stats()
    assert naive == 952522          # sum dv*(dv-1) for the N=64800 code
    assert ops < naive


def test_guarded_minsum_equals_plain():
    """core.cuh: the guard-bit min-sum used by the vectorised CN kernel == the plain SWAR min-sum."""
    assert _emul.lib().emul_cng_selftest(200000, 12345) == 0
