"""GF(2) set-up on the device: rank of H and LDPC_Generator_Systematic::construct as Gauss-Jordan
elimination on the GPU, the bit-packed encoder, and a decoder for the column-permuted code.  Checked against what the
reference computed (nchk_lin_indep of the golden bundles) and against the code itself (every codeword satisfies H, the
code is systematic, decoding statistics do not depend on the codeword)."""
import numpy as np
import pytest

import _golden
from _alist import write_alist

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cfg", ["c2_irr_n500", "c1b_reg36_n1000", "c1a_reg632_wide", "c3_irr_n10000", "c5_n64800"])
def test_systematic_generator_on_device(cfg):
    import lut_ldpc_b200 as L
    src = L.Decoder(_golden.bundle_path(cfg))
    sysd, perm = src.systematic()
    # rank as the reference's dense GF2mat::row_rank found it ((6,32) N=2048: 384 checks, rank 325)
    assert sysd.nchk_lin_indep == src.nchk_lin_indep and sysd.ninfo == src.ninfo
    assert sorted(perm.tolist()) == list(range(src.nvar)) and sysd.has_gen
    assert (sysd.nvar, sysd.nchk, sysd.num_edges) == (src.nvar, src.nchk, src.num_edges)
    rng = np.random.default_rng(5)
    n = 300 if src.nvar <= 10000 else 64
    info = rng.integers(0, 2, size=(n, sysd.ninfo), dtype=np.uint8)
    cw = sysd.encode(info)
    assert np.array_equal(cw[:, :sysd.ninfo], info)               # systematic: codeword = [u, P*u]
    assert sysd.syndrome_check(cw).all()                          # ... of the permuted code
    inv = np.argsort(perm)                                        # old column of every new column
    assert src.syndrome_check(cw[:, perm]).all()                  # and, columns moved back, of the original H
    assert not sysd.syndrome_check(cw ^ np.eye(1, src.nvar, 3, dtype=np.uint8)).any()
    assert inv.shape == perm.shape
    # the zero word encodes to zero, and encoding is linear
    z = sysd.encode(np.zeros((1, sysd.ninfo), np.uint8))
    assert not z.any()
    assert np.array_equal(sysd.encode(info[:1] ^ info[1:2]), cw[:1] ^ cw[1:2])


def test_random_codeword_simulation_on_permuted_decoder():
    """Same FER with the all-zero word on the original decoder and with random codewords on the systematic one."""
    import lut_ldpc_b200 as L
    from test_gpu_ber import _agree
    src = L.Decoder(_golden.bundle_path("c1b_reg36_n1000"))
    sysd, _ = src.systematic()
    for d in (src, sysd):
        d.set_exit_conditions(d.max_iters, True, True)
    a = src.simulate_snr(2.0, 30000, seed=5, zero_codeword=True)
    b = sysd.simulate_snr(2.0, 30000, seed=6, zero_codeword=False)
    assert _agree(a.frame_errors, a.frames, b.frame_errors, b.frames, z=3.29)
    assert abs(a.sum_iters / a.frames - b.sum_iters / b.frames) < 1.0
    assert b.frame_errors > 100


def test_rank_from_alist_on_device(tmp_path):
    from lut_ldpc_b200 import capi
    for cfg in ("c4_reg632_high", "c5_n64800"):
        path = str(tmp_path / (cfg + ".alist"))
        orc = write_alist(cfg, path)
        g = capi.alist_graph(path, rank_mode=2)
        assert g["nchk_lin_indep"] == orc.nvar - orc.ninfo
        assert np.array_equal(g["cn_msg_idx"], orc.cn_msg_idx)
