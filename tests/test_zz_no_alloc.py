"""lutldpc_syndrome_check_device_noalloc: the allocation-free variant (static __device__ scratch) against the existing
entry point lutldpc_syndrome_check, through the C ABI, on two calls in a row.  Named test_zz_* so that pytest
collects it last."""
import numpy as np
import pytest

import _golden

pytestmark = pytest.mark.gpu


def test_noalloc_syndrome_check_matches_existing_entry_point():
    import torch
    import lut_ldpc_b200 as L
    dec = L.Decoder(_golden.bundle_path("c2_irr_n500"))
    rng = np.random.default_rng(2024)
    stream = torch.cuda.current_stream()
    dec.set_stream(stream.cuda_stream)
    # two calls in a row; the second is smaller and ragged, so stale rows / flags of the first would show
    for call, n in enumerate((1500, 333)):
        info = rng.integers(0, 2, size=(n, dec.ninfo), dtype=np.uint8)
        bits = dec.encode(info)
        bad = rng.choice(n, size=n // 3, replace=False)
        bits[bad, rng.integers(0, dec.nvar, size=bad.size)] ^= 1
        want = dec.syndrome_check(bits)                      # existing entry point (host buffers)
        assert want.sum() == n - bad.size
        d_bits = torch.from_numpy(bits).cuda()
        d_ok = torch.full((n,), 7, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        dec.syndrome_check_device_noalloc(n, d_bits.data_ptr(), d_ok.data_ptr())
        torch.cuda.synchronize()
        got = d_ok.cpu().numpy()
        assert np.array_equal(got.astype(bool), want), "call %d" % call   # integer work: exact
        assert set(np.unique(got)) <= {0, 1}
    # beyond the static array: a clear error, nothing launched
    too_many = dec.noalloc_max_frames + 1
    assert dec.noalloc_max_frames >= 256
    with pytest.raises(L.LutLdpcError) as e:
        dec.syndrome_check_device_noalloc(too_many, d_bits.data_ptr(), d_ok.data_ptr())
    assert e.value.code == -1 and "static array" in str(e.value)
    dec.set_stream(0)
