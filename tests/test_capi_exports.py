"""The C-ABI library loads without a GPU and exports every entry point include/lutldpc.h declares;
creating a decoder without a CUDA device fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

import _golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "lutldpc.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(lutldpc_[a-z_0-9]+)\s*\(", txt)))


def test_header_symbols_are_exported():
    import lut_ldpc_b200 as L
    if not os.path.exists(L.library_path()):
        L.build_library()
    lib = ctypes.CDLL(L.library_path())
    names = _declared()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.lutldpc_abi_version() == 1
    from lut_ldpc_b200 import capi
    assert set(capi.EXPORTS) <= set(names)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import lut_ldpc_b200 as L
    with pytest.raises(L.LutLdpcError) as e:
        L.Decoder(_golden.bundle_path("c1b_reg36_n1000"))
    assert "no CPU fallback" in str(e.value)


def test_bad_bundle_is_rejected(tmp_path):
    import lut_ldpc_b200 as L
    p = tmp_path / "bad.bundle"
    p.write_bytes(b"LUTLDPC1" + b"\x00" * 16)
    with pytest.raises(L.LutLdpcError):
        L.Decoder(str(p))
    with pytest.raises(L.LutLdpcError):
        L.Decoder(str(tmp_path / "missing.bundle"))
