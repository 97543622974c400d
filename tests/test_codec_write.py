"""lutldpc_save_code / lutldpc_convert_codec == LDPC_Code_LUT::save_code (src/LDPC_Code_LUT.cpp:643-697): the written
file is the reference's own format.  Checked three ways: byte-identical with the codec files the reference itself wrote
for the same decoders (tests/golden/*.codec.it.xz); the unmodified reference reads the written file back into the same
decoder (oracle/_ref/lutldpc_ref bundle); and the library's own reader gives the same decoder again."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import _golden
from lut_ldpc_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "lutldpc_ref")
CFGS = sorted(_golden.MANIFEST)


def _convert(src, dst):
    L = capi.load_library()
    L.lutldpc_convert_codec.argtypes = [C.c_char_p, C.c_char_p]
    rc = L.lutldpc_convert_codec(os.fsencode(src), os.fsencode(dst))
    assert rc == 0, (L.lutldpc_last_error(None) or b"").decode()


def _same_decoder(a, b):
    for k in a:
        if isinstance(a[k], np.ndarray):
            assert np.array_equal(a[k], b[k]), k
        else:
            assert a[k] == b[k], k


@pytest.mark.parametrize("cfg", CFGS)
def test_written_codec_file(cfg, tmp_path):
    src, out = _golden.bundle_path(cfg), str(tmp_path / "lut_codec.it")
    _convert(src, out)
    want = _golden.read_bundle(src)
    if "codec_sha256" in _golden.MANIFEST[cfg]:  # the reference's own save_code output for this decoder
        assert open(out, "rb").read() == open(_golden.codec_path(cfg), "rb").read()
    # own reader: codec file -> codec file is a fixed point
    again = str(tmp_path / "again.it")
    _convert(out, again)
    assert open(again, "rb").read() == open(out, "rb").read()
    if os.path.exists(REF):  # the unmodified reference loads it (load_code) and dumps what it holds
        back = str(tmp_path / "back.bundle")
        subprocess.run([REF, "bundle", out, back], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        _same_decoder(want, _golden.read_bundle(back))


def test_convert_reports_errors(tmp_path):
    L = capi.load_library()
    L.lutldpc_convert_codec.argtypes = [C.c_char_p, C.c_char_p]
    assert L.lutldpc_convert_codec(b"/nonexistent/file", os.fsencode(tmp_path / "x.it")) != 0
    assert b"cannot open" in L.lutldpc_last_error(None)
    assert L.lutldpc_convert_codec(os.fsencode(_golden.bundle_path("c2_irr_n500")), b"/nonexistent/dir/x.it") != 0


# ---- bp_codec.it (IT++ LDPC_Code::save_code, the [BP] path's codec file)
BP_SHA = __import__("json").load(open(os.path.join(_golden.GOLD, "bp_codec_sha256.json")))


@pytest.mark.parametrize("cfg", sorted(k for k in BP_SHA if k != "c2_gen"))
def test_bp_codec_file_equals_the_reference_written_one(cfg, tmp_path):
    """alist -> bp_codec.it: C, V, sumX1, sumX2, iind, jind as LDPC_Code::decoder_parameterization lays them out; the file
    has the hash of the one the reference wrote for the same matrix (tests/golden/make_bp_golden.py)"""
    import hashlib
    import _alist
    alist, out = str(tmp_path / "h.alist"), str(tmp_path / "bp_codec.it")
    _alist.write_alist(cfg, alist)
    capi.bp_convert_alist(alist, out)
    assert hashlib.sha256(open(out, "rb").read()).hexdigest() == BP_SHA[cfg]
    if os.path.exists(REF):  # and the unmodified reference loads it and writes the same bytes again
        back = str(tmp_path / "back.it")
        subprocess.run([REF, "bpreload", out, back], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        assert open(back, "rb").read() == open(out, "rb").read()
