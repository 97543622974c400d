"""The reference-side binding of INTEGRATION.md, compiled against the UNMODIFIED reference and run: oracle/_ref/
b200_binding_test holds an `itpp::Channel_Code* C` the way LDPC_BER_Sim does (src/LDPC_BER_Sim.hpp:159) and runs the
reference's frame loop (src/LDPC_BER_Sim.cpp:260-290) twice with the same IT++ seed -- C = the reference's
LDPC_Code_LUT, then C = LDPC_Code_LUT_B200 (oracle/b200_binding/LDPC_Code_LUT_B200.hpp over liblutldpc_b200.so) -- for
three exit configurations, and requires bit-identical bitsout for every frame.  Also drives the C++ facade
lut_ldpc_b200/host/ldpc_code_lut.hpp (decode / encode / lut_decode / syndrome_check) against a reference fixture."""
import json
import os
import subprocess

import numpy as np
import pytest

import _golden
import pyoracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "b200_binding_test")


from _alist import write_alist  # noqa: E402


@pytest.mark.parametrize("cfg,args", [
    # (3,6) N=1000, all-zero codeword (LDPC.zero_codeword=true)
    ("c1b_reg36_n1000", "--design-thr 0.88 --max-iter 50 --qcha 4 --qmsg 4 --snr 1.8 --frames 24 --seed 5"),
    # irregular N=500 as params/ber.ini.irregular.example: random codewords through the systematic generator
    ("c2_irr_n500", "--design-thr 0.88 --max-iter 50 --qcha 4 --qmsg 4 --gen --snr 1.5 --frames 24 --seed 6"),
])
def test_reference_frame_loop_through_the_channel_code_binding(tmp_path, cfg, args):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/b200_binding_test not built (needs /root/reference at build time)")
    alist = str(tmp_path / (cfg + ".alist"))
    write_alist(cfg, alist)
    p = subprocess.run([BIN, alist] + args.split(), capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-2000:])
    r = json.loads(p.stdout.strip().splitlines()[-1])
    assert r["frames"] == 72 and r["mismatched_frames"] == 0
    assert r["random_codewords"] == (1 if "--gen" in args else 0)
    assert r["cpu_data_bit_errors"] > 0  # the operating point is noisy enough to exercise failing frames


def test_cpp_facade_against_reference_fixture(tmp_path):
    """host/ldpc_code_lut.hpp: Channel_Code::decode(vec), encode, lut_decode, syndrome_check, one frame per call like the
    reference's callers, against a reference fixture (random codewords)."""
    cfg, tag = "c2_irr_n500", "snr1.0"
    exe = str(tmp_path / "facade_check")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-o", exe, os.path.join(ROOT, "tests", "facade_check.cpp"),
                           "-I" + ROOT, "-L" + os.path.join(ROOT, "lut_ldpc_b200"), "-llutldpc_b200",
                           "-Wl,-rpath," + os.path.join(ROOT, "lut_ldpc_b200")])
    p = subprocess.run([exe, _golden.bundle_path(cfg), _golden.fixture_path(cfg, tag)], capture_output=True, text=True,
                       timeout=600)
    assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-2000:])
    r = json.loads(p.stdout.strip().splitlines()[-1])
    assert r["frames"] == 8 and r["mismatches"] == 0 and r["checked"] >= 5 * 8
