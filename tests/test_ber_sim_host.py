"""Host logic of the Monte-Carlo harness (lut_ldpc_b200/ber_sim.py): .ini surface, frame sharding,
global stop rules, and the world_size-2 path over gloo (CPU)."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from lut_ldpc_b200 import ber_sim as bs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

INI = """
; comment
[Sim]
   SNRdB = 0:.5:2
   Nframes = 1e3
   Nfers = 20
   codec_filename = x.bundle
[LDPC]
   parity_filename = rate0.50_dv02-17_dc08-09_lut_q4_N500
   zero_codeword = false
[LUT]
   max_iter = 50
   design_thr = 0.88 ; trailing comment
   initial_message_mode = from_quantized_channel_llrs
"""


def test_parse_vec_itpp_syntax():
    assert bs.parse_vec("0:.5:4") == [0, .5, 1, 1.5, 2, 2.5, 3, 3.5, 4]
    assert bs.parse_vec("3:.5:6")[-1] == 6.0
    assert bs.parse_vec("[1 2.5, 4]") == [1, 2.5, 4]
    assert bs.parse_vec("1:3") == [1, 2, 3]


def test_ini_defaults_and_keys(tmp_path):
    f = tmp_path / "p.ini"
    f.write_text(INI)
    p = bs.SimParams(str(f))
    assert p.SNRdB == [0, .5, 1, 1.5, 2] and p.Nframes == 1000 and p.Nfers == 20
    assert p.ber_min == 1e-7 and p.fer_min == 1e-5 and p.rand_seed_offset == 0          # reference defaults
    assert p.zero_codeword is False and p.parity_check_iter is True and p.max_iter == 50
    assert p.initial_message_mode == "from_quantized_channel_llrs" and p.has_lut_section


def test_split_frames_is_a_partition():
    for todo in (0, 1, 7, 8192, 10001):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                lo, n = bs.split_frames(todo, 500, r, world)
                got.extend(range(lo, lo + n))
            assert got == list(range(500, 500 + todo))


def _fake(fer):
    """Deterministic per-frame outcome keyed by the global frame index (what Philox gives the GPU path)."""
    def sim(snr, n, first):
        a = np.zeros(bs.N_COUNTERS, np.int64)
        idx = np.arange(first, first + n, dtype=np.uint64)
        h = (idx * np.uint64(2654435761) + np.uint64(int(snr * 1000))) % np.uint64(1000)
        err = h < np.uint64(int(fer(snr) * 1000))
        a[0], a[1], a[2], a[3] = n, err.sum(), n * 100, 3 * err.sum()
        a[4], a[5] = n * 200, n * 7
        return a
    return sim


def test_stop_rules_single_rank():
    p = bs.SimParams(None, SNRdB=[1.0, 2.0, 3.0, 4.0], Nframes=4000, Nfers=50, ber_min=1e-7, fer_min=1e-5)
    sim = bs.LDPC_BER_Sim_LUT(p, simulate=_fake(lambda s: {1.0: 0.5, 2.0: 0.01, 3.0: 0.0}.get(s, 0.0)), chunk_frames=256)
    r = sim.run(verbose=False)
    assert r["sim_frame_errors"][0] > 50 and r["sim_Nframes"][0] < 4000      # Nfers: '>' at chunk granularity
    assert r["sim_Nframes"][1] == 4000                                       # ran to Nframes
    assert r["sim_Nframes"][2] == 4000 and r["sim_frame_errors"][2] == 0     # FER < fer_min recorded, then stop
    assert r["sim_Nframes"][3] == 0                                          # later points stay empty


WORKER = textwrap.dedent("""
    import json, os, sys
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
    import torch.distributed as dist
    from lut_ldpc_b200 import ber_sim as bs
    from test_ber_sim_host import _fake
    dist.init_process_group("gloo")
    p = bs.SimParams(None, SNRdB=[1.0, 2.0, 3.0], Nframes=3000, Nfers=40, ber_min=1e-7, fer_min=1e-5)
    sim = bs.LDPC_BER_Sim_LUT(p, simulate=_fake(lambda s: {1.0: 0.3, 2.0: 0.02}.get(s, 0.0)), chunk_frames=128, dist=dist)
    r = sim.run(verbose=False)
    if dist.get_rank() == 0:
        print("RESULT " + json.dumps(r))
    dist.destroy_process_group()
""")


def test_world_size_2_matches_world_size_1(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % (ROOT, ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29611")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29611", str(script)],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    got = json.loads([l for l in out.stdout.splitlines() if l.startswith("RESULT ")][0][7:])
    p = bs.SimParams(None, SNRdB=[1.0, 2.0, 3.0], Nframes=3000, Nfers=40, ber_min=1e-7, fer_min=1e-5)
    # a single rank that consumes frames in the same global chunks (2 x 128) sees the same frames
    one = bs.LDPC_BER_Sim_LUT(p, simulate=_fake(lambda s: {1.0: 0.3, 2.0: 0.02}.get(s, 0.0)), chunk_frames=256)
    want = one.run(verbose=False)
    assert got == want


def test_chunks_follow_the_fer_estimate_so_the_nfers_stop_overshoots_little():
    """The reference stops at the first frame with more than Nfers frame errors (src/LDPC_BER_Sim.cpp:289); the GPU
    path can only stop between chunks, so chunks are sized from the running FER estimate."""
    assert bs.next_chunk(0, 0, 20, 8192) == 1344                 # small first chunk: 64 x (Nfers + 1)
    assert bs.next_chunk(0, 0, 10 ** 15, 8192) == 8192           # no Nfers rule: always the maximal chunk
    assert bs.next_chunk(1344, 700, 20, 8192) == 1024            # already past the limit at high FER: minimal chunk
    assert bs.next_chunk(1344, 10, 20, 8192) == 1849             # 11 errors missing at FER 10/1344 -> 1.25 x 1478
    assert bs.next_chunk(8192, 0, 20, 8192) == 8192              # no error seen yet: maximal chunk
    # a simulated point: FER 0.5, Nfers 20 -> the reference would stop after ~42 frames; one small chunk here
    p = bs.SimParams(None, SNRdB=[1.0], Nframes=10 ** 6, Nfers=20, fer_min=0.0, ber_min=0.0)
    sim = bs.LDPC_BER_Sim_LUT(p, simulate=_fake(lambda s: 0.5), chunk_frames=8192)
    acc, used, _ = sim.sim_snr_point(1.0, 0)
    assert acc[1] > 20 and used == 1344


BP_INI = """
[Sim]
   SNRdB = 1:1:3
[LDPC]
   parity_filename = somecode
[BP]
   max_iter = 40
   qllr_table_size = 0   ; min-sum
"""


def test_bp_section_selects_the_bp_harness(tmp_path):
    """prog/ber_sim.cpp:136-147: [LUT] wins, else [BP], else it_error; [BP] defaults of src/LDPC_BER_Sim.cpp:74-78"""
    f = tmp_path / "bp.ini"
    f.write_text(BP_INI)
    p = bs.SimParams(str(f))
    assert p.has_bp_section and not p.has_lut_section
    assert (p.bp_max_iter, p.qllr_scale_res, p.qllr_table_size, p.qllr_spacing_res, p.qllr_total_res) == (40, 12, 0, 7, 28)
    assert type(bs.make_sim(p)) is bs.LDPC_BER_Sim
    f.write_text(INI)
    assert type(bs.make_sim(bs.SimParams(str(f)))) is bs.LDPC_BER_Sim_LUT
    f.write_text("[Sim]\nSNRdB = 1\n")
    with pytest.raises(RuntimeError, match="type of decoder"):
        bs.make_sim(bs.SimParams(str(f)))
    f.write_text("[Sim]\nSNRdB = 1\ncodec_type = BP\n")
    assert type(bs.make_sim(bs.SimParams(str(f)))) is bs.LDPC_BER_Sim
    # the shared run loop works for the BP harness too (stop rules, chunking)
    f.write_text(BP_INI)
    sim = bs.make_sim(bs.SimParams(str(f), Nframes=3000, Nfers=10), simulate=_fake(lambda s: {1.0: 0.2}.get(s, 0.0)), chunk_frames=256)
    r = sim.run(verbose=False)
    assert r["sim_frame_errors"][0] > 10 and r["sim_Nframes"][1] == 3000 and r["sim_Nframes"][2] == 0
