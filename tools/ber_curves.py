#!/usr/bin/env python3
"""BER/FER curve of one golden decoder over an SNR grid with the multi-GPU Monte-Carlo harness (ber_sim.py: disjoint Philox
frame ranges per rank, lutldpc_simulate_snr, NCCL all-reduce of the counters after every chunk, global stop rules as
LDPC_BER_Sim::run / sim_snr_point).  Launch with torchrun (one rank per GPU) or plain python (one GPU).
  ber_curves.py <cfg> <SNRdB vec, e.g. 0:.5:4> <Nframes> <Nfers> <ber_min> <fer_min> <chunk per GPU> <out.json> [bp [max_iter]]
A trailing `bp` runs the BP comparison decoder of the [BP] section on the same code and the same noise (same seed and
frame indices) instead of the LUT decoder."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _golden  # noqa: E402
from lut_ldpc_b200 import ber_sim as bs  # noqa: E402


def main():
    cfg, grid = sys.argv[1], sys.argv[2]
    nframes, nfers, ber_min, fer_min = int(float(sys.argv[3])), int(float(sys.argv[4])), float(sys.argv[5]), float(sys.argv[6])
    chunk, out = int(sys.argv[7]), sys.argv[8]
    local, world = int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    man = _golden.MANIFEST[cfg]
    p = bs.SimParams(None, SNRdB=bs.parse_vec(grid), Nframes=nframes, Nfers=nfers, ber_min=ber_min, fer_min=fer_min,
                     codec_filename=_golden.bundle_path(cfg), zero_codeword=True, max_iter=0,
                     initial_message_mode="from_quantized_channel_llrs" if man["init_mode"] == "qcha" else "from_continuous_input")
    sim = bs.LDPC_BER_Sim_LUT(p, rand_seed=2026, chunk_frames=chunk, dist=dist if world > 1 else None,
                              device="cuda" if world > 1 else None)
    from lut_ldpc_b200.capi import BpDecoder, Decoder, INIT_CONT, INIT_QCHA
    use_bp = len(sys.argv) > 9 and sys.argv[9] == "bp"
    if use_bp:
        b = _golden.read_bundle(p.codec_filename)
        sim.dec = BpDecoder(dict(dv_vec=b["dv_vec"], dc_vec=b["dc_vec"], cn_msg_idx=b["cn_msg_idx"],
                                 nchk_lin_indep=b["nchk_lin_indep"]),
                            max_iters=int(sys.argv[10]) if len(sys.argv) > 10 else b["max_iters"])
    else:
        sim.dec = Decoder(p.codec_filename)
        sim.dec.set_exit_conditions(sim.dec.max_iters, True, True)
        sim.dec.set_initial_message_mode(INIT_QCHA if man["init_mode"] == "qcha" else INIT_CONT)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = sim.run(verbose=int(os.environ.get("RANK", 0)) == 0)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    r.update(cfg=cfg, decoder="bp (IT++ bp_decode, QLLR 12/300/7/28)" if use_bp else "lut", design=man["design"], world=world, wall_s=round(wall, 3), Nframes=nframes, Nfers=nfers, ber_min=ber_min,
             fer_min=fer_min, chunk_frames_per_gpu=chunk, allreduces=sim.n_allreduce,
             allreduce_s=round(sim.t_allreduce, 4), simulate_s=round(sim.t_simulate, 3),
             ninfo=sim.dec.ninfo, nvar=sim.dec.nvar,
             ber=[(e / b if b else None) for e, b in zip(r["sim_data_bit_errors"], r["sim_Ndatabits"])],
             fer=[(e / n if n else None) for e, n in zip(r["sim_frame_errors"], r["sim_Nframes"])],
             info_gbit_s=round(sum(r["sim_Ndatabits"]) / wall / 1e9, 3))
    if world > 1:
        dist.destroy_process_group()
    if int(os.environ.get("RANK", 0)) == 0:
        os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
        json.dump(r, open(out, "w"), indent=1)
        print(json.dumps({k: r[k] for k in ("cfg", "world", "wall_s", "info_gbit_s", "sim_SNRdB", "sim_Nframes", "sim_frame_errors", "ber")}))


if __name__ == "__main__":
    main()
