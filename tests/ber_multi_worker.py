"""Worker of tests/test_gpu_ber.py::test_multi_gpu_harness_is_split_invariant, launched by torchrun with one rank per
GPU: the ber_sim.py harness over NCCL (Philox frame ranges per rank + lutldpc_simulate_snr + all_reduce of the
counters + global stop rules).  Rank 0 prints the results as one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _golden  # noqa: E402
from lut_ldpc_b200 import ber_sim as bs  # noqa: E402


def main():
    cfg, nframes, nfers, chunk = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    snrs = [float(x) for x in sys.argv[5].split(",")]
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    p = bs.SimParams(None, SNRdB=snrs, Nframes=nframes, Nfers=nfers, codec_filename=_golden.bundle_path(cfg),
                     zero_codeword=True, max_iter=50, fer_min=0.0, ber_min=0.0)
    sim = bs.LDPC_BER_Sim_LUT(p, rand_seed=11, chunk_frames=chunk, dist=dist if world > 1 else None,
                              device="cuda" if world > 1 else None).load()
    r = sim.run(verbose=False)
    r["world"] = world
    r["allreduces"] = sim.n_allreduce
    if world > 1:
        dist.destroy_process_group()
    if int(os.environ.get("RANK", 0)) == 0:
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
