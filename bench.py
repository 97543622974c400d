#!/usr/bin/env python3
"""bench.py -- decoded information Gbit/s of the batched LUT-LDPC decoder on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  (N > 1: launched by torchrun, one rank per GPU; frames shard, no data-path collective)

A step = one pass of the hot path (LDPC_Code_LUT::lut_decode for every frame of one batch) over one
batch of synthetic frames: all-zero codeword + BPSK/AWGN at --snr, quantised by the decoder's own
thresholds.  `value` is timed with the quantised inputs resident in HBM; `e2e` is the same batch
through the host-buffer C-ABI call (pinned host memory, H2D/D2H inside the timed region).
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name -> (golden cfg, default snr dB, default frames per step per GPU, description)
    "c5_n64800": ("c5_n64800", 1.3, 32768,
                  "rate0.50_dv02-08_dc07-08_lut_q4_N64800, LUT q=4/4, max_iter 50, psc=pisc=1"),
    "c3_irr_n10000": ("c3_irr_n10000", 1.3, 32768, "rate0.50_dv02-17_dc08-09_lut_q4_N10000, LUT q=4/4, max_iter 50"),
    "c1b_reg36_n1000": ("c1b_reg36_n1000", 2.0, 65536, "rate0.50_dv03_dc06_N1000 (3,6), LUT q=4/4, max_iter 50"),
    "c2_irr_n500": ("c2_irr_n500", 2.0, 65536, "rate0.50_dv02-17_dc08-09_lut_q4_N500, LUT q=4/4, max_iter 50 (params/ber.ini.irregular.example)"),
    "c4_reg632_high": ("c4_reg632_high", 4.5, 65536, "rate0.84_reg_v6c32_N2048 with trees/6_32_high.ini, LUT q=4/2, max_iter 8"),
    "c1a_reg632_wide": ("c1a_reg632_wide", 4.5, 65536, "rate0.84_reg_v6c32_N2048 with trees/6_32_wide.ini, LUT q=4/3, QCHA, max_iter 8 (params/ber.ini.regular.example)"),
}
SEC_PER_FRAME_EXTRA = {"c2_irr_n500": 0.013, "c4_reg632_high": 0.0064, "c1a_reg632_wide": 0.0115}
METRIC, UNIT = "decoded_info_gbit_per_s", "Gbit/s"


def common_config(workload, desc, snr):
    """`config` of the JSON line: identical in both arms (--impl b200 / reference) for the same command line."""
    return {"workload": "%s: %s" % (workload, desc), "snr_db": snr, "codeword": "all-zero (LDPC.zero_codeword=true)",
            "l2": "no flush needed: on the GPU arm the labels and message state of one step (frames x (2N + E/2) bytes, "
                  "several GB) are far larger than the 126 MB L2; the CPU arm decodes one frame per core at a time"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in o.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=3)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_inputs(torch, dec, F, snr_db, seed):
    """All-zero codeword over BPSK/AWGN, LLR = 4x/N0 (src/LDPC_BER_Sim.cpp:248-278), quantised on the
    device by the decoder's own quant_nonlin kernel.  Returns device uint8 [F][N] labels (cha, msg)."""
    N = dec.nvar
    N0 = 10.0 ** (-snr_db / 10.0) / dec.rate
    sigma = (N0 / 2.0) ** 0.5
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    cha = torch.empty((F, N), dtype=torch.uint8, device="cuda")
    msg = torch.empty((F, N), dtype=torch.uint8, device="cuda")
    chunk = max(1, min(F, (64 << 20) // N))
    for f0 in range(0, F, chunk):
        n = min(chunk, F - f0)
        llr = (1.0 + sigma * torch.randn((n, N), dtype=torch.float64, device="cuda", generator=g)) * (4.0 / N0)
        torch.cuda.current_stream().synchronize()
        dec.quantize_device(n, llr.data_ptr(), cha[f0:].data_ptr(), msg[f0:].data_ptr())
        torch.cuda.synchronize()
    return cha, msg


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import _golden
    import lut_ldpc_b200 as L

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg, snr_def, frames_def, desc = WORKLOADS[args.workload]
    snr = args.snr if args.snr is not None else snr_def
    F = args.frames or frames_def
    dec = L.Decoder(_golden.bundle_path(cfg))
    dec.set_exit_conditions(dec.max_iters, True, True)
    if _golden.MANIFEST[cfg]["init_mode"] == "qcha":
        dec.set_initial_message_mode(L.INIT_QCHA)
    if os.environ.get("LUTLDPC_POLL"):
        dec.set_poll_interval(int(os.environ["LUTLDPC_POLL"]))
    stream = torch.cuda.Stream()
    dec.set_stream(stream.cuda_stream)
    K, N, E = dec.ninfo, dec.nvar, dec.num_edges

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    with torch.cuda.stream(stream):
        cha, msg = make_inputs(torch, dec, F, snr, 1000 + rank)
    out = torch.empty((F, N), dtype=torch.uint8, device="cuda")
    its = torch.empty((F,), dtype=torch.int32, device="cuda")

    def step_device():
        dec.decode_quantized_device(F, cha.data_ptr(), msg.data_ptr(), out.data_ptr(), its.data_ptr())

    # ---- device-resident timing (value): K steps, no per-launch events
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    clocks = sampler.summary()
    launches = (dec.last_call_stats()["kernel_launches"] + 1) * args.steps  # + unpack; every step decodes the same batch
    # ---- the same K steps again with a CUDA event after every launch: per-kernel times for the roofline
    # (the events cost ~1 % of the step, so they stay out of the region `value` is taken from)
    dec.set_profiling(True)
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(stream)
    for _ in range(args.steps):
        step_device()
    p1.record(stream)
    torch.cuda.synchronize()
    profiled_ms = p0.elapsed_time(p1)
    ms = max_over_ranks(e0.elapsed_time(e1))
    prof = dec.get_profile()
    dec.set_profiling(False)
    it_host = its.cpu().numpy()
    wrong = out.sum(dim=1, dtype=torch.int64).cpu().numpy() != 0   # frames not decoded to the sent (zero) codeword
    iters_used = np.abs(it_host).astype(np.int64)
    sum_iters = int(iters_used.sum())
    vn_passes = int(np.where(it_host > 0, np.minimum(it_host, dec.max_iters - 1), np.where(it_host < 0, dec.max_iters - 1, 0)).sum())
    sum_iters_all = sum_over_ranks(sum_iters)
    frames_total = sum_over_ranks(F) * args.steps
    value = frames_total * K / (ms * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (algorithmic bytes / CUDA-event time, this rank)
    peak, peak_src = peaks()
    b_iter = 2 * E + N / 2.0
    dom = max(prof, key=lambda k: prof[k][0])
    per_frame_iter = {"cn_minsum": float(E)}
    for d, n in zip(dec.vn_class_degree, dec.vn_class_nodes):
        per_frame_iter["vn_dv%d" % d] = n * (d + 0.5)   # dv nibbles in + out, channel nibble (E + N/2 per pass)
    # the fused launch runs every specialised class: all their nodes in one kernel
    per_frame_iter["vn_fused"] = sum(v for k, v in per_frame_iter.items() if k.startswith("vn_dv"))
    units = {"cn_minsum": sum_iters}
    alg = per_frame_iter.get(dom, 0.0) * units.get(dom, vn_passes) * args.steps
    dom_ms = prof[dom][0]
    achieved = alg / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        # measured per launch with every frame of the profiled batch live: scale to this run's batch
        traffic = tj.get(dom)
        if traffic is not None and tj.get("frames_per_step"):
            traffic = int(traffic * F / tj["frames_per_step"])
    step_gbs = sum_iters * b_iter * args.steps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    if dom == "onchip":
        # on-chip regime (second bound): the same algorithmic message bytes, 2E + N/2 per frame-iteration,
        # move through shared memory instead of HBM; peak = 128 B per clock and SM at the maximum SM clock
        smem_peak = 148 * 128 * (clocks["sm_max_mhz"] or 1965.0) * 1e6 / 1e9
        alg = b_iter * sum_iters * args.steps
        achieved = alg / (dom_ms * 1e-3) / 1e9
        roofline = {"bound": "smem", "kernel": "k_onchip (whole lut_decode of a batch, state in shared memory)",
                    "achieved": round(achieved, 1), "peak": round(smem_peak, 1), "unit": "GB/s",
                    "frac": round(achieved / smem_peak, 4), "traffic": None,
                    "peak_source": "148 SMs x 128 B/clk shared-memory bandwidth x max SM clock (B200_PROFILING.md)",
                    "algorithmic_bytes": "SURVEY 8(d): 2E + N/2 message bytes per frame-iteration, through shared memory; "
                                         "LUT look-ups (one LDS.U8 per frame-look-up) come on top and are the actual limit: "
                                         "see profiles/ for lsu wavefronts and issue utilisation",
                    "hbm_bytes_per_frame": 3 * N + 4, "kernel_ms_per_launch": round(dom_ms / max(1, prof[dom][1]), 4),
                    "profile_ms": {k: round(v[0], 3) for k, v in prof.items()},
                    "kernel_share_of_step": {k: round(v[0] / sum(x[0] for x in prof.values()), 4) for k, v in prof.items()}}
    else:
        roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                  "frac": round(achieved / peak, 4), "traffic": traffic,
                  "traffic_source": "profiles/traffic.json (ncu --set full capture of one all-frames-live launch, static; "
                                    "scaled to this run's frames per step)",
                  "algorithmic_bytes": "SURVEY 8(d): check pass E, variable pass E + N/2 bytes per frame-pass "
                                       "(the N/8 hard-decision bytes a variable pass also writes are not counted)",
                  "peak_source": peak_src,
                  "kernel_ms_per_launch": round(dom_ms / max(1, prof[dom][1]), 4),
                  "whole_step_algorithmic_gbs": round(step_gbs, 1), "whole_step_frac": round(step_gbs / peak, 4),
                  "profiled_pass_ms_per_step": round(profiled_ms / args.steps, 3),
                  "profile_ms": {k: round(v[0], 3) for k, v in prof.items()},
                  "kernel_share_of_step": {k: round(v[0] / sum(x[0] for x in prof.values()), 4) for k, v in prof.items()}}

    # ---- end to end through the host-buffer C-ABI call (pinned host memory)
    cha_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True)
    msg_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True)
    cha_h.copy_(cha)
    msg_h.copy_(msg)
    torch.cuda.synchronize()
    del cha, msg, out
    cha_np, msg_np = cha_h.numpy(), msg_h.numpy()
    out_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True)
    its_h = torch.empty((F,), dtype=torch.int32, pin_memory=True)
    o_np, i_np = out_h.numpy(), its_h.numpy()
    e2e_steps = max(1, min(args.steps, 3))
    dec.decode_quantized(cha_np, msg_np, out=o_np, iters=i_np)  # warm-up (allocates the staging buffers)
    barrier()
    t0 = time.perf_counter()
    e0.record(stream)
    for _ in range(e2e_steps):
        dec.decode_quantized(cha_np, msg_np, out=o_np, iters=i_np)
    e1.record(stream)
    barrier()
    e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3))
    assert np.array_equal(i_np, it_host), "host-buffer path and device path disagree"
    e2e_value = sum_over_ranks(F) * e2e_steps * K / (e2e_ms * 1e-3) / 1e9

    # ---- the fused Monte-Carlo step, i.e. what ber_sim runs per SNR point (LDPC_BER_Sim::sim_snr_point,
    # src/LDPC_BER_Sim.cpp:246-311): Philox source/AWGN -> LLR -> quantise -> lut_decode -> BERC/BLERC counting, all on
    # the device through lutldpc_simulate_snr, driven by the ber_sim.py harness: every step ends in the all-reduce of
    # the 264 counters (NCCL for N > 1) and the global stop-rule test.  Same workload, SNR and frames per step.
    from lut_ldpc_b200.ber_sim import LDPC_BER_Sim_LUT, SimParams
    sp = SimParams(SNRdB=[snr], Nframes=F * world, Nfers=1 << 60, ber_min=0.0, fer_min=0.0, zero_codeword=True)
    harness = LDPC_BER_Sim_LUT(sp, rand_seed=20261105, decoder=dec, chunk_frames=F, dist=dist if world > 1 else None,
                               device="cuda" if world > 1 else None)
    first = 0
    for _ in range(max(1, min(args.warmup, 2))):
        harness.sim_snr_point(snr, first)
        first += F * world
    harness.t_simulate = harness.t_allreduce = 0.0
    sim_steps = max(1, min(args.steps, 3))
    barrier()
    t0 = time.perf_counter()
    sim_acc = None
    sim_dev_ms = 0.0
    for _ in range(sim_steps):
        acc, used, _stop = harness.sim_snr_point(snr, first)
        first += used
        sim_acc = acc if sim_acc is None else sim_acc + acc
        sim_dev_ms += dec.last_call_stats()["device_ms"]
    barrier()
    sim_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    dec.set_profiling(True)
    harness.sim_snr_point(snr, first)
    sim_prof = dec.get_profile()
    dec.set_profiling(False)
    sim = {"value": round(float(sim_acc[0]) * K / (sim_ms * 1e-3) / 1e9, 4), "unit": UNIT,
           "frames_per_s": round(float(sim_acc[0]) / (sim_ms * 1e-3), 1), "ms_per_step": round(sim_ms / sim_steps, 3),
           "device_ms_per_step_rank0": round(sim_dev_ms / sim_steps, 3), "steps": sim_steps,
           "fer": float(sim_acc[1]) / float(sim_acc[0]), "ber": float(sim_acc[3]) / float(sim_acc[2]),
           "mean_iterations": round(float(sim_acc[6]) / float(sim_acc[0]), 2),
           "allreduce_ms_per_step": round(1e3 * harness.t_allreduce / sim_steps, 3),
           "allreduce_share": round(1e3 * harness.t_allreduce / sim_ms, 5),
           "profile_ms_one_step": {k: round(v[0], 3) for k, v in sim_prof.items()},
           "call": "ber_sim.py LDPC_BER_Sim_LUT.sim_snr_point -> lutldpc_simulate_snr (Philox4x32-10 AWGN, quantiser, "
                   "lut_decode, error counting on the device) + all_reduce(SUM) of 264 int64 counters per step"}

    # ---- random codewords (north_star: "throughput on synthetic random-codeword frames").  The generator comes from
    # lutldpc_create_systematic: GF(2) Gauss-Jordan elimination of H on the device (LDPC_Generator_Systematic::construct;
    # the reference needs minutes for N=64800), a decoder for the column-permuted code, a bit-packed encoder.  Same
    # Monte-Carlo step as `sim` with LDPC.zero_codeword=false: source bits -> encode -> BPSK/AWGN -> ... -> count.
    t0 = time.perf_counter()
    sysd, _perm = dec.systematic()
    torch.cuda.synchronize()
    construct_s = time.perf_counter() - t0
    sysd.set_exit_conditions(sysd.max_iters, True, True)
    if _golden.MANIFEST[cfg]["init_mode"] == "qcha":
        sysd.set_initial_message_mode(L.INIT_QCHA)
    sysd.simulate_snr(snr, F, seed=99, first_frame=0, zero_codeword=False, batch_frames=F)
    rc_ms, rc_acc = 0.0, None
    for i in range(sim_steps):
        c = sysd.simulate_snr(snr, F, seed=99, first_frame=(i + 1) * F + rank * (sim_steps + 1) * F, zero_codeword=False,
                              batch_frames=F)
        rc_ms += sysd.last_call_stats()["device_ms"]
        rc_acc = c.as_array() if rc_acc is None else rc_acc + c.as_array()
    sysd.set_profiling(True)
    sysd.simulate_snr(snr, F, seed=98, zero_codeword=False, batch_frames=F)
    rc_prof = sysd.get_profile()
    sysd.set_profiling(False)
    rc_ms = max_over_ranks(rc_ms)
    random_cw = {"value": round(sum_over_ranks(F) * sim_steps * K / (rc_ms * 1e-3) / 1e9, 4), "unit": UNIT,
                 "ms_per_step": round(rc_ms / sim_steps, 3), "steps": sim_steps,
                 "fer": float(rc_acc[1]) / float(rc_acc[0]), "mean_iterations": round(float(rc_acc[6]) / float(rc_acc[0]), 2),
                 "generator_construction_s": round(construct_s, 3),
                 "profile_ms_one_step": {k: round(v[0], 3) for k, v in rc_prof.items() if v[0] >= 0.5},
                 "call": "lutldpc_create_systematic (rank + generator by Gauss-Jordan elimination on the device) + "
                         "lutldpc_simulate_snr with zero_codeword = 0 (device timing, rank-local frame ranges)"}
    sysd.close()

    # ---- the same batch through the packed host entry point (one byte per symbol in, one bit per decision out)
    both_h = torch.empty((F, N), dtype=torch.uint8, pin_memory=True)
    both_h.copy_(cha_h | (msg_h << 4))
    NB = (N + 7) // 8
    pk_h = torch.empty((F, NB), dtype=torch.uint8, pin_memory=True)
    both_np, pk_np = both_h.numpy(), pk_h.numpy()
    dec.decode_packed(both_np, out=pk_np, iters=i_np)  # warm-up
    barrier()
    t0 = time.perf_counter()
    e0.record(stream)
    for _ in range(e2e_steps):
        dec.decode_packed(both_np, out=pk_np, iters=i_np)
    e1.record(stream)
    barrier()
    pk_ms = max_over_ranks(max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3))
    assert np.array_equal(i_np, it_host), "packed host path and device path disagree"
    assert np.array_equal(np.packbits(o_np[:64], axis=1, bitorder="little"), pk_np[:64])
    e2e_packed = {"value": round(sum_over_ranks(F) * e2e_steps * K / (pk_ms * 1e-3) / 1e9, 4), "unit": UNIT,
                  "h2d_bytes_per_step": int(F * N) * world, "d2h_bytes_per_step": int(F * NB + 4 * F) * world,
                  "steps": e2e_steps,
                  "call": "lutldpc_decode_packed (pinned host: cha | msg << 4 per symbol in, bit-packed decisions + "
                          "iteration counts out): the same decode with 2.65x fewer bytes across PCIe"}

    # ---- the BP comparison decoder of ber_sim's [BP] section on the same code and SNR (IT++ LDPC_Code::bp_decode,
    # int32 QLLR messages, [BP] defaults except max_iter = the LUT decoder's): Monte-Carlo step on the device
    from lut_ldpc_b200 import capi as _capi
    _b = _golden.read_bundle(_golden.bundle_path(cfg))  # the graph of the same code
    bp_graph = dict(dv_vec=_b["dv_vec"], dc_vec=_b["dc_vec"], cn_msg_idx=_b["cn_msg_idx"],
                    nchk_lin_indep=N - K)  # same rate as the LUT decoder of this bundle
    bpd = _capi.BpDecoder(bp_graph, max_iters=dec.max_iters)
    Fbp = min(F, 8192)
    bpd.simulate_snr(snr, Fbp, seed=5, batch_frames=Fbp)
    bp_ms, bp_acc, bp_fi = 0.0, None, 0
    for i in range(sim_steps):
        c = bpd.simulate_snr(snr, Fbp, seed=5, first_frame=(i + 1) * Fbp + rank * (sim_steps + 1) * Fbp, batch_frames=Fbp)
        st = bpd.last_call_stats()
        bp_ms += st["device_ms"]
        bp_fi += st["frame_iterations"]
        bp_acc = c.as_array() if bp_acc is None else bp_acc + c.as_array()
    bp_ms = max_over_ranks(bp_ms)
    bp_bytes = float(bp_fi) * (4 * E + N) * 4  # per frame-iteration: every edge message read and written in both passes + the channel value
    bp = {"value": round(sum_over_ranks(Fbp) * sim_steps * K / (bp_ms * 1e-3) / 1e9, 4), "unit": UNIT,
          "ms_per_step": round(bp_ms / sim_steps, 3), "frames_per_step_per_gpu": Fbp, "steps": sim_steps,
          "fer": float(bp_acc[1]) / float(bp_acc[0]), "mean_iterations": round(float(bp_acc[6]) / float(bp_acc[0]), 2),
          "algorithmic_GBps_rank0": round(bp_bytes / (bp_ms * 1e-3) / 1e9, 1),
          "frac_of_hbm_peak": round(bp_bytes / (bp_ms * 1e-3) / 1e9 / peak, 4),
          "call": "lutldpc_bp_simulate_snr (Philox AWGN, to_qllr, bp_decode with per-frame early return and batch "
                  "narrowing, error counting on the device); (4 E + N) * 4 algorithmic bytes per frame-iteration"}
    bpd.close()

    line = {"metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u4 (8 codes per u32, SWAR)", "data": "synthetic",
            "config": common_config(args.workload, desc, snr),
            "step": {"frames_per_step_per_gpu": F, "frames_per_step": F * world,
                     "resident_bytes": "%.0f MB labels + %.0f MB messages per step and GPU" % (2e-6 * F * N, 1e-6 * F * E / 2),
                     "mean_iterations": round(sum_iters / F, 2), "decode_failures": int((it_host < 0).sum()),
                     "frame_errors": int(wrong.sum()), "undetected_frame_errors": int((wrong & (it_host >= 0)).sum()),
                     "edge_updates_per_s": round(2.0 * E * sum_iters_all * args.steps / (ms * 1e-3), 1)},
            "roofline": roofline, "sim": sim, "random_codeword": random_cw, "e2e_packed": e2e_packed, "bp": bp, "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": round(e2e_value, 4), "unit": UNIT, "h2d_bytes_per_step": int(2 * F * N) * world,
                    "d2h_bytes_per_step": int(F * N + 4 * F) * world, "steps": e2e_steps,
                    "call": "lutldpc_decode_quantized (pinned host u8 labels in, bits + iteration counts out; chunks pipelined over copy and decode streams)"}}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.workload, snr, seconds=args.cpu_seconds)
    if world > 1:
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# CPU side: the reference's own decoder (oracle/_ref/lutldpc_ref, compiled unmodified from the
# reference sources) running the sim_snr_point frame loop, one process per host core with seeds
# 0..P-1 (the reference's own way of using several cores, prog/ber_sim.cpp:67).  If the binary is not
# there, the plain-C port in oracle/ is timed instead ("kind": "port").
# ------------------------------------------------------------------------------------------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "lutldpc_ref")
SEC_PER_FRAME = {"c5_n64800": 1.9, "c3_irr_n10000": 0.7, "c1b_reg36_n1000": 0.025, **SEC_PER_FRAME_EXTRA}  # SURVEY.md 6.2, one core


def _ref_sample(workload, snr, frames_per_proc, procs, seed0=0):
    import _golden
    cfg = WORKLOADS[workload][0]
    codec = _golden.codec_path(cfg)
    cmds = [[REF_BIN, "sim", codec, "--snr", str(snr), "--frames", str(frames_per_proc), "--seed", str(seed0 + i),
             "--count-iters"] for i in range(procs)]
    ps = [subprocess.Popen(c, stdout=subprocess.PIPE, text=True) for c in cmds]
    rows = []
    for p in ps:
        o, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("lutldpc_ref sim failed")
        rows.append(json.loads(o.strip().splitlines()[-1]))
    frames = sum(r["frames"] for r in rows)
    wall = max(r["wall_s"] for r in rows)  # frame loops only; codec load is reported separately
    return frames, wall, rows


def _port_worker(a):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import pyoracle
    bundle, snr, frames, seed, qcha = a
    orc = pyoracle.Oracle(bundle)
    orc.set_exit_conditions(orc.max_iters, True, True)
    rate = orc.ninfo / orc.nvar
    N0 = 10.0 ** (-snr / 10.0) / rate
    rng = np.random.default_rng(seed)
    llr = (1.0 + (N0 / 2) ** 0.5 * rng.standard_normal((frames, orc.nvar))) * 4.0 / N0
    cha, msg = orc.quantise(llr, pyoracle.INIT_QCHA if qcha else pyoracle.INIT_CONT)
    t0 = time.perf_counter()
    orc.lut_decode_batch(cha, msg)
    return time.perf_counter() - t0


def _port_sample(workload, snr, frames_per_proc, procs, seed0=0):
    import multiprocessing as mp
    import _golden
    bundle = _golden.bundle_path(WORKLOADS[workload][0])
    qcha = _golden.MANIFEST[WORKLOADS[workload][0]]["init_mode"] == "qcha"
    with mp.get_context("spawn").Pool(procs) as pool:
        walls = pool.map(_port_worker, [(bundle, snr, frames_per_proc, seed0 + i, qcha) for i in range(procs)])
    return frames_per_proc * procs, max(walls), []


def _cpu_kind(workload):
    """The compiled reference when it and the workload's reference-written codec file are there, else the C port."""
    import _golden
    have = os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK) and "codec_sha256" in _golden.MANIFEST[WORKLOADS[workload][0]]
    return ("reference", _ref_sample) if have else ("port", _port_sample)


def _ninfo(workload):
    import _golden
    return _golden.MANIFEST[WORKLOADS[workload][0]]["design_report"]["ninfo"]


def cpu_baseline(workload, snr, seconds=15.0):
    kind, fn = _cpu_kind(workload)
    procs = os.cpu_count() or 1
    per = max(1, int(seconds / SEC_PER_FRAME.get(workload, 1.0)))
    frames, wall, _ = fn(workload, snr, per, procs)
    return {"value": round(frames * _ninfo(workload) / wall / 1e9, 9), "unit": UNIT, "cores": procs, "kind": kind,
            "sample": "%d frames (%d per process, seeds 0..%d) of the same workload at %.2f dB, frame loop wall time %.2f s"
                      % (frames, per, procs - 1, snr, wall)}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores."""
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    cfg, snr_def, frames_def, desc = WORKLOADS[args.workload]
    snr = args.snr if args.snr is not None else snr_def
    kind, fn = _cpu_kind(args.workload)
    procs = os.cpu_count() or 1
    per = max(1, int(args.ref_seconds_per_step / SEC_PER_FRAME.get(args.workload, 1.0)))
    for i in range(args.warmup):
        fn(args.workload, snr, 1, procs, seed0=1000 * (i + 1))
    frames = wall = 0.0
    for i in range(args.steps):
        f, w, _ = fn(args.workload, snr, per, procs, seed0=100 * i)
        frames += f
        wall += w
    value = frames * _ninfo(args.workload) / wall / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 9), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * wall / args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32 (IT++ ivec)",
            "data": "synthetic",
            "config": common_config(args.workload, desc, snr),
            "step": {"frames_per_step": per * procs},
            "cpu_baseline": {"value": round(value, 9), "unit": UNIT, "cores": procs, "kind": kind,
                             "sample": "%d frames per step (%d per process x %d processes, distinct seeds), "
                                       "frame-loop wall time only" % (per * procs, per, procs)},
            "e2e": {"value": round(value, 9), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c5_n64800", choices=sorted(WORKLOADS))
    ap.add_argument("--snr", type=float, default=None)
    ap.add_argument("--frames", type=int, default=0, help="frames per step per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--ref-seconds-per-step", type=float, default=6.0)
    args = ap.parse_args()
    # ONE JSON line on stdout: libraries that write to file descriptor 1 themselves (NCCL prints its version banner
    # there) are sent to stderr; the line goes to the saved descriptor.
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(saved, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
