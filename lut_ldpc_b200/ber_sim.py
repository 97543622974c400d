"""Host-side mirror of the reference's Monte-Carlo harness for the LUT decoder
(LDPC_BER_Sim / LDPC_BER_Sim_LUT, src/LDPC_BER_Sim.{hpp,cpp}), with the frame loop of
sim_snr_point (:246-311) running on the GPU through lutldpc_simulate_snr and the frames of every SNR
point sharded over the ranks of a torch.distributed job (one process per GPU):

  * rank r simulates disjoint Philox frame ranges -- frames are independent trials, no data moves;
  * after every chunk ONE all_reduce(SUM) of the counter vector (NCCL over NVLink on GPUs, gloo in the
    CPU tests) makes the stop rules of the reference global:
        frames >= Nframes, frame_errors > Nfers           (:260, :289)
        BER < ber_min or FER < fer_min ends the sweep      (:307)
    Every rank sees the same sums, so every rank takes the same decision without further messages.

Params come from the reference's own .ini (sections [Sim], [LDPC], [LUT]; keys and defaults as
src/LDPC_BER_Sim.cpp:53-78,386-413).  LUT design is not part of this path: Sim.codec_filename names a
designed decoder, i.e. the lut_codec.it a reference run saved (LDPC_Code_LUT::save_code) or a bundle.
"""
import configparser
import math
import os
import time

import numpy as np

from .capi import ITER_HIST_MAX, SimCounters

N_COUNTERS = 8 + ITER_HIST_MAX


def parse_vec(s):
    """IT++ vec string syntax used by Sim.SNRdB: 'a:step:b', 'a:b', '[a b c]', 'a,b,c'."""
    out = []
    for tok in s.replace("[", " ").replace("]", " ").replace(",", " ").replace(";", " ").split():
        if ":" not in tok:
            out.append(float(tok))
            continue
        p = [float(x) for x in tok.split(":")]
        a, step, b = (p[0], 1.0, p[1]) if len(p) == 2 else p
        n = int(math.floor((b - a) / step + 1e-9)) if step else 0
        out.extend(a + i * step for i in range(n + 1))
    return out


def _flag(v):
    return str(v).strip().lower() in ("1", "true")


class SimParams:
    """[Sim]/[LDPC]/[LUT] keys with the reference's defaults."""

    def __init__(self, path=None, **over):
        cp = configparser.ConfigParser(inline_comment_prefixes=(";",), comment_prefixes=(";", "#"), strict=False)
        cp.optionxform = str
        if path:
            with open(path) as f:
                cp.read_file(f)
        g = lambda s, k, d: cp.get(s, k, fallback=d) if cp.has_section(s) else d
        self.SNRdB = parse_vec(g("Sim", "SNRdB", "")) if g("Sim", "SNRdB", "") else []
        self.Nframes = int(float(g("Sim", "Nframes", 1e5)))
        self.Nfers = int(float(g("Sim", "Nfers", 20)))
        self.ber_min = float(g("Sim", "ber_min", 1e-7))
        self.fer_min = float(g("Sim", "fer_min", 1e-5))
        self.rand_seed_offset = int(float(g("Sim", "rand_seed_offset", 0)))
        self.results_prefix = g("Sim", "results_prefix", "RES")
        self.results_dir = g("Sim", "results_dir", "results")
        self.custom_name = g("Sim", "custom_name", "")
        self.codec_filename = g("Sim", "codec_filename", "")
        self.save_codec = int(float(g("Sim", "save_codec", 0)))
        self.zero_codeword = _flag(g("LDPC", "zero_codeword", "true"))
        self.parity_check_iter = _flag(g("LDPC", "parity_check_iter", "true"))
        self.max_iter = int(float(g("LUT", "max_iter", 30)))
        # design keys (src/LDPC_BER_Sim.cpp:386-413), used when no codec file is named
        self.codes_dir = g("Sim", "codes_dir", "codes")
        self.parity_filename = g("LDPC", "parity_filename", "")
        self.design_thr = float(g("LUT", "design_thr", 0)) if g("LUT", "design_thr", None) is not None else None
        self.design_SNRdB = float(g("LUT", "design_SNRdB", 0)) if g("LUT", "design_SNRdB", None) is not None else None
        self.qbits_channel = int(float(g("LUT", "qbits_channel", 4)))
        self.qbits_message_uniform = int(float(g("LUT", "qbits_message_uniform", 3)))
        self.qbits_messages = [int(x) for x in parse_vec(g("LUT", "qbits_messages", ""))] if g("LUT", "qbits_messages", "") else None
        self.reuse_lut = [int(x) for x in parse_vec(g("LUT", "reuse_lut", ""))] if g("LUT", "reuse_lut", "") else None
        self.tree_mode = g("LUT", "tree_mode", "auto_bin_balanced")
        self.trees_dir = g("LUT", "trees_dir", "trees")
        self.trees_filename = g("LUT", "trees_filename", "")
        self.min_lut = _flag(g("LUT", "min_lut", "true"))
        self.initial_message_mode = g("LUT", "initial_message_mode", "from_continuous_input")
        self.has_lut_section = cp.has_section("LUT") or g("Sim", "codec_type", "none") == "LUT"
        # [BP]: the comparison decoder (src/LDPC_BER_Sim.cpp:74-78)
        self.bp_max_iter = int(float(g("BP", "max_iter", 30)))
        self.qllr_scale_res = int(float(g("BP", "qllr_scale_res", 12)))
        self.qllr_table_size = int(float(g("BP", "qllr_table_size", 300)))
        self.qllr_spacing_res = int(float(g("BP", "qllr_spacing_res", 7)))
        self.qllr_total_res = int(float(g("BP", "qllr_total_res", 28)))
        self.has_bp_section = cp.has_section("BP") or g("Sim", "codec_type", "none") == "BP"
        for k, v in over.items():
            setattr(self, k, v)


def next_chunk(frames, frame_errors, nfers, chunk_max, chunk_min=1024):
    """Frames of the next chunk of an SNR point.  The reference stops at the first frame after which more than Nfers
    frame errors have been seen (src/LDPC_BER_Sim.cpp:289); chunks are sized from the running FER estimate so that the
    GPU path overshoots that point by a fraction of a chunk instead of a whole maximal one: a small first chunk, then
    about 1.25 x the frames expected to bring the missing frame errors."""
    if nfers >= (1 << 40):
        return chunk_max
    if frames == 0:
        return min(chunk_max, max(chunk_min, 64 * (nfers + 1)))
    fer = max(frame_errors, 0.5) / frames
    want = int(1.25 * (nfers + 1 - frame_errors) / fer) + 1
    return max(min(chunk_min, chunk_max), min(chunk_max, want))


def split_frames(todo, first_frame, rank, world):
    """Contiguous, disjoint share of `todo` frames starting at global index first_frame."""
    per = (todo + world - 1) // world
    lo = min(todo, rank * per)
    hi = min(todo, lo + per)
    return first_frame + lo, hi - lo


class LDPC_BER_Sim_LUT:
    """load() / run() / results like the reference class; `simulate` is the per-rank GPU call
    (snr_db, n_frames, first_frame) -> int64 counter vector, by default Decoder.simulate_snr."""

    def __init__(self, params, rand_seed=0, decoder=None, simulate=None, chunk_frames=8192, dist=None, device=None):
        self.p, self.rand_seed, self.dec = params, int(rand_seed), decoder
        self.chunk, self.dist, self.device = int(chunk_frames), dist, device
        self._sim = simulate
        self.points = []
        self.t_simulate = self.t_allreduce = 0.0  # host seconds spent in the per-rank GPU calls / the counter all-reduce
        self.n_allreduce = 0
        self.rank = dist.get_rank() if dist is not None and dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1

    def load(self, basedir="."):
        """LDPC_BER_Sim_LUT::load (src/LDPC_BER_Sim.cpp:434-550) minus the design step."""
        from .capi import Decoder, INIT_CONT, INIT_QCHA, alist_graph, design_luts
        p = self.p
        if p.codec_filename:
            path = p.codec_filename if os.path.isabs(p.codec_filename) else os.path.join(basedir, p.codec_filename)
            self.dec = Decoder(path)
        else:
            # alist -> graph (+ rank) -> design_luts -> decoder, and for LDPC.zero_codeword = false the systematic
            # generator: LDPC_BER_Sim_LUT::load, src/LDPC_BER_Sim.cpp:434-507
            if not p.parity_filename:
                raise RuntimeError("LDPC_BER_Sim_LUT::load(): LDPC.parity_filename (or Sim.codec_filename) is required")
            if p.design_thr is None and p.design_SNRdB is None:
                raise RuntimeError("LDPC_BER_Sim_LUT::load(): Either design_thr or design_SNRdB must be specified")
            alist = os.path.join(basedir, p.codes_dir, p.parity_filename + ".alist")
            nvar = alist_graph(alist)["nvar"]
            graph = alist_graph(alist, rank_mode=0 if nvar >= 100000 else (2 if nvar > 20000 else 1))
            rank = graph["nchk_lin_indep"] if graph["nchk_lin_indep"] >= 0 else graph["nchk"]
            rate = 1.0 - rank / graph["nvar"]
            sigma2 = p.design_thr ** 2 if p.design_thr is not None else 10.0 ** (-p.design_SNRdB / 10) / (2 * rate)
            nq = [1 << q for q in (p.qbits_messages or [p.qbits_message_uniform] * p.max_iter)]
            method = "auto_bin_balanced"  # the other auto modes behave like it in the reference's ber_sim (:487-488)
            if p.tree_mode == "file":
                method = "filename=" + os.path.join(basedir, p.trees_dir, p.trees_filename)
            design = design_luts(graph["dv_vec"], graph["dc_vec"], p.max_iter, 1 << p.qbits_channel, nq, sigma2,
                                 reuse_vec=p.reuse_lut, min_lut=p.min_lut, tree_method=method)
            self.dec = Decoder.from_design(graph, design, p.max_iter, 1 << p.qbits_channel, nq, reuse_vec=p.reuse_lut,
                                           min_lut=p.min_lut)
            if not p.zero_codeword:
                self.dec = self.dec.systematic()[0]
            if self.rand_seed == p.save_codec and self.rank == 0:  # src/LDPC_BER_Sim.cpp:511-520
                d = os.path.join(basedir, p.results_dir, self.gen_filename())
                os.makedirs(d, exist_ok=True)
                self.dec.save_code(os.path.join(d, "lut_codec.it"))
        self.dec.set_exit_conditions(self.p.max_iter, self.p.parity_check_iter, self.p.parity_check_iter)  # :500
        self.dec.set_initial_message_mode(INIT_QCHA if self.p.initial_message_mode == "from_quantized_channel_llrs"
                                          else INIT_CONT)
        return self

    def gen_filename(self):
        """LDPC_BER_Sim_LUT::gen_filename, src/LDPC_BER_Sim.cpp:553-568 (the rate as an ostream prints a double)"""
        p = self.p
        return "%s_N%d_R%s_maxIter%d_zcw%d_frames%d%s%s" % (p.results_prefix, self.dec.nvar, "%g" % self.dec.rate, p.max_iter,
                                                           int(p.zero_codeword), p.Nframes, "_minLUT" if p.min_lut else "_LUT",
                                                           p.custom_name)

    def _simulate(self, snr, n, first):
        if self._sim is not None:
            return np.asarray(self._sim(snr, n, first), dtype=np.int64)
        if n <= 0:
            return np.zeros(N_COUNTERS, np.int64)
        c = self.dec.simulate_snr(snr, n, seed=self.rand_seed + self.p.rand_seed_offset, first_frame=first,
                                  zero_codeword=self.p.zero_codeword, batch_frames=self.chunk)
        return c.as_array()

    def _allreduce(self, a):
        if self.world == 1:
            return a
        import torch
        t0 = time.perf_counter()
        t = torch.from_numpy(a.copy())
        if self.device is not None:
            t = t.to(self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        out = t.cpu().numpy()  # synchronises: the stop rules below need the sums on the host
        self.t_allreduce += time.perf_counter() - t0
        self.n_allreduce += 1
        return out

    def sim_snr_point(self, snr, first_frame):
        """One SNR point; returns (counters, frames consumed, stop_sweep)."""
        acc = np.zeros(N_COUNTERS, np.int64)
        used = 0
        while acc[0] < self.p.Nframes and acc[1] <= self.p.Nfers:
            todo = int(min(self.p.Nframes - acc[0], next_chunk(acc[0], acc[1], self.p.Nfers, self.chunk * self.world)))
            lo, n = split_frames(todo, first_frame + used, self.rank, self.world)
            t0 = time.perf_counter()
            mine = self._simulate(snr, n, lo)
            self.t_simulate += time.perf_counter() - t0
            acc += self._allreduce(mine)
            used += todo
        ber = acc[3] / acc[2] if acc[2] else 0.0
        fer = acc[1] / acc[0] if acc[0] else 0.0
        return acc, used, (ber < self.p.ber_min or fer < self.p.fer_min)

    def run(self, verbose=True):
        """LDPC_BER_Sim::run (src/LDPC_BER_Sim.cpp:121-155)."""
        self.points, first, stop = [], 0, False
        for snr in self.p.SNRdB:
            if stop:  # later points are stored with zero counts (:142-149)
                self.points.append((snr, np.zeros(N_COUNTERS, np.int64)))
                continue
            acc, used, stop = self.sim_snr_point(snr, first)
            first += used
            self.points.append((snr, acc))
            if verbose and self.rank == 0:
                ubr = acc[5] / acc[4] if acc[4] else 0.0
                print("SNR = %g  Simulated %d frames and %d data bits. Obtained %d data bit errors.  Data BER: %g "
                      "Uncoded BER: %g FER: %g" % (snr, acc[0], acc[2], acc[3], acc[3] / acc[2] if acc[2] else 0.0, ubr,
                                                   acc[1] / acc[0] if acc[0] else 0.0), flush=True)
        return self.results()

    def results(self):
        """The variables the reference stores (src/LDPC_BER_Sim.cpp:347-358)."""
        P = self.points
        return {"sim_SNRdB": [s for s, _ in P], "sim_Nframes": [int(a[0]) for _, a in P],
                "sim_frame_errors": [int(a[1]) for _, a in P], "sim_Ndatabits": [int(a[2]) for _, a in P],
                "sim_data_bit_errors": [int(a[3]) for _, a in P], "sim_uncoded_bit_errors": [int(a[5]) for _, a in P],
                "sim_sum_iterations": [int(a[6]) for _, a in P], "sim_decode_failures": [int(a[7]) for _, a in P]}


class LDPC_BER_Sim(LDPC_BER_Sim_LUT):
    """The same harness around the BP comparison decoder: LDPC_BER_Sim::load (src/LDPC_BER_Sim.cpp:157-244) builds
    IT++'s LDPC_Code from the alist with the LLR_calc_unit of the [BP] section; the frame loop is
    lutldpc_bp_simulate_snr.  (In the reference this is the base class and the LUT harness derives from it; here the
    shared part -- run, sim_snr_point, the stop rules and the all-reduce -- lives in LDPC_BER_Sim_LUT.)"""

    def load(self, basedir="."):
        from .capi import BpDecoder, alist_graph
        p = self.p
        kw = dict(max_iters=p.bp_max_iter, psc=p.parity_check_iter, pisc=p.parity_check_iter,
                  qllr_scale_res=p.qllr_scale_res, qllr_table_size=p.qllr_table_size,
                  qllr_spacing_res=p.qllr_spacing_res, qllr_total_res=p.qllr_total_res)
        if p.codec_filename:  # codes_path / codec_filename, src/LDPC_BER_Sim.cpp:219-239
            path = os.path.join(basedir, p.codes_dir, p.codec_filename)
            if os.path.isabs(p.codec_filename) or not os.path.exists(path):
                path = p.codec_filename if os.path.isabs(p.codec_filename) else os.path.join(basedir, p.codec_filename)
            if not os.path.exists(path):
                raise RuntimeError("Codec file" + path + " does not exist!")
            self.dec = BpDecoder.from_file(path, **kw)
            if not p.zero_codeword and not self.dec.has_gen:
                raise RuntimeError("LDPC_BER_Sim::load(): LDPC.zero_codeword = false needs a codec file with a generator")
            return self
        if not p.parity_filename:
            raise RuntimeError("LDPC_BER_Sim::load(): LDPC.parity_filename is required")
        alist = os.path.join(basedir, p.codes_dir, p.parity_filename + ".alist")
        if not os.path.exists(alist):
            raise RuntimeError("Parity file" + alist + " does not exist!")  # :162-164
        self.dec = BpDecoder(alist_graph(alist), with_generator=not p.zero_codeword, **kw)
        if self.rand_seed == p.save_codec and self.rank == 0:  # :207-216
            d = os.path.join(basedir, p.results_dir, self.gen_filename())
            os.makedirs(d, exist_ok=True)
            self.dec.save_code(os.path.join(d, "bp_codec.it"))
        return self

    def gen_filename(self):
        """LDPC_BER_Sim::gen_filename, src/LDPC_BER_Sim.cpp:104-114"""
        p = self.p
        return "%s_N%d_R%s_maxIter%d_zcw%d_frames%d%s" % (p.results_prefix, self.dec.nvar, "%g" % self.dec.rate, p.bp_max_iter,
                                                         int(p.zero_codeword), p.Nframes, p.custom_name)


def make_sim(params, **kw):
    """prog/ber_sim.cpp:136-147: a [LUT] section (or Sim.codec_type = LUT) selects the LUT decoder, else a [BP]
    section the BP decoder, else it_error."""
    if params.has_lut_section:
        return LDPC_BER_Sim_LUT(params, **kw)
    if params.has_bp_section:
        return LDPC_BER_Sim(params, **kw)
    raise RuntimeError("You must specify the type of decoder in the params file by either explicitly setting at least one "
                       "parameter in the [LUT] or [BP] section of the parameter file or loading a codec via its filename "
                       "and specifying its type")


def counters_from_array(a):
    return SimCounters.from_array(a)
