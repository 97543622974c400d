"""ctypes binding of include/lutldpc.h (liblutldpc_b200.so).  Fails loudly when the library is absent."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
INIT_CONT, INIT_QCHA = 0, 1
TRACE_INIT, TRACE_CN, TRACE_VN = 0, 1, 2
ITER_HIST_MAX = 256


class LutLdpcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("lutldpc error %d: %s" % (code, msg))
        self.code = code


def library_path():
    # LUTLDPC_LIB: another build of the same library (kernel tuning experiments)
    return os.environ.get("LUTLDPC_LIB") or os.path.join(HERE, "liblutldpc_b200.so")


def build_library(verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a build of csrc/ (see csrc/Makefile)."""
    subprocess.check_call(["make", "-C", os.path.join(HERE, "csrc"), "-j8"] + ([] if verbose else ["-s"]))
    return library_path()


class _Info(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("nvar", "nchk", "nchk_lin_indep", "ninfo", "num_edges", "max_iters", "Nq_Cha",
                                         "min_lut", "has_gen", "n_degree_classes")] + \
               [("lookups_per_frame_iter", C.c_int64), ("lookups_per_frame_iter_naive", C.c_int64), ("rate", C.c_double),
                ("n_var_programs", C.c_int32), ("n_fast_var_programs", C.c_int32),
                ("onchip", C.c_int32), ("onchip_smem_bytes", C.c_int32)]


class _Graph(C.Structure):
    _fields_ = [("nvar", C.c_int32), ("nchk", C.c_int32), ("num_edges", C.c_int32), ("nchk_lin_indep", C.c_int32),
                ("dv_vec", C.POINTER(C.c_int32)), ("dc_vec", C.POINTER(C.c_int32)), ("cn_msg_idx", C.POINTER(C.c_int32))]


def alist_graph(path, rank_mode=0):
    """LDPC_Parity(alist) + LDPC_Code_LUT::decoder_parameterization: dict with nvar, nchk, num_edges, nchk_lin_indep
    (-1 unless rank_mode 1 = host / 2 = device) and the numpy arrays dv_vec, dc_vec, cn_msg_idx."""
    L = load_library()
    g = _Graph()
    rc = L.lutldpc_alist_graph(os.fsencode(path), int(rank_mode), C.byref(g))
    if rc != 0:
        raise LutLdpcError(rc, (L.lutldpc_last_error(None) or b"").decode())
    try:
        return dict(nvar=g.nvar, nchk=g.nchk, num_edges=g.num_edges, nchk_lin_indep=g.nchk_lin_indep,
                    dv_vec=np.ctypeslib.as_array(g.dv_vec, (g.nvar,)).copy(),
                    dc_vec=np.ctypeslib.as_array(g.dc_vec, (g.nchk,)).copy(),
                    cn_msg_idx=np.ctypeslib.as_array(g.cn_msg_idx, (g.num_edges,)).copy())
    finally:
        L.lutldpc_graph_free(C.byref(g))


class _Desc(C.Structure):
    _fields_ = [("nvar", C.c_int32), ("nchk", C.c_int32), ("nchk_lin_indep", C.c_int32), ("num_edges", C.c_int32),
                ("max_iters", C.c_int32), ("Nq_Cha", C.c_int32), ("min_lut", C.c_int32),
                ("dv_vec", C.c_void_p), ("dc_vec", C.c_void_p), ("cn_msg_idx", C.c_void_p), ("Nq_Msg", C.c_void_p),
                ("reuse_vec", C.c_void_p), ("Nq_Cha_2_Nq_Msg_map", C.c_void_p), ("qb_Cha", C.c_void_p), ("qb_Msg", C.c_void_p),
                ("n_qb_Msg", C.c_int32), ("var_tree_string", C.c_char_p), ("var_tree_len", C.c_size_t),
                ("chk_tree_string", C.c_char_p), ("chk_tree_len", C.c_size_t), ("gen_parity", C.c_void_p),
                ("gen_rows", C.c_int32), ("gen_cols", C.c_int32)]


class _DesignParams(C.Structure):
    _fields_ = [("dv_vec", C.POINTER(C.c_int32)), ("nvar", C.c_int32), ("dc_vec", C.POINTER(C.c_int32)), ("nchk", C.c_int32),
                ("max_iters", C.c_int32), ("Nq_Cha", C.c_int32), ("min_lut", C.c_int32), ("Nq_Msg", C.POINTER(C.c_int32)),
                ("reuse_vec", C.POINTER(C.c_int32)), ("sigma2", C.c_double), ("tree_method", C.c_char_p)]


class _DesignResult(C.Structure):
    _fields_ = [("qb_Cha", C.POINTER(C.c_double)), ("n_qb_Cha", C.c_int32), ("qb_Msg", C.POINTER(C.c_double)),
                ("n_qb_Msg", C.c_int32), ("map", C.POINTER(C.c_int32)), ("n_map", C.c_int32),
                ("var_tree_string", C.POINTER(C.c_char)), ("var_tree_len", C.c_size_t),
                ("chk_tree_string", C.POINTER(C.c_char)), ("chk_tree_len", C.c_size_t)]


def design_luts(dv, dc, max_iters, Nq_Cha, Nq_Msg, sigma2, reuse_vec=None, min_lut=True, tree_method="auto_bin_balanced"):
    """LDPC_Code_LUT::design_luts on the host (no device needed): dict with qb_Cha, qb_Msg, Nq_Cha_2_Nq_Msg_map and
    the var / chk tree strings (bytes, the reference's text serialisation)."""
    L = load_library()
    L.lutldpc_design_luts.argtypes = [C.POINTER(_DesignParams), C.POINTER(_DesignResult)]
    L.lutldpc_design_free.argtypes = [C.POINTER(_DesignResult)]
    L.lutldpc_design_free.restype = None
    dv = np.ascontiguousarray(dv, np.int32)
    dc = np.ascontiguousarray(dc, np.int32)
    nq = np.ascontiguousarray(Nq_Msg, np.int32)
    assert nq.shape == (max_iters,)
    ru = None if reuse_vec is None else np.ascontiguousarray(reuse_vec, np.int32)
    i32p = C.POINTER(C.c_int32)
    p = _DesignParams(dv.ctypes.data_as(i32p), len(dv), dc.ctypes.data_as(i32p), len(dc), int(max_iters), int(Nq_Cha),
                      int(bool(min_lut)), nq.ctypes.data_as(i32p), ru.ctypes.data_as(i32p) if ru is not None else None,
                      float(sigma2), os.fsencode(tree_method))
    r = _DesignResult()
    rc = L.lutldpc_design_luts(C.byref(p), C.byref(r))
    if rc != 0:
        raise LutLdpcError(rc, (L.lutldpc_last_error(None) or b"").decode())
    try:
        return dict(qb_Cha=np.ctypeslib.as_array(r.qb_Cha, (r.n_qb_Cha,)).copy(),
                    qb_Msg=np.ctypeslib.as_array(r.qb_Msg, (r.n_qb_Msg,)).copy(),
                    Nq_Cha_2_Nq_Msg_map=np.ctypeslib.as_array(r.map, (r.n_map,)).copy(),
                    var_tree_string=C.string_at(r.var_tree_string, r.var_tree_len),
                    chk_tree_string=C.string_at(r.chk_tree_string, r.chk_tree_len))
    finally:
        L.lutldpc_design_free(C.byref(r))


class _SimParams(C.Structure):
    _fields_ = [("snr_db", C.c_double), ("n_frames", C.c_int64), ("seed", C.c_uint64), ("first_frame", C.c_uint64),
                ("zero_codeword", C.c_int32), ("batch_frames", C.c_int32)]


class SimCounters(C.Structure):
    """lutldpc_sim_counters: the BERC/BLERC counters of LDPC_BER_Sim::sim_snr_point."""
    _fields_ = [(n, C.c_uint64) for n in ("frames", "frame_errors", "data_bits", "data_bit_errors", "uncoded_bits",
                                          "uncoded_bit_errors", "sum_iters", "decode_failures")] + \
               [("iter_hist", C.c_uint64 * ITER_HIST_MAX)]

    NAMES = ("frames", "frame_errors", "data_bits", "data_bit_errors", "uncoded_bits", "uncoded_bit_errors",
             "sum_iters", "decode_failures")

    def as_array(self):
        return np.array([getattr(self, n) for n in self.NAMES] + list(self.iter_hist), dtype=np.int64)

    @classmethod
    def from_array(cls, a):
        c = cls()
        for i, n in enumerate(cls.NAMES):
            setattr(c, n, int(a[i]))
        for i in range(ITER_HIST_MAX):
            c.iter_hist[i] = int(a[len(cls.NAMES) + i])
        return c

    def as_dict(self):
        return {n: int(getattr(self, n)) for n in self.NAMES}


_lib = None
EXPORTS = ["lutldpc_abi_version", "lutldpc_device_count", "lutldpc_set_device", "lutldpc_last_error", "lutldpc_create", "lutldpc_create_from_bundle",
           "lutldpc_destroy", "lutldpc_get_info", "lutldpc_set_exit_conditions", "lutldpc_set_initial_message_mode",
           "lutldpc_set_stream", "lutldpc_set_poll_interval", "lutldpc_decode_quantized",
           "lutldpc_decode_quantized_device", "lutldpc_decode_llr", "lutldpc_decode_llr_device", "lutldpc_quantize",
           "lutldpc_decode_quantized_traced", "lutldpc_encode", "lutldpc_syndrome_check", "lutldpc_simulate_snr",
           "lutldpc_last_call_stats", "lutldpc_set_profiling", "lutldpc_get_profile", "lutldpc_quantize_device",
           "lutldpc_set_repack", "lutldpc_get_repack_count", "lutldpc_syndrome_check_device_noalloc",
           "lutldpc_noalloc_max_frames", "lutldpc_set_onchip", "lutldpc_decode_packed", "lutldpc_decode_packed_device",
           "lutldpc_alist_graph", "lutldpc_graph_free", "lutldpc_create_systematic",
           "lutldpc_design_luts", "lutldpc_design_free", "lutldpc_simulate_snr_stimuli"]
PROFILE_KINDS = 64
PROFILE_NAMES = {0: "reset", 1: "pack", 2: "channel", 3: "cn_minsum", 4: "syndrome", 5: "finalize", 6: "unpack",
                 7: "count", 8: "encode", 10: "repack", 11: "retire", 12: "onchip"}


class _BpParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("max_iters", "psc", "pisc", "qllr_scale_res", "qllr_table_size",
                                         "qllr_spacing_res", "qllr_total_res")]


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    p = library_path()
    if not os.path.exists(p):
        raise LutLdpcError(-3, "%s is missing: build it with lut_ldpc_b200.build_library() "
                               "(nvcc, sm_100a); there is no CPU fallback" % p)
    L = C.CDLL(p)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.lutldpc_last_error.restype = C.c_char_p
    L.lutldpc_last_error.argtypes = [vp]
    L.lutldpc_create_from_bundle.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.lutldpc_destroy.argtypes = [vp]
    L.lutldpc_destroy.restype = None
    L.lutldpc_get_info.argtypes = [vp, C.POINTER(_Info)]
    L.lutldpc_set_exit_conditions.argtypes = [vp, i32, i32, i32]
    L.lutldpc_set_initial_message_mode.argtypes = [vp, i32]
    L.lutldpc_set_stream.argtypes = [vp, vp]
    L.lutldpc_set_poll_interval.argtypes = [vp, i32]
    L.lutldpc_set_repack.argtypes = [vp, i32, C.c_double]
    L.lutldpc_get_repack_count.argtypes = [vp, vp]
    L.lutldpc_decode_quantized.argtypes = [vp, i32, vp, vp, vp, vp]
    L.lutldpc_decode_quantized_device.argtypes = [vp, i32, vp, vp, vp, vp]
    L.lutldpc_decode_llr.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_decode_llr_device.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_quantize.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_decode_quantized_traced.argtypes = [vp, i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, vp]
    L.lutldpc_encode.argtypes = [vp, i32, vp, vp]
    L.lutldpc_syndrome_check.argtypes = [vp, i32, vp, vp]
    L.lutldpc_simulate_snr.argtypes = [vp, C.POINTER(_SimParams), C.POINTER(SimCounters)]
    L.lutldpc_simulate_snr_stimuli.argtypes = [vp, C.POINTER(_SimParams), C.POINTER(SimCounters), vp, vp]
    L.lutldpc_last_call_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64), C.POINTER(i64)]
    L.lutldpc_quantize_device.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_set_profiling.argtypes = [vp, i32]
    L.lutldpc_syndrome_check_device_noalloc.argtypes = [vp, i32, vp, vp]
    L.lutldpc_noalloc_max_frames.argtypes = [vp]
    L.lutldpc_set_onchip.argtypes = [vp, i32]
    L.lutldpc_decode_packed.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_alist_graph.argtypes = [C.c_char_p, i32, C.POINTER(_Graph)]
    L.lutldpc_graph_free.argtypes = [C.POINTER(_Graph)]
    L.lutldpc_graph_free.restype = None
    L.lutldpc_create_systematic.argtypes = [vp, C.POINTER(vp), vp]
    L.lutldpc_decode_packed_device.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_get_profile.argtypes = [vp, vp, vp, vp, vp]
    L.lutldpc_save_code.argtypes = [vp, C.c_char_p]
    L.lutldpc_convert_codec.argtypes = [C.c_char_p, C.c_char_p]
    L.lutldpc_bp_create.argtypes = [C.POINTER(_Graph), C.POINTER(_BpParams), i32, C.POINTER(vp), vp]
    L.lutldpc_bp_create_from_file.argtypes = [C.c_char_p, C.POINTER(_BpParams), C.POINTER(vp)]
    L.lutldpc_bp_save_code.argtypes = [vp, C.c_char_p]
    L.lutldpc_bp_convert_alist.argtypes = [C.c_char_p, C.c_char_p]
    L.lutldpc_bp_destroy.argtypes = [vp]
    L.lutldpc_bp_destroy.restype = None
    L.lutldpc_bp_last_error.restype = C.c_char_p
    L.lutldpc_bp_last_error.argtypes = [vp]
    L.lutldpc_bp_get_info.argtypes = [vp, C.POINTER(_Info)]
    L.lutldpc_bp_decode_llr.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_bp_decode_qllr.argtypes = [vp, i32, vp, vp, vp]
    L.lutldpc_bp_encode.argtypes = [vp, i32, vp, vp]
    L.lutldpc_bp_simulate_snr.argtypes = [vp, C.POINTER(_SimParams), C.POINTER(SimCounters)]
    L.lutldpc_bp_last_call_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64), C.POINTER(i64)]
    L.lutldpc_bp_set_repack.argtypes = [vp, i32]
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Decoder:
    """One lutldpc_decoder handle (an LDPC_Code_LUT bound to the current CUDA device)."""

    def __init__(self, bundle_path, _handle=None):
        """bundle_path: the reference's codec file (lut_codec.it, IT++ it_file) or a decoder bundle."""
        self._L = load_library()
        h = C.c_void_p()
        if _handle is not None:
            h = _handle
        else:
            rc = self._L.lutldpc_create_from_bundle(os.fsencode(bundle_path), C.byref(h))
            if rc != 0:
                raise LutLdpcError(rc, (self._L.lutldpc_last_error(None) or b"").decode())
        self._h = h
        self._refresh_info()

    @classmethod
    def from_design(cls, graph, design, max_iters, Nq_Cha, Nq_Msg, reuse_vec=None, min_lut=True):
        """lutldpc_create from the graph arrays of alist_graph() and the tables of design_luts(): the decoder
        LDPC_BER_Sim_LUT::load builds when the params name an alist instead of a codec file."""
        L = load_library()
        keep = [np.ascontiguousarray(graph[k], np.int32) for k in ("dv_vec", "dc_vec", "cn_msg_idx")]
        nq = np.ascontiguousarray(Nq_Msg, np.int32)
        ru = np.ascontiguousarray(reuse_vec if reuse_vec is not None else np.zeros(max_iters), np.int32)
        mp = np.ascontiguousarray(design["Nq_Cha_2_Nq_Msg_map"], np.int32)
        qc, qm = np.ascontiguousarray(design["qb_Cha"], np.float64), np.ascontiguousarray(design["qb_Msg"], np.float64)
        rank = graph["nchk_lin_indep"] if graph["nchk_lin_indep"] >= 0 else graph["nchk"]
        d = _Desc(graph["nvar"], graph["nchk"], rank, graph["num_edges"], int(max_iters), int(Nq_Cha), int(bool(min_lut)),
                  keep[0].ctypes.data, keep[1].ctypes.data, keep[2].ctypes.data, nq.ctypes.data, ru.ctypes.data,
                  mp.ctypes.data if len(mp) == Nq_Cha else None, qc.ctypes.data, qm.ctypes.data, len(qm),
                  design["var_tree_string"], len(design["var_tree_string"]),
                  design["chk_tree_string"], len(design["chk_tree_string"]), None, 0, 0)
        h = C.c_void_p()
        L.lutldpc_create.argtypes = [C.POINTER(_Desc), C.POINTER(C.c_void_p)]
        rc = L.lutldpc_create(C.byref(d), C.byref(h))
        if rc != 0:
            raise LutLdpcError(rc, (L.lutldpc_last_error(None) or b"").decode())
        return cls(None, _handle=h)

    def systematic(self):
        """LDPC_Generator_Systematic::construct on the device: (decoder of the column-permuted code with the generator
        attached, col_perm = new index of every old column)."""
        h = C.c_void_p()
        perm = np.zeros(self.nvar, np.int32)
        rc = self._L.lutldpc_create_systematic(self._h, C.byref(h), _p(perm))
        if rc != 0:
            raise LutLdpcError(rc, (self._L.lutldpc_last_error(None) or b"").decode())
        return Decoder(None, _handle=h), perm

    def _refresh_info(self):
        info = _Info()
        self._ck(self._L.lutldpc_get_info(self._h, C.byref(info)))
        self.info = info
        for n, _ in _Info._fields_:
            setattr(self, n, getattr(info, n))

    from_bundle = classmethod(lambda cls, path: cls(path))

    def close(self):
        if getattr(self, "_h", None):
            self._L.lutldpc_destroy(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise LutLdpcError(rc, (self._L.lutldpc_last_error(self._h) or b"").decode())

    # ---- configuration (LDPC_Code_LUT::set_exit_conditions / set_initial_message_mode)
    def save_code(self, path):
        """LDPC_Code_LUT::save_code: the decoder as the reference's codec file (lut_codec.it)."""
        self._ck(self._L.lutldpc_save_code(self._h, os.fsencode(path)))

    def set_exit_conditions(self, max_iters, psc=True, pisc=False):
        self._ck(self._L.lutldpc_set_exit_conditions(self._h, int(max_iters), int(psc), int(pisc)))

    def set_initial_message_mode(self, mode):
        self._ck(self._L.lutldpc_set_initial_message_mode(self._h, int(mode)))

    def set_stream(self, cuda_stream_ptr):
        self._ck(self._L.lutldpc_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def set_poll_interval(self, every):
        self._ck(self._L.lutldpc_set_poll_interval(self._h, int(every)))

    def set_repack(self, on, alpha=0.7):
        """Frame repacking (gather of the unfinished frames into a narrower layout); results do not depend on it."""
        self._ck(self._L.lutldpc_set_repack(self._h, int(bool(on)), float(alpha)))

    def set_onchip(self, on):
        """On-chip regime (one persistent kernel per batch, state in shared memory) on/off; results do not depend on it."""
        self._ck(self._L.lutldpc_set_onchip(self._h, int(bool(on))))
        self.onchip = int(bool(on))

    @property
    def repack_count(self):
        n = C.c_int64(0)
        self._ck(self._L.lutldpc_get_repack_count(self._h, C.byref(n)))
        return int(n.value)

    # ---- decode (host buffers)
    def decode_quantized(self, cha, msg, out=None, iters=None):
        """lut_decode for a batch of host label arrays.  `out` [F, nvar] u8 / `iters` [F] i32 may be supplied
        (e.g. views of pinned memory, so that the copies overlap the decoder); else they are allocated."""
        cha = np.ascontiguousarray(cha, np.uint8).reshape(-1, self.nvar)
        msg = np.ascontiguousarray(msg, np.uint8).reshape(-1, self.nvar)
        assert cha.shape == msg.shape
        out = np.empty(cha.shape, np.uint8) if out is None else out
        it = np.empty(cha.shape[0], np.int32) if iters is None else iters
        assert out.dtype == np.uint8 and out.shape == cha.shape and out.flags.c_contiguous
        assert it.dtype == np.int32 and it.shape == (cha.shape[0],) and it.flags.c_contiguous
        self._ck(self._L.lutldpc_decode_quantized(self._h, cha.shape[0], _p(cha), _p(msg), _p(out), _p(it)))
        return out, it

    def decode_packed(self, cha_msg, out=None, iters=None):
        """lut_decode on packed host buffers: cha_msg [F, nvar] u8 = cha label | msg label << 4; returns the bit-packed
        decisions [F, ceil(nvar/8)] (numpy.unpackbits(..., bitorder="little") gives LLRout) and the returns."""
        cha_msg = np.ascontiguousarray(cha_msg, np.uint8).reshape(-1, self.nvar)
        nb = (self.nvar + 7) // 8
        out = np.empty((cha_msg.shape[0], nb), np.uint8) if out is None else out
        it = np.empty(cha_msg.shape[0], np.int32) if iters is None else iters
        assert out.dtype == np.uint8 and out.shape == (cha_msg.shape[0], nb) and out.flags.c_contiguous
        assert it.dtype == np.int32 and it.shape == (cha_msg.shape[0],) and it.flags.c_contiguous
        self._ck(self._L.lutldpc_decode_packed(self._h, cha_msg.shape[0], _p(cha_msg), _p(out), _p(it)))
        return out, it

    def decode_packed_device(self, n_frames, d_cha_msg, d_out, d_iters=0):
        self._ck(self._L.lutldpc_decode_packed_device(self._h, int(n_frames), C.c_void_p(d_cha_msg), C.c_void_p(d_out),
                                                      C.c_void_p(d_iters or None)))

    def decode_quantized_traced(self, cha, msg, trace_frame, max_rec):
        cha = np.ascontiguousarray(cha, np.uint8).reshape(-1, self.nvar)
        msg = np.ascontiguousarray(msg, np.uint8).reshape(-1, self.nvar)
        out = np.empty(cha.shape, np.uint8)
        it = np.empty(cha.shape[0], np.int32)
        tm = np.zeros((max_rec, self.num_edges), np.uint8)
        tk = np.zeros(max_rec, np.int32)
        ti = np.zeros(max_rec, np.int32)
        n = C.c_int32(0)
        self._ck(self._L.lutldpc_decode_quantized_traced(self._h, cha.shape[0], _p(cha), _p(msg), _p(out), _p(it),
                                                         int(trace_frame), _p(tm), _p(tk), _p(ti), int(max_rec),
                                                         C.byref(n)))
        return out, it, [(int(tk[i]), int(ti[i]), tm[i].copy()) for i in range(n.value)]

    def decode_llr(self, llr):
        llr = np.ascontiguousarray(llr, np.float64).reshape(-1, self.nvar)
        out = np.empty(llr.shape, np.uint8)
        it = np.empty(llr.shape[0], np.int32)
        self._ck(self._L.lutldpc_decode_llr(self._h, llr.shape[0], _p(llr), _p(out), _p(it)))
        return out, it

    def quantize(self, llr):
        llr = np.ascontiguousarray(llr, np.float64).reshape(-1, self.nvar)
        cha = np.empty(llr.shape, np.uint8)
        msg = np.empty(llr.shape, np.uint8)
        self._ck(self._L.lutldpc_quantize(self._h, llr.shape[0], _p(llr), _p(cha), _p(msg)))
        return cha, msg

    def encode(self, info):
        info = np.ascontiguousarray(info, np.uint8).reshape(-1, self.ninfo)
        cw = np.empty((info.shape[0], self.nvar), np.uint8)
        self._ck(self._L.lutldpc_encode(self._h, info.shape[0], _p(info), _p(cw)))
        return cw

    def syndrome_check(self, bits):
        bits = np.ascontiguousarray(bits, np.uint8).reshape(-1, self.nvar)
        ok = np.empty(bits.shape[0], np.uint8)
        self._ck(self._L.lutldpc_syndrome_check(self._h, bits.shape[0], _p(bits), _p(ok)))
        return ok.astype(bool)

    def syndrome_check_device_noalloc(self, n_frames, d_bits, d_ok):
        """Allocation-free syndrome check on device pointers (static __device__ scratch; asynchronous on the stream)."""
        self._ck(self._L.lutldpc_syndrome_check_device_noalloc(self._h, int(n_frames), C.c_void_p(d_bits),
                                                               C.c_void_p(d_ok)))

    @property
    def noalloc_max_frames(self):
        return int(self._L.lutldpc_noalloc_max_frames(self._h))

    # ---- decode (device pointers, e.g. torch tensors' data_ptr())
    def decode_quantized_device(self, n_frames, d_cha, d_msg, d_out, d_iters=0):
        self._ck(self._L.lutldpc_decode_quantized_device(self._h, int(n_frames), C.c_void_p(d_cha), C.c_void_p(d_msg),
                                                         C.c_void_p(d_out), C.c_void_p(d_iters or None)))

    def quantize_device(self, n_frames, d_llr, d_cha, d_msg):
        self._ck(self._L.lutldpc_quantize_device(self._h, int(n_frames), C.c_void_p(d_llr), C.c_void_p(d_cha),
                                                 C.c_void_p(d_msg)))

    def decode_llr_device(self, n_frames, d_llr, d_out, d_iters=0):
        self._ck(self._L.lutldpc_decode_llr_device(self._h, int(n_frames), C.c_void_p(d_llr), C.c_void_p(d_out),
                                                   C.c_void_p(d_iters or None)))

    # ---- fused Monte-Carlo step (LDPC_BER_Sim::sim_snr_point's frame loop)
    def simulate_snr(self, snr_db, n_frames, seed=0, first_frame=0, zero_codeword=True, batch_frames=0, acc=None):
        p = _SimParams(float(snr_db), int(n_frames), int(seed), int(first_frame), int(bool(zero_codeword)),
                       int(batch_frames))
        acc = acc if acc is not None else SimCounters()
        self._ck(self._L.lutldpc_simulate_snr(self._h, C.byref(p), C.byref(acc)))
        return acc

    def simulate_snr_stimuli(self, snr_db, n_frames, seed=0, first_frame=0, zero_codeword=True):
        """One batch of the fused step, returning (counters, llr_in_cha labels [F, nvar], LLRout [F, nvar]): the pairs
        the reference prints as "Stimuli Pair" at output_verbosity > 0."""
        p = _SimParams(float(snr_db), int(n_frames), int(seed), int(first_frame), int(bool(zero_codeword)), 0)
        acc = SimCounters()
        cha = np.zeros((n_frames, self.nvar), np.uint8)
        out = np.zeros((n_frames, self.nvar), np.uint8)
        self._ck(self._L.lutldpc_simulate_snr_stimuli(self._h, C.byref(p), C.byref(acc), _p(cha), _p(out)))
        return acc, cha, out

    def set_profiling(self, on):
        self._ck(self._L.lutldpc_set_profiling(self._h, int(bool(on))))

    def get_profile(self):
        """{kernel name: (device ms, launches)} since profiling was switched on; VN classes by degree."""
        ms = np.zeros(PROFILE_KINDS, np.float64)
        n = np.zeros(PROFILE_KINDS, np.int64)
        deg = np.zeros(16, np.int32)
        cnt = np.zeros(16, np.int32)
        self._ck(self._L.lutldpc_get_profile(self._h, _p(ms), _p(n), _p(deg), _p(cnt)))
        out = {}
        for k in range(PROFILE_KINDS):
            if n[k] == 0:
                continue
            if k < 16:
                name = PROFILE_NAMES.get(k, "kind%d" % k)
            else:
                name = "%s_c%d" % (("vn", "dec", "chk")[k // 16 - 1], k % 16)
                if k == 31:
                    name = "vn_fused"
                elif k < 48:
                    name = "%s_dv%d" % (("vn", "dec")[k // 16 - 1], deg[k % 16])
            out[name] = (float(ms[k]), int(n[k]))
        self.vn_class_degree = [int(d) for d in deg if d]
        self.vn_class_nodes = [int(c) for d, c in zip(deg, cnt) if d]
        return out

    def last_call_stats(self):
        ms, nl, fi = C.c_double(0), C.c_int64(0), C.c_int64(0)
        self._ck(self._L.lutldpc_last_call_stats(self._h, C.byref(ms), C.byref(nl), C.byref(fi)))
        return dict(device_ms=ms.value, kernel_launches=nl.value, frame_iterations=fi.value)


class BpDecoder:
    """One lutldpc_bp handle: IT++'s LDPC_Code with the LLR_calc_unit of ber_sim's [BP] section, batched
    (itpp/itpp/comm/ldpc.cpp:1398-1622; src/LDPC_BER_Sim.cpp:74-78,197-201)."""

    def __init__(self, graph, max_iters=30, psc=True, pisc=None, qllr_scale_res=12, qllr_table_size=300,
                 qllr_spacing_res=7, qllr_total_res=28, with_generator=False, codec_file=None):
        """graph: the dict of alist_graph() (or None with codec_file = a bp_codec.it, see from_file).  pisc defaults to
        psc, as LDPC_BER_Sim::load passes parity_check_iter twice."""
        self._L = load_library()
        pisc = psc if pisc is None else pisc
        par = _BpParams(int(max_iters), int(bool(psc)), int(bool(pisc)), int(qllr_scale_res), int(qllr_table_size),
                        int(qllr_spacing_res), int(qllr_total_res))
        h = C.c_void_p()
        if codec_file is not None:
            self.col_perm = None
            rc = self._L.lutldpc_bp_create_from_file(os.fsencode(codec_file), C.byref(par), C.byref(h))
        else:
            dv = np.ascontiguousarray(graph["dv_vec"], np.int32)
            dc = np.ascontiguousarray(graph["dc_vec"], np.int32)
            idx = np.ascontiguousarray(graph["cn_msg_idx"], np.int32)
            i32p = C.POINTER(C.c_int32)
            g = _Graph(int(dv.size), int(dc.size), int(idx.size), int(graph.get("nchk_lin_indep", -1)),
                       dv.ctypes.data_as(i32p), dc.ctypes.data_as(i32p), idx.ctypes.data_as(i32p))
            self.col_perm = np.empty(dv.size, np.int32)
            rc = self._L.lutldpc_bp_create(C.byref(g), C.byref(par), int(bool(with_generator)), C.byref(h), _p(self.col_perm))
        if rc != 0:
            raise LutLdpcError(rc, (self._L.lutldpc_bp_last_error(None) or b"").decode())
        self._h = h
        info = _Info()
        self._ck(self._L.lutldpc_bp_get_info(self._h, C.byref(info)))
        self.nvar, self.nchk, self.ninfo, self.num_edges = info.nvar, info.nchk, info.ninfo, info.num_edges
        self.nchk_lin_indep, self.max_iters, self.has_gen, self.rate = info.nchk_lin_indep, info.max_iters, bool(info.has_gen), info.rate

    @classmethod
    def from_file(cls, codec_file, **kw):
        """LDPC_Code(filename, G): IT++'s bp_codec.it (LDPC_Code::save_code, with the generator if it was saved)."""
        return cls(None, codec_file=codec_file, **kw)

    def save_code(self, path):
        """LDPC_Code::save_code (+ LDPC_Generator_Systematic::save when the handle carries the generator)."""
        self._ck(self._L.lutldpc_bp_save_code(self._h, os.fsencode(path)))

    def close(self):
        if getattr(self, "_h", None):
            self._L.lutldpc_bp_destroy(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise LutLdpcError(rc, (self._L.lutldpc_bp_last_error(self._h) or b"").decode())

    def set_repack(self, on):
        self._ck(self._L.lutldpc_bp_set_repack(self._h, int(bool(on))))

    def decode_llr(self, llr):
        """LDPC_Code::decode for [n, nvar] LLRs -> (hard decisions of all nvar bits [n, nvar], bp_decode returns [n])."""
        llr = np.ascontiguousarray(llr, np.float64).reshape(-1, self.nvar)
        n = llr.shape[0]
        bits = np.empty((n, self.nvar), np.uint8)
        iters = np.empty(n, np.int32)
        self._ck(self._L.lutldpc_bp_decode_llr(self._h, n, _p(llr), _p(bits), _p(iters)))
        return bits, iters

    def decode_qllr(self, qllr):
        """LDPC_Code::bp_decode for [n, nvar] QLLRs -> (LLRout [n, nvar] int32, returns [n])."""
        q = np.ascontiguousarray(qllr, np.int32).reshape(-1, self.nvar)
        n = q.shape[0]
        out = np.empty((n, self.nvar), np.int32)
        iters = np.empty(n, np.int32)
        self._ck(self._L.lutldpc_bp_decode_qllr(self._h, n, _p(q), _p(out), _p(iters)))
        return out, iters

    def encode(self, info):
        info = np.ascontiguousarray(info, np.uint8).reshape(-1, self.ninfo)
        cw = np.empty((info.shape[0], self.nvar), np.uint8)
        self._ck(self._L.lutldpc_bp_encode(self._h, info.shape[0], _p(info), _p(cw)))
        return cw

    def simulate_snr(self, snr_db, n_frames, seed=0, first_frame=0, zero_codeword=True, batch_frames=0, acc=None):
        p = _SimParams(float(snr_db), int(n_frames), int(seed), int(first_frame), int(bool(zero_codeword)),
                       int(batch_frames))
        acc = acc if acc is not None else SimCounters()
        self._ck(self._L.lutldpc_bp_simulate_snr(self._h, C.byref(p), C.byref(acc)))
        return acc

    def last_call_stats(self):
        ms, nl, fi = C.c_double(), C.c_int64(), C.c_int64()
        self._ck(self._L.lutldpc_bp_last_call_stats(self._h, C.byref(ms), C.byref(nl), C.byref(fi)))
        return dict(device_ms=ms.value, kernel_launches=nl.value, frame_iterations=fi.value)


def bp_convert_alist(alist_path, out_it_path):
    """alist -> bp_codec.it as IT++'s LDPC_Code(H).save_code() writes it (host only)."""
    L = load_library()
    rc = L.lutldpc_bp_convert_alist(os.fsencode(alist_path), os.fsencode(out_it_path))
    if rc != 0:
        raise LutLdpcError(rc, (L.lutldpc_bp_last_error(None) or b"").decode())
