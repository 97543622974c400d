"""TEST INFRASTRUCTURE ONLY: ctypes view of oracle/liblutldpc_oracle.so (the plain-C restatement
of the reference's LUT decode path, see lutldpc_oracle.h) plus a reader for the fixture files the
compiled reference writes (oracle/ref_driver.cpp, `fixture`).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import this module; the product package never does."""
import ctypes as C
import os
import struct
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblutldpc_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "lutldpc_ref")

(NVAR, NCHK, NCHK_LIN_INDEP, NUM_EDGES, MAX_ITERS, NQ_CHA, MIN_LUT, NINFO, HAS_GEN, N_ITER_SLOTS,
 N_DEG_SLOTS) = range(11)
INIT_CONT, INIT_QCHA = 0, 1
TRACE_INIT, TRACE_CN, TRACE_VN = 0, 1, 2

_lib = None


def build():
    """Compile the C restatement (gcc only; needs no reference tree)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_load_bundle.restype = C.c_void_p
        L.orc_load_bundle.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_info.argtypes = [C.c_void_p, C.c_int]
        L.orc_set_exit_conditions.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_quant_nonlin.argtypes = [C.c_double, C.c_void_p, C.c_int]
        L.orc_quantise.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_lut_decode.argtypes = [C.c_void_p] + [C.c_void_p] * 6 + [C.c_int, C.c_void_p]
        L.orc_lut_decode_batch.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.orc_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_syndrome_ok.argtypes = [C.c_void_p, C.c_void_p]
        for f in (L.orc_get_dv, L.orc_get_dc, L.orc_get_cn_msg_idx, L.orc_get_nq_msg):
            f.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_qb.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """One decoder bundle loaded into the C restatement."""

    def __init__(self, bundle_path):
        err = C.create_string_buffer(256)
        self.h = lib().orc_load_bundle(os.fsencode(bundle_path), err, 256)
        if not self.h:
            raise RuntimeError("oracle: %s (%s)" % (err.value.decode(), bundle_path))
        g = lambda w: lib().orc_info(self.h, w)
        self.nvar, self.nchk, self.num_edges = g(NVAR), g(NCHK), g(NUM_EDGES)
        self.max_iters, self.nq_cha, self.ninfo = g(MAX_ITERS), g(NQ_CHA), g(NINFO)
        self.min_lut, self.has_gen = g(MIN_LUT), g(HAS_GEN)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_free(self.h)
            self.h = None

    def _ivec(self, fn, n):
        a = np.zeros(n, np.int32)
        fn(self.h, _p(a))
        return a

    @property
    def dv(self):
        return self._ivec(lib().orc_get_dv, self.nvar)

    @property
    def dc(self):
        return self._ivec(lib().orc_get_dc, self.nchk)

    @property
    def cn_msg_idx(self):
        return self._ivec(lib().orc_get_cn_msg_idx, self.num_edges)

    @property
    def nq_msg(self):
        return self._ivec(lib().orc_get_nq_msg, self.max_iters)

    def qb(self, which):
        a = np.zeros(64, np.float64)
        n = lib().orc_get_qb(self.h, which, _p(a))
        return a[:n].copy()

    def set_exit_conditions(self, max_iters, psc=True, pisc=False):
        lib().orc_set_exit_conditions(self.h, int(max_iters), int(psc), int(pisc))
        self.max_iters = int(max_iters)

    def quantise(self, llr, init_mode=INIT_CONT):
        llr = np.ascontiguousarray(llr, np.float64).reshape(-1, self.nvar)
        cha = np.zeros(llr.shape, np.uint8)
        msg = np.zeros(llr.shape, np.uint8)
        for f in range(llr.shape[0]):
            lib().orc_quantise(self.h, _p(llr[f]), init_mode, _p(cha[f]), _p(msg[f]))
        return cha, msg

    def lut_decode(self, cha, msg, max_trace=0):
        """One frame. Returns (bits, ret) or (bits, ret, [(kind, iter, msgs)...]) with max_trace > 0."""
        cha = np.ascontiguousarray(cha, np.uint8)
        msg = np.ascontiguousarray(msg, np.uint8)
        out = np.zeros(self.nvar, np.uint8)
        if max_trace <= 0:
            r = lib().orc_lut_decode(self.h, _p(cha), _p(msg), _p(out), None, None, None, 0, None)
            return out, r
        tm = np.zeros((max_trace, self.num_edges), np.uint8)
        tk = np.zeros(max_trace, np.int32)
        ti = np.zeros(max_trace, np.int32)
        n = C.c_int32(0)
        r = lib().orc_lut_decode(self.h, _p(cha), _p(msg), _p(out), _p(tm), _p(tk), _p(ti), max_trace,
                                 C.byref(n))
        return out, r, [(int(tk[i]), int(ti[i]), tm[i].copy()) for i in range(n.value)]

    def lut_decode_batch(self, cha, msg):
        cha = np.ascontiguousarray(cha, np.uint8).reshape(-1, self.nvar)
        msg = np.ascontiguousarray(msg, np.uint8).reshape(-1, self.nvar)
        out = np.zeros(cha.shape, np.uint8)
        ret = np.zeros(cha.shape[0], np.int32)
        lib().orc_lut_decode_batch(self.h, cha.shape[0], _p(cha), _p(msg), _p(out), _p(ret))
        return out, ret

    def encode(self, info):
        info = np.ascontiguousarray(info, np.uint8)
        cw = np.zeros(self.nvar, np.uint8)
        if lib().orc_encode(self.h, _p(info), _p(cw)) != 0:
            raise RuntimeError("bundle has no generator")
        return cw

    def syndrome_ok(self, bits):
        return bool(lib().orc_syndrome_ok(self.h, _p(np.ascontiguousarray(bits, np.uint8))))


def read_fixture(path):
    """Parse a LUTFIX01 file written by `lutldpc_ref fixture` (layout: oracle/ref_driver.cpp), or its LUTFIX02 form
    in which tests/golden/make_golden.py replaced the message array of every trace record by its SHA-256 digest
    (the record's third element is then a 32-byte `bytes`; compare with same_msgs)."""
    b = open(path, "rb").read()
    assert b[:8] in (b"LUTFIX01", b"LUTFIX02"), path
    hashed = b[:8] == b"LUTFIX02"
    n, e, nf, nt, with_llr, zero_cw = struct.unpack("<6i", b[8:32])
    (snr,) = struct.unpack("<d", b[32:40])
    off = 40
    fx = dict(nvar=n, num_edges=e, snr_db=snr, zero_cw=bool(zero_cw))
    llr = np.zeros((nf, n), np.float64) if with_llr else None
    cw, cha, msg, out = (np.zeros((nf, n), np.uint8) for _ in range(4))
    ret = np.zeros(nf, np.int32)
    for f in range(nf):
        if with_llr:
            llr[f] = np.frombuffer(b, "<f8", n, off)
            off += 8 * n
        for a in (cw, cha, msg, out):
            a[f] = np.frombuffer(b, np.uint8, n, off)
            off += n
        ret[f] = struct.unpack_from("<i", b, off)[0]
        off += 4
    traces = []
    for _ in range(nt):
        (nr,) = struct.unpack_from("<i", b, off)
        off += 4
        recs = []
        for _ in range(nr):
            kind, it = struct.unpack_from("<2i", b, off)
            off += 8
            if hashed:
                recs.append((kind, it, bytes(b[off:off + 32])))
                off += 32
            else:
                recs.append((kind, it, np.frombuffer(b, np.uint8, e, off).copy()))
                off += e
        traces.append(recs)
    assert off == len(b), (off, len(b))
    fx.update(llr=llr, cw=cw, cha=cha, msg=msg, out=out, ret=ret, traces=traces)
    return fx


def same_msgs(ref_msgs, mine):
    """A fixture trace record's messages (array, or SHA-256 digest in LUTFIX02 files) against a computed array."""
    mine = np.ascontiguousarray(mine, np.uint8)
    if isinstance(ref_msgs, bytes):
        import hashlib
        return hashlib.sha256(mine.tobytes()).digest() == ref_msgs
    return np.array_equal(ref_msgs, mine)


# ---- BP comparison decoder (oracle/bp_oracle.c: LDPC_Code::bp_decode, itpp/itpp/comm/ldpc.cpp:1398-1622)
class BpParams:
    """max_iter, exit conditions and the four LLR_calc_unit parameters of the [BP] section (src/LDPC_BER_Sim.cpp:74-78)."""

    def __init__(self, max_iter=30, psc=True, pisc=True, d1=12, d2=300, d3=7, d4=28):
        self.max_iter, self.psc, self.pisc, self.d1, self.d2, self.d3, self.d4 = max_iter, bool(psc), bool(pisc), d1, d2, d3, d4


def bp_to_qllr(llr, d1=12, d4=28):
    L = lib()
    llr = np.ascontiguousarray(llr, np.float64)
    out = np.empty(llr.shape, np.int32)
    L.bp_oracle_to_qllr(C.c_int(d1), C.c_int(d4), _p(llr), C.c_int(llr.size), _p(out))
    return out


def bp_logexp_table(d1=12, d2=300, d3=7, d4=28):
    out = np.empty(d2, np.int32)
    lib().bp_oracle_logexp_table(C.c_int(d1), C.c_int(d2), C.c_int(d3), C.c_int(d4), _p(out))
    return out


def bp_decode(dv, dc, cn_msg_idx, qllr_in, par):
    """One frame or [n][nvar] frames of QLLR through bp_decode -> (qllr_out, ret)."""
    L = lib()
    L.bp_oracle_decode.restype = C.c_int
    dv = np.ascontiguousarray(dv, np.int32)
    dc = np.ascontiguousarray(dc, np.int32)
    idx = np.ascontiguousarray(cn_msg_idx, np.int32)
    q = np.ascontiguousarray(qllr_in, np.int32)
    single = q.ndim == 1
    q2 = q.reshape(-1, dv.size)
    out = np.empty_like(q2)
    ret = np.empty(q2.shape[0], np.int32)
    for f in range(q2.shape[0]):
        row = np.ascontiguousarray(q2[f])
        o = np.empty(dv.size, np.int32)
        ret[f] = L.bp_oracle_decode(C.c_int(dv.size), C.c_int(dc.size), _p(dv), _p(dc), _p(idx), C.c_int(par.max_iter),
                                    C.c_int(par.psc), C.c_int(par.pisc), C.c_int(par.d1), C.c_int(par.d2), C.c_int(par.d3),
                                    C.c_int(par.d4), _p(row), _p(o))
        out[f] = o
    return (out[0], int(ret[0])) if single else (out, ret)


def read_bp_fixture(path):
    """BPFIX001 (oracle/ref_driver.cpp cmd_bpfixture): the reference's LLRs, to_qllr, bp_decode output/return, decode() bits."""
    import lzma
    import struct
    raw = (lzma.open(path).read() if str(path).endswith(".xz") else open(path, "rb").read())
    assert raw[:8] == b"BPFIX001"
    N, nf, mi, psc, pisc, d1, d2, d3, d4, K, with_llr = struct.unpack("<11i", raw[8:52])
    snr, scale = struct.unpack("<2d", raw[52:68])
    off = 68
    fx = dict(nvar=N, ninfo=K, par=BpParams(mi, psc, pisc, d1, d2, d3, d4), snr=snr, scale=scale,
              llr=np.empty((nf, N)) if with_llr else None, qin=np.empty((nf, N), np.int32), qout=np.empty((nf, N), np.int32),
              ret=np.empty(nf, np.int32), syst=np.empty((nf, K), np.uint8))
    for f in range(nf):
        if with_llr:
            fx["llr"][f] = np.frombuffer(raw, "<f8", N, off); off += 8 * N
        fx["qin"][f] = np.frombuffer(raw, "<i4", N, off); off += 4 * N
        fx["qout"][f] = np.frombuffer(raw, "<i4", N, off); off += 4 * N
        fx["ret"][f] = struct.unpack("<i", raw[off:off + 4])[0]; off += 4
        fx["syst"][f] = np.frombuffer(raw, "u1", K, off); off += K
    assert off == len(raw)
    return fx
