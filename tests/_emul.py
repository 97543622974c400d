"""TEST INFRASTRUCTURE ONLY: builds and binds tests/host_emul.cpp (the kernels' per-thread arithmetic
and the tree compiler run on the CPU).  Not a product path."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BUILD = os.path.join(HERE, "_build")
LIB = os.path.join(BUILD, "libhost_emul.so")
SRCS = [os.path.join(HERE, "host_emul.cpp"), os.path.join(ROOT, "lut_ldpc_b200", "csrc", "model.cpp")]
DEPS = SRCS + [os.path.join(ROOT, "lut_ldpc_b200", "csrc", n) for n in ("core.cuh", "model.hpp")]
_lib = None


def lib():
    global _lib
    if _lib is None:
        os.makedirs(BUILD, exist_ok=True)
        if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in DEPS):
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-x", "c++", "-o", LIB] + SRCS)
        L = C.CDLL(LIB)
        L.emul_create.restype = C.c_void_p
        L.emul_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.emul_destroy.argtypes = [C.c_void_p]
        L.emul_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.emul_decode.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.emul_cng_selftest.argtypes = [C.c_int, C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Emul:
    def __init__(self, bundle):
        err = C.create_string_buffer(512)
        self.h = lib().emul_create(os.fsencode(bundle), err, 512)
        if not self.h:
            raise RuntimeError(err.value.decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().emul_destroy(self.h)
            self.h = None

    def stats(self):
        s = np.zeros(2, np.int64)
        lib().emul_stats(self.h, _p(s))
        return int(s[0]), int(s[1])

    def decode(self, cha, msg, psc=1, pisc=1, trace_frame=-1, num_edges=0, max_rec=0):
        cha = np.ascontiguousarray(cha, np.uint8)
        msg = np.ascontiguousarray(msg, np.uint8)
        F, N = cha.shape
        out = np.zeros((F, N), np.uint8)
        it = np.zeros(F, np.int32)
        tm = np.zeros((max(max_rec, 1), max(num_edges, 1)), np.uint8)
        tk = np.zeros(max(max_rec, 1), np.int32)
        ti = np.zeros(max(max_rec, 1), np.int32)
        n = C.c_int32(0)
        lib().emul_decode(self.h, F, _p(cha), _p(msg), int(psc), int(pisc), _p(out), _p(it), trace_frame, _p(tm),
                          _p(tk), _p(ti), max_rec, C.byref(n))
        recs = [(int(tk[i]), int(ti[i]), tm[i].copy()) for i in range(n.value)]
        return out, it, recs
