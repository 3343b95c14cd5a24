"""ctypes binding of libsw_cuda.so (the C ABI in include/sw_cuda.h).

Mirrors the reference's Smith-Waterman interface (tools/GKL/src/main/native/smithwaterman/
IntelSmithWatermanCSource.cpp: initNative :22, align :45; Java side IntelSmithWaterman.java:84-104:
`align(ref, alt, SWParameters, SWOverhangStrategy) -> (cigar, alignment_offset)`) on batches of
pairs.  There is no CPU fallback: a missing library or no visible B200 raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .sw_pairs import SwPairs, from_lists

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsw_cuda.so")

SOFTCLIP, INDEL, LEADING_INDEL, IGNORE = 9, 10, 11, 12      # IntelSmithWaterman.java:106-121
ERRORS = {1: "SW_ERR_ARG", 2: "SW_ERR_EMPTY_SEQ", 3: "SW_ERR_TOO_LONG", 4: "SW_ERR_CUDA", 5: "SW_ERR_NOMEM"}
EXPORTS = ["sw_abi_version", "sw_create", "sw_destroy", "sw_last_error", "sw_align_flat", "sw_align_one",
           "sw_workload_create", "sw_workload_run", "sw_workload_results", "sw_workload_destroy",
           "initNative", "align"]


class SwError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s (%d): %s" % (ERRORS.get(code, "SW_ERR"), code, msg))
        self.code = code


class CFlatPairs(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("ref_len", C.c_void_p), ("ref_off", C.c_void_p), ("ref_pool", C.c_void_p),
                ("alt_len", C.c_void_p), ("alt_off", C.c_void_p), ("alt_pool", C.c_void_p)]


class CParams(C.Structure):
    _fields_ = [("match", C.c_int32), ("mismatch", C.c_int32), ("open", C.c_int32), ("extend", C.c_int32)]


class CStats(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("n_cells", C.c_int64), ("n_overflow", C.c_int64), ("h2d_bytes", C.c_int64),
                ("d2h_bytes", C.c_int64), ("bt_bytes", C.c_int64), ("n_launches", C.c_int32), ("n_classes", C.c_int32),
                ("ms_kernels", C.c_float), ("ms_total", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def load_library():
    """Load libsw_cuda.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SwError(4, "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        lib.sw_last_error.restype = C.c_char_p
        lib.sw_last_error.argtypes = [C.c_void_p]
        _lib = lib
    return _lib


def _view(pairs: SwPairs):
    z = np.zeros(1, np.uint8)
    vp = lambda a: (a if a.size else z).ctypes.data_as(C.c_void_p)
    return CFlatPairs(pairs.n_pairs, vp(pairs.ref_len), vp(pairs.ref_off), vp(pairs.ref_pool),
                      vp(pairs.alt_len), vp(pairs.alt_off), vp(pairs.alt_pool)), z


class SwResult:
    def __init__(self, n, stride):
        self.score = np.zeros(max(1, n), np.int32)[:n]
        self.offset = np.zeros(max(1, n), np.int32)[:n]
        self.cigar_len = np.zeros(max(1, n), np.int32)[:n]
        self.buf = np.zeros(max(1, n) * stride, np.uint8)
        self.stride = stride
        self.stats = {}
        self._patched = {}

    @property
    def cigars(self):
        out = [self.buf[p * self.stride:p * self.stride + n].tobytes().decode() if n >= 0 else None
               for p, n in enumerate(self.cigar_len)]
        for p, s in self._patched.items():
            out[p] = s
        return out


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a.size else None


class Workload:
    """Device-resident pairs (sw_workload_*)."""

    def __init__(self, eng: "SmithWaterman", pairs: SwPairs, stride: int):
        self.eng, self.n, self.stride = eng, pairs.n_pairs, stride
        self._h = C.c_void_p()
        view, keep = _view(pairs)
        eng._check(eng.lib.sw_workload_create(eng._h, C.byref(view), C.c_int32(stride), C.byref(self._h)))

    def run(self, params, strategy) -> dict:
        st = CStats()
        pm = CParams(*[int(x) for x in params])
        self.eng._check(self.eng.lib.sw_workload_run(self.eng._h, self._h, C.byref(pm), C.c_int(strategy), C.byref(st)))
        return st.as_dict()

    def results(self) -> SwResult:
        r = SwResult(self.n, self.stride)
        self.eng._check(self.eng.lib.sw_workload_results(self.eng._h, self._h, _p(r.score), _p(r.offset),
                                                         r.buf.ctypes.data_as(C.c_void_p), _p(r.cigar_len)))
        return r

    def close(self):
        if self._h:
            self.eng.lib.sw_workload_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SmithWaterman:
    """One context on one B200 (sw_create = the reference's initNative)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self._h = C.c_void_p()
        rc = self.lib.sw_create(C.c_int(device), C.byref(self._h))
        if rc:
            raise SwError(rc, (self.lib.sw_last_error(None) or b"").decode())

    def _check(self, rc):
        if rc:
            raise SwError(rc, (self.lib.sw_last_error(self._h) or b"").decode())

    def workload(self, pairs: SwPairs, stride: int = 64) -> Workload:
        return Workload(self, pairs, stride)

    def align_pairs(self, pairs: SwPairs, params, strategy, stride: int = 64, retry: bool = True) -> SwResult:
        """sw_align_flat; pairs whose CIGAR does not fit `stride` are re-run with the stride they need."""
        r = SwResult(pairs.n_pairs, stride)
        view, keep = _view(pairs)
        pm = CParams(*[int(x) for x in params])
        st = CStats()
        self._check(self.lib.sw_align_flat(self._h, C.byref(view), C.byref(pm), C.c_int(strategy), _p(r.score),
                                           _p(r.offset), r.buf.ctypes.data_as(C.c_void_p), C.c_int32(stride),
                                           _p(r.cigar_len), C.byref(st)))
        r.stats = st.as_dict()
        over = np.nonzero(r.cigar_len < 0)[0]
        if retry and len(over):
            need = int((-1 - r.cigar_len[over]).max()) + 1
            sub = self.align_pairs(pairs.select(over), params, strategy, stride=need, retry=False)
            for k, p in enumerate(over):
                r._patched[int(p)] = sub.cigars[k]
        return r

    def align(self, ref: bytes, alt: bytes, params, strategy):
        """IntelSmithWaterman.align: (cigar, alignment offset), plus the score the Java side drops."""
        cap = 2 * max(len(ref), len(alt)) + 16          # the Java side allocates 2 * max(len) (:99)
        cig = C.create_string_buffer(cap)
        sc, off = C.c_int32(), C.c_int32()
        pm = CParams(*[int(x) for x in params])
        self._check(self.lib.sw_align_one(self._h, ref, C.c_int32(len(ref)), alt, C.c_int32(len(alt)), C.byref(pm),
                                          C.c_int(strategy), cig, C.c_int32(cap), C.byref(sc), C.byref(off)))
        return cig.value.decode(), off.value, sc.value

    def close(self):
        if self._h:
            self.lib.sw_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


__all__ = ["SmithWaterman", "SwPairs", "SwError", "from_lists", "load_library"]
