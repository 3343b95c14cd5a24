"""TEST INFRASTRUCTURE -- ctypes bindings of the Smith-Waterman oracle (oracle/sw_oracle.c, kind
"port") and of the reference's own AVX2 / AVX-512 objects (oracle/_ref/libsw_ref.so, kind
"reference").  Same access rule as the rest of oracle/: tests, smoke() and the CPU-baseline legs of
the benchmarks only."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import REF_DIR, _load, build  # noqa: F401

SOFTCLIP, INDEL, LEADING_INDEL, IGNORE = 9, 10, 11, 12


def have_ref() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "libsw_ref.so"))


class SwResult:
    def __init__(self, n, stride):
        self.score = np.zeros(n, np.int32)
        self.offset = np.zeros(n, np.int32)
        self.cigar_len = np.zeros(n, np.int32)
        self.buf = np.zeros(max(1, n) * stride, np.uint8)
        self.stride = stride

    @property
    def cigars(self):
        # None where the CIGAR did not fit the stride (cigar_len = -1 - needed), as the product reports it
        return [self.buf[p * self.stride:p * self.stride + n].tobytes().decode() if n >= 0 else None
                for p, n in enumerate(self.cigar_len)]


def _run(fn, extra, pairs, params, strategy, nthreads, stride):
    n = pairs.n_pairs
    res = SwResult(n, stride)
    if n == 0:
        return res
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = fn(C.c_int64(n), vp(pairs.ref_pool), vp(pairs.ref_off), vp(pairs.ref_len), vp(pairs.alt_pool),
            vp(pairs.alt_off), vp(pairs.alt_len), *[C.c_int(int(x)) for x in params], C.c_int(strategy),
            vp(res.score), vp(res.offset), vp(res.buf), C.c_int(stride), vp(res.cigar_len),
            C.c_int(nthreads or (os.cpu_count() or 1)), *extra)
    res.rc = rc
    return res


def run_port(pairs, params, strategy, nthreads=0, stride=0):
    stride = stride or int(4 * (pairs.ref_len.astype(np.int64) + pairs.alt_len).max() + 16) if pairs.n_pairs else 16
    lib = _load("libsw_oracle.so")
    lib.sw_oracle_run.restype = None
    return _run(lib.sw_oracle_run, (), pairs, params, strategy, nthreads, stride)


def run_ref(pairs, params, strategy, nthreads=0, isa=0, stride=0):
    """The reference's own objects.  score holds the int16 the reference returns (sign-extended)."""
    stride = stride or int(4 * (pairs.ref_len.astype(np.int64) + pairs.alt_len).max() + 16) if pairs.n_pairs else 16
    lib = _load("libsw_ref.so")
    lib.sw_ref_run.restype = C.c_int
    return _run(lib.sw_ref_run, (C.c_int(isa),), pairs, params, strategy, nthreads, stride)
