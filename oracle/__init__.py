"""TEST INFRASTRUCTURE -- ctypes bindings for the CPU oracle of the phmm hot path.

Two libraries, both built by `oracle/build_ref.sh` into `oracle/_ref/` (git-ignored, but shipped
to the GPU box):

* `liboracle.so`     plain-C restatement (`oracle/phmm_oracle.c`), kind "port";
* `libphmm_ref.so`   the reference's own AVX/AVX-512 sources compiled in place from
                     /root/reference + `oracle/ref_harness.cpp`, kind "reference"
                     (`libphmm_ref_fma.so`: same with g++'s default FP contraction).

Only tests/, `__graft_entry__.smoke()` and bench.py's cpu_baseline / `--impl reference` legs
may import this package.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
MM_SIZE = (255 * 256) >> 1

_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(force: bool = False) -> None:
    """Compile the oracle (and the reference, when /root/reference exists)."""
    need = force or not os.path.exists(os.path.join(REF_DIR, "liboracle.so"))
    if os.path.isdir("/root/reference") and not os.path.exists(os.path.join(REF_DIR, "libphmm_ref.so")):
        need = True
    if need:
        subprocess.run(["bash", os.path.join(HERE, "build_ref.sh")], check=True,
                       stdout=subprocess.DEVNULL)


def have_ref(fma: bool = False) -> bool:
    return os.path.exists(os.path.join(REF_DIR, "libphmm_ref_fma.so" if fma else "libphmm_ref.so"))


_libs: dict = {}


def _load(name: str):
    if name not in _libs:
        path = os.path.join(REF_DIR, name)
        if not os.path.exists(path):
            build()
        _libs[name] = C.CDLL(path)
    return _libs[name]


class Result:
    def __init__(self, n):
        self.out = np.zeros(n, np.float64)       # log10 likelihoods, reference order
        self.raw32 = np.zeros(n, np.float32)     # FP32 sums (scaled by 2^120)
        self.raw64 = np.zeros(n, np.float64)     # FP64 sums (scaled by 2^1020), rescued pairs only
        self.rescued = np.zeros(n, np.uint8)


def _pools(fb):
    return [np.ascontiguousarray(x, np.uint8) for x in (fb.rs, fb.q, fb.i, fb.d, fb.c, fb.hap)]


def _run(fn, extra, fb, pairs, mode, nthreads):
    rsl, hpl, rso, hpo = pairs if pairs is not None else fb.pairs()
    n = len(rsl)
    res = Result(n)
    rs, q, i, d, c, hap = _pools(fb)
    # the kernels read exactly rslen/haplen bytes; pad pools so empty pools still have an address
    pad = lambda a: a if len(a) else np.zeros(1, np.uint8)
    fn(C.c_int64(n), np.ascontiguousarray(rsl, np.int32), np.ascontiguousarray(hpl, np.int32),
       np.ascontiguousarray(rso, np.int64), np.ascontiguousarray(hpo, np.int64),
       pad(rs), pad(q), pad(i), pad(d), pad(c), pad(hap),
       res.out, res.raw32.ctypes.data_as(C.c_void_p), res.raw64.ctypes.data_as(C.c_void_p),
       res.rescued.ctypes.data_as(C.c_void_p), C.c_int(mode), C.c_int(nthreads), *extra)
    return res


_COMMON = [C.c_int64, _i32p, _i32p, _i64p, _i64p, _u8p, _u8p, _u8p, _u8p, _u8p, _u8p,
           _f64p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]


def run_port(fb, pairs=None, mode=0, nthreads=0) -> Result:
    """The C restatement.  mode 0 = FP32 + FP64 rescue (the benchmark), 1 = FP32 only, 2 = FP64 only."""
    lib = _load("liboracle.so")
    lib.phmm_oracle_run.argtypes = _COMMON
    lib.phmm_oracle_run.restype = None
    return _run(lib.phmm_oracle_run, (), fb, pairs, mode, nthreads)


def run_ref(fb, pairs=None, mode=0, nthreads=0, isa=0, ftz=0, fma=False) -> Result:
    """The reference's own kernels.  isa 0 = its cpuid dispatch, 1 = AVX-256, 2 = AVX-512."""
    lib = _load("libphmm_ref_fma.so" if fma else "libphmm_ref.so")
    lib.phmm_ref_run.argtypes = _COMMON + [C.c_int, C.c_int]
    lib.phmm_ref_run.restype = None
    return _run(lib.phmm_ref_run, (C.c_int(isa), C.c_int(ftz)), fb, pairs, mode, nthreads)


def ref_isa() -> int:
    lib = _load("libphmm_ref.so")
    lib.phmm_ref_isa.restype = C.c_int
    return int(lib.phmm_ref_isa())


def _tables(fn):
    t = {"ph2pr_f": np.zeros(128, np.float32), "mm_f": np.zeros(MM_SIZE, np.float32),
         "ph2pr_d": np.zeros(128, np.float64), "mm_d": np.zeros(MM_SIZE, np.float64)}
    sc_f = (C.c_float * 2)()
    sc_d = (C.c_double * 2)()
    fn.restype = None
    fn(t["ph2pr_f"].ctypes.data_as(C.c_void_p), t["mm_f"].ctypes.data_as(C.c_void_p),
       t["ph2pr_d"].ctypes.data_as(C.c_void_p), t["mm_d"].ctypes.data_as(C.c_void_p),
       C.byref(sc_f, 0), C.byref(sc_f, 4), C.byref(sc_d, 0), C.byref(sc_d, 8))
    t["init_f"], t["log10_init_f"] = np.float32(sc_f[0]), np.float32(sc_f[1])
    t["init_d"], t["log10_init_d"] = np.float64(sc_d[0]), np.float64(sc_d[1])
    return t


def port_tables():
    return _tables(_load("liboracle.so").phmm_oracle_tables)


def ref_tables():
    return _tables(_load("libphmm_ref.so").phmm_ref_tables)


def ref_kat(rs: bytes, q: bytes, i: bytes, d: bytes, c: bytes, hap: bytes):
    """Single pair through the reference's real C entry points
    (computelikelihoodsfloat/double, IntelPairHmmCSource.cpp:89-114)."""
    lib = _load("libphmm_ref.so")
    f = C.c_float()
    dd = C.c_double()
    lib.phmm_ref_kat.restype = None
    lib.phmm_ref_kat(C.c_int(len(rs)), C.c_int(len(hap)), rs, q, i, d, c, hap, C.byref(f), C.byref(dd))
    return float(f.value), float(dd.value)
