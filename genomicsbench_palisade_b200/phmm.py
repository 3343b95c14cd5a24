"""ctypes binding of libphmm_cuda.so (the C ABI in include/phmm_cuda.h).

Mirrors the reference's library interface for the phmm path
(tools/GKL/src/main/native/pairhmm/IntelPairHmmCSource.cpp: initPairHMM :29,
computelikelihoodsboth :61, computelikelihoodsfloat :89, computelikelihoodsdouble :103) on top
of flat batches.  There is no CPU fallback: if the shared library is missing, or no B200 is
visible, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .batches import FlatBatches

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libphmm_cuda.so")

OPT_PRECISION, OPT_FP32_KERNEL, OPT_FP64_KERNEL, OPT_SEG_COLS, OPT_FORCE_CLASS, OPT_CHUNK_CELLS, OPT_PLAN = 1, 2, 3, 4, 6, 7, 8
PLAN_AUTO, PLAN_THROUGHPUT, PLAN_LATENCY = 0, 1, 2
BOTH, FLOAT_ONLY, DOUBLE_ONLY = 0, 1, 2
ERRORS = {1: "PHMM_ERR_ARG", 2: "PHMM_ERR_EMPTY_SEQ", 3: "PHMM_ERR_TOO_LONG", 4: "PHMM_ERR_CUDA",
          5: "PHMM_ERR_NOMEM", 6: "PHMM_ERR_PARSE", 7: "PHMM_ERR_IO"}


class PhmmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s (%d): %s" % (ERRORS.get(code, "PHMM_ERR"), code, msg))
        self.code = code


class CFlatBatches(C.Structure):
    _fields_ = [("n_batches", C.c_int64), ("batch_num_reads", C.c_void_p), ("batch_num_haps", C.c_void_p),
                ("n_reads", C.c_int64), ("read_len", C.c_void_p), ("read_off", C.c_void_p),
                ("rs", C.c_void_p), ("q", C.c_void_p), ("i", C.c_void_p), ("d", C.c_void_p), ("c", C.c_void_p),
                ("n_haps", C.c_int64), ("hap_len", C.c_void_p), ("hap_off", C.c_void_p), ("hap", C.c_void_p)]


class CStats(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("n_cells", C.c_int64), ("n_rescued", C.c_int64),
                ("n_rescued_cells", C.c_int64), ("n_guard", C.c_int64), ("n_tasks", C.c_int64),
                ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64), ("n_launches", C.c_int32),
                ("n_classes", C.c_int32), ("ms_fp32", C.c_float), ("ms_fp64", C.c_float),
                ("ms_total", C.c_float), ("ms_h2d", C.c_float), ("ms_d2h", C.c_float),
                ("n_spec_pairs", C.c_int64), ("n_spec_cells", C.c_int64), ("n_spec_direct", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class CTestcase(C.Structure):
    _fields_ = [("rslen", C.c_int), ("haplen", C.c_int), ("q", C.c_void_p), ("i", C.c_void_p),
                ("d", C.c_void_p), ("c", C.c_void_p), ("hap", C.c_void_p), ("rs", C.c_void_p)]


EXPORTS = ["phmm_abi_version", "phmm_device_count", "phmm_create", "phmm_destroy", "phmm_set_option",
           "phmm_get_option", "phmm_last_error", "phmm_compute_pairs", "phmm_compute_batches",
           "phmm_compute_flat", "phmm_submit_pairs", "phmm_submit_flat", "phmm_flatten_pairs", "phmm_wait", "phmm_host_register", "phmm_host_unregister", "phmm_workload_create", "phmm_workload_run", "phmm_workload_results",
           "phmm_workload_destroy", "phmm_partition_lpt", "phmm_batch_costs", "phmm_parse_file", "phmm_parse_buffer",
           "phmm_parsed_batches", "phmm_parsed_free", "phmm_microbench", "phmm_get_tables",
           "phmm_plan_summary", "phmm_cache_write", "phmm_cache_open", "phmm_gatk_init",
           "phmm_gatk_compute_likelihoods", "phmm_gatk_compute_regions", "phmm_gatk_done",
           "phmm_gatk_last_error", "phmm_gatk_context", "phmm_testdata_parse_buffer",
           "phmm_testdata_parse_file", "phmm_testdata_expected", "phmm_forward_region_noalloc"]
LEGACY_EXPORTS = ["_Z11initPairHMMv", "_Z22computelikelihoodsbothP8testcasePdi",
                  "_Z23computelikelihoodsfloatP8testcasePf", "_Z24computelikelihoodsdoubleP8testcasePd",
                  "_ZN11ConvertChar15conversionTableE"]

_lib = None


def load_library():
    """Load libphmm_cuda.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PhmmError(4, "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'`"
                               % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        lib.phmm_last_error.restype = C.c_char_p
        lib.phmm_last_error.argtypes = [C.c_void_p]
        lib.phmm_get_option.restype = C.c_int64
        lib.phmm_parsed_batches.restype = C.POINTER(CFlatBatches)
        lib.phmm_parsed_batches.argtypes = [C.c_void_p]
        lib.phmm_parsed_free.argtypes = [C.c_void_p]
        lib.phmm_parsed_free.restype = None
        lib.phmm_gatk_last_error.restype = C.c_char_p
        lib.phmm_gatk_last_error.argtypes = [C.c_void_p]
        lib.phmm_gatk_context.restype = C.c_void_p
        lib.phmm_gatk_context.argtypes = [C.c_void_p]
        lib.phmm_gatk_done.argtypes = [C.c_void_p]
        lib.phmm_testdata_expected.restype = C.c_int64
        lib.phmm_testdata_expected.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        lib.phmm_forward_region_noalloc.restype = C.c_int
        lib.phmm_forward_region_noalloc.argtypes = (
            [C.c_int, C.c_void_p] + [C.c_void_p] * 6 + [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32,
                                                         C.c_int32, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t])
        _lib = lib
    return _lib


# stated maxima of phmm_forward_region_noalloc (include/phmm_cuda.h)
NOALLOC_MAX_READS, NOALLOC_MAX_HAPS, NOALLOC_MAX_PAIRS = 1024, 128, 32768
NOALLOC_MAX_READ_LEN, NOALLOC_MAX_HAP_BYTES = 255, 65536


def forward_region_noalloc(device, stream, d_rs, d_q, d_i, d_d, d_c, d_read_off, n_reads, max_read_len,
                           d_hap, d_hap_off, n_haps, hap_bytes, d_out, d_rescued=None):
    """phmm_forward_region_noalloc (no counterpart in the reference; include/phmm_cuda.h): every pointer is a DEVICE
    address given as an int (e.g. torch.Tensor.data_ptr()), `stream` a cudaStream_t as an int or None.  Queues the
    work and returns; raises PhmmError with the library's message when a stated maximum is exceeded."""
    err = C.create_string_buffer(256)
    rc = load_library().phmm_forward_region_noalloc(
        device, stream, d_rs, d_q, d_i, d_d, d_c, d_read_off, n_reads, max_read_len, d_hap, d_hap_off, n_haps,
        hap_bytes, d_out, d_rescued, err, len(err))
    if rc != 0:
        raise PhmmError(rc, err.value.decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class _Keep:
    """A CFlatBatches plus the numpy arrays it points into."""

    def __init__(self, fb: FlatBatches):
        z = np.zeros(1, np.uint8)
        self.arrays = [fb.batch_num_reads, fb.batch_num_haps, fb.read_len, fb.read_off, fb.rs, fb.q,
                       fb.i, fb.d, fb.c, fb.hap_len, fb.hap_off, fb.hap, z]
        pool = lambda a: (a if a.size else z).ctypes.data_as(C.c_void_p)
        self.c = CFlatBatches(fb.n_batches, _ptr(fb.batch_num_reads), _ptr(fb.batch_num_haps),
                              len(fb.read_len), _ptr(fb.read_len), _ptr(fb.read_off),
                              pool(fb.rs), pool(fb.q), pool(fb.i), pool(fb.d), pool(fb.c),
                              len(fb.hap_len), _ptr(fb.hap_len), _ptr(fb.hap_off), pool(fb.hap))


def flat_from_c(cf: CFlatBatches) -> FlatBatches:
    """Copy a C-side phmm_flat_batches (e.g. from the parser) into numpy arrays."""
    def arr(p, n, dt):
        if not n:
            return np.zeros(0, dt)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(np.ctypeslib.as_ctypes_type(dt))), (n,)).copy()
    rl = arr(cf.read_len, cf.n_reads, np.int32)
    ro = arr(cf.read_off, cf.n_reads, np.int64)
    hl = arr(cf.hap_len, cf.n_haps, np.int32)
    ho = arr(cf.hap_off, cf.n_haps, np.int64)
    nrb = int((ro + rl).max()) if len(rl) else 0
    nhb = int((ho + hl).max()) if len(hl) else 0
    return FlatBatches(arr(cf.batch_num_reads, cf.n_batches, np.int32),
                       arr(cf.batch_num_haps, cf.n_batches, np.int32), rl, ro,
                       arr(cf.rs, nrb, np.uint8), arr(cf.q, nrb, np.uint8), arr(cf.i, nrb, np.uint8),
                       arr(cf.d, nrb, np.uint8), arr(cf.c, nrb, np.uint8), hl, ho, arr(cf.hap, nhb, np.uint8))


def parse_text(data: bytes) -> FlatBatches:
    """The library's parser of the reference's text format (no device needed)."""
    lib = load_library()
    h = C.c_void_p()
    err = C.create_string_buffer(256)
    rc = lib.phmm_parse_buffer(data, C.c_size_t(len(data)), C.byref(h), err, C.c_size_t(256))
    if rc:
        raise PhmmError(rc, err.value.decode())
    try:
        return flat_from_c(lib.phmm_parsed_batches(h).contents)
    finally:
        lib.phmm_parsed_free(h)


def cache_write(fb: FlatBatches, path: str) -> None:
    """phmm_cache_write: the binary cache of packed batches (no device needed)."""
    lib = load_library()
    keep = _Keep(fb)
    err = C.create_string_buffer(256)
    rc = lib.phmm_cache_write(C.byref(keep.c), os.fsencode(path), err, C.c_size_t(256))
    if rc:
        raise PhmmError(rc, err.value.decode())


def cache_open(path: str) -> FlatBatches:
    """phmm_cache_open (mmap + checksums), copied out into numpy arrays."""
    lib = load_library()
    h = C.c_void_p()
    err = C.create_string_buffer(256)
    rc = lib.phmm_cache_open(os.fsencode(path), C.byref(h), err, C.c_size_t(256))
    if rc:
        raise PhmmError(rc, err.value.decode())
    try:
        return flat_from_c(lib.phmm_parsed_batches(h).contents)
    finally:
        lib.phmm_parsed_free(h)


def parse_testdata(data: bytes):
    """GATK regression-vector lines (PairHmmUnitTest.java:166-185) -> (1 x 1 batches, expected)."""
    lib = load_library()
    h = C.c_void_p()
    err = C.create_string_buffer(256)
    rc = lib.phmm_testdata_parse_buffer(data, C.c_size_t(len(data)), C.byref(h), err, C.c_size_t(256))
    if rc:
        raise PhmmError(rc, err.value.decode())
    try:
        n = lib.phmm_testdata_expected(h, None, 0)
        exp = np.zeros(max(1, n), np.float64)
        lib.phmm_testdata_expected(h, _ptr(exp), n)
        return flat_from_c(lib.phmm_parsed_batches(h).contents), exp[:n]
    finally:
        lib.phmm_parsed_free(h)


class CGatkRead(C.Structure):
    _fields_ = [("readBases", C.c_void_p), ("readQuals", C.c_void_p), ("insertionGOP", C.c_void_p),
                ("deletionGOP", C.c_void_p), ("overallGCP", C.c_void_p), ("length", C.c_int32)]


class CGatkHap(C.Structure):
    _fields_ = [("haplotypeBases", C.c_void_p), ("length", C.c_int32)]


class CGatkRegion(C.Structure):
    _fields_ = [("reads", C.c_void_p), ("haps", C.c_void_p), ("likelihoods", C.c_void_p),
                ("n_reads", C.c_int32), ("n_haps", C.c_int32)]


class GatkPairHMM:
    """IntelPairHmm.initialize / computeLikelihoods / done (IntelPairHmm.java:75-125) over
    phmm_gatk_*.  A read is a 5-tuple of bytes objects (readBases, readQuals, insertionGOP,
    deletionGOP, overallGCP) with raw Phred values, a haplotype a bytes object."""

    def __init__(self, device: int = 0, use_double: bool = False, max_threads: int = 1, fast=None):
        """fast: None = the library's default (bit-exact kernels unless PHMM_GATK_FAST=1), True / False = set the
        kernels through the underlying context as a C caller would."""
        self.lib = load_library()
        self._h = C.c_void_p()
        rc = self.lib.phmm_gatk_init(C.c_int(device), C.c_int(int(use_double)), C.c_int(max_threads),
                                     C.c_int(0), C.byref(self._h))
        if rc:
            raise PhmmError(rc, (self.lib.phmm_gatk_last_error(None) or b"").decode())
        if fast is not None:
            ctx = C.c_void_p(self.lib.phmm_gatk_context(self._h))
            for opt in (OPT_FP32_KERNEL, OPT_FP64_KERNEL):
                self._check(self.lib.phmm_set_option(ctx, C.c_int(opt), C.c_int64(int(bool(fast)))))

    def _check(self, rc):
        if rc:
            raise PhmmError(rc, (self.lib.phmm_gatk_last_error(self._h) or b"").decode())

    @staticmethod
    def _holders(reads, haps, keep):
        cr = (CGatkRead * max(1, len(reads)))()
        for k, rd in enumerate(reads):
            bufs = [C.create_string_buffer(bytes(x), len(x)) if len(x) else C.create_string_buffer(1) for x in rd]
            keep.extend(bufs)
            cr[k] = CGatkRead(*[C.cast(b, C.c_void_p) for b in bufs], len(rd[0]))
        ch = (CGatkHap * max(1, len(haps)))()
        for k, hp in enumerate(haps):
            b = C.create_string_buffer(bytes(hp), len(hp)) if len(hp) else C.create_string_buffer(1)
            keep.append(b)
            ch[k] = CGatkHap(C.cast(b, C.c_void_p), len(hp))
        keep.extend([cr, ch])
        return cr, ch

    def compute_likelihoods(self, reads, haps) -> np.ndarray:
        keep = []
        cr, ch = self._holders(reads, haps, keep)
        out = np.zeros(max(1, len(reads) * len(haps)), np.float64)
        self._check(self.lib.phmm_gatk_compute_likelihoods(self._h, cr, C.c_int32(len(reads)), ch,
                                                           C.c_int32(len(haps)), _ptr(out)))
        return out[:len(reads) * len(haps)]

    def compute_regions(self, regions):
        """regions: list of (reads, haps); one library call for all of them."""
        keep, outs = [], []
        cg = (CGatkRegion * max(1, len(regions)))()
        for k, (reads, haps) in enumerate(regions):
            cr, ch = self._holders(reads, haps, keep)
            out = np.zeros(max(1, len(reads) * len(haps)), np.float64)
            outs.append(out)
            cg[k] = CGatkRegion(C.cast(cr, C.c_void_p), C.cast(ch, C.c_void_p), out.ctypes.data,
                                len(reads), len(haps))
        self._check(self.lib.phmm_gatk_compute_regions(self._h, cg, C.c_int64(len(regions))))
        return [o[:len(r) * len(h)] for o, (r, h) in zip(outs, regions)]

    def done(self):
        if self._h:
            self.lib.phmm_gatk_done(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.done()
        except Exception:
            pass


def partition_lpt(cost, parts: int) -> np.ndarray:
    lib = load_library()
    cost = np.ascontiguousarray(cost, np.int64)
    owner = np.zeros(len(cost), np.int32)
    rc = lib.phmm_partition_lpt(_ptr(cost), C.c_int64(len(cost)), C.c_int(parts), _ptr(owner))
    if rc:
        raise PhmmError(rc, "phmm_partition_lpt")
    return owner


def batch_costs(fb: FlatBatches, fast: bool = True, precision: int = BOTH, predicted=False):
    """Predicted device time per batch (phmm_batch_costs): the cost to hand to partition_lpt."""
    lib = load_library()
    keep = _Keep(fb)
    cost = np.zeros(max(1, fb.n_batches), np.int64)
    pred = np.zeros(max(1, len(fb.read_len)), np.uint8)
    rc = lib.phmm_batch_costs(C.byref(keep.c), C.c_int(int(fast)), C.c_int(precision), _ptr(cost), _ptr(pred))
    if rc:
        raise PhmmError(rc, "phmm_batch_costs")
    cost = cost[:fb.n_batches]
    return (cost, pred[:len(fb.read_len)]) if predicted else cost


def plan_summary(fb: FlatBatches, precision=0, seg_cols=0, cover=True, fast=False, fp64_first=False):
    """fp64_first (with fast): plan as the library does by default for the fast kernels -- reads predicted to fail the
    FP32 pass against their whole batch get FP64-first tasks instead of a place in an FP32 class; "auto" = only when the
    library's per-call rule (plan_spec: work per read, share of such reads in a sample) says so."""
    lib = load_library()
    keep = _Keep(fb)
    info = np.zeros(8 + 64 + 32 + 2, np.int64)
    cov = np.zeros(max(1, fb.n_pairs), np.uint8) if cover else None
    rc = lib.phmm_plan_summary(C.byref(keep.c), C.c_int(precision | (0x100 if fast else 0) | (0x400 if fp64_first == "auto" else 0x200 if fp64_first else 0)),
                               C.c_int(seg_cols),
                               _ptr(info), _ptr(cov) if cover else None)
    if rc:
        raise PhmmError(rc, "phmm_plan_summary")
    ncls = int(info[2])
    classes = [tuple(int(x) for x in info[8 + 4 * c: 12 + 4 * c]) for c in range(min(ncls, 16))]
    fill = [(int(info[72 + 2 * c]), int(info[73 + 2 * c])) for c in range(min(ncls, 16))]
    return {"pairs": int(info[0]), "cells": int(info[1]), "n_classes": ncls, "stream_bytes": int(info[3]),
            "max_rslen": int(info[4]), "max_haplen": int(info[5]), "tasks": int(info[6]),
            "tile_cells": int(info[7]), "class_cells": fill,
            "classes": classes, "cover": cov[:fb.n_pairs] if cover else None,
            "fp64_first_pairs": int(info[104]), "fp64_first_cells": int(info[105])}


def get_tables():
    lib = load_library()
    t = {"ph2pr_f": np.zeros(128, np.float32), "mm_f": np.zeros(8256, np.float32),
         "ph2pr_d": np.zeros(128, np.float64), "mm_d": np.zeros(8256, np.float64)}
    cf = np.zeros(2, np.float32)
    cd = np.zeros(2, np.float64)
    n = lib.phmm_get_tables(_ptr(t["ph2pr_f"]), _ptr(t["mm_f"]), _ptr(t["ph2pr_d"]), _ptr(t["mm_d"]),
                            _ptr(cf), _ptr(cd))
    assert n == 8256
    t.update(init_f=cf[0], log10_init_f=cf[1], init_d=cd[0], log10_init_d=cd[1])
    return t


class Workload:
    """Device-resident planned workload (phmm_workload_*)."""

    def __init__(self, engine: "PairHMM", fb: FlatBatches):
        self.engine = engine
        self.n_pairs = fb.n_pairs
        self._keep = _Keep(fb)
        self._h = C.c_void_p()
        engine._check(engine.lib.phmm_workload_create(engine._h, C.byref(self._keep.c), C.byref(self._h)))
        self._keep = None

    def run(self) -> dict:
        st = CStats()
        self.engine._check(self.engine.lib.phmm_workload_run(self.engine._h, self._h, C.byref(st)))
        return st.as_dict()

    def results(self, raw=False):
        n = self.n_pairs
        out = np.zeros(n, np.float64)
        if not raw:
            self.engine._check(self.engine.lib.phmm_workload_results(self.engine._h, self._h, _ptr(out), None, None, None))
            return out
        raw32 = np.zeros(n, np.float32)
        raw64 = np.zeros(n, np.float64)
        resc = np.zeros(n, np.uint8)
        self.engine._check(self.engine.lib.phmm_workload_results(self.engine._h, self._h, _ptr(out), _ptr(raw32),
                                                               _ptr(raw64), _ptr(resc)))
        return out, raw32, raw64, resc

    def close(self):
        if self._h:
            self.engine.lib.phmm_workload_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PairHMM:
    """One context on one B200 (phmm_create = the reference's initPairHMM)."""

    def __init__(self, device: int = 0, precision: int = BOTH, fast: bool = False, fast64: bool = False,
                 seg_cols: int = 0, force_class=None, chunk_cells: int = 0, plan: int = PLAN_AUTO):
        self.lib = load_library()
        self._h = C.c_void_p()
        rc = self.lib.phmm_create(C.c_int(device), C.byref(self._h))
        if rc:
            raise PhmmError(rc, (self.lib.phmm_last_error(None) or b"").decode())
        self.set_option(OPT_PRECISION, precision)
        self.set_option(OPT_FP32_KERNEL, int(fast))
        self.set_option(OPT_FP64_KERNEL, int(fast64))
        self.set_option(OPT_SEG_COLS, seg_cols)
        if force_class:
            self.set_option(OPT_FORCE_CLASS, (force_class[0] << 8) | force_class[1])
        if chunk_cells:
            self.set_option(OPT_CHUNK_CELLS, chunk_cells)
        if plan:
            self.set_option(OPT_PLAN, plan)

    def _check(self, rc):
        if rc:
            raise PhmmError(rc, (self.lib.phmm_last_error(self._h) or b"").decode())

    def set_option(self, opt, value):
        self._check(self.lib.phmm_set_option(self._h, C.c_int(opt), C.c_int64(int(value))))

    def workload(self, fb: FlatBatches) -> Workload:
        return Workload(self, fb)

    def compute(self, fb: FlatBatches, stats=False, out=None):
        """Host buffers in, log10 likelihoods out (phmm_compute_flat).  `out`: caller-owned float64
        array of fb.n_pairs to fill (as the reference driver reuses its results buffer)."""
        keep = _Keep(fb)
        if out is None:
            out = np.zeros(fb.n_pairs, np.float64)
        assert out.dtype == np.float64 and out.size >= fb.n_pairs and out.flags.c_contiguous
        st = CStats()
        self._check(self.lib.phmm_compute_flat(self._h, C.byref(keep.c), _ptr(out), C.byref(st)))
        return (out, st.as_dict()) if stats else out

    def pin(self, *arrays):
        """phmm_host_register on caller-owned numpy arrays that live across many calls."""
        for a in arrays:
            if a is not None and a.nbytes:
                self._check(self.lib.phmm_host_register(self._h, C.c_void_p(a.ctypes.data), C.c_size_t(a.nbytes)))

    def unpin(self, *arrays):
        for a in arrays:
            if a is not None and a.nbytes:
                self._check(self.lib.phmm_host_unregister(self._h, C.c_void_p(a.ctypes.data)))

    def compute_pairs(self, fb: FlatBatches):
        """Through the reference's own argument list: a testcase[] of pointers (computelikelihoodsboth)."""
        rsl, hpl, rso, hpo = fb.pairs()
        n = len(rsl)
        tcs = (CTestcase * max(1, n))()
        base = {k: getattr(fb, k).ctypes.data for k in ("rs", "q", "i", "d", "c", "hap")}
        for p in range(n):
            t = tcs[p]
            t.rslen, t.haplen = int(rsl[p]), int(hpl[p])
            ro, ho = int(rso[p]), int(hpo[p])
            t.rs, t.q, t.i, t.d, t.c = (base["rs"] + ro, base["q"] + ro, base["i"] + ro,
                                         base["d"] + ro, base["c"] + ro)
            t.hap = base["hap"] + ho
        out = np.zeros(n, np.float64)
        self._check(self.lib.phmm_compute_pairs(self._h, tcs, _ptr(out), C.c_int64(n)))
        return out

    def microbench(self, which: int):
        ops = C.c_double()
        ms = C.c_float()
        self._check(self.lib.phmm_microbench(self._h, C.c_int(which), C.byref(ops), C.byref(ms)))
        return ops.value, ms.value

    def close(self):
        if self._h:
            self.lib.phmm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
