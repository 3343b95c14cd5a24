"""GPU (-m gpu): libphmm_cuda through its C ABI against the CPU oracle and the golden vectors.

Bars (BASELINE.json north_star): FP64-rescue set identical; FP32-path log10 within 1e-4 relative;
rescued pairs within 1e-9 relative.  The bit-exact kernels are held to more: raw FP32 and FP64
sums bit-identical to the reference's AVX code."""
import ctypes as C

import numpy as np
import pytest

import oracle
from conftest import GOLDEN_NAMES, load_golden, max_rel, same_or_close
from genomicsbench_palisade_b200 import phmm, synth
from genomicsbench_palisade_b200.batches import from_lists

pytestmark = pytest.mark.gpu

RTOL_FP32 = 1e-4      # north_star: <= 1e-4 relative on the FP32 path
RTOL_FP64 = 1e-9      # north_star: <= 1e-9 relative on FP64-rescued pairs


def run(fb, **kw):
    eng = phmm.PairHMM(**kw)
    w = eng.workload(fb)
    st = w.run()
    res = w.results(raw=True)
    w.close()
    eng.close()
    return res, st


def check(res, ref_out, ref_raw32, ref_raw64, ref_resc, bit_exact32, bit_exact64):
    out, raw32, raw64, resc = res
    assert np.array_equal(resc, ref_resc), "FP64-rescue set differs"
    r = ref_resc.astype(bool)
    assert same_or_close(out[~r], ref_out[~r], RTOL_FP32), max_rel(out[~r], ref_out[~r])
    assert same_or_close(out[r], ref_out[r], RTOL_FP64), max_rel(out[r], ref_out[r])
    if bit_exact32:
        assert np.array_equal(raw32.view(np.uint32), ref_raw32.view(np.uint32))
    if bit_exact64:
        assert np.array_equal(raw64[r].view(np.uint64), ref_raw64[r].view(np.uint64))


@pytest.mark.parametrize("name", GOLDEN_NAMES)
@pytest.mark.parametrize("variant", ["exact", "fast", "fast-throughput-plan", "fast-latency-plan", "fast64"])
def test_golden_vectors(name, variant):
    fb, g = load_golden(name)
    # the golden sets are small, so "fast" alone takes the plan of a small call; the other two force each plan
    kw = {"exact": dict(), "fast": dict(fast=True), "fast-throughput-plan": dict(fast=True, plan=phmm.PLAN_THROUGHPUT),
          "fast-latency-plan": dict(fast=True, plan=phmm.PLAN_LATENCY), "fast64": dict(fast64=True)}[variant]
    res, _ = run(fb, **kw)
    check(res, g["out"], g["raw32"], g["raw64"], g["rescued"], variant in ("exact", "fast64"),
          variant != "fast64")


def test_known_answer_through_the_legacy_symbols():
    # PairHmmUnitTest.java:42-49,55 through the reference's own entry points
    lib = phmm.load_library()
    lib["_Z11initPairHMMv"]()
    bufs = [C.create_string_buffer(b, 4) for b in (b"ACGT", b"++++", b"++++", b"++++", b"++++", b"ACGT")]
    tc = phmm.CTestcase(4, 4, *[C.cast(b, C.c_void_p) for b in (bufs[1], bufs[2], bufs[3], bufs[4], bufs[5], bufs[0])])
    out = C.c_double()
    lib["_Z22computelikelihoodsbothP8testcasePdi"](C.byref(tc), C.byref(out), C.c_int(1))
    assert abs(out.value - (-6.022797e-01)) < 1e-5
    f = C.c_float()
    lib["_Z23computelikelihoodsfloatP8testcasePf"](C.byref(tc), C.byref(f))
    assert abs(f.value - (-6.022797e-01)) < 1e-5
    d = C.c_double()
    lib["_Z24computelikelihoodsdoubleP8testcasePd"](C.byref(tc), C.byref(d))
    assert abs(d.value - (-6.022776589e-01)) < 1e-9


def _testcases(fb):
    rsl, hpl, rso, hpo = fb.pairs()
    n = len(rsl)
    tcs = (phmm.CTestcase * max(1, n))()
    base = {k: getattr(fb, k).ctypes.data for k in ("rs", "q", "i", "d", "c", "hap")}
    for p in range(n):
        t = tcs[p]
        t.rslen, t.haplen = int(rsl[p]), int(hpl[p])
        ro, ho = int(rso[p]), int(hpo[p])
        t.rs, t.q, t.i, t.d, t.c = (base["rs"] + ro, base["q"] + ro, base["i"] + ro, base["d"] + ro, base["c"] + ro)
        t.hap = base["hap"] + ho
    return tcs, n


def test_legacy_symbol_coalesces_concurrent_callers():
    # the benchmark driver's call pattern (PairHMMUnitTest_orig.cpp:271-299): every OpenMP thread calls
    # computelikelihoodsboth with one batch at a time; behind the symbol one dispatcher merges whatever calls are
    # pending into one planned chunk.  Eight threads x several batches each, every result checked against the oracle
    import threading
    lib = phmm.load_library()
    lib["_Z11initPairHMMv"]()
    fn = lib["_Z22computelikelihoodsbothP8testcasePdi"]
    fb = synth.generic(123, 48, (2, 40), (1, 9), (20, 260), (30, 500), q_rng=(6, 45), subst=0.05, foreign_frac=0.2)
    ref = oracle.run_port(fb)
    bos = fb.batch_out_start
    subs = [fb.select_batches(np.array([b])) for b in range(fb.n_batches)]
    packed = [_testcases(s) for s in subs]
    outs = [np.full(max(1, n), np.nan) for _, n in packed]
    errs = []

    def worker(tid):
        try:
            for b in range(tid, len(subs), 8):
                tcs, n = packed[b]
                fn(tcs, outs[b].ctypes.data_as(C.c_void_p), C.c_int(n))
        except Exception as e:      # pragma: no cover
            errs.append(e)

    for rep in range(2):
        th = [threading.Thread(target=worker, args=(k,)) for k in range(8)]
        [t.start() for t in th]
        [t.join() for t in th]
    assert not errs
    for b, (_, n) in enumerate(packed):
        want, resc = ref.out[bos[b]:bos[b + 1]], ref.rescued[bos[b]:bos[b + 1]].astype(bool)
        got = outs[b][:n]
        assert same_or_close(got[~resc], want[~resc], RTOL_FP32) and same_or_close(got[resc], want[resc], RTOL_FP64), b


def test_async_submit_and_wait():
    # phmm_submit_pairs / phmm_wait: several requests per submission, several submissions in flight, waited for
    # out of order
    lib = phmm.load_library()

    class CReq(C.Structure):
        _fields_ = [("pairs", C.c_void_p), ("out", C.c_void_p), ("n", C.c_int64)]

    fb = synth.generic(321, 14, (2, 30), (1, 8), (20, 250), (30, 400), q_rng=(6, 45), subst=0.05, foreign_frac=0.2)
    ref = oracle.run_port(fb)
    bos = fb.batch_out_start
    packed = [_testcases(fb.select_batches(np.array([b]))) for b in range(fb.n_batches)]
    outs = [np.full(max(1, n), np.nan) for _, n in packed]
    eng = phmm.PairHMM()
    tickets = []
    for g in range(0, fb.n_batches, 2):          # seven submissions of two requests each: more than the four slots
        reqs = (CReq * 2)()
        for k in range(2):
            tcs, n = packed[g + k]
            reqs[k] = CReq(C.cast(tcs, C.c_void_p), outs[g + k].ctypes.data, n)
        t = C.c_int64()
        eng._check(lib.phmm_submit_pairs(eng._h, reqs, C.c_int(2), C.byref(t)))
        tickets.append(t.value)
    assert tickets == sorted(tickets) and len(set(tickets)) == len(tickets)
    st = phmm.CStats()
    eng._check(lib.phmm_wait(eng._h, C.c_int64(tickets[-1]), C.byref(st)))
    assert st.n_pairs == fb.n_pairs and st.n_cells == fb.n_cells and st.n_rescued == int(ref.rescued.sum())
    eng._check(lib.phmm_wait(eng._h, C.c_int64(tickets[0]), None))        # already retired: returns at once
    # the same through phmm_flatten_pairs + phmm_submit_flat (what the legacy symbols' dispatcher does)
    class CFlatReq(C.Structure):
        _fields_ = [("batches", C.c_void_p), ("out", C.c_void_p)]

    lib.phmm_parsed_batches.restype = C.c_void_p
    outs2 = [np.full(max(1, n), np.nan) for _, n in packed]
    handles = []
    for tcs, n in packed:
        h = C.c_void_p()
        assert lib.phmm_flatten_pairs(tcs, C.c_int64(n), C.byref(h), None, C.c_size_t(0)) == 0
        handles.append(h)
    for g in range(0, fb.n_batches, 7):
        reqs = (CFlatReq * 7)()
        for k in range(7):
            reqs[k] = CFlatReq(lib.phmm_parsed_batches(handles[g + k]), outs2[g + k].ctypes.data)
        t = C.c_int64()
        eng._check(lib.phmm_submit_flat(eng._h, reqs, C.c_int(7), C.byref(t)))
    eng._check(lib.phmm_wait(eng._h, t, None))
    lib.phmm_parsed_batches.restype = C.POINTER(phmm.CFlatBatches)
    for h in handles:
        lib.phmm_parsed_free(h)
    eng.close()
    for b, (_, n) in enumerate(packed):
        assert np.array_equal(outs2[b][:n], outs[b][:n])
    for b, (_, n) in enumerate(packed):
        want, resc = ref.out[bos[b]:bos[b + 1]], ref.rescued[bos[b]:bos[b + 1]].astype(bool)
        assert same_or_close(outs[b][:n][~resc], want[~resc], RTOL_FP32) and same_or_close(outs[b][:n][resc], want[resc], RTOL_FP64)


@pytest.mark.parametrize("precision", [phmm.FLOAT_ONLY, phmm.DOUBLE_ONLY])
def test_single_precision_modes(precision):
    fb, g = load_golden("c1_small")
    eng = phmm.PairHMM(precision=precision)
    out = eng.compute(fb)
    eng.close()
    want = g["out_f32only"] if precision == phmm.FLOAT_ONLY else g["out_f64only"]
    assert same_or_close(out, want, RTOL_FP32 if precision == phmm.FLOAT_ONLY else RTOL_FP64)


def test_host_entry_points_agree():
    fb, g = load_golden("c2_large")
    eng = phmm.PairHMM()
    a = eng.compute(fb)                 # phmm_compute_flat
    b = eng.compute_pairs(fb)           # computelikelihoodsboth's testcase[] argument list
    # phmm_compute_batches: the driver's Batch structs
    keep, batches = [], []
    brs, bhs, bos = fb.batch_read_start, fb.batch_hap_start, fb.batch_out_start
    results = np.zeros(fb.n_pairs, np.float64)

    class CRead(C.Structure):
        _fields_ = [("length", C.c_int), ("bases", C.c_void_p), ("q", C.c_void_p), ("i", C.c_void_p),
                    ("d", C.c_void_p), ("c", C.c_void_p)]

    class CHap(C.Structure):
        _fields_ = [("length", C.c_int), ("bases", C.c_void_p)]

    class CBatch(C.Structure):
        _fields_ = [("num_reads", C.c_int), ("num_haps", C.c_int), ("reads", C.c_void_p), ("haps", C.c_void_p),
                    ("results", C.c_void_p)]

    arr = (CBatch * fb.n_batches)()
    for k in range(fb.n_batches):
        nr, nh = int(fb.batch_num_reads[k]), int(fb.batch_num_haps[k])
        reads = (CRead * nr)()
        for j in range(nr):
            o, n = int(fb.read_off[brs[k] + j]), int(fb.read_len[brs[k] + j])
            reads[j] = CRead(n, *[getattr(fb, f).ctypes.data + o for f in ("rs", "q", "i", "d", "c")])
        haps = (CHap * nh)()
        for j in range(nh):
            haps[j] = CHap(int(fb.hap_len[bhs[k] + j]), fb.hap.ctypes.data + int(fb.hap_off[bhs[k] + j]))
        keep += [reads, haps]
        arr[k] = CBatch(nr, nh, C.cast(reads, C.c_void_p), C.cast(haps, C.c_void_p),
                        results.ctypes.data + 8 * int(bos[k]))
    eng._check(eng.lib.phmm_compute_batches(eng._h, arr, C.c_int64(fb.n_batches)))
    eng.close()
    assert np.array_equal(a, b) and np.array_equal(a, results)
    assert same_or_close(a, g["out"], RTOL_FP32)


def test_host_entry_point_pipelines_chunks():
    # phmm_compute_flat cuts the batches into chunks and overlaps plan + upload of one with the kernels
    # of the other: any chunk size must give the results of the single device-resident workload (bit for bit
    # with the bit-exact kernels)
    fb = synth.generic(77, 23, (5, 40), (1, 9), (20, 300), (30, 500), q_rng=(6, 45), subst=0.05, foreign_frac=0.2)
    ref = oracle.run_port(fb)
    for fast in (False, True):
        res, _ = run(fb, fast=fast)
        for chunk in (1, 200_000, 3_000_000, 0):
            eng = phmm.PairHMM(fast=fast, chunk_cells=chunk)
            out, st = eng.compute(fb, stats=True)
            out2 = eng.compute(fb, out=np.empty(fb.n_pairs))        # arenas reused
            eng.close()
            r = ref.rescued.astype(bool)
            # the low bits of the fast FP32 kernels depend on the class a read runs in; the plan (throughput or
            # latency) is decided once per call and the guard band per pair, so the cut into chunks changes nothing
            assert np.array_equal(out, res[0])
            if fast:
                assert same_or_close(out[~r], ref.out[~r], RTOL_FP32)
            assert np.array_equal(out2, out)
            assert st["n_pairs"] == fb.n_pairs and st["n_cells"] == fb.n_cells
            assert st["n_rescued"] == int(ref.rescued.sum())
    assert same_or_close(res[0][ref.rescued == 0], ref.out[ref.rescued == 0], RTOL_FP32)
    # a shard that references scattered, out-of-order pieces of shared pools (what the driver hands each
    # GPU): only the referenced segments are uploaded
    sub = fb.select_batches(np.array([19, 2, 11, 4, 3]))
    rs = oracle.run_port(sub)
    eng = phmm.PairHMM(chunk_cells=1_000_000)
    out, st = eng.compute(sub, stats=True)
    eng.close()
    r = rs.rescued.astype(bool)
    assert same_or_close(out[~r], rs.out[~r], RTOL_FP32) and same_or_close(out[r], rs.out[r], RTOL_FP64)
    assert st["h2d_bytes"] < 5 * int(sub.read_len.sum()) + 64 * len(sub.read_len) + int(sub.hap_len.sum()) * 2 + 200_000


def test_random_workloads_against_the_oracle():
    for fb in (synth.generic(31, 12, (1, 60), (1, 20), (1, 255), (1, 700), q_rng=(6, 45), subst=0.08,
                             foreign_frac=0.15),
               synth.c1_small(0.04), synth.c4_stress(800), synth.c5_skewed(4000)):
        ref = oracle.run_port(fb)
        for kw, e32 in ((dict(), True), (dict(plan=phmm.PLAN_THROUGHPUT), True),
                        (dict(fast=True, fast64=True, plan=phmm.PLAN_THROUGHPUT), False),
                        (dict(fast=True, fast64=True, plan=phmm.PLAN_LATENCY), False)):
            res, _ = run(fb, **kw)
            check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, e32, e32)


def cpu_reference(fb):
    """The reference's own AVX code when oracle/_ref holds it (it does on the GPU box), else the C restatement."""
    return oracle.run_ref(fb) if oracle.have_ref() else oracle.run_port(fb)


def test_c3_active_regions_against_the_oracle():
    # BASELINE.json configs[2]: every read 250 bp, all pairs of a region (IntelPairHmm.cc:99-145) -- every read
    # in one kernel class with both haplotype runs full, a shape no other set produces.  Both plans, the exact
    # and the fast kernels, and the GATK-shaped entry the upstream caller of this shape uses.
    fb = synth.c3_regions(24, seed=synth.SEED0 + 33)
    ref = cpu_reference(fb)
    assert fb.n_pairs > 30_000 and set(fb.read_len.tolist()) == {250}
    for kw, e32 in ((dict(plan=phmm.PLAN_THROUGHPUT), True), (dict(plan=phmm.PLAN_LATENCY), True),
                    (dict(fast=True, fast64=True, plan=phmm.PLAN_THROUGHPUT), False),
                    (dict(fast=True, fast64=True, plan=phmm.PLAN_LATENCY), False)):
        res, st = run(fb, **kw)
        check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, e32, e32)
    # with a quarter of the reads unrelated to the region, so that the rescue path of this shape runs too
    fb2, _ = load_golden("c3_regions")
    ref2 = cpu_reference(fb2)
    assert 0.05 < ref2.rescued.mean() < 0.95
    for kw, e32 in ((dict(), True), (dict(fast=True, fast64=True, plan=phmm.PLAN_THROUGHPUT), False)):
        res, st = run(fb2, **kw)
        check(res, ref2.out, ref2.raw32, ref2.raw64, ref2.rescued, e32, e32)
    g = phmm.GatkPairHMM()
    together = np.concatenate(g.compute_regions(_regions(fb2)))
    g.done()
    r = ref2.rescued.astype(bool)
    assert same_or_close(together[~r], ref2.out[~r], RTOL_FP32) and same_or_close(together[r], ref2.out[r], RTOL_FP64)


@pytest.mark.parametrize("name", ["c2", "c5"])
def test_benchmarked_path_on_a_large_slice(name):
    # what bench.py times -- fast FP32 + fast FP64 kernels, throughput plan, device-resident and through the
    # chunked host-buffer call -- on >= 1e5 pairs of the benchmark's own generators against the reference
    fb = synth.c2_large(200_000) if name == "c2" else synth.c5_skewed(100_000)
    assert fb.n_pairs >= 100_000
    ref = cpu_reference(fb)
    eng = phmm.PairHMM(fast=True, fast64=True)
    w = eng.workload(fb)
    st = w.run()
    res = w.results(raw=True)
    w.close()
    check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, False, False)
    assert st["n_rescued"] == int(ref.rescued.sum()) > 0
    out = eng.compute(fb)
    eng.close()
    assert np.array_equal(out, res[0])


def test_every_kernel_class_is_bit_exact():
    # pin each register-tile height / lane-group width in turn on reads that fit it; odd read counts
    # leave f32x2 halves and lane groups idle
    for lanes in (32, 16):
        for kh in range(1, 9):
            cap = kh * lanes - 1
            fb = synth.generic(1000 + kh * 100 + lanes, 3, (5, 11), (2, 6), (max(1, cap - 40), cap), (10, 300),
                               q_rng=(6, 45), subst=0.05, foreign_frac=0.2)
            ref = oracle.run_port(fb)
            for fast in (False, True):
                res, st = run(fb, force_class=(kh, lanes), fast=fast)
                assert st["n_classes"] == 1
                check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, not fast, True)


def test_every_haplotype_pair_class():
    # the fast FP32 kernels: one read per lane group against two runs of haplotypes; batches with
    # one haplotype leave the second run idle, odd read counts leave lane groups idle
    for lanes in (32, 16, 8):
        for kh in range(1, 17):
            cap = kh * lanes - 1
            fb = synth.generic(2000 + kh * 100 + lanes, 4, (3, 9), (1, 7), (max(1, cap - 40), cap), (10, 300),
                               q_rng=(6, 45), subst=0.05, foreign_frac=0.2)
            ref = oracle.run_port(fb)
            for seg in (0, 512):
                res, st = run(fb, force_class=(kh, lanes), fast=True, seg_cols=seg)
                assert st["n_classes"] == 1
                check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, False, True)


def test_long_reads_multi_pass():
    # reads far beyond the largest register tile; haplotypes shorter and longer than the reads
    fb = synth.generic(51, 3, (3, 6), (2, 5), (256, 1500), (50, 2500), q_rng=(6, 45), subst=0.05, foreign_frac=0.3)
    ref = oracle.run_port(fb)
    for kw, exact in ((dict(), True), (dict(fast=True), False)):
        res, _ = run(fb, **kw)
        check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, exact, True)
    eng = phmm.PairHMM(precision=phmm.DOUBLE_ONLY)
    out = eng.compute(fb)
    eng.close()
    assert same_or_close(out, oracle.run_port(fb, mode=2).out, RTOL_FP64)


def test_fast_kernel_keeps_the_rescue_set_near_the_threshold():
    # many pairs whose FP32 sum lands around MIN_ACCEPTED = 1e-28f: the fast kernel must send exactly
    # the reference's set to FP64 (guard band + bit-exact re-decision)
    fb = synth.generic(61, 40, (30, 30), (6, 6), (100, 150), (150, 300), q_rng=(20, 40), subst=0.15)
    ref = oracle.run_port(fb)
    frac = ref.rescued.mean()
    assert 0.02 < frac < 0.98
    res, st = run(fb, fast=True)
    assert np.array_equal(res[3], ref.rescued)
    near = (ref.raw32 > 0.99e-28) & (ref.raw32 < 1.01e-28)
    assert st["n_guard"] >= int(near.sum())


def test_fast_kernel_with_exotic_gap_penalties():
    """The fast FP32 kernel runs a rescaled recurrence (X / prod(pXX) inside groups of up to four rows);
    gap-continuation and gap-open qualities far from GATK's constants make that scale tiny or the gap
    terms large.  Overflowing pairs must come back through the bit-exact kernel, everything else stay
    inside the FP32 bar, and the rescue set stay identical."""
    rng = np.random.default_rng(7)
    guard_total = 0
    for case in range(4):
        fb = synth.generic(700 + case, 6, (4, 12), (2, 6), (20, 250), (30, 400), q_rng=(6, 45), subst=0.03,
                           foreign_frac=0.1)
        n = fb.c.size
        if case == 0:      # GCP anywhere in the Phred range, incl. 0 (pGAPM = 0) and very high
            fb.c[:] = rng.integers(0, 94, n)
        elif case == 1:    # high constant GCP: scale 1e-12 per group
            fb.c[:] = 40
            fb.i[:] = rng.integers(0, 10, n)
            fb.d[:] = rng.integers(0, 10, n)
        elif case == 2:    # cheap gaps everywhere
            fb.c[:] = rng.integers(1, 6, n)
            fb.i[:] = rng.integers(0, 46, n)
            fb.d[:] = rng.integers(0, 46, n)
        else:              # unmasked bytes: the kernel masks qualities with & 127 (:110-112,125)
            fb.c[:] = rng.integers(0, 256, n)
            fb.i[:] = rng.integers(0, 256, n)
            fb.d[:] = rng.integers(0, 256, n)
        ref = oracle.run_port(fb)
        res, st = run(fb, fast=True)
        check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, False, True)
        guard_total += st["n_guard"]
        eng = phmm.PairHMM(precision=phmm.FLOAT_ONLY, fast=True)
        out = eng.compute(fb)
        eng.close()
        assert same_or_close(out, oracle.run_port(fb, mode=1).out, RTOL_FP32)
        # the fast FP64 kernels use the same rescaling; their overflows (X / sigma beyond 1.8e308 with
        # the 2^1020 initial constant: GCP above ~Q60) come back through the bit-exact FP64 kernel
        res64, _ = run(fb, fast=True, fast64=True)
        check(res64, ref.out, ref.raw32, ref.raw64, ref.rescued, False, False)
        eng = phmm.PairHMM(precision=phmm.DOUBLE_ONLY, fast64=True)
        out = eng.compute(fb)
        eng.close()
        assert same_or_close(out, oracle.run_port(fb, mode=2).out, RTOL_FP64)
    assert guard_total > 0      # at least one case must have exercised the overflow path


def _regions(fb):
    """FlatBatches -> [(reads, haps)] in the GATK holder shape (bytes objects, raw Phred values)."""
    brs, bhs = fb.batch_read_start, fb.batch_hap_start
    regs = []
    for b in range(fb.n_batches):
        reads = []
        for r in range(brs[b], brs[b + 1]):
            o, n = int(fb.read_off[r]), int(fb.read_len[r])
            reads.append(tuple(getattr(fb, f)[o:o + n].tobytes() for f in ("rs", "q", "i", "d", "c")))
        haps = [fb.hap[int(fb.hap_off[h]):int(fb.hap_off[h]) + int(fb.hap_len[h])].tobytes()
                for h in range(bhs[b], bhs[b + 1])]
        regs.append((reads, haps))
    return regs


def test_gatk_shaped_api():
    # IntelPairHmm.computeLikelihoods (IntelPairHmm.cc:103-156): all reads x all haplotypes, read-major
    for use_double, fast in ((False, None), (True, None), (False, True)):
        g = phmm.GatkPairHMM(use_double=use_double, fast=fast)
        # the reference's known-answer test through this entry (PairHmmUnitTest.java:42-55): raw
        # (un-normalised) '+' bytes, tolerance 1e-5
        kat = g.compute_likelihoods([(b"ACGT", b"++++", b"++++", b"++++", b"++++")], [b"ACGT"])
        assert abs(kat[0] - (-6.022797e-01)) < 1e-5
        fb = synth.generic(91, 9, (1, 30), (1, 12), (10, 260), (20, 500), q_rng=(6, 45), subst=0.05, foreign_frac=0.25)
        ref = oracle.run_port(fb, mode=2 if use_double else 0)
        regs = _regions(fb)
        one_by_one = np.concatenate([g.compute_likelihoods(r, h) for r, h in regs])
        together = np.concatenate(g.compute_regions(regs))
        if use_double:
            assert np.array_equal(one_by_one, together)
            assert same_or_close(together, ref.out, RTOL_FP64)
        else:
            r = ref.rescued.astype(bool)
            assert r.any() and (~r).any()
            # default = bit-exact kernels: a pair's likelihood does not depend on what shares its call.  The fast
            # kernels are an opt-in: their low bits depend on the class a read runs in, which the plan picks by the
            # size of the call (a single region is planned for latency: 32 lanes per read, and every read -- its
            # FP64 rescue included -- in the class of the region's longest read)
            if not fast:
                assert np.array_equal(one_by_one, together)
            assert same_or_close(one_by_one[r], together[r], RTOL_FP64)
            assert same_or_close(one_by_one[~r], ref.out[~r], RTOL_FP32)
            assert same_or_close(together[~r], ref.out[~r], RTOL_FP32) and same_or_close(together[r], ref.out[r], RTOL_FP64)
        # empty region, and the errors the reference leaves undefined
        assert len(g.compute_likelihoods([], [b"ACGT"])) == 0
        with pytest.raises(phmm.PhmmError) as e:
            g.compute_likelihoods([(b"", b"", b"", b"", b"")], [b"ACGT"])
        assert e.value.code == 2
        g.done()
    with pytest.raises(phmm.PhmmError):
        lib = phmm.load_library()
        h = C.c_void_p()
        rc = lib.phmm_gatk_init(0, 0, 1, 1, C.byref(h))          # use_fpga
        raise phmm.PhmmError(rc, "use_fpga") if rc else AssertionError("use_fpga accepted")


_GATK_LARGE_CALL = r"""
import sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from genomicsbench_palisade_b200 import phmm, synth
sys.path.insert(0, sys.argv[1] + "/tests")
from test_gpu_parity import _regions
fb = synth.generic(synth.SEED0 + 61, 224, (40, 120), (4, 12), (100, 250), (200, 500), foreign_frac=0.03)
regs = _regions(fb)
regs.insert(100, ([], [b"ACGT"]))                      # an empty region in the middle of a group
eng = phmm.PairHMM(fast=False)
want = eng.compute(fb)
eng.close()
g = phmm.GatkPairHMM()
got = np.concatenate(g.compute_regions(regs))
assert got.shape == want.shape and np.array_equal(got.view(np.uint64), want.view(np.uint64)), "results differ"
# a malformed region late in the call: reported, nothing left in flight, the next call is fine
bad = list(regs)
bad[200] = ([(b"", b"", b"", b"", b"")], [b"ACGT"])
try:
    g.compute_regions(bad)
    raise SystemExit("empty read accepted")
except phmm.PhmmError as e:
    assert e.code == 2, e
again = np.concatenate(g.compute_regions(regs))
assert np.array_equal(again.view(np.uint64), want.view(np.uint64))
g.done()
# useDoublePrecision (IntelPairHmm.cc:123-124) through the same pipeline: every pair in FP64
eng = phmm.PairHMM(precision=phmm.DOUBLE_ONLY, fast=False)
want64 = eng.compute(fb)
eng.close()
g = phmm.GatkPairHMM(use_double=True)
got64 = np.concatenate(g.compute_regions(regs))
g.done()
assert np.array_equal(got64.view(np.uint64), want64.view(np.uint64)), "FP64-only results differ"
print("OK", fb.n_pairs, fb.n_cells)
"""


def test_gatk_large_call_is_staged_in_groups_under_the_kernels():
    # hundreds of regions in one phmm_gatk_compute_regions call: groups of regions staged by helper threads while
    # the previous groups compute (phmm_gatk.cpp: compute_regions_pipelined); same bits as phmm_compute_flat
    import os
    import subprocess
    import sys
    from conftest import ROOT
    env = dict(os.environ, PHMM_TRACE="1")
    pr = subprocess.run([sys.executable, "-c", _GATK_LARGE_CALL, ROOT], capture_output=True, text=True, timeout=600, env=env)
    assert pr.returncode == 0, pr.stderr[-3000:]
    assert pr.stdout.strip().startswith("OK")
    assert "gatk pipelined call" in pr.stderr, "the large call did not take the pipelined path"


_FP64_FIRST = r"""
import sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from genomicsbench_palisade_b200 import phmm, synth
fb = synth.generic(synth.SEED0 + 81, 40, (20, 40), (40, 64), (100, 250), (150, 473), foreign_frac=0.2)
eng = phmm.PairHMM(fast=True, fast64=True)
w = eng.workload(fb)
st = w.run()
st2 = w.run()                                   # the lists' heads are restored for every run
out, raw32, raw64, resc = w.results(raw=True)
w.close()
flat = eng.compute(fb)                          # and the chunked host-buffer call returns the same
eng.close()
assert st["n_rescued"] == st2["n_rescued"] and np.array_equal(flat.view(np.uint64), out.view(np.uint64))
np.savez(sys.argv[2], out=out, resc=resc, raw64=raw64, n_rescued=st["n_rescued"], spec=st["n_spec_pairs"],
         direct=st["n_spec_direct"], launches=st["n_launches"])
"""


def test_fp64_first_reads_keep_results_and_rescue_set(tmp_path):
    # Reads predicted to fail the FP32 pass against their whole batch run FP64 first (PlanOptions::spec).  Default,
    # every such pair forced through the band list and the bit-exact FP32 decision (PHMM_SPEC_BAND huge), and the feature
    # off must agree: same rescue set as the oracle, rescued pairs bit for bit, the rest within the FP32 bar.
    import os
    import subprocess
    import sys
    from conftest import ROOT
    fb = synth.generic(synth.SEED0 + 81, 40, (20, 40), (40, 64), (100, 250), (150, 473), foreign_frac=0.2)
    ref = oracle.run_port(fb)
    r = ref.rescued.astype(bool)
    res = {}
    for name, env in (("default", {}), ("band", {"PHMM_SPEC_BAND": "1e9"}), ("off", {"PHMM_SPEC64": "0"})):
        path = str(tmp_path / (name + ".npz"))
        pr = subprocess.run([sys.executable, "-c", _FP64_FIRST, ROOT, path], capture_output=True, text=True, timeout=600,
                            env=dict(os.environ, **env))
        assert pr.returncode == 0, pr.stderr[-3000:]
        z = res[name] = np.load(path)
        assert np.array_equal(z["resc"], ref.rescued), name
        assert int(z["n_rescued"]) == int(r.sum()), name
        assert same_or_close(z["out"][~r], ref.out[~r], RTOL_FP32) and same_or_close(z["out"][r], ref.out[r], RTOL_FP64), name
    assert int(res["default"]["spec"]) > 0.1 * fb.n_pairs and int(res["band"]["spec"]) == int(res["default"]["spec"])
    assert int(res["off"]["spec"]) == 0
    assert int(res["band"]["launches"]) > int(res["default"]["launches"])          # the decision pass ran
    for name in ("band", "off"):
        assert np.array_equal(res[name]["out"][r].view(np.uint64), res["default"]["out"][r].view(np.uint64)), name
        assert np.array_equal(res[name]["raw64"][r].view(np.uint64), res["default"]["raw64"][r].view(np.uint64)), name


def test_driver_cli_with_binary_cache(tmp_path):
    # the rewritten driver: text input, then the same input mapped from its binary cache
    import os
    import subprocess
    from conftest import ROOT
    fb = synth.generic(5, 6, (2, 12), (1, 6), (20, 200), (30, 400), q_rng=(6, 45), foreign_frac=0.2)
    ref = oracle.run_port(fb)
    path = str(tmp_path / "in.txt")
    open(path, "wb").write(fb.to_text())
    exe = os.path.join(ROOT, "genomicsbench_palisade_b200", "phmm")
    outs = []
    for args in (["-p"], ["-p", "-c"], ["-p", "-c"]):
        pr = subprocess.run([exe, "-f", path] + args, capture_output=True, text=True, timeout=300)
        assert pr.returncode == 0, pr.stderr
        lines = pr.stdout.strip().split("\n")
        assert lines[-1].startswith("PairHMM completed. Kernel runtime:")
        outs.append((np.array([float(x) for x in lines[:-1]]), pr.stderr))
    assert os.path.exists(path + ".phmmsoa")
    assert "parsed from text" in outs[1][1] and "mapped from cache" in outs[2][1]
    for o, _ in outs:
        assert len(o) == fb.n_pairs and np.allclose(o, ref.out, rtol=1e-4, atol=2e-6)     # "%lf": 6 decimals
    assert np.array_equal(outs[0][0], outs[2][0])


def test_driver_cli_pulls_many_chunks_from_the_queue(tmp_path):
    # the CLI's shared queue with chunks small enough that a single GPU keeps three in flight and retires many:
    # results land at each batch's own place, with the fast kernels behind it
    import os
    import subprocess
    from conftest import ROOT
    fb = synth.generic(15, 40, (2, 20), (1, 8), (20, 250), (30, 500), q_rng=(6, 45), foreign_frac=0.2)
    ref = oracle.run_port(fb)
    path = str(tmp_path / "in.txt")
    open(path, "wb").write(fb.to_text())
    exe = os.path.join(ROOT, "genomicsbench_palisade_b200", "phmm")
    for fast in ("0", "1"):
        pr = subprocess.run([exe, "-f", path, "-p", "-l", "2"], capture_output=True, text=True, timeout=300,
                            env=dict(os.environ, PHMM_CLI_CHUNK_COST="3e6", PHMM_FAST=fast))
        assert pr.returncode == 0, pr.stderr
        lines = pr.stdout.strip().split("\n")
        out = np.array([float(x) for x in lines[:-1]])
        assert len(out) == fb.n_pairs and np.allclose(out, ref.out, rtol=1e-4, atol=2e-6)
        n_chunks = int(pr.stderr.split("phmm: ")[1].split(" chunks")[0])
        assert n_chunks >= 8, pr.stderr
        assert "%d FP64-rescued" % int(ref.rescued.sum()) in pr.stderr


def test_page_locked_caller_buffers():
    # phmm_host_register: the same call from buffers the caller pinned once gives the same bits
    fb = synth.generic(19, 10, (5, 40), (2, 9), (20, 250), (30, 500), q_rng=(6, 45), foreign_frac=0.2)
    eng = phmm.PairHMM(fast=True, chunk_cells=2_000_000)
    a = eng.compute(fb)
    o = np.empty(fb.n_pairs)
    pools = (fb.rs, fb.q, fb.i, fb.d, fb.c, fb.hap, o)
    eng.pin(*pools)
    b = eng.compute(fb, out=o).copy()
    c = eng.compute(fb, out=o).copy()
    eng.unpin(*pools)
    d = eng.compute(fb)
    eng.close()
    assert np.array_equal(a, b) and np.array_equal(a, c) and np.array_equal(a, d)


def test_unmodified_reference_driver_linked_against_libphmm_cuda(tmp_path):
    """INTEGRATION.md section 1 for real: the reference's own driver source
    (tools/GKL/src/examples/PairHMM/PairHMMUnitTest.cpp, compiled unmodified by oracle/build_ref.sh in the
    build container) linked against libphmm_cuda.so prints the likelihoods its twin linked against the
    reference's own objects prints."""
    import os
    import subprocess
    from conftest import ROOT
    cuda = os.path.join(ROOT, "oracle", "_ref", "gkl_example_driver_cuda")
    ref = os.path.join(ROOT, "oracle", "_ref", "gkl_example_driver_ref")
    # built by oracle/build_ref.sh in the build container (where /root/reference exists) and shipped with the
    # tree; on a box with a GPU their absence is a failure of the build, not a reason to skip the boundary test
    assert os.path.exists(cuda) and os.path.exists(ref), "oracle/_ref/gkl_example_driver_{cuda,ref} missing: run __graft_entry__.build() where /root/reference exists"
    fb = synth.generic(17, 5, (2, 10), (1, 6), (20, 200), (30, 400), q_rng=(6, 45), foreign_frac=0.3)
    path = str(tmp_path / "in.txt")
    open(path, "wb").write(fb.to_text())
    outs = []
    for exe in (cuda, ref):
        pr = subprocess.run([exe, path], capture_output=True, text=True, timeout=300,
                            env=dict(os.environ, CHECK="1", OMP_NUM_THREADS="4"))
        assert pr.returncode == 0 and "PairHMM completed." in pr.stdout, pr.stderr
        vals = []
        for ln in pr.stdout.split("\n"):
            try:
                vals.append(float(ln))          # "%lf" lines of check_results (:158-161); the fork's own
            except ValueError:                  # "i: .. result_final" chatter and headers are skipped
                pass
        outs.append(np.array(vals))
    assert len(outs[0]) == len(outs[1]) == fb.n_pairs
    # default = bit-exact kernels: the FP32 / FP64 sums are identical; the printed value goes through
    # log10f / log10 (IntelPairHmmCSource.cpp:75,78), CUDA's on one side and glibc's on the other, which
    # differ in the last float bit now and then.  log10f(sum) lies in [8, 36.2) before the float subtraction
    # of log10f(2^120) (:78), so one ulp there is 3.8e-6: two of them + the "%lf" rounding, far inside 1e-4
    tol = 2 * 2.0 ** -18 + 1.1e-6 + 3 * 2.0 ** -23 * np.abs(outs[1])
    assert np.all(np.abs(outs[0] - outs[1]) <= tol), np.max(np.abs(outs[0] - outs[1]))
    assert np.all(np.abs(outs[0] - oracle.run_port(fb).out) <= tol)


def test_full_size_properties():
    """BASELINE config 2 at full size (2 M pairs): size-independent checks instead of the oracle."""
    fb = synth.c2_large()
    eng = phmm.PairHMM()
    w = eng.workload(fb)
    w.run()
    out, raw32, raw64, resc = w.results(raw=True)
    w.run()
    out2 = w.results()
    w.close()
    assert np.array_equal(out, out2)                               # deterministic / idempotent
    assert np.isfinite(out).all() and (out < 0).all()              # log10 of a probability
    # a different tiling (every read in the widest class) is a different program: identical raw sums
    eng_s = phmm.PairHMM(force_class=(8, 32))
    w = eng_s.workload(fb)
    w.run()
    _, raw32_s, raw64_s, resc_s = w.results(raw=True)
    w.close()
    eng_s.close()
    assert np.array_equal(raw32.view(np.uint32), raw32_s.view(np.uint32))
    assert np.array_equal(raw64.view(np.uint64), raw64_s.view(np.uint64)) and np.array_equal(resc, resc_s)
    # permuting the haplotypes of every batch permutes the results and nothing else
    rng = np.random.default_rng(1)
    bhs, bos = fb.batch_hap_start, fb.batch_out_start
    perm = np.concatenate([bhs[b] + rng.permutation(int(fb.batch_num_haps[b])) for b in range(fb.n_batches)])
    fb_p = type(fb)(fb.batch_num_reads, fb.batch_num_haps, fb.read_len, fb.read_off, fb.rs, fb.q, fb.i, fb.d,
                    fb.c, fb.hap_len[perm], fb.hap_off[perm], fb.hap)
    out_p = eng.compute(fb_p)
    eng.close()
    idx = []
    for b in range(fb.n_batches):
        nr, nh = int(fb.batch_num_reads[b]), int(fb.batch_num_haps[b])
        local = perm[bhs[b]:bhs[b] + nh] - bhs[b]
        idx.append((bos[b] + np.arange(nr)[:, None] * nh + local[None, :]).ravel())
    assert np.array_equal(out_p, out[np.concatenate(idx)])
    # a seeded sample against the oracle
    sub = fb.select_batches(np.array([0]))
    ref = oracle.run_port(sub)
    n = sub.n_pairs
    assert np.array_equal(raw32[:n].view(np.uint32), ref.raw32.view(np.uint32))
    assert np.array_equal(resc[:n], ref.rescued)


def test_red_zones_behind_device_buffers_stay_intact(monkeypatch):
    # compute-sanitizer is closed on the GPU pool, so out-of-bounds writes are looked for with the library's
    # own red zones (PHMM_CANARY=1: 0xA5 behind every device buffer of a workload, verified after each run)
    # over every kernel family: register-tile classes, haplotype-pair classes, multi-pass, FP64 rescue,
    # guard band, both plans, and the chunked host-buffer entry point
    monkeypatch.setenv("PHMM_CANARY", "1")
    fbs = [synth.generic(61, 8, (1, 40), (1, 12), (1, 255), (1, 700), q_rng=(6, 45), subst=0.08, foreign_frac=0.2),
           synth.generic(62, 3, (3, 6), (2, 5), (256, 1500), (50, 2500), q_rng=(6, 45), subst=0.05, foreign_frac=0.3),
           synth.c4_stress(300)]
    for fb in fbs:
        ref = oracle.run_port(fb)
        for kw, e32 in ((dict(), True), (dict(fast=True, fast64=True, plan=phmm.PLAN_THROUGHPUT), False),
                        (dict(fast=True, fast64=True, plan=phmm.PLAN_LATENCY), False)):
            res, _ = run(fb, **kw)
            check(res, ref.out, ref.raw32, ref.raw64, ref.rescued, e32, e32)
        eng = phmm.PairHMM(fast=True, chunk_cells=max(1, fb.n_cells // 5))
        out = eng.compute(fb)
        eng.close()
        r = ref.rescued.astype(bool)
        assert same_or_close(out[~r], ref.out[~r], RTOL_FP32) and same_or_close(out[r], ref.out[r], RTOL_FP64)
    # the check itself: PHMM_CANARY=2 damages one zone on purpose and the run must report it
    monkeypatch.setenv("PHMM_CANARY", "2")
    eng = phmm.PairHMM()
    w = eng.workload(fbs[0])
    with pytest.raises(phmm.PhmmError, match="PHMM_CANARY"):
        w.run()
    w.close()
    eng.close()
