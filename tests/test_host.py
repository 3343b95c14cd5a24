"""CPU: host-side logic of libphmm_cuda (no device): exports, tables, planner, parser, scheduler."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import oracle
from conftest import ROOT, load_golden
from genomicsbench_palisade_b200 import phmm, synth
from genomicsbench_palisade_b200.batches import FlatBatches, from_lists


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "phmm_cuda.h")).read()
    body = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(phmm_[a-z0-9_]+)\s*\(", body))
    assert {"phmm_create", "phmm_compute_pairs", "phmm_compute_flat", "phmm_workload_run"} <= declared
    lib = phmm.load_library()
    for name in sorted(declared):
        assert hasattr(lib, name), "libphmm_cuda.so does not export " + name
    assert declared == set(phmm.EXPORTS)
    # the reference's own C++-linkage symbols (IntelPairHmmCSource.cpp:29,61,89,103; pairhmm_common.cc:3)
    for name in phmm.LEGACY_EXPORTS:
        assert hasattr(lib, name), name
    assert lib.phmm_abi_version() == 5


def test_no_cpu_fallback_without_a_device():
    lib = phmm.load_library()
    if lib.phmm_device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(phmm.PhmmError) as e:
        phmm.PairHMM()
    assert e.value.code == 4 and "no CPU fallback" in str(e.value)


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "genomicsbench_palisade_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in txt and "liboracle" not in txt and "libphmm_ref" not in txt, f
                assert "oracle/" not in txt, f


def test_uploaded_tables_equal_the_oracle_tables():
    t, o = phmm.get_tables(), oracle.port_tables()
    assert np.array_equal(t["ph2pr_f"].view(np.uint32), o["ph2pr_f"].view(np.uint32))
    assert np.array_equal(t["ph2pr_d"].view(np.uint64), o["ph2pr_d"].view(np.uint64))
    assert np.array_equal(t["mm_f"].view(np.uint32), o["mm_f"][:8256].view(np.uint32))
    assert np.array_equal(t["mm_d"].view(np.uint64), o["mm_d"][:8256].view(np.uint64))
    for k in ("init_f", "log10_init_f", "init_d", "log10_init_d"):
        assert t[k] == o[k], k


@pytest.mark.parametrize("precision", [0, 2])
def test_plan_covers_every_pair_exactly_once(precision):
    for fb in (synth.edge_cases(), synth.c1_small(0.03), synth.c2_large(60000), synth.c4_stress(800),
               synth.c5_skewed(6000)):
        s = phmm.plan_summary(fb, precision=precision)
        assert s["pairs"] == fb.n_pairs and s["cells"] == fb.n_cells
        assert (s["cover"] == 1).all()
        for kh, lanes, vl, ntasks in s["classes"]:
            assert lanes in (16, 32) and 1 <= kh <= 8 and vl == (2 if precision != 2 else 1) and ntasks > 0
    s = phmm.plan_summary(synth.c2_large(60000), seg_cols=700)
    assert (s["cover"] == 1).all()


def test_fast_plan_pairs_runs_of_haplotypes():
    # PHMM_OPT_FP32_KERNEL 1: haplotype-pair tasks (one read x two runs of haplotypes)
    for fb in (synth.edge_cases(), synth.c1_small(0.03), synth.c2_large(60000), synth.c4_stress(800),
               synth.c5_skewed(6000)):
        for seg in (0, 300, 700):
            s = phmm.plan_summary(fb, fast=True, seg_cols=seg)
            assert s["pairs"] == fb.n_pairs and s["cells"] == fb.n_cells
            assert (s["cover"] == 1).all()
            for kh, lanes, vl, ntasks in s["classes"]:
                assert lanes in (8, 16, 32) and 1 <= kh <= 16 and vl == 2 and ntasks > 0


def test_fp64_first_reads_are_planned_once_and_only_when_predicted():
    # PlanOptions::spec: reads the k-mer predictor expects to fail the FP32 pass against their whole batch get FP64-first
    # tasks (heads of the rescue lists) instead of a place in an FP32 class; every pair is still produced exactly once
    for fb in (synth.c1_small(0.03), synth.c2_large(60000), synth.c4_stress(800), synth.c5_skewed(6000), synth.edge_cases()):
        s = phmm.plan_summary(fb, fast=True, fp64_first=True)
        assert (s["cover"] == 1).all() and s["pairs"] == fb.n_pairs and s["cells"] == fb.n_cells
        _, pred = phmm.batch_costs(fb, fast=True, predicted=True)
        sure = (pred & 2) != 0
        nh_of_read = np.repeat(fb.batch_num_haps, fb.batch_num_reads)
        # (small calls are planned for latency and keep every read in the FP32 classes)
        assert s["fp64_first_pairs"] in (0, int(nh_of_read[sure & (fb.read_len <= 255)].sum()))
        assert (s["fp64_first_pairs"] > 0) == (s["fp64_first_cells"] > 0)
        assert sum(c for c, _ in s["class_cells"]) + s["fp64_first_cells"] == fb.n_cells
        assert not ((pred & 2) != 0)[(pred & 1) == 0].any()          # "all pairs, by a factor of ten" implies "most pairs"
    assert phmm.plan_summary(synth.c2_large(60000), fast=True, fp64_first=True)["fp64_first_pairs"] > 0
    clean = synth.generic(5, 200, (40, 40), (8, 8), (100, 150), (150, 300), foreign_frac=0.0)
    assert phmm.plan_summary(clean, fast=True, fp64_first=True)["fp64_first_pairs"] == 0
    dirty = synth.generic(5, 200, (40, 40), (8, 8), (100, 150), (150, 300), foreign_frac=0.9)
    assert phmm.plan_summary(dirty, fast=True, fp64_first=True)["fp64_first_pairs"] > 0.5 * dirty.n_pairs
    # the per-call rule (plan_spec): only where a read is enough device work to pay for its prediction, and only when a
    # sample of the reads has such reads at all
    many_haps = dict(reads_rng=(20, 40), haps_rng=(40, 64), rlen_rng=(100, 250), hlen_rng=(150, 473))
    assert phmm.plan_summary(synth.generic(6, 40, foreign_frac=0.2, **many_haps), fast=True, fp64_first="auto")["fp64_first_pairs"] > 0
    short_reads = dict(many_haps, rlen_rng=(60, 120))          # none of these hangs over its haplotype: nothing fails the FP32 pass
    assert phmm.plan_summary(synth.generic(6, 60, foreign_frac=0.0, **short_reads), fast=True, fp64_first="auto")["fp64_first_pairs"] == 0
    assert phmm.plan_summary(dirty, fast=True, fp64_first="auto")["fp64_first_pairs"] == 0       # 8 short haplotypes per batch
    assert phmm.plan_summary(dirty, fast=False, fp64_first="auto")["fp64_first_pairs"] == 0
    # a one-wave call (one region) keeps its one class: no FP64-first tasks there
    assert phmm.plan_summary(synth.generic(5, 1, (20, 20), (6, 6), (100, 150), (150, 300), foreign_frac=0.9),
                             fast=True, fp64_first=True)["fp64_first_pairs"] == 0


def test_random_shapes_every_pair_planned_exactly_once():
    # property: whatever the batch shapes, lengths and share of unrelated reads, and whichever plan (bit-exact classes,
    # haplotype-pair classes, FP64-first reads forced on or by the library's rule, short runs), every pair is produced once
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=80, deadline=None, derandomize=True)
    @given(seed=st.integers(0, 10_000), nb=st.integers(1, 12), rmax=st.integers(1, 40), hmax=st.integers(1, 30),
           rlen=st.integers(1, 520), hlen=st.integers(1, 900), ff=st.sampled_from([0.0, 0.3, 1.0]),
           mode=st.sampled_from([(False, False, 0), (True, False, 0), (True, True, 0), (True, "auto", 0), (True, True, 300)]))
    def check(seed, nb, rmax, hmax, rlen, hlen, ff, mode):
        fb = synth.generic(seed, nb, (1, rmax), (1, hmax), (1, rlen), (max(1, hlen // 3), hlen), foreign_frac=ff)
        fast, spec, seg = mode
        s = phmm.plan_summary(fb, fast=fast, fp64_first=spec, seg_cols=seg)
        assert s["pairs"] == fb.n_pairs and s["cells"] == fb.n_cells and (s["cover"] == 1).all()
        assert sum(c for c, _ in s["class_cells"]) + s["fp64_first_cells"] == fb.n_cells or s["n_classes"] > 16

    check()


def test_one_wave_calls_are_planned_with_one_class():
    # a call whose tasks fit the device in one wave (one active region): every read in the 32-lane class of the
    # longest read -- one launch per phase instead of one per class (phmm_plan.cpp: plan_one_class)
    region = synth.generic(synth.SEED0 + 71, 1, (60, 60), (8, 8), (100, 250), (200, 500))
    assert len(set(int(x + 32) // 32 for x in region.read_len)) > 2          # the reads span several class heights
    for fast in (False, True):
        s = phmm.plan_summary(region, fast=fast)
        assert (s["cover"] == 1).all()
        assert [(kh, lanes) for kh, lanes, _, _ in s["classes"]] == [((int(region.read_len.max()) + 32) // 32, 32)]
    # hundreds of regions in one call do not fit one wave: classes by read length again
    many = synth.generic(synth.SEED0 + 72, 64, (40, 80), (6, 10), (100, 250), (200, 500))
    assert len(phmm.plan_summary(many, fast=True)["classes"]) > 1 and len(phmm.plan_summary(many)["classes"]) > 1


def test_plan_rejects_bad_input():
    with pytest.raises(phmm.PhmmError) as e:
        phmm.plan_summary(from_lists([([(b"", b"", b"", b"", b"")], [b"ACGT"])]))
    assert e.value.code == 2                      # PHMM_ERR_EMPTY_SEQ: rslen == 0
    with pytest.raises(phmm.PhmmError) as e:
        phmm.plan_summary(from_lists([([(b"A", b"5", b"5", b"5", b"5")], [b""])]))
    assert e.value.code == 2                      # haplen == 0: the reference divides by it
    empty = from_lists([])
    s = phmm.plan_summary(empty)
    assert s["pairs"] == 0 and s["tasks"] == 0
    s = phmm.plan_summary(from_lists([([(b"ACGT", b"5555", b"5555", b"5555", b"5555")], [])]))
    assert s["pairs"] == 0


def test_parser_matches_the_reference_driver_semantics():
    fb = synth.generic(5, 7, (1, 9), (1, 5), (1, 80), (1, 120), q_rng=(6, 45), foreign_frac=0.2)
    txt = fb.to_text()
    a = phmm.parse_text(txt)
    b = FlatBatches.from_text(txt)                   # python mirror of read_batch
    for k in ("batch_num_reads", "batch_num_haps", "read_len", "hap_len", "rs", "q", "i", "d", "c", "hap"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
        assert np.array_equal(getattr(a, k), getattr(fb, k)), k
    # normalize(): max(floor, byte - 33) with floor 6 for q and 0 for i/d/c (_orig.cpp:88-92,120-123)
    t = phmm.parse_text(b"1 1\nACGT !!+5 !!!+ \"#$% ++++\nAC\n")
    assert t.q.tolist() == [6, 6, 10, 20] and t.i.tolist() == [0, 0, 0, 10] and t.d.tolist() == [1, 2, 3, 4]
    # any whitespace separates tokens, trailing whitespace ends the file (operator>> + ws)
    t2 = phmm.parse_text(b"  1\t1 ACGT\n!!+5\r\n!!!+ \"#$% ++++ AC \n\n")
    assert np.array_equal(t2.q, t.q) and t2.hap.tobytes() == b"AC"
    assert phmm.parse_text(b"").n_batches == 0
    for bad in (b"1 1\nACGT !!! !!!! !!!! !!!!\nAC\n", b"1 1\nACGT !!!! !!!! !!!!\n", b"x 1\n", b"2 1\nA ! ! ! !\nAC\n"):
        with pytest.raises(phmm.PhmmError) as e:
            phmm.parse_text(bad)
        assert e.value.code == 6


def test_binary_cache_round_trip_and_corruption(tmp_path):
    fb = synth.generic(11, 9, (1, 12), (1, 6), (1, 200), (1, 300), q_rng=(6, 45), foreign_frac=0.2)
    path = str(tmp_path / "b.phmmsoa")
    phmm.cache_write(fb, path)
    a = phmm.cache_open(path)
    for k in ("batch_num_reads", "batch_num_haps", "read_len", "read_off", "hap_len", "hap_off",
              "rs", "q", "i", "d", "c", "hap"):
        assert np.array_equal(getattr(a, k), getattr(fb, k)), k
    # a shard that points into larger pools is stored packed
    shard = fb.select_batches(np.arange(1, fb.n_batches, 3))
    phmm.cache_write(shard, path)
    b = phmm.cache_open(path)
    assert len(b.rs) == int(shard.read_len.sum()) and len(b.hap) == int(shard.hap_len.sum())
    ra, rb = oracle.run_port(shard), oracle.run_port(b)
    assert np.array_equal(ra.out.view(np.uint64), rb.out.view(np.uint64))
    # empty set of batches
    phmm.cache_write(fb.select_batches(np.arange(0)), path)
    assert phmm.cache_open(path).n_batches == 0
    # corruption, truncation and foreign files are refused, not read
    phmm.cache_write(fb, path)
    raw = bytearray(open(path, "rb").read())
    for pos in (3, 60, 130, len(raw) - 5):
        bad = bytearray(raw)
        bad[pos] ^= 0x40
        open(path, "wb").write(bad)
        with pytest.raises(phmm.PhmmError) as e:
            phmm.cache_open(path)
        assert e.value.code == 6, pos
    open(path, "wb").write(raw[:len(raw) // 2])
    with pytest.raises(phmm.PhmmError):
        phmm.cache_open(path)
    # a crafted header whose counts wrap when multiplied (n_reads = 2^61 -> 0 bytes of read_len) together with a
    # section table rewritten to match (the FNV sums are no integrity boundary: an attacker recomputes them)
    import struct

    def fnv(b):
        h, k = 0xcbf29ce484222325, 0
        while k + 8 <= len(b):
            h = ((h ^ int.from_bytes(b[k:k + 8], "little")) * 0x100000001b3) & (2 ** 64 - 1)
            k += 8
        for x in b[k:]:
            h = ((h ^ x) * 0x100000001b3) & (2 ** 64 - 1)
        return h

    evil = bytearray(raw)
    struct.pack_into("<q", evil, 24, 2 ** 61)                       # Header::n_reads
    tab = bytearray(evil[128:128 + 12 * 24])
    for sec in (2, 3):                                              # read_len, read_off: sizes wrap to 0
        off, _, _ = struct.unpack_from("<QQQ", tab, sec * 24)
        struct.pack_into("<QQQ", tab, sec * 24, off, 0, fnv(b""))
    evil[128:128 + 12 * 24] = tab
    if fnv(bytes(raw[128:128 + 12 * 24])) == struct.unpack_from("<Q", raw, 56)[0]:     # the test's FNV is the file's
        struct.pack_into("<Q", evil, 56, fnv(bytes(tab)))
        open(path, "wb").write(evil)
        with pytest.raises(phmm.PhmmError) as e:
            phmm.cache_open(path)
        assert e.value.code == 6 and "count larger than the file" in str(e.value)
    else:
        raise AssertionError("test FNV does not reproduce the file's table checksum")
    with pytest.raises(phmm.PhmmError) as e:
        phmm.cache_open(str(tmp_path / "missing"))
    assert e.value.code == 7


def test_gatk_testdata_loader():
    # the line format of GATK's pairhmm-testdata.txt as dataFileTest reads it
    # (PairHmmUnitTest.java:166-185); the file itself is not in the reference checkout, so the
    # vectors here are the reference's own simpleTest (:42-55) and SURVEY.md 8c's tiny.in pairs
    txt = (b"# hap read quals ins del gcp expected\n"
           b"ACGT ACGT LLLL LLLL LLLL LLLL -6.022797e-01\n"
           b"#comment line\n"
           b"  TTACGTACGTACTT\tACGTACGTAC IIIIIIIIII NNNNNNNNNN NNNNNNNNNN ++++++++++ -1.192551\n"
           b"ACGT ACGT !!+5 !!!+ \"#$% ++++ 0\n")
    fb, exp = phmm.parse_testdata(txt)
    assert fb.n_batches == 3 and fb.n_pairs == 3 and exp.tolist() == [-6.022797e-01, -1.192551, 0.0]
    assert fb.hap.tobytes() == b"ACGTTTACGTACGTACTTACGT" and fb.read_len.tolist() == [4, 10, 4]
    # normalize(): byte - 33, read quals floored at 6 (:268-278)
    assert fb.q[:4].tolist() == [43] * 4 and fb.q[-4:].tolist() == [6, 6, 10, 20]
    assert fb.i[-4:].tolist() == [0, 0, 0, 10] and fb.d[-4:].tolist() == [1, 2, 3, 4]
    r = oracle.run_port(fb)
    assert abs(r.out[0] - exp[0]) < 1e-5 and abs(r.out[1] - exp[1]) < 1e-5     # the test's own bar
    assert phmm.parse_testdata(b"# only a comment\n")[0].n_batches == 0
    for bad in (b"ACGT ACGT LLLL LLLL LLLL LLLL\n", b"ACGT ACGT LLL LLLL LLLL LLLL -1\n",
                b"ACGT ACGT LLLL LLLL LLLL LLLL x\n"):
        with pytest.raises(phmm.PhmmError) as e:
            phmm.parse_testdata(bad)
        assert e.value.code == 6


def test_lpt_partition():
    rng = np.random.default_rng(3)
    cost = rng.integers(1, 10**9, 500)
    for parts in (1, 2, 4, 8):
        owner = phmm.partition_lpt(cost, parts)
        assert owner.min() >= 0 and owner.max() < parts
        load = np.bincount(owner, weights=cost, minlength=parts)
        assert load.max() - load.min() <= cost.max()          # LPT bound
        assert load.max() / load.mean() < 1.01                # < 1 % imbalance
        assert np.array_equal(owner, phmm.partition_lpt(cost, parts))   # deterministic across ranks


def test_batch_costs_and_rescue_prediction():
    # phmm_batch_costs: the cost the scheduler balances -- register-tile geometry + predicted FP64 work
    for fb, lo in ((synth.c2_large(60_000), 0.97), (synth.c1_small(0.05), 0.97), (synth.c5_skewed(20_000), 0.97),
                   (synth.c4_stress(2_000), 0.80)):
        cost, pred = phmm.batch_costs(fb, fast=True, predicted=True)
        assert cost.shape == (fb.n_batches,) and (cost > 0).all()
        ref = oracle.run_port(fb)
        bos, brs = fb.batch_out_start, fb.batch_read_start
        act = np.concatenate([ref.rescued[bos[b]:bos[b + 1]].reshape(int(fb.batch_num_reads[b]), -1).mean(1) >= 0.5
                              for b in range(fb.n_batches)])
        assert (pred.astype(bool) == act).mean() >= lo, (pred.mean(), act.mean())
        # never cheaper than the cells themselves, and FP64-only / FP32-only never dearer than both ... than sensible
        c32 = phmm.batch_costs(fb, fast=True, precision=phmm.FLOAT_ONLY)
        assert (c32 <= cost).all() and (c32 >= fb.batch_cells).all()
    # the all-pairs cost of a batch with many rescued reads exceeds that of its clean twin
    clean = synth.generic(5, 4, (20, 20), (6, 6), (100, 150), (150, 300), foreign_frac=0.0)
    dirty = synth.generic(5, 4, (20, 20), (6, 6), (100, 150), (150, 300), foreign_frac=0.9)
    assert phmm.batch_costs(dirty).sum() > 1.5 * phmm.batch_costs(clean).sum()
    # LPT over predicted time balances predicted time
    fb = synth.c2_large(400_000)
    cost = phmm.batch_costs(fb)
    owner = phmm.partition_lpt(cost, 2)
    load = np.bincount(owner, weights=cost, minlength=2)
    assert load.max() / load.mean() < 1.05
    empty = synth.generic(1, 0, (1, 2), (1, 2), (10, 20), (10, 20))
    assert len(phmm.batch_costs(empty)) == 0
    # the planner's own account of padding: cells against the register-tile cells the FP32 kernels compute
    for fast in (False, True):
        p = phmm.plan_summary(fb, fast=fast, cover=False)
        assert sum(c for c, _ in p["class_cells"]) == p["cells"] == fb.n_cells
        assert sum(t for _, t in p["class_cells"]) == p["tile_cells"] and all(t >= c for c, t in p["class_cells"])
        assert 0.85 < p["cells"] / p["tile_cells"] < 1.0


def test_flatten_pairs_folds_the_testcase_slab_back_into_batches():
    # phmm_flatten_pairs: what a caller thread of the legacy symbol does before the dispatcher merges the calls
    import ctypes as C
    lib = phmm.load_library()
    fb = synth.generic(77, 5, (2, 9), (1, 6), (10, 120), (20, 200), q_rng=(6, 45), foreign_frac=0.2)
    rsl, hpl, rso, hpo = fb.pairs()
    n = len(rsl)
    tcs = (phmm.CTestcase * n)()
    base = {k: getattr(fb, k).ctypes.data for k in ("rs", "q", "i", "d", "c", "hap")}
    for p in range(n):
        t = tcs[p]
        t.rslen, t.haplen = int(rsl[p]), int(hpl[p])
        t.rs, t.q, t.i, t.d, t.c = (base[k] + int(rso[p]) for k in ("rs", "q", "i", "d", "c"))
        t.hap = base["hap"] + int(hpo[p])
    h = C.c_void_p()
    err = C.create_string_buffer(160)
    assert lib.phmm_flatten_pairs(tcs, C.c_int64(n), C.byref(h), err, C.c_size_t(160)) == 0
    got = phmm.flat_from_c(lib.phmm_parsed_batches(h).contents)
    lib.phmm_parsed_free(h)
    # the all-pairs slabs fold back into the reads x haplotypes batches they came from, every string stored once
    assert got.batch_num_reads.tolist() == fb.batch_num_reads.tolist()
    assert got.batch_num_haps.tolist() == fb.batch_num_haps.tolist()
    assert len(got.rs) == int(fb.read_len.sum()) and len(got.hap) == int(fb.hap_len.sum())
    a, b = oracle.run_port(fb), oracle.run_port(got)
    assert np.array_equal(a.out.view(np.uint64), b.out.view(np.uint64))
    tcs[3].rslen = 0
    assert lib.phmm_flatten_pairs(tcs, C.c_int64(n), C.byref(h), err, C.c_size_t(160)) == 2 and b"rslen" in err.value


def test_same_shape_parts_of_a_weak_scaling_run():
    a = synth.c2_large(60_000)
    b = synth.same_shape(a, 7)
    c = synth.same_shape(a, 8)
    for x in (b, c):
        assert x.n_cells == a.n_cells and x.n_pairs == a.n_pairs
        assert np.array_equal(x.read_len, a.read_len) and np.array_equal(x.hap_len, a.hap_len)
        assert np.array_equal(x.batch_num_reads, a.batch_num_reads)
    assert (b.rs != c.rs).mean() > 0.5 and (b.hap != a.hap).mean() > 0.5
    ra, rb = oracle.run_port(a), oracle.run_port(b)
    assert abs(ra.rescued.mean() - rb.rescued.mean()) < 0.05 and rb.rescued.any()
    assert np.array_equal(synth.same_shape(a, 7).rs, b.rs)       # deterministic


def test_bench_parity_block_and_config():
    import bench
    fb, g = load_golden("c2_large")
    ref = oracle.run_port(fb)
    ok = bench.parity_block(ref, (ref.out.copy(), ref.rescued.copy()), "fast")
    assert ok["ok"] and ok["rescue_symdiff"] == 0 and ok["pairs"] == fb.n_pairs and ok["max_rel_fp32"] == 0.0
    out = ref.out.copy()
    k = int(np.nonzero(ref.rescued == 0)[0][0])
    out[k] *= 1 + 5e-4
    assert not bench.parity_block(ref, (out, ref.rescued.copy()), "fast")["ok"]
    resc = ref.rescued.copy()
    resc[k] = 1
    bad = bench.parity_block(ref, (ref.out.copy(), resc), "fast")
    assert not bad["ok"] and bad["rescue_symdiff"] == 1
    out = ref.out.copy()
    k = int(np.nonzero(ref.rescued == 1)[0][0])
    out[k] *= 1 + 1e-8
    assert not bench.parity_block(ref, (out, ref.rescued.copy()), "fast")["ok"]
    # a prefix: the reference ran on the first pairs only
    half = fb.select_batches(np.arange(1))
    rh = oracle.run_port(half)
    assert bench.parity_block(rh, (ref.out, ref.rescued), "exact")["pairs"] == half.n_pairs
    # both arms print the same config block
    tiny = bench.make_workload("tiny", 1)
    assert bench.config_block("tiny", tiny, 1) == bench.config_block("tiny", bench.make_workload("tiny", 1), 1)
    assert bench.config_block("tiny", bench.make_workload("tiny", 2), 2)["cells"] == 2 * tiny.n_cells


def test_batches_helpers():
    fb, g = load_golden("c1_small")
    rsl, hpl, _, _ = fb.pairs()
    assert int((rsl.astype(np.int64) * hpl).sum()) == fb.n_cells == int(fb.batch_cells.sum())
    sub, idx = fb.select_reads(fb.read_len <= 120)
    r = oracle.run_port(sub)
    assert np.array_equal(r.out, g["out"][idx])
    half = fb.select_batches(np.arange(0, fb.n_batches, 2))
    r2 = oracle.run_port(half)
    bos = fb.batch_out_start
    want = np.concatenate([g["out"][bos[b]:bos[b + 1]] for b in range(0, fb.n_batches, 2)])
    assert np.array_equal(r2.out, want)


def test_two_rank_sharding_gloo(tmp_path):
    """N > 1 path on CPU: two gloo ranks build the same LPT map, take disjoint batch sets that
    cover the workload, and the all-reduced cell counts add up (bench.py's sharding logic)."""
    script = tmp_path / "rank.py"
    script.write_text(
        "import os, sys, json\n"
        "sys.path.insert(0, %r)\n"
        "import numpy as np, torch, torch.distributed as dist\n"
        "import bench\n"
        "dist.init_process_group('gloo')\n"
        "rank, world = dist.get_rank(), dist.get_world_size()\n"
        "fb = bench.make_workload('tiny', world)\n"
        "mine, idx = bench.shard_of(fb, rank, world, True)\n"
        "t = torch.tensor([mine.n_cells, mine.n_pairs, mine.n_batches], dtype=torch.int64)\n"
        "dist.all_reduce(t)\n"
        "own = torch.zeros(fb.n_batches, dtype=torch.int64)\n"
        "own[idx] = 1\n"
        "dist.all_reduce(own)\n"
        "ok = bool((own == 1).all()) and int(t[0]) == fb.n_cells and int(t[1]) == fb.n_pairs\n"
        "ok = ok and fb.n_batches == 24 and fb.n_cells == 2 * bench.make_workload('tiny', 1).n_cells\n"
        "if rank == 0: print(json.dumps({'ok': ok, 'cells': int(t[0]), 'world': world}))\n"
        "dist.destroy_process_group()\n" % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29517", str(script)],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    import json
    res = json.loads(line)
    assert res["ok"] and res["world"] == 2
