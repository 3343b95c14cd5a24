"""Smith-Waterman with backtrack: oracle against the reference-generated golden
vectors (CPU), libsw_cuda through its C ABI against the oracle and the golden vectors (-m gpu).
Bar: bit-exact -- the same score, alignment offset and CIGAR string for every pair."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT
from genomicsbench_palisade_b200 import sw, sw_pairs
from oracle import sw as osw

SW_GOLDEN = ["sw_edge", "sw_haplotypes", "sw_divergent"]
STRATEGIES = [sw.SOFTCLIP, sw.INDEL, sw.LEADING_INDEL, sw.IGNORE]


def load_sw_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    ps = sw_pairs.SwPairs(z["ref_len"], z["ref_off"], z["ref_pool"], z["alt_len"], z["alt_off"], z["alt_pool"])
    return ps, z


def golden_cases(z):
    for k, pm in enumerate(z["params"]):
        for st in STRATEGIES:
            yield tuple(int(x) for x in pm), st, z["score16_%d_%d" % (k, st)], z["offset_%d_%d" % (k, st)], \
                z["cigar_%d_%d" % (k, st)].tobytes().decode().split("\n")


def assert_same(res, score16, offset, cigars):
    assert res.cigars == cigars
    assert np.array_equal(res.offset, offset)
    # the reference hands its int32 score back through an int16_t* (PairWiseSW.h:441)
    assert np.array_equal(res.score.astype(np.int16), score16)


# ---- CPU -------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", SW_GOLDEN)
def test_oracle_matches_reference_generated_golden(name):
    ps, z = load_sw_golden(name)
    for pm, st, score16, offset, cigars in golden_cases(z):
        assert_same(osw.run_port(ps, pm, st), score16, offset, cigars)


@pytest.mark.skipif(not osw.have_ref(), reason="reference library only exists in the build container")
def test_oracle_matches_live_reference():
    ps = sw_pairs.synth(99, 120, ref_rng=(10, 500), snp=0.05, indel=0.02, read_like=0.3)
    for pm in ((200, -150, -260, -11), (2, -3, -5, -1)):
        for st in STRATEGIES:
            a, b = osw.run_port(ps, pm, st), osw.run_ref(ps, pm, st)
            assert_same(a, b.score.astype(np.int16), b.offset, b.cigars)


def test_oracle_hand_checked_alignments():
    one = lambda r, a, st=sw.SOFTCLIP, pm=(200, -150, -260, -11): (
        lambda x: (x.cigars[0], int(x.offset[0]), int(x.score[0])))(osw.run_port(sw_pairs.from_lists([(r, a)]), pm, st))
    assert one(b"ACGTACGT", b"ACGTACGT") == ("8M", 0, 1600)
    assert one(b"TTTTACGTACGTTTTT", b"ACGTACGT") == ("8M", 4, 1600)            # offset = start on the reference
    assert one(b"ACGTACGT", b"GGGGACGTACGTGGGG") == ("4S8M4S", 0, 1600)        # overhangs soft-clipped
    assert one(b"ACGTACGTACGTACGT", b"ACGTACGTCCACGTACGT")[0] == "8M2I8M"
    assert one(b"ACGTACGTCCACGTACGT", b"ACGTACGTACGTACGT")[0] == "8M2D8M"
    assert one(b"ACGTACGT", b"GGGGACGTACGTGGGG", sw.INDEL)[0] == "4I8M4I"


def test_sw_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "sw_cuda.h")).read()
    body = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(sw_[a-z0-9_]+|initNative|align)\s*\(", body))
    lib = sw.load_library()
    for name in sorted(declared):
        assert hasattr(lib, name), "libsw_cuda.so does not export " + name
    assert declared == set(sw.EXPORTS)
    assert lib.sw_abi_version() == 1


def test_sw_no_cpu_fallback_without_a_device():
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from genomicsbench_palisade_b200 import sw\n"
            "try:\n    sw.SmithWaterman()\n    print('CREATED')\n"
            "except sw.SwError as e:\n    print('ERR', e.code, e)\n") % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300).stdout
    if "CREATED" in out:
        pytest.skip("a GPU is visible")
    assert "ERR 4" in out and "no CPU fallback" in out


# ---- GPU -------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", SW_GOLDEN)
def test_gpu_matches_golden(name):
    ps, z = load_sw_golden(name)
    eng = sw.SmithWaterman()
    for pm, st, score16, offset, cigars in golden_cases(z):
        assert_same(eng.align_pairs(ps, pm, st), score16, offset, cigars)
    eng.close()


@pytest.mark.gpu
def test_gpu_matches_oracle_on_random_pairs():
    eng = sw.SmithWaterman()
    sets = [sw_pairs.synth(5, 400, ref_rng=(1, 300), snp=0.03, indel=0.01, read_like=0.4),
            sw_pairs.synth(6, 60, ref_rng=(600, 1024), alts_per_ref=(2, 6), snp=0.02, indel=0.01),
            sw_pairs.synth(7, 150, ref_rng=(30, 400), snp=0.25, indel=0.08),
            sw_pairs.synth(8, 100, ref_rng=(5, 200), alphabet=b"AAAC"),          # low complexity: many ties
            sw_pairs.full_length_cases(),
            sw_pairs.from_lists([(b"A" * 1024, b"C" * 1024), (b"A" * 1023, b"C" * 1024)])]
    for ps in sets:
        # the last two parameter sets select the kernels that keep the MATRIX_MIN_CUTOFF clamp (sw_api.cu:
        # cutoff_can_bind); with the last one the clamp decides cells of the all-mismatch 1024 x 1024 pair
        for pm in ((200, -150, -260, -11), (1, -1, -1, -1), (25, -50, -110, -6), (200, -100000, -260, -11),
                   (100, -100000, -100000, -100000)):
            for st in STRATEGIES:
                ref = osw.run_port(ps, pm, st)
                got = eng.align_pairs(ps, pm, st)
                assert got.cigars == ref.cigars
                assert np.array_equal(got.offset, ref.offset) and np.array_equal(got.score, ref.score)
    eng.close()


@pytest.mark.gpu
def test_gpu_entry_points_agree_and_errors():
    eng = sw.SmithWaterman()
    ps = sw_pairs.synth(11, 50, ref_rng=(20, 300), read_like=0.3)
    pm = (200, -150, -260, -11)
    ref = osw.run_port(ps, pm, sw.SOFTCLIP)
    # device-resident workload, run twice (idempotent), other strategy on the same upload
    w = eng.workload(ps, stride=64)
    st = w.run(pm, sw.SOFTCLIP)
    a = w.results()
    w.run(pm, sw.SOFTCLIP)
    b = w.results()
    assert a.cigars == b.cigars == ref.cigars and st["n_cells"] == ps.n_cells and st["n_launches"] >= 1
    w.run(pm, sw.INDEL)
    assert w.results().cigars == osw.run_port(ps, pm, sw.INDEL).cigars
    w.close()
    # one pair at a time: alignNative's argument list
    for p in range(0, ps.n_pairs, 7):
        r, alt = ps.pair(p)
        cig, off, sc = eng.align(r, alt, pm, sw.SOFTCLIP)
        assert (cig, off, sc) == (ref.cigars[p], int(ref.offset[p]), int(ref.score[p]))
    # a stride too small for some CIGARs: flagged, nothing written, retried by the binding
    div = sw_pairs.synth(12, 40, ref_rng=(100, 300), snp=0.3, indel=0.1)
    rd = osw.run_port(div, pm, sw.SOFTCLIP)
    tight = eng.align_pairs(div, pm, sw.SOFTCLIP, stride=8, retry=False)
    assert (tight.cigar_len < 0).any() and tight.stats["n_overflow"] == int((tight.cigar_len < 0).sum())
    assert np.array_equal(-1 - tight.cigar_len[tight.cigar_len < 0],
                          np.array([len(c) for c in rd.cigars])[tight.cigar_len < 0])
    assert eng.align_pairs(div, pm, sw.SOFTCLIP, stride=8).cigars == rd.cigars
    # the reference's own C entry points (IntelSmithWatermanCSource.h)
    lib = sw.load_library()
    lib.initNative()
    buf = C.create_string_buffer(256)
    sc, off = C.c_int(-7), C.c_int(-7)
    lib.align(b"TTTTACGTACGTTTTT", b"ACGTACGT", buf, 200, -150, -260, -11, 9, C.byref(sc), C.byref(off))
    assert (buf.value, off.value, sc.value) == (b"8M", 4, 1600)
    # errors the reference leaves undefined
    for bad, code in (((b"", b"A"), 2), ((b"A" * 1025, b"A"), 3)):
        with pytest.raises(sw.SwError) as e:
            eng.align_pairs(sw_pairs.from_lists([bad]), pm, sw.SOFTCLIP)
        assert e.value.code == code
    with pytest.raises(sw.SwError):
        eng.align_pairs(ps, pm, 13)
    with pytest.raises(sw.SwError):
        eng.align_pairs(ps, (200, -150, -260, -10**6), sw.SOFTCLIP)
    assert eng.align_pairs(ps.select([]), pm, sw.SOFTCLIP).cigars == []
    eng.close()


@pytest.mark.gpu
def test_gpu_full_size_properties():
    """A region-scale batch (20 000 pairs): size-independent checks plus a seeded sample against the oracle."""
    ps = sw_pairs.synth(2026, 20000, ref_rng=(100, 600), read_like=0.2)
    pm = (200, -150, -260, -11)
    eng = sw.SmithWaterman()
    res = eng.align_pairs(ps, pm, sw.SOFTCLIP)
    cig = res.cigars
    # a CIGAR consumes exactly the alt sequence (M, I, S) and stays inside the reference (offset + M + D)
    pat = re.compile(r"(\d+)([MIDS])")
    for p in range(0, ps.n_pairs, 37):
        ops = pat.findall(cig[p])
        assert "".join(n + o for n, o in ops) == cig[p]
        assert sum(int(n) for n, o in ops if o in "MIS") == int(ps.alt_len[p])
        assert int(res.offset[p]) + sum(int(n) for n, o in ops if o in "MD") <= int(ps.ref_len[p])
    # permuting the pairs permutes the results
    perm = np.random.default_rng(1).permutation(ps.n_pairs)
    rp = eng.align_pairs(ps.select(perm), pm, sw.SOFTCLIP)
    assert rp.cigars == [cig[k] for k in perm] and np.array_equal(rp.score, res.score[perm])
    idx = np.arange(0, ps.n_pairs, 101)
    ref = osw.run_port(ps.select(idx), pm, sw.SOFTCLIP)
    assert ref.cigars == [cig[k] for k in idx] and np.array_equal(ref.score, res.score[idx])
    eng.close()


@pytest.mark.gpu
def test_gpu_red_zones_behind_device_buffers_stay_intact(monkeypatch):
    # compute-sanitizer is closed on the GPU pool: SW_CANARY=1 puts 0xA5 red zones behind every device buffer
    # (pools, results, backtrack words, CIGAR elements) and sw_workload_results verifies them after the kernels
    monkeypatch.setenv("SW_CANARY", "1")
    eng = sw.SmithWaterman()
    sets = [sw_pairs.synth(15, 300, ref_rng=(1, 300), snp=0.03, indel=0.01, read_like=0.4),
            sw_pairs.synth(16, 40, ref_rng=(600, 1024), alts_per_ref=(2, 6), snp=0.02, indel=0.01),
            sw_pairs.synth(17, 100, ref_rng=(30, 400), snp=0.25, indel=0.08),
            sw_pairs.full_length_cases(),
            sw_pairs.from_lists([(b"A" * 1024, b"C" * 1024), (b"A" * 1023, b"C" * 1024)])]
    for ps in sets:
        for pm in ((200, -150, -260, -11), (100, -100000, -100000, -100000)):
            for st in STRATEGIES:
                ref = osw.run_port(ps, pm, st)
                got = eng.align_pairs(ps, pm, st, stride=16)        # a short stride: overflow + retry path too
                assert got.cigars == ref.cigars
                assert np.array_equal(got.offset, ref.offset) and np.array_equal(got.score, ref.score)
    # the check itself: SW_CANARY=2 damages one zone on purpose and the call must report it
    monkeypatch.setenv("SW_CANARY", "2")
    with pytest.raises(sw.SwError, match="SW_CANARY"):
        eng.align_pairs(sets[0], (200, -150, -260, -11), STRATEGIES[0])
    eng.close()
