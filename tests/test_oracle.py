"""CPU: the oracle (plain-C restatement) against the reference's golden vectors.

Pins the oracle three ways: the reference's only in-repo known-answer test, golden vectors
written by the reference's own AVX code (tests/golden/make_golden.py), and -- when the reference
library was built here -- the live reference on fresh random pairs."""
import numpy as np
import pytest

import oracle
from conftest import GOLDEN_NAMES, load_golden
from genomicsbench_palisade_b200 import synth
from genomicsbench_palisade_b200.batches import FlatBatches, from_lists


def test_known_answer_vector():
    # tools/GKL/src/test/java/com/intel/gkl/pairhmm/PairHmmUnitTest.java:42-49,55:
    # hap ACGT, read ACGT, all four quality strings the raw bytes "++++", expected
    # -6.022797e-01 +- 1e-5 (FP32 path)
    fb = from_lists([([(b"ACGT", b"++++", b"++++", b"++++", b"++++")], [b"ACGT"])])
    r = oracle.run_port(fb)
    assert abs(r.out[0] - (-6.022797e-01)) < 1e-5
    assert r.rescued[0] == 0
    # the FP64-only path gives -6.022776589e-01: the KAT also pins "finalise in float"
    d = oracle.run_port(fb, mode=2)
    assert abs(d.out[0] - (-6.022776589e-01)) < 1e-9


def test_tiny_in_vectors():
    # output of the reference's example driver on tiny.in
    txt = (b"2 2\nACGT LLLL NNNN NNNN ++++\nACGTACGTAC IIIIIIIIII NNNNNNNNNN NNNNNNNNNN ++++++++++\n"
           b"ACGT\nTTACGTACGTACTT\n")
    r = oracle.run_port(FlatBatches.from_text(txt))
    np.testing.assert_allclose(r.out, [-0.647972, -0.890999, -10.098480, -1.192551], atol=1e-6)


def test_table_check_values():
    # SURVEY.md 8c table check values (reference Context.h compiled in the build container)
    t = oracle.port_tables()
    bits = lambda x: int(np.float32(x).view(np.uint32))
    assert bits(t["log10_init_f"]) == 0x42107e91
    assert np.float64(t["log10_init_d"]).view(np.uint64) == 0x407330cf3d4eda85
    for q, want in [(6, 0x3e809bcc), (10, 0x3dcccccd), (20, 0x3c23d70a), (30, 0x3a83126f),
                    (40, 0x38d1b717), (43, 0x3852368c), (45, 0x3804a2b3), (127, 0x2a60a58c)]:
        assert bits(t["ph2pr_f"][q]) == want, q
    mm = lambda a, b: max(a, b) * (max(a, b) + 1) // 2 + min(a, b)
    for (a, b), want in [((45, 45), 0x3f7ffbdb), ((43, 43), 0x3f7ff96e), ((40, 45), 0x3f7ff760),
                         ((10, 20), 0x3f63d70a), ((0, 0), 0), ((127, 127), 0x3f800000)]:
        assert bits(t["mm_f"][mm(a, b)]) == want, (a, b)
    assert np.float64(t["mm_d"][mm(45, 45)]).view(np.uint64) == 0x3fefff7b5d4d5b21
    assert np.float64(t["mm_d"][mm(127, 127)]).view(np.uint64) == 0x3feffffffffff1f6
    # the float tables are not the rounded double tables
    assert np.count_nonzero(t["mm_f"][:8256] != t["mm_d"][:8256].astype(np.float32)) > 100


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_port_matches_reference_golden(name):
    fb, g = load_golden(name)
    r = oracle.run_port(fb)
    assert np.array_equal(r.rescued, g["rescued"])
    assert np.array_equal(r.raw32.view(np.uint32), g["raw32"].view(np.uint32))     # bit-exact FP32 sums
    assert np.array_equal(r.raw64.view(np.uint64), g["raw64"].view(np.uint64))     # bit-exact FP64 sums
    assert np.array_equal(r.out.view(np.uint64), g["out"].view(np.uint64))
    assert np.array_equal(oracle.run_port(fb, mode=1).out.view(np.uint64), g["out_f32only"].view(np.uint64))
    assert np.array_equal(oracle.run_port(fb, mode=2).out.view(np.uint64), g["out_f64only"].view(np.uint64))


def test_golden_covers_the_interesting_cases():
    fb, g = load_golden("edge")
    assert g["rescued"].sum() > 0 and (g["rescued"] == 0).sum() > 0
    assert fb.read_len.min() == 1 and fb.hap_len.min() == 1
    rsl, hpl, _, _ = fb.pairs()
    assert (rsl > hpl).any()                      # R > H
    assert (fb.rs == ord("N")).any() and (fb.hap == ord("N")).any()
    assert (fb.q > 127).any()                     # bit 7 set: masked by the kernel


@pytest.mark.skipif(not oracle.have_ref(), reason="reference library only exists in the build container")
def test_port_matches_live_reference():
    for fb in (synth.generic(777, 4, (5, 20), (3, 9), (1, 300), (1, 600), q_rng=(6, 45), subst=0.1,
                             foreign_frac=0.2), synth.edge_cases(seed=4242)):
        a, b = oracle.run_port(fb), oracle.run_ref(fb)
        assert np.array_equal(a.raw32.view(np.uint32), b.raw32.view(np.uint32))
        assert np.array_equal(a.raw64.view(np.uint64), b.raw64.view(np.uint64))
        assert np.array_equal(a.rescued, b.rescued)
        assert np.array_equal(a.out.view(np.uint64), b.out.view(np.uint64))
        # AVX-256 and AVX-512 instantiations agree bit for bit (contraction off)
        c = oracle.run_ref(fb, isa=1)
        assert np.array_equal(b.raw32.view(np.uint32), c.raw32.view(np.uint32))


@pytest.mark.skipif(not oracle.have_ref(), reason="reference library only exists in the build container")
def test_reference_tables_match_port():
    a, b = oracle.port_tables(), oracle.ref_tables()
    for k in ("ph2pr_f", "mm_f", "ph2pr_d", "mm_d"):
        assert np.array_equal(a[k], b[k]), k
    assert a["log10_init_f"] == b["log10_init_f"] and a["log10_init_d"] == b["log10_init_d"]


@pytest.mark.skipif(not oracle.have_ref(fma=True), reason="reference library only exists in the build container")
def test_default_flag_reference_build_stays_in_tolerance():
    # g++'s default -ffp-contract=fast puts FMAs into the AVX-512 TU only (SURVEY.md 7 "hard
    # parts"): results move in the last bits, far inside the 1e-4 tolerance
    fb = synth.generic(99, 3, (10, 20), (4, 8), (100, 250), (150, 473), foreign_frac=0.05)
    a, b = oracle.run_ref(fb), oracle.run_ref(fb, fma=True)
    keep = (a.rescued == 0) & (b.rescued == 0)
    assert np.max(np.abs(a.out[keep] - b.out[keep]) / np.abs(a.out[keep])) < 1e-5


@pytest.mark.skipif(not oracle.have_ref(), reason="reference library only exists in the build container")
def test_reference_with_ftz_equals_reference_without():
    # the GATK entry (IntelPairHmm.cc:66-70) turns x86 flush-to-zero on, the benchmark's C entry
    # does not; include/phmm_cuda.h claims the outputs agree on anything but pathological pairs
    for fb in (synth.generic(31, 4, (10, 20), (4, 8), (100, 250), (150, 473), foreign_frac=0.3),
               load_golden("c4_stress")[0], load_golden("c1_small")[0]):
        a, b = oracle.run_ref(fb, ftz=0), oracle.run_ref(fb, ftz=1)
        assert np.array_equal(a.rescued, b.rescued)
        assert np.array_equal(a.out.view(np.uint64), b.out.view(np.uint64))
