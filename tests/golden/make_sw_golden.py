"""Generates tests/golden/sw_*.npz from the REFERENCE's own Smith-Waterman objects
(oracle/_ref/libsw_ref.so = tools/GKL/src/main/native/smithwaterman/avx2_impl.cc + avx512_impl.cc
compiled in place by oracle/build_ref.sh).  Run in the build container only:
    python tests/golden/make_sw_golden.py
The reference has no golden vectors of its own for this path: its only test feeds a file that is
not in the checkout and asserts nothing (SmithWatermanUnitTest.java:21,45-63)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from genomicsbench_palisade_b200 import sw_pairs  # noqa: E402
from oracle import sw as osw  # noqa: E402

PARAM_SETS = [(200, -150, -260, -11),      # SmithWatermanUnitTest.java:48 (GATK's NEW_SW_PARAMETERS)
              (3, -1, -4, -3),             # :49, commented-out alternative
              (25, -50, -110, -6),         # GATK ALIGNMENT_TO_BEST_HAPLOTYPE_SW_PARAMETERS
              (1, -1, -1, -1)]             # many ties
STRATEGIES = [9, 10, 11, 12]


def dump(name, ps):
    out = {"ref_len": ps.ref_len, "ref_off": ps.ref_off, "ref_pool": ps.ref_pool, "alt_len": ps.alt_len,
           "alt_off": ps.alt_off, "alt_pool": ps.alt_pool}
    for k, pm in enumerate(PARAM_SETS):
        for st in STRATEGIES:
            a = osw.run_ref(ps, pm, st)
            b = osw.run_ref(ps, pm, st, isa=1)
            assert a.cigars == b.cigars and np.array_equal(a.offset, b.offset) and np.array_equal(a.score, b.score)
            out["score16_%d_%d" % (k, st)] = a.score.astype(np.int16)
            out["offset_%d_%d" % (k, st)] = a.offset
            out["cigar_%d_%d" % (k, st)] = np.frombuffer("\n".join(a.cigars).encode(), np.uint8)
    out["params"] = np.asarray(PARAM_SETS, np.int32)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **out)
    print(name, ps.n_pairs, "pairs", ps.n_cells, "cells")


if __name__ == "__main__":
    dump("sw_edge", sw_pairs.edge_cases())
    dump("sw_haplotypes", sw_pairs.synth(20261104, 160, ref_rng=(60, 700), read_like=0.25))
    dump("sw_divergent", sw_pairs.synth(20261105, 80, ref_rng=(20, 400), snp=0.2, indel=0.06, read_like=0.3))
