#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE's own code (oracle/_ref/libphmm_ref.so, i.e.
/root/reference/tools/GKL/src/main/native/pairhmm compiled in place by oracle/build_ref.sh with
-ffp-contract=off).  Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Each file holds the flat input batches and, per pair in reference order, the reference's log10
likelihood (`out`), raw FP32 sum (`raw32`), raw FP64 sum of rescued pairs (`raw64`) and the
FP64-rescue flag (`rescued`), plus FP32-only and FP64-only outputs of the single-pair entry
points' arithmetic (`out_f32only`, `out_f64only`)."""
import os, sys
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import numpy as np
import oracle
from genomicsbench_palisade_b200 import synth
from genomicsbench_palisade_b200.batches import FlatBatches, from_lists

FIELDS = ["batch_num_reads", "batch_num_haps", "read_len", "read_off", "rs", "q", "i", "d", "c",
          "hap_len", "hap_off", "hap"]


def save(name, fb):
    assert oracle.have_ref(), "reference library not built"
    r = oracle.run_ref(fb, mode=0, isa=0)
    r256 = oracle.run_ref(fb, mode=0, isa=1)
    assert np.array_equal(r.raw32.view(np.uint32), r256.raw32.view(np.uint32)), "AVX-512 != AVX-256"
    f32 = oracle.run_ref(fb, mode=1)
    f64 = oracle.run_ref(fb, mode=2)
    d = {k: getattr(fb, k) for k in FIELDS}
    d.update(out=r.out, raw32=r.raw32, raw64=r.raw64, rescued=r.rescued, out_f32only=f32.out,
             out_f64only=f64.out)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **d)
    print("%-14s pairs %6d cells %.2e rescued %5d  -> %s (%d KB)" % (
        name, fb.n_pairs, fb.n_cells, int(r.rescued.sum()), os.path.basename(path), os.path.getsize(path) // 1024))


def main():
    # the reference's one known-answer vector (PairHmmUnitTest.java:42-49): raw '+' quality bytes
    save("kat", from_lists([([(b"ACGT", b"++++", b"++++", b"++++", b"++++")], [b"ACGT"])]))
    # SURVEY.md 8c tiny.in (reference text format, parsed with the driver's normalisation)
    tiny = (b"2 2\nACGT LLLL NNNN NNNN ++++\nACGTACGTAC IIIIIIIIII NNNNNNNNNN NNNNNNNNNN ++++++++++\n"
            b"ACGT\nTTACGTACGTACTT\n")
    save("tiny", FlatBatches.from_text(tiny))
    save("edge", synth.edge_cases())
    save("c1_small", synth.generic(synth.SEED0 + 21, 6, (4, 30), (2, 12), (100, 150), (100, 500),
                                   foreign_frac=0.08, name="golden-c1"))
    save("c2_large", synth.generic(synth.SEED0 + 22, 2, (40, 60), (16, 24), (100, 250), (150, 473),
                                   foreign_frac=0.04, name="golden-c2"))
    # C3 shape (active regions, 250 bp reads x all candidate haplotypes) with 25 % unrelated reads so that the
    # FP64-rescue path of this shape is pinned too
    c3 = synth.c3_regions(3, seed=synth.SEED0 + 23)
    gen = dict(c3.meta["gen"], foreign_frac=0.25)
    rng = np.random.Generator(np.random.PCG64(synth.SEED0 + 123))
    save("c3_regions", synth._assemble(rng, c3.batch_num_reads, c3.batch_num_haps, c3.hap_len, c3.hap,
                                       c3.read_len.astype(np.int64), {"workload": "golden-c3"}, **gen))
    save("c4_stress", synth.generic(synth.SEED0 + 24, 3, (6, 6), (4, 4), (200, 250), (1000, 5000),
                                    q_rng=(6, 15), subst=0.20, name="golden-c4"))


if __name__ == "__main__":
    main()
