#!/usr/bin/env python3
"""One active region through the GATK-shaped API, N calls (for launch lists: ncu --metrics gpu__time_duration.sum)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from genomicsbench_palisade_b200 import phmm, synth
from tools.gatk_bench import regions_of

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
fb = synth.generic(synth.SEED0 + 60, 4, (20, 120), (2, 12), (100, 250), (200, 500), foreign_frac=0.03, name="gatk-regions")
reads, haps = regions_of(fb)[1]
g = phmm.GatkPairHMM()
for k in range(n):
    t = time.perf_counter()
    out = g.compute_likelihoods(reads, haps)
    print("call %d: %d reads x %d haps, %.3f ms" % (k, len(reads), len(haps), 1e3 * (time.perf_counter() - t)))
g.done()
