#!/usr/bin/env python3
"""Two-second check of the FP64-first path on a B200: the fast kernels (FP64-first reads planned) against the bit-exact
kernels on the same 62 595 pairs -- same rescue set, outputs within the 1e-4 / 1e-9 bars, same rescued count.
(What tests/test_gpu_parity.py::test_fp64_first_reads_keep_results_and_rescue_set checks against the oracle, without it.)"""
import sys, time
t0=time.time()
sys.path.insert(0, '.')
import numpy as np
from genomicsbench_palisade_b200 import phmm, synth
fb = synth.generic(synth.SEED0 + 81, 40, (20, 40), (40, 64), (100, 250), (150, 473), foreign_frac=0.2)
print("gen %.2f s" % (time.time()-t0), flush=True)
res = {}
for fast in (True, False):
    eng = phmm.PairHMM(fast=fast, fast64=fast)
    w = eng.workload(fb); st = w.run(); out, raw32, raw64, resc = w.results(raw=True); w.close(); eng.close()
    res[fast] = (out, resc, st)
a, b = res[True], res[False]
assert np.array_equal(a[1], b[1]), "rescue sets differ"
r = b[1].astype(bool)
assert np.all(np.abs(a[0][~r]-b[0][~r]) <= 1e-4*np.abs(b[0][~r])) and np.all(np.abs(a[0][r]-b[0][r]) <= 1e-9*np.abs(b[0][r]))
print("OK spec pairs", a[2]["n_spec_pairs"], "direct", a[2]["n_spec_direct"], "rescued", a[2]["n_rescued"], b[2]["n_rescued"], "%.2f s" % (time.time()-t0))
