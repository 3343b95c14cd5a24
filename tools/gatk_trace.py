#!/usr/bin/env python3
"""PHMM_TRACE=1 timeline of one large phmm_gatk_compute_regions call (the workload of tools/gatk_bench.py)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
os.environ.setdefault("PHMM_TRACE", "1")
import ctypes as C  # noqa: E402
import numpy as np  # noqa: E402
from genomicsbench_palisade_b200 import phmm, synth  # noqa: E402
from gatk_bench import regions_of  # noqa: E402

fb = synth.generic(synth.SEED0 + 60, 1024, (20, 120), (2, 12), (100, 250), (200, 500), foreign_frac=0.03, name="gatk-regions")
regs = regions_of(fb)
g = phmm.GatkPairHMM(fast=(len(sys.argv) > 1 and sys.argv[1] == "fast"))
keep, outs = [], []
cg = (phmm.CGatkRegion * len(regs))()
for k, (reads, haps) in enumerate(regs):
    cr, ch = g._holders(reads, haps, keep)
    o = np.zeros(max(1, len(reads) * len(haps)), np.float64)
    outs.append(o)
    cg[k] = phmm.CGatkRegion(C.cast(cr, C.c_void_p), C.cast(ch, C.c_void_p), o.ctypes.data_as(C.c_void_p), len(reads), len(haps))
for it in range(3):
    sys.stderr.flush()
    print("---- call %d" % it, file=sys.stderr, flush=True)
    t = time.perf_counter()
    rc = g.lib.phmm_gatk_compute_regions(g._h, cg, C.c_int64(len(regs)))
    print("wall %.3f ms rc %d" % ((time.perf_counter() - t) * 1e3, rc), file=sys.stderr, flush=True)
