#!/usr/bin/env python3
"""Device-resident GCUPS of a workload against the planner's run length (PHMM_OPT_SEG_COLS): shorter runs = more,
shorter tasks (better tails) but more wavefront fills.  python tools/seg_cols_sweep.py c1 [c2 ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from genomicsbench_palisade_b200 import phmm

for name in sys.argv[1:] or ["c1"]:
    fb = bench.make_workload(name, 1)
    for seg in (0, 4096, 2048, 1024, 512, 256):
        eng = phmm.PairHMM(fast=True, fast64=True, seg_cols=seg)
        w = eng.workload(fb)
        best = None
        for _ in range(6):
            st = w.run()
            if best is None or st["ms_total"] < best["ms_total"]:
                best = st
        w.close(); eng.close()
        print("%s seg_cols %4d: %.3f ms (fp32 %.3f fp64 %.3f) %.0f GCUPS, %d tasks" % (
            name, seg, best["ms_total"], best["ms_fp32"], best["ms_fp64"], fb.n_cells / best["ms_total"] / 1e6, best["n_tasks"]), flush=True)
