#!/usr/bin/env python3
"""Cell rate of every FP32 kernel class on reads that fill its register tile exactly (no padded rows):
haplotype-pair classes (fast) against the two-reads classes (fast, PHMM_NO_HP=1 and exact).
Development tool: python tools/class_sweep.py [hp|classic|exact]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genomicsbench_palisade_b200 import phmm, synth

def rate(kh, lanes, **kw):
    cap = kh * lanes - 1
    nb = max(8, min(64, 6_000_000 // (cap * 64)))
    fb = synth.generic(900 + cap, nb, (192, 192), (32, 32), (cap, cap), (290, 330), q_rng=(10, 40), subst=0.01)
    eng = phmm.PairHMM(precision=phmm.FLOAT_ONLY, force_class=(kh, lanes), **kw)
    w = eng.workload(fb)
    best = None
    for _ in range(3):
        st = w.run()
        if best is None or st["ms_fp32"] < best["ms_fp32"]:
            best = st
    w.close(); eng.close()
    return fb.n_cells / best["ms_fp32"] / 1e6, best

which = sys.argv[1] if len(sys.argv) > 1 else "hp"
if which == "hp":
    for lanes in (8, 16, 32):
        for kh in (4, 6, 8, 9, 10, 12, 13, 14, 16):
            g, st = rate(kh, lanes, fast=True)
            print("HP      kh %2d lanes %2d  rows %3d  %8.1f GCUPS  (%.3f ms, %d tasks, %d classes)" % (
                kh, lanes, kh * lanes, g, st["ms_fp32"], st["n_tasks"], st["n_classes"]), flush=True)
else:
    if which == "classic":
        os.environ["PHMM_NO_HP"] = "1"
    for lanes in (16, 32):
        for kh in (4, 6, 8):
            g, st = rate(kh, lanes, fast=(which == "classic"))
            print("%-7s kh %2d lanes %2d  rows %3d  %8.1f GCUPS  (%.3f ms, %d tasks)" % (
                which, kh, lanes, kh * lanes, g, st["ms_fp32"], st["n_tasks"]), flush=True)
