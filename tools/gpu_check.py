#!/usr/bin/env python3
"""First-contact GPU script: pipe microbenchmarks, parity against the CPU oracle, kernel-class sweep."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from genomicsbench_palisade_b200 import phmm, synth

def parity(name, eng, fb, exact=True):
    ref = oracle.run_port(fb)
    w = eng.workload(fb)
    st = w.run()
    out, raw32, raw64, resc = w.results(raw=True)
    w.close()
    ok_resc = np.array_equal(resc, ref.rescued)
    raw_eq = np.array_equal(raw32.view(np.uint32), ref.raw32.view(np.uint32))
    nr = ref.rescued == 0
    rel32 = np.abs(out[nr] - ref.out[nr]) / np.maximum(np.abs(ref.out[nr]), 1e-30)
    rel64 = np.abs(out[~nr] - ref.out[~nr]) / np.maximum(np.abs(ref.out[~nr]), 1e-30)
    raw64_eq = np.array_equal(raw64[~nr], ref.raw64[~nr])
    print("PARITY %-28s pairs %7d resc %6d set_equal %s raw32_biteq %s (mismatch %d) raw64_biteq %s max_rel32 %.3g max_rel64 %.3g guard %d" % (
        name, fb.n_pairs, int(resc.sum()), ok_resc, raw_eq, int((raw32.view(np.uint32) != ref.raw32.view(np.uint32)).sum()),
        raw64_eq, rel32.max() if rel32.size else 0, rel64.max() if rel64.size else 0, st["n_guard"]), flush=True)
    return ok_resc and (raw_eq or not exact)

def main():
    eng = phmm.PairHMM()
    names = ["FFMA", "FFMA2", "FMUL+FADD", "FMUL2+FADD", "DFMA"]
    for k, n in enumerate(names):
        ops, ms = eng.microbench(k)
        print("MICRO %-12s %.3f T lane-ops/s  (%.2f ms)" % (n, ops / 1e12, ms), flush=True)
    eng.close()

    edge = synth.edge_cases(max_read_len=255)
    c1 = synth.c1_small(0.05)
    c4 = synth.c4_stress(1200)
    ok = True
    for label, kw in [("packed-exact", dict()), ("packed-fast", dict(fast=True)),
                      ("packed-exact-f64fast", dict(fast64=True))]:
        eng = phmm.PairHMM(**kw)
        exact = "exact" in label and "f64fast" not in label
        for nm, fb in [("edge", edge), ("c1", c1), ("c4", c4)]:
            ok &= parity(label + ":" + nm, eng, fb, exact="fast" not in label.split("-")[1][:4])
        eng.close()
    for prec, nm in [(1, "float-only"), (2, "double-only")]:
        eng = phmm.PairHMM(precision=prec)
        ref = oracle.run_port(c1, mode=prec)
        out = eng.compute(c1)
        print("MODE %s max abs diff %.3g" % (nm, np.abs(out - ref.out).max()), flush=True)
        eng.close()
    # host-buffer entry points
    eng = phmm.PairHMM()
    ref = oracle.run_port(edge)
    o1 = eng.compute(edge)
    o2 = eng.compute_pairs(edge)
    print("ENTRY compute_flat vs pairs equal:", np.array_equal(o1, o2), "max rel vs oracle %.3g" % (np.abs(o1 - ref.out) / np.abs(ref.out)).max(), flush=True)
    eng.close()

    # kernel-class sweep on a C2-shaped workload
    fb = synth.c2_large(400_000)
    print("SWEEP workload pairs %d cells %.3e" % (fb.n_pairs, fb.n_cells), flush=True)
    for label, kw in [("auto exact", dict()), ("auto fast", dict(fast=True)),
                      ("kh8 l32 exact", dict(force_class=(8, 32))), ("kh8 l32 fast", dict(force_class=(8, 32), fast=True))]:
        try:
            eng = phmm.PairHMM(**kw)
            w = eng.workload(fb)
            best = None
            for rep in range(3):
                st = w.run()
                if best is None or st["ms_total"] < best["ms_total"]:
                    best = st
            print("SWEEP %-20s ms_fp32 %8.3f ms_fp64 %8.3f total %8.3f  GCUPS %8.1f  classes %d tasks %d resc %d launches %d" % (
                label, best["ms_fp32"], best["ms_fp64"], best["ms_total"], fb.n_cells / best["ms_total"] / 1e6,
                best["n_classes"], best["n_tasks"], best["n_rescued"], best["n_launches"]), flush=True)
            w.close(); eng.close()
        except Exception as e:
            print("SWEEP %-20s FAILED %s" % (label, e), flush=True)
    print("ALL_OK" if ok else "PARITY_FAILED")

if __name__ == "__main__":
    main()
