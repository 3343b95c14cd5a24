#!/usr/bin/env python3
"""Measurement of row (f1), the GATK-shaped batch API (phmm_gatk_*), on one B200:

    python tools/gatk_bench.py [--regions N] [--no-cpu]

HaplotypeCaller calls computeLikelihoods once per active region (tens of reads x a few haplotypes).
Reported: latency of one phmm_gatk_compute_likelihoods call per region, and throughput when G regions
are handed over together through phmm_gatk_compute_regions (G = 1, 16, 256, all), in regions/s and
GCUPS, next to the reference's own AVX code (oracle/_ref, FTZ on as initNative sets it,
IntelPairHmm.cc:66-70) looping over the same regions one at a time with all host threads.
Host buffers in, host results out: everything is end to end.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from genomicsbench_palisade_b200 import phmm, synth  # noqa: E402


def regions_of(fb):
    brs, bhs = fb.batch_read_start, fb.batch_hap_start
    regs = []
    for b in range(fb.n_batches):
        reads = []
        for r in range(brs[b], brs[b + 1]):
            o, n = int(fb.read_off[r]), int(fb.read_len[r])
            reads.append(tuple(getattr(fb, f)[o:o + n].tobytes() for f in ("rs", "q", "i", "d", "c")))
        haps = [fb.hap[int(fb.hap_off[h]):int(fb.hap_off[h]) + int(fb.hap_len[h])].tobytes()
                for h in range(bhs[b], bhs[b + 1])]
        regs.append((reads, haps))
    return regs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--regions", type=int, default=1024)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--kernels", default="exact", choices=["exact", "fast"],
                    help="exact = the entry's default (bit-exact kernels); fast = the opt-in (PHMM_OPT_FP32/FP64_KERNEL 1)")
    a = ap.parse_args()
    print(json.dumps(measure(a.regions, a.kernels, a.no_cpu)))


def measure(n_regions=1024, kernels="exact", no_cpu=False, device=0, sizes=(1, 16, 256, 0)):
    """sizes: regions per call, 0 = all in one call."""
    import types
    a = types.SimpleNamespace(regions=n_regions, kernels=kernels, no_cpu=no_cpu)
    # active regions: 20-120 reads of 100-250 bp x 2-12 haplotypes of 200-500 bp
    fb = synth.generic(synth.SEED0 + 60, a.regions, (20, 120), (2, 12), (100, 250), (200, 500), foreign_frac=0.03,
                       name="gatk-regions")
    regs = regions_of(fb)
    cells = fb.batch_cells.astype(np.int64)
    out = {"metric": "GATK-shaped batch API (phmm_gatk_*), end to end", "regions": len(regs), "pairs": fb.n_pairs,
           "cells": int(cells.sum()), "mean_pairs_per_region": fb.n_pairs / len(regs), "kernels": a.kernels, "groups": []}
    g = phmm.GatkPairHMM(device=device, fast=(a.kernels == "fast"))
    # ctypes marshalling of thousands of small byte strings is Python time, not library time: build the C
    # holder arrays once per group outside the timed region
    import ctypes as C

    def marshal(group):
        keep, outs = [], []
        cg = (phmm.CGatkRegion * len(group))()
        for k, (reads, haps) in enumerate(group):
            cr, ch = phmm.GatkPairHMM._holders(reads, haps, keep)
            o = np.zeros(max(1, len(reads) * len(haps)), np.float64)
            outs.append(o)
            cg[k] = phmm.CGatkRegion(C.cast(cr, C.c_void_p), C.cast(ch, C.c_void_p), o.ctypes.data, len(reads), len(haps))
        return cg, keep, outs

    ref_out = None
    for G in [x if x else len(regs) for x in sizes]:
        groups = [regs[k:k + G] for k in range(0, len(regs), G)]
        if G == 1:
            groups = groups[:256]                      # latency sample
        packed = [marshal(gr) for gr in groups]
        # warm-up (arenas, module load; large calls: the staging pools of the pipelined path are allocated by the first
        # such call and page-locked by the second, ~20 ms once per handle)
        for cg, _, _ in (packed[:2] if G < 256 else packed + packed + packed[:1]):
            g._check(g.lib.phmm_gatk_compute_regions(g._h, cg, C.c_int64(len(cg))))
        lat = []
        t0 = time.perf_counter()
        for cg, _, _ in packed:
            t = time.perf_counter()
            g._check(g.lib.phmm_gatk_compute_regions(g._h, cg, C.c_int64(len(cg))))
            lat.append(time.perf_counter() - t)
        dt = time.perf_counter() - t0
        nreg = sum(len(gr) for gr in groups)
        ncell = int(cells[:nreg].sum())
        out["groups"].append({"regions_per_call": G, "calls": len(groups), "regions_per_s": nreg / dt,
                              "gcups": ncell / dt / 1e9, "ms_per_call_median": 1e3 * float(np.median(lat)),
                              "ms_per_call_p95": 1e3 * float(np.percentile(lat, 95))})
        if G == len(regs):
            ref_out = np.concatenate([o[:len(r) * len(h)] for (_, _, outs), gr in zip(packed, groups)
                                      for o, (r, h) in zip(outs, gr)])
    g.done()
    if not a.no_cpu:
        import oracle
        kind = "reference" if oracle.have_ref() else "port"
        subs = [fb.select_batches(np.array([b])) for b in range(min(len(regs), 256))]
        t0 = time.perf_counter()
        for s in subs:
            (oracle.run_ref(s, nthreads=0, ftz=1) if kind == "reference" else oracle.run_port(s, nthreads=0))
        dt = time.perf_counter() - t0
        whole = oracle.run_ref(fb, nthreads=0, ftz=1) if kind == "reference" else oracle.run_port(fb, nthreads=0)
        r = whole.rescued.astype(bool)
        ok = bool(np.all(np.abs(ref_out[~r] - whole.out[~r]) <= 1e-4 * np.abs(whole.out[~r])) and
                  np.all(np.abs(ref_out[r] - whole.out[r]) <= 1e-9 * np.abs(whole.out[r])))
        out["cpu_baseline"] = {"kind": kind, "cores": os.cpu_count(), "ftz": True, "regions_per_s": len(subs) / dt,
                               "gcups": int(cells[:len(subs)].sum()) / dt / 1e9,
                               "ms_per_region": 1e3 * dt / len(subs),
                               "sample": "first %d regions, one call per region, OpenMP over the pairs of a region" % len(subs),
                               "gpu_results_within_tolerance": ok}
    return out


if __name__ == "__main__":
    main()
