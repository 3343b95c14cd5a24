#!/usr/bin/env python3
"""The rewritten driver (`phmm` CLI: one process, one host thread per GPU pulling chunks of batches from a shared
queue) on N GPUs of one box: GCUPS from its own "Kernel runtime" line (plan + copies + kernels, as the reference
driver times its compute phase), parsing outside the timed region as in the reference.

    python tools/cli_bench.py [--gpus 1,2,4,8] [--workloads c2w,c5] [--loops 5]

c2w: N x the C2 set (weak scaling, same-shape parts); c5: a 2.5 M-pair slice of the C5 generator, the same file at
every N (strong scaling)."""
import argparse, json, os, re, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from genomicsbench_palisade_b200 import phmm, synth

EXE = os.path.join(ROOT, "genomicsbench_palisade_b200", "phmm")


def run(path, gpus, loops):
    pr = subprocess.run([EXE, "-f", path, "-c", "-t", str(gpus), "-l", str(loops)], capture_output=True, text=True,
                        env=dict(os.environ, PHMM_FAST="1"), timeout=1800)
    m = re.search(r"Kernel runtime: ([0-9.]+) sec", pr.stdout)
    if pr.returncode != 0 or not m:
        raise RuntimeError(pr.stderr[-800:])
    return float(m.group(1)), pr.stderr.strip().splitlines()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--workloads", default="c2w,c5")
    ap.add_argument("--loops", type=int, default=10)
    a = ap.parse_args()
    gpus = [int(x) for x in a.gpus.split(",")]
    tmp = tempfile.mkdtemp()
    out = {"what": "phmm CLI (csrc/driver/phmm_main.cpp), PHMM_FAST=1, -l %d; GCUPS = cells x loops / its 'Kernel runtime'" % a.loops,
           "runs": []}
    for w in a.workloads.split(","):
        base = None
        for n in gpus:
            if w == "c2w":
                fb = bench.make_workload("c2", n)
            elif base is None:
                fb = base = synth.c5_skewed(2_500_000)
            else:
                fb = base
            path = os.path.join(tmp, "%s_%d.in" % (w, n if w == "c2w" else 0))
            cache = path + ".phmmsoa"
            if not os.path.exists(cache):
                phmm.cache_write(fb, cache)             # the CLI maps the cache when it is newer than the text
                open(path, "wb").write(b"")             # (an empty text file older than the cache)
                os.utime(path, (time.time() - 100, time.time() - 100))
            run(path, n, 1)                             # warm-up: module load, arenas
            rt, log = run(path, n, a.loops)
            r = {"workload": w, "gpus": n, "pairs": fb.n_pairs, "cells": fb.n_cells, "loops": a.loops,
                 "kernel_runtime_s": rt, "gcups": fb.n_cells * a.loops / max(rt, 1e-3) / 1e9, "log": log[-2:]}
            out["runs"].append(r)
            print(json.dumps(r), file=sys.stderr, flush=True)
    for w in a.workloads.split(","):
        rs = [r for r in out["runs"] if r["workload"] == w]
        one = [r for r in rs if r["gpus"] == 1]
        if one:
            for r in rs:
                r["vs_1gpu"] = r["gcups"] / one[0]["gcups"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
