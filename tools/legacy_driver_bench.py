#!/usr/bin/env python3
"""The link-time drop-in, timed: the benchmark's own driver (benchmarks/phmm/PairHMMUnitTest_orig.cpp, compiled
UNMODIFIED by oracle/build_ref.sh) linked against libphmm_cuda.so, next to its twin linked against the reference's
own objects, on C1- and C2-shaped text inputs.  GCUPS = cells x loops / the driver's own "Kernel runtime" line
(:327; two decimals, hence the loops).  PHMM_FAST=0/1 picks the bit-exact or the fast kernels behind the legacy
symbols; -t is the driver's OpenMP thread count (one computelikelihoodsboth call per batch and thread, :271-299).

    python tools/legacy_driver_bench.py [--workloads c1,c2] [--threads 1,16] [--no-ref]
"""
import argparse, json, os, re, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from genomicsbench_palisade_b200 import synth

REF = os.path.join(ROOT, "oracle", "_ref")


def run(exe, path, threads, loops, env=None):
    t = time.perf_counter()
    pr = subprocess.run([exe, "-f", path, "-t", str(threads), "-l", str(loops)], stdout=subprocess.PIPE,
                        stderr=subprocess.PIPE, text=True, env=dict(os.environ, **(env or {})), timeout=1800)
    wall = time.perf_counter() - t
    m = re.search(r"Kernel runtime: ([0-9.]+) sec", pr.stdout[-400:])
    if pr.returncode != 0 or not m:
        raise RuntimeError("%s failed: %s" % (exe, (pr.stderr or pr.stdout)[-500:]))
    return float(m.group(1)), wall


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="c1,c2")
    ap.add_argument("--threads", default="1,16")
    ap.add_argument("--no-ref", action="store_true")
    a = ap.parse_args()
    out = {"what": "benchmarks/phmm/PairHMMUnitTest_orig.cpp, unmodified, -l loops over the input; GCUPS from its own "
                   "'Kernel runtime' line", "host_cores": os.cpu_count(), "runs": []}
    tmp = tempfile.mkdtemp()
    for w in a.workloads.split(","):
        fb = synth.c1_small() if w == "c1" else synth.c2_large()
        path = os.path.join(tmp, w + ".in")
        open(path, "wb").write(fb.to_text())
        # a prefix for the CPU twin: ~6e9 cells per loop
        nb = int(np.searchsorted(np.cumsum(fb.batch_cells), 6e9)) + 1
        sub = fb.select_batches(np.arange(min(nb, fb.n_batches)))
        spath = os.path.join(tmp, w + "_ref.in")
        open(spath, "wb").write(sub.to_text())
        for th in [int(x) for x in a.threads.split(",")]:
            for fast in (0, 1):
                loops = 10 if w == "c2" else 40
                for coalesce in (("0", "1") if th > 1 else ("0",)):
                    rt, wall = run(os.path.join(REF, "phmm_orig_driver_cuda"), path, th, loops,
                                   {"PHMM_FAST": str(fast), "PHMM_LEGACY_COALESCE": coalesce})
                    out["runs"].append({"workload": w, "lib": "libphmm_cuda", "threads": th, "PHMM_FAST": fast,
                                        "coalesce": int(coalesce), "loops": loops, "kernel_runtime_s": rt,
                                        "gcups": fb.n_cells * loops / max(rt, 1e-3) / 1e9, "pairs": fb.n_pairs,
                                        "wall_s": wall})
                    print(json.dumps(out["runs"][-1]), file=sys.stderr, flush=True)
            if not a.no_ref:
                rt, wall = run(os.path.join(REF, "phmm_orig_driver_ref"), spath, th, 1)
                out["runs"].append({"workload": w, "lib": "reference objects (libphmm_ref, FTZ off, fork printf on)",
                                    "threads": th, "loops": 1, "kernel_runtime_s": rt,
                                    "gcups": sub.n_cells / max(rt, 1e-3) / 1e9, "pairs": sub.n_pairs, "wall_s": wall})
                print(json.dumps(out["runs"][-1]), file=sys.stderr, flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
