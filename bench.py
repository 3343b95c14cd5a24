#!/usr/bin/env python3
"""Benchmark of the phmm hot path: pair-HMM forward GCUPS on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--kernel exact|fast] [--workload c2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU path on the host cores

One "step" = one pass of the hot path (FP32 forward kernels + FP64 rescue kernels) over the
whole synthetic workload of this rank.  `value` is device-timed with the inputs resident in HBM;
`e2e` goes through the host-buffer C-ABI call (plan + H2D + kernels + D2H inside the timed
region).  Cells of FP64-rescued pairs are counted once.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from genomicsbench_palisade_b200 import synth  # noqa: E402

METRIC = "pair-HMM forward GCUPS (FP32 + FP64 rescue)"
UNIT = "GCUPS"
FLOP_PER_CELL = 12          # 8 mul + 4 add, avx-pairhmm-template.h:188,194,197


# ---------------------------------------------------------------------------------------------
# workload and sharding
# ---------------------------------------------------------------------------------------------
def make_workload(name: str, world: int):
    """Global workload for `world` GPUs: `world` independently seeded parts (weak scaling)."""
    parts = []
    for k in range(world):
        if name == "c2":        # BASELINE configs[1]: 'large'-shaped, ~2 M pairs per GPU
            parts.append(synth.c2_large(2_000_000, seed=synth.SEED0 + 2 + 100 * k))
        elif name == "c1":
            parts.append(synth.c1_small(1.0, seed=synth.SEED0 + 1 + 100 * k))
        elif name == "c3":
            parts.append(synth.c3_regions(1024, seed=synth.SEED0 + 3 + 100 * k))
        elif name == "c4":
            parts.append(synth.c4_stress(200_000, seed=synth.SEED0 + 4 + 100 * k))
        elif name == "c5":
            parts.append(synth.c5_skewed(10_000_000 // 8, seed=synth.SEED0 + 5 + 100 * k))
        elif name == "tiny":
            parts.append(synth.generic(synth.SEED0 + 50 + k, 12, (5, 30), (2, 10), (100, 250), (150, 473),
                                       foreign_frac=0.05, name="tiny"))
        else:
            raise SystemExit("unknown workload " + name)
    return concat(parts) if world > 1 else parts[0]


def concat(parts):
    from genomicsbench_palisade_b200.batches import FlatBatches
    ro, ho, acc_r, acc_h = [], [], 0, 0
    for p in parts:
        ro.append(p.read_off + acc_r)
        ho.append(p.hap_off + acc_h)
        acc_r += len(p.rs)
        acc_h += len(p.hap)
    cat = lambda k: np.concatenate([getattr(p, k) for p in parts])
    return FlatBatches(cat("batch_num_reads"), cat("batch_num_haps"), cat("read_len"), np.concatenate(ro),
                       cat("rs"), cat("q"), cat("i"), cat("d"), cat("c"), cat("hap_len"), np.concatenate(ho),
                       cat("hap"), dict(parts[0].meta))


def my_batches(fb, rank: int, world: int):
    """Length-balanced scheduler: LPT over batches by exact cell count (phmm_partition_lpt)."""
    if world == 1:
        return np.arange(fb.n_batches)
    from genomicsbench_palisade_b200 import phmm
    owner = phmm.partition_lpt(fb.batch_cells, world)
    return np.nonzero(owner == rank)[0]


def make_shard(name: str, rank: int, world: int):
    fb = make_workload(name, world)
    return fb, (fb if world == 1 else fb.select_batches(my_batches(fb, rank, world), compact=True))


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        mhz, mx, reasons = [], None, set()
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7 or not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            try:
                mhz.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(mhz)) if mhz else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(mhz)}


# ---------------------------------------------------------------------------------------------
# CPU reference arm
# ---------------------------------------------------------------------------------------------
def cpu_sample(fb, target_cells):
    """First batches of the workload up to ~target_cells cells."""
    cells = np.cumsum(fb.batch_cells)
    nb = int(np.searchsorted(cells, target_cells)) + 1
    return fb.select_batches(np.arange(min(nb, fb.n_batches)))


def host_threads():
    # every core this process may run on; passed explicitly because torchrun exports OMP_NUM_THREADS=1 to its
    # ranks when there is more than one
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def run_cpu(fb, kind):
    import oracle
    t = time.perf_counter()
    if kind == "reference":
        oracle.run_ref(fb, nthreads=host_threads())
    else:
        oracle.run_port(fb, nthreads=host_threads())
    return time.perf_counter() - t


def cpu_variants(fb, seconds=4.0):
    """the reference's AVX-512 and AVX-256 instantiations, FTZ off (as the benchmark ships)
    and on (as the JNI path sets it), all host threads and one thread; a few seconds each."""
    import oracle
    out = {"cpu_model": "", "variants": []}
    try:
        out["cpu_model"] = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        pass
    probe = cpu_sample(fb, 2e8)
    rate = probe.n_cells / max(run_cpu(probe, "reference"), 1e-6)
    cores = os.cpu_count() or 1
    for isa, isa_name in ((2, "avx512"), (1, "avx256")):
        for ftz in (0, 1):
            for nt in (cores, 1):
                sample = cpu_sample(fb, rate * seconds * nt / cores)
                t = time.perf_counter()
                r = oracle.run_ref(sample, nthreads=nt, isa=isa, ftz=ftz)
                dt = time.perf_counter() - t
                out["variants"].append({"isa": isa_name, "ftz": bool(ftz), "threads": nt, "gcups": sample.n_cells / dt / 1e9,
                                        "cells": sample.n_cells, "rescued_frac": float(r.rescued.mean())})
    return out


def cpu_baseline(fb, seconds=12.0):
    import oracle
    oracle.build()
    kind = "reference" if oracle.have_ref() else "port"
    cores = host_threads()
    probe = cpu_sample(fb, 2e8)
    dt = run_cpu(probe, kind)
    rate = probe.n_cells / max(dt, 1e-6)
    sample = cpu_sample(fb, rate * seconds)
    dt = run_cpu(sample, kind)
    ftz_on = None
    if kind == "reference":
        # the same code with x86 flush-to-zero on, as GKL's JNI entry sets it (IntelPairHmm.cc:66-70): the
        # benchmark's C entry leaves it off and pays for denormals in the FP32 pass
        t = time.perf_counter()
        oracle.run_ref(sample, nthreads=host_threads(), ftz=1)
        ftz_on = sample.n_cells / (time.perf_counter() - t) / 1e9
    return {"value": sample.n_cells / dt / 1e9, "unit": UNIT, "cores": cores, "kind": kind, "value_ftz_on": ftz_on,
            "sample": "first %d batches of the workload: %d pairs, %.3e cells, %.1f s; %s, OpenMP schedule(dynamic,1) over "
                      "pairs as IntelPairHmmCSource.cpp:64-66, FTZ off as the benchmark ships" % (
                          sample.n_batches, sample.n_pairs, sample.n_cells, dt,
                          "reference AVX sources compiled in place (oracle/_ref), cpuid-dispatched" if kind == "reference"
                          else "plain-C restatement (oracle/phmm_oracle.c)"),
            "isa": ("avx512" if oracle.ref_isa() else "avx") if kind == "reference" else "scalar"}


def reference_arm(args, rank):
    if rank != 0:
        return
    fb = make_workload(args.workload, 1)
    import oracle
    oracle.build()
    kind = "reference" if oracle.have_ref() else "port"
    probe = cpu_sample(fb, 2e8)
    rate = probe.n_cells / max(run_cpu(probe, kind), 1e-6)
    budget = 150.0 / max(1, args.steps + args.warmup)          # whole run within a few minutes
    sample = cpu_sample(fb, rate * min(budget, 15.0))
    for _ in range(args.warmup):
        run_cpu(sample, kind)
    t = [run_cpu(sample, kind) for _ in range(args.steps)]
    ms = 1e3 * float(np.mean(t))
    val = sample.n_cells / (ms * 1e-3) / 1e9
    cores = host_threads()
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32+f64", "data": "synthetic",
        "config": {"workload": workload_name(args.workload), "sample_pairs": sample.n_pairs, "sample_cells": sample.n_cells},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "each step = first %d batches (%d pairs, %.3e cells) of the workload" % (
                             sample.n_batches, sample.n_pairs, sample.n_cells)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def workload_name(w):
    return {"c2": "C2 GenomicsBench phmm 'large'-shaped synthetic batch (~2 M read x haplotype pairs per GPU, reads 100-250 bp, "
                  "haplotypes 150-473 bp, <= 50 000 pairs per batch)",
            "c1": "C1 'small'-shaped (550 batches)", "c3": "C3 active regions", "c4": "C4 low-quality long-haplotype stress",
            "c5": "C5 length-skewed sweep", "tiny": "tiny"}[w]


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--kernel", default=os.environ.get("PHMM_BENCH_KERNEL", "fast"), choices=["exact", "fast"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-variants", action="store_true",
                    help="add the reference's AVX-512 / AVX-256 x FTZ off / on x all threads / 1 thread table to cpu_baseline")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    from genomicsbench_palisade_b200 import phmm
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    fb_all, fb = make_shard(args.workload, rank, world)
    fast = args.kernel == "fast"
    eng = phmm.PairHMM(device=local, fast=fast, fast64=fast)
    peak_ops, _ = eng.microbench(0)            # FFMA lane-ops/s measured on this device, now
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    # ---- device-resident timing (value) ---------------------------------------------------------
    w = eng.workload(fb)
    for _ in range(args.warmup):
        flush.zero_()
        torch.cuda.synchronize()
        st = w.run()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    t0 = time.perf_counter()
    ms_total = ms32 = ms64 = 0.0
    for _ in range(args.steps):
        flush.zero_()                          # L2 flush between timed iterations
        torch.cuda.synchronize()
        st = w.run()                           # CUDA events on the library's launch stream
        ms_total += st["ms_total"]
        ms32 += st["ms_fp32"]
        ms64 += st["ms_fp64"]
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1) if sampler else None
    dev_ms = max_over_ranks(ms_total / args.steps)
    cells_all = sum_over_ranks(fb.n_cells)
    pairs_all = sum_over_ranks(fb.n_pairs)
    out, raw32, raw64, resc = w.results(raw=True)
    resc_cells = int(fb.pair_cells()[resc.astype(bool)].sum()) if fb.n_pairs < 50_000_000 else 0
    w.close()

    # ---- end to end through the host-buffer C ABI (e2e) --------------------------------------------
    o = np.empty(fb.n_pairs, np.float64)          # the caller's results buffer, reused as the driver does
    for _ in range(0 if args.no_e2e else 2):
        eng.compute(fb, out=o)
    o.fill(0.0)
    barrier()
    te = time.perf_counter()
    h2d = d2h = 0
    for _ in range(0 if args.no_e2e else args.steps):
        o, s2 = eng.compute(fb, stats=True, out=o)
        h2d, d2h = s2["h2d_bytes"], s2["d2h_bytes"]
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - te) * 1e3 / args.steps)
    e2e_ok = bool(np.array_equal(o, out)) if not args.no_e2e else None
    h2d_all, d2h_all = sum_over_ranks(h2d), sum_over_ranks(d2h)

    if rank == 0:
        value = cells_all / (dev_ms * 1e-3) / 1e9
        ach = FLOP_PER_CELL * fb.n_cells / (ms32 / args.steps * 1e-3) / 1e12
        peak = 2 * peak_ops / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32+f64", "data": "synthetic",
            "config": {"workload": workload_name(args.workload), "pairs": int(pairs_all), "cells": int(cells_all),
                       "batches": fb_all.n_batches, "kernel": args.kernel,
                       "sharding": "LPT over batches by cells (phmm_partition_lpt), no collective on the data path",
                       "l2": "256 MB memset between timed iterations", "seed": synth.SEED0 + 2},
            "clocks": clocks,
            "e2e": {"value": cells_all / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d_all),
                    "d2h_bytes_per_step": int(d2h_all), "ms_per_step": e2e_ms, "matches_resident_run": e2e_ok,
                    "includes": "host planning + H2D from pageable host buffers + kernels + D2H of the 8 B/pair results into the caller's buffer; chunks of batches pipelined (plan + upload of one under the kernels of the other)"},
            "gpu_launches": int(st["n_launches"]) * args.steps,
            "roofline": {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                         "traffic": ncu_traffic("phmm") if (fast and args.workload == "c2") else None,
                         "hbm": {"algorithmic_bytes_per_step": int(h2d + 8 * fb.n_pairs),
                                 "note": "inputs as uploaded + 8 B per result; `traffic` = dram__bytes_read.sum + "
                                         "dram__bytes_write.sum over the same launches of one step from the ncu --set full "
                                         "capture in profiles/ (C2, one GPU); both are ~1e-3 of what HBM moves in the step time"},
                         "kernel": ("phmm_forward_hp_kernel<KH, LANES>" if fast else "phmm_forward_kernel<float, KH, LANES>")
                                   + " (all FP32 class instantiations of the step together, rank 0)",
                         "how": "12 flop/cell x cells / CUDA-event time of the FP32 launches; peak = 2 x FFMA lane-ops/s "
                                "measured by phmm_microbench on this device in this run (MEASURED_PEAKS.json has no FP32 entry)",
                         # FP32 lane-operations issued per cell / FFMA lane-ops peak: 6.25 packed instructions per
                         # cell pair in the fast kernels (scaled recurrence, 16-row tiles), 12 unfused in the exact ones
                         "fma_pipe_slot_frac": ach / peak * 2 * ((6.25 if fast else 12.0) / 12.0),
                         "ms_fp32": ms32 / args.steps, "ms_fp64": ms64 / args.steps},
            "rescue": {"pairs": int(resc.sum()), "frac": float(resc.mean()) if resc.size else 0.0, "cells": resc_cells,
                       "guard_band_pairs": int(st["n_guard"])},
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(fb)
            if args.cpu_variants and line["cpu_baseline"]["kind"] == "reference":
                line["cpu_baseline"].update(cpu_variants(fb))
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def ncu_traffic(tag):
    """DRAM bytes per step of the kernel set `achieved` covers, from the committed ncu --set full capture."""
    try:
        return int(json.load(open(os.path.join(ROOT, "profiles", "r1_v9_dram_traffic.json")))[tag]["traffic_bytes_per_step"])
    except Exception:
        return None


if __name__ == "__main__":
    main()
