#!/usr/bin/env python3
"""Benchmark of the phmm hot path: pair-HMM forward GCUPS on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--kernel exact|fast] [--workload c2]
                    [--configs c1,c3,c4,c5 | none]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU path on the host cores

One "step" = one pass of the hot path (FP32 forward kernels + FP64 rescue kernels) over the
whole synthetic workload of this rank.  `value` is device-timed with the inputs resident in HBM;
`e2e` goes through the host-buffer C-ABI call (plan + H2D + kernels + D2H inside the timed
region).  Cells of FP64-rescued pairs are counted once.

The headline line is BASELINE.json configs[1] (C2, weak scaling: every GPU gets a set of the shape
of the N = 1 set).  Its `configs` block carries the other four configs of BASELINE.json, a few
steps each: C1, C3, C4 (every rank runs the same set: "replicas") and C5, ONE 10 M-pair set at
every N, sharded by the scheduler (strong scaling).  Every block has a `parity` entry: the
reference's own CPU code run on a prefix of the very pairs that were timed, compared with what
the timed kernels produced.
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
TOL_FP32, TOL_FP64 = 1e-4, 1e-9      # BASELINE.json north_star: relative, per pair

# name -> (generator, scaling at N > 1, description)
WORKLOADS = {
    "c1": (lambda: synth.c1_small(1.0), "replicas",
           "C1 GenomicsBench phmm 'small'-shaped synthetic testcases (550 batches, <= 110 reads x <= 37 haplotypes, "
           "reads 100-150 bp, haplotypes 100-500 bp)"),
    "c2": (lambda: synth.c2_large(2_000_000), "weak",
           "C2 GenomicsBench phmm 'large'-shaped synthetic batch (~2 M read x haplotype pairs per GPU, reads 100-250 bp, "
           "haplotypes 150-473 bp, <= 50 000 pairs per batch)"),
    "c3": (lambda: synth.c3_regions(4096), "replicas",
           "C3 HaplotypeCaller-style active regions (4096 regions, 64-256 reads of 250 bp x 4-32 candidate haplotypes of "
           "300-500 bp)"),
    "c4": (lambda: synth.c4_stress(200_000), "replicas",
           "C4 low-quality / long-haplotype stress set (200 000 pairs, reads 200-250 bp, haplotypes 1-5 kb, Q 6-15)"),
    "c5": (lambda: synth.c5_skewed(10_000_000), "strong",
           "C5 length-skewed sweep: ONE 10 M-pair set (reads 100/150/250 bp, haplotypes log-uniform 100-5000 bp) "
           "sharded over the GPUs by the length-balanced scheduler"),
    "skew": (lambda: synth.skewed_rescue(4_000_000), "strong",
             "scheduler test: C2-shaped, 4 M pairs, every third batch 60 % unrelated reads (rescued in FP64): equal "
             "cells are not equal time"),
    "tiny": (lambda: synth.generic(synth.SEED0 + 50, 12, (5, 30), (2, 10), (100, 250), (150, 473), foreign_frac=0.05,
                                   name="tiny"), "weak", "tiny"),
}


def workload_name(w):
    return WORKLOADS[w][2]


# ---------------------------------------------------------------------------------------------
# workload and sharding
# ---------------------------------------------------------------------------------------------
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


def make_workload(name: str, world: int):
    """The global workload of a run on `world` GPUs.
    weak: `world` sets of exactly the shape of the N = 1 set (same batch sizes, lengths and cell count, different
    sequences and qualities: synth.same_shape), so per-GPU work does not drift with N;  strong / replicas: the
    N = 1 set itself."""
    gen, scaling, _ = WORKLOADS[name]
    base = gen()
    if world == 1 or scaling != "weak":
        return base
    return concat([base] + [synth.same_shape(base, int(base.meta.get("seed", synth.SEED0)) + 100 * k)
                            for k in range(1, world)])


def shard_of(fb, rank: int, world: int, fast: bool, by: str = "time"):
    """Length-balanced scheduler: whole batches (keeps read / haplotype reuse) assigned longest-processing-time
    first (phmm_partition_lpt) by predicted device time (phmm_batch_costs: register-tile geometry of each read's
    kernel class + predicted FP64-rescue work), or by exact cell count (`by` = "cells", what round 1 did)."""
    if world == 1:
        return fb, np.arange(fb.n_batches)
    from genomicsbench_palisade_b200 import phmm
    cost = fb.batch_cells if by == "cells" else phmm.batch_costs(fb, fast=fast)
    owner = phmm.partition_lpt(cost, world)
    mine = np.nonzero(owner == rank)[0]
    return fb.select_batches(mine, compact=True), mine


def config_block(name, fb_all, world):
    """Identical in the GPU arm and the reference arm (the driver compares the two)."""
    scaling = WORKLOADS[name][1]
    return {"workload": workload_name(name), "pairs": int(fb_all.n_pairs * (world if scaling == "replicas" else 1)),
            "cells": int(fb_all.n_cells * (world if scaling == "replicas" else 1)), "batches": int(fb_all.n_batches),
            "seed": int(fb_all.meta.get("seed", 0)),
            "sample_of": "GPU arm: every step = the whole workload; reference arm (CPU): every step = a prefix of the "
                         "workload's batches sized to the time budget (cpu_baseline.sample); GCUPS is a rate",
            "sharding": "LPT over whole batches by predicted device time (phmm_batch_costs + phmm_partition_lpt), "
                        "no collective on the data path",
            "l2": "GPU arm: 256 MB memset between timed iterations"}


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
# CPU reference: baseline, parity, reference arm
# ---------------------------------------------------------------------------------------------
def cpu_sample(fb, target_cells):
    """First batches of the workload up to ~target_cells cells (a prefix, so its results are a prefix too)."""
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


def cpu_kind():
    import oracle
    oracle.build()
    return "reference" if oracle.have_ref() else "port"


def run_cpu(fb, kind, ftz=0):
    """-> (seconds, oracle.Result)"""
    import oracle
    t = time.perf_counter()
    if kind == "reference":
        res = oracle.run_ref(fb, nthreads=host_threads(), ftz=ftz)
    else:
        res = oracle.run_port(fb, nthreads=host_threads())
    return time.perf_counter() - t, res


def cpu_rate(fb, kind):
    probe = cpu_sample(fb, 2e8)
    dt, _ = run_cpu(probe, kind)
    return probe.n_cells / max(dt, 1e-6)


def cpu_variants(fb, seconds=4.0):
    """the reference's AVX-512 and AVX-256 instantiations, FTZ off (as the benchmark ships)
    and on (as the JNI path sets it), all host threads and one thread; a few seconds each."""
    import oracle
    out = {"cpu_model": "", "variants": []}
    try:
        out["cpu_model"] = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        pass
    rate = cpu_rate(fb, "reference")
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


def parity_block(ref, got, kernel, plan="throughput"):
    """the reference's results for the first len(ref.out) pairs of the timed workload
    against what the timed GPU path produced for them.  got = (out, rescued) of the whole workload."""
    n = len(ref.out)
    out, resc = got[0][:n], got[1][:n].astype(bool)
    r = ref.rescued.astype(bool)
    both32, both64 = ~r & ~resc, r & resc
    rel = np.abs(out - ref.out) / np.maximum(np.abs(ref.out), 1e-300)
    m32 = float(rel[both32].max()) if both32.any() else 0.0
    m64 = float(rel[both64].max()) if both64.any() else 0.0
    bad = int((r != resc).sum())
    ok = bad == 0 and m32 <= TOL_FP32 and m64 <= TOL_FP64 and bool(np.all(np.isfinite(out)))
    return {"pairs": int(n), "rescued_pairs": int(r.sum()), "rescue_symdiff": bad, "max_rel_fp32": m32,
            "max_rel_rescued": m64, "bars": {"fp32": TOL_FP32, "rescued": TOL_FP64}, "kernel": kernel, "plan": plan,
            "against": "reference AVX code (oracle/_ref), FTZ off, on the same pairs", "ok": ok}


def cpu_baseline(fb, kind, seconds=12.0):
    """-> (cpu_baseline block, oracle.Result of the sample)"""
    import oracle
    cores = host_threads()
    sample = cpu_sample(fb, cpu_rate(fb, kind) * seconds)
    dt, res = run_cpu(sample, kind)
    ftz_on = None
    if kind == "reference":
        # the same code with x86 flush-to-zero on, as GKL's JNI entry sets it (IntelPairHmm.cc:66-70): the
        # benchmark's C entry leaves it off and pays for denormals in the FP32 pass
        dt1, _ = run_cpu(sample, kind, ftz=1)
        ftz_on = sample.n_cells / dt1 / 1e9
    return {"value": sample.n_cells / dt / 1e9, "unit": UNIT, "cores": cores, "kind": kind, "value_ftz_on": ftz_on,
            "sample": "first %d batches of the workload: %d pairs, %.3e cells, %.1f s; %s, OpenMP schedule(dynamic,1) over "
                      "pairs as IntelPairHmmCSource.cpp:64-66, FTZ off as the benchmark ships" % (
                          sample.n_batches, sample.n_pairs, sample.n_cells, dt,
                          "reference AVX sources compiled in place (oracle/_ref), cpuid-dispatched" if kind == "reference"
                          else "plain-C restatement (oracle/phmm_oracle.c)"),
            "isa": ("avx512" if oracle.ref_isa() else "avx") if kind == "reference" else "scalar"}, res


def reference_arm(args, rank, world):
    if rank != 0:
        return
    fb_all = make_workload(args.workload, world)
    kind = cpu_kind()
    rate = cpu_rate(fb_all, kind)
    budget = 150.0 / max(1, args.steps + args.warmup)          # whole run within a few minutes
    sample = cpu_sample(fb_all, rate * min(budget, 15.0))
    for _ in range(args.warmup):
        run_cpu(sample, kind)
    t = [run_cpu(sample, kind)[0] for _ in range(args.steps)]
    ms = 1e3 * float(np.mean(t))
    val = sample.n_cells / (ms * 1e-3) / 1e9
    cores = host_threads()
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32+f64", "data": "synthetic",
        "config": config_block(args.workload, fb_all, world),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "each step = first %d batches (%d pairs, %.3e cells) of the workload; FTZ off as the "
                                   "benchmark ships (with FTZ on, as GKL's JNI entry sets it, the same code runs ~10x "
                                   "faster: cpu_baseline.value_ftz_on of the GPU arm's line)" % (
                                       sample.n_batches, sample.n_pairs, sample.n_cells)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
class Ranks:
    """torch.distributed plumbing: barrier and small reductions for the timing, nothing on the data path."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def flush_l2(self):
        self.flush_buf.zero_()
        self.torch.cuda.synchronize()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def gather(self, x):
        """x (float) of every rank, on every rank."""
        if self.world == 1:
            return [float(x)]
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    def max(self, x):
        return max(self.gather(x))

    def sum(self, x):
        return sum(self.gather(x))

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def measure(R, eng, fb, steps, warmup, e2e=True, clocks=False):
    """Times `steps` passes over this rank's batches `fb`: device-resident (CUDA events on the library's launch
    stream, L2 flushed before every pass) and end to end through phmm_compute_flat.  Barriers on both sides of each
    timed loop; times are reduced over the ranks by the caller."""
    w = eng.workload(fb)
    for _ in range(warmup):
        R.flush_l2()
        st = w.run()
    R.barrier()
    sampler = ClockSampler(R.local) if (clocks and R.rank == 0) else None
    t0 = time.perf_counter()
    ms_total = ms32 = ms64 = 0.0
    for _ in range(steps):
        R.flush_l2()                           # L2 flush between timed iterations
        st = w.run()                           # CUDA events on the library's launch stream
        ms_total += st["ms_total"]
        ms32 += st["ms_fp32"]
        ms64 += st["ms_fp64"]
    R.barrier()
    t1 = time.perf_counter()
    m = {"ms": ms_total / steps, "ms_fp32": ms32 / steps, "ms_fp64": ms64 / steps, "st": st,
         "clocks": sampler.stop(t0, t1) if sampler else None}
    out, raw32, raw64, resc = w.results(raw=True)
    w.close()
    m["out"], m["rescued"] = out, resc
    m["resc_cells"] = int(fb.pair_cells()[resc.astype(bool)].sum())
    if e2e:
        o = np.empty(fb.n_pairs, np.float64)          # the caller's results buffer, reused as the driver does
        for _ in range(2):
            eng.compute(fb, out=o)
        o.fill(0.0)
        R.barrier()
        te = time.perf_counter()
        for _ in range(steps):
            o, s2 = eng.compute(fb, stats=True, out=o)
        R.barrier()
        m["e2e_ms"] = (time.perf_counter() - te) * 1e3 / steps
        m["h2d"], m["d2h"] = s2["h2d_bytes"], s2["d2h_bytes"]
        m["e2e_ok"] = bool(np.array_equal(o, out))
        # the same call from buffers the caller page-locked once (phmm_host_register): what a driver that owns its
        # pools for the whole run gets; `e2e` itself stays the pageable drop-in case
        pools = (fb.rs, fb.q, fb.i, fb.d, fb.c, fb.hap, o)
        eng.pin(*pools)
        eng.compute(fb, out=o)
        o.fill(0.0)
        R.barrier()
        te = time.perf_counter()
        for _ in range(steps):
            eng.compute(fb, out=o)
        R.barrier()
        m["e2e_pinned_ms"] = (time.perf_counter() - te) * 1e3 / steps
        m["e2e_ok"] = m["e2e_ok"] and bool(np.array_equal(o, out))
        eng.unpin(*pools)
    return m


def check_parity(fb, m, kernel, seconds, kind):
    """Reference CPU code on a prefix of this rank's batches (~`seconds` of host time) against the timed results."""
    if kind != "reference":
        return {"ok": None, "note": "oracle/_ref not built: reference unavailable on this box"}
    sample = cpu_sample(fb, cpu_rate(fb, kind) * seconds)
    _, ref = run_cpu(sample, kind)
    return parity_block(ref, (m["out"], m["rescued"]), kernel)


def sub_config(R, eng, name, fast, kind, peak, steps=3, warmup=3):
    """One of BASELINE.json's other configs, a few steps: same measurements as the headline, in short."""
    scaling = WORKLOADS[name][1]
    fb_all = make_workload(name, R.world)
    fb, _ = shard_of(fb_all, R.rank, R.world, fast) if scaling != "replicas" else (fb_all, None)
    blk = {"workload": workload_name(name), "scaling": scaling if R.world > 1 else "n/a (1 GPU)", "steps": steps,
           "warmup": warmup}
    n1_ms = None
    if scaling == "strong" and R.world > 1:
        # the whole set on rank 0 alone, in this run: the denominator of the strong-scaling speed-up
        if R.rank == 0:
            n1_ms = measure(_Solo(R), eng, fb_all, 2, 1, e2e=False)["ms"]
        R.barrier()
    m = measure(R, eng, fb, steps, warmup)
    per_rank = R.gather(m["ms"])
    dev_ms, e2e_ms, e2e_pin_ms = max(per_rank), R.max(m["e2e_ms"]), R.max(m["e2e_pinned_ms"])
    cells, pairs = R.sum(fb.n_cells), R.sum(fb.n_pairs)
    ok_all = R.sum(0.0 if m["e2e_ok"] else 1.0) == 0
    if R.rank == 0:
        # FP64-first reads never enter the FP32 kernels: their cells are not FP32 work
        ach = FLOP_PER_CELL * (fb.n_cells - int(m["st"].get("n_spec_cells", 0))) / (m["ms_fp32"] * 1e-3) / 1e12
        blk.update({"pairs": int(pairs), "cells": int(cells), "value": cells / (dev_ms * 1e-3) / 1e9, "unit": UNIT,
                    "ms_per_step": dev_ms, "ms_fp32": m["ms_fp32"], "ms_fp64": m["ms_fp64"],
                    "e2e": {"value": cells / (e2e_ms * 1e-3) / 1e9, "ms_per_step": e2e_ms,
                            "h2d_bytes_per_step": int(R.world * m["h2d"]) if scaling == "replicas" else None,
                            "matches_resident_run": ok_all, "pinned_value": cells / (e2e_pin_ms * 1e-3) / 1e9},
                    "frac": ach / peak, "tile_fill": tile_fill(fb, fast), "rescue_frac": float(m["rescued"].mean()),
                    "rescue_cells_frac": m["resc_cells"] / max(1, fb.n_cells),
                    "fp64_first": {"pairs": int(m["st"].get("n_spec_pairs", 0)), "cells": int(m["st"].get("n_spec_cells", 0)),
                                   "passed_after_all": int(m["st"].get("n_spec_direct", 0))},
                    "imbalance": max(per_rank) / (sum(per_rank) / len(per_rank)),
                    "gpu_launches": int(m["st"]["n_launches"]) * steps,
                    "parity": check_parity(fb, m, "fast" if fast else "exact", 3.0, kind)})
        if n1_ms:
            blk["n1_ms_per_step"] = n1_ms
            blk["speedup_vs_n1"] = n1_ms / dev_ms
    return blk


class _Solo:
    """The Ranks interface for a measurement one rank makes alone (no barrier with the others)."""

    def __init__(self, R):
        self.R, self.rank, self.local, self.world = R, R.rank, R.local, 1

    def flush_l2(self):
        self.R.flush_l2()

    def barrier(self):
        self.R.torch.cuda.synchronize()


def hbm_peak():
    """Measured copy bandwidth of this pool's B200s (driver-written MEASURED_PEAKS.json), else the profiling guide's
    fallback."""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6500.0


def tile_fill(fb, fast):
    """cells / tile cells of rank 0's plan: the share of the FP32 kernels' row steps that are real read rows (a read
    of R bases occupies kh x lanes rows of its class's register tile; separator columns count as tile cells too)."""
    from genomicsbench_palisade_b200 import phmm
    # planned as the library plans it (FP64-first reads, where its per-call rule has them, are in no FP32 tile)
    p = phmm.plan_summary(fb, cover=False, fast=fast, fp64_first="auto" if fast else False)
    return {"real_over_tile": (p["cells"] - p["fp64_first_cells"]) / max(1, p["tile_cells"]),
            "classes": [{"kh": k[0], "lanes": k[1], "real_over_tile": round(c / max(1, t), 4)}
                        for k, (c, t) in zip(p["classes"], p["class_cells"])]}


def sw_block(R, no_cpu):
    """tools/sw_bench.py's measurement (device-resident, end to end through sw_align_flat, reference on the host
    cores with a bit-for-bit comparison of score, offset and CIGAR) as a block of this line."""
    import types
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import sw_bench
    from genomicsbench_palisade_b200 import sw
    a = types.SimpleNamespace(steps=5, warmup=3, pairs=100_000, strategy=sw.SOFTCLIP, no_cpu=no_cpu, no_e2e=False)
    return sw_bench.measure(a, device=R.local, flush=R.flush_buf, cpu_seconds=4.0)


def gatk_block(R, no_cpu):
    """tools/gatk_bench.py's measurement of SURVEY.md 8f row 1 (the GATK-shaped batch API, host holders in, host
    likelihoods out) as a block of this line: the entry's default bit-exact kernels at 1 / 16 / 256 / 1024 regions
    per call with the reference (FTZ on, as its JNI entry sets it) one region per call beside them, and the opt-in
    fast kernels for the single-region and the all-regions call."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gatk_bench
    b = gatk_bench.measure(1024, "exact", no_cpu, device=R.local)
    f = gatk_bench.measure(1024, "fast", True, device=R.local, sizes=(1, 0))
    b["fast_kernels"] = f["groups"]
    return b


def ncu_traffic():
    """DRAM bytes per step of the FP32 launches (dram__bytes_read.sum + dram__bytes_write.sum) from the newest
    committed ncu --set full capture of this benchmark (C2, one GPU); None when there is none."""
    best = None
    pdir = os.path.join(ROOT, "profiles")
    for f in sorted(os.listdir(pdir)) if os.path.isdir(pdir) else []:
        if f.endswith("_dram_traffic.json") and "phmm" in json.load(open(os.path.join(pdir, f))):
            best = f
    if not best:
        return None, None
    try:
        d = json.load(open(os.path.join(pdir, best)))["phmm"]
        return int(d["traffic_bytes_per_step"]), "profiles/" + best + " (" + d.get("build", "build not recorded") + ")"
    except Exception:
        return None, None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--configs", default=None,
                    help="other BASELINE.json configs to add to the line, a few steps each (default: c1,c3,c4,c5 when the "
                         "headline workload is c2; 'none' to skip)")
    ap.add_argument("--kernel", default=os.environ.get("PHMM_BENCH_KERNEL", "fast"), choices=["exact", "fast"])
    ap.add_argument("--shard-by", default="time", choices=["time", "cells"], help="cost the scheduler balances")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and parity legs")
    ap.add_argument("--cpu-variants", action="store_true",
                    help="add the reference's AVX-512 / AVX-256 x FTZ off / on x all threads / 1 thread table to cpu_baseline")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.configs is None:
        args.configs = "c1,c3,c4,c5,sw,gatk" if args.workload == "c2" else "none"
    subs = [c for c in args.configs.split(",") if c and c not in ("none", "sw", "gatk")]
    with_sw = "sw" in args.configs.split(",")
    with_gatk = "gatk" in args.configs.split(",")

    if args.impl == "reference":
        reference_arm(args, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return

    from genomicsbench_palisade_b200 import phmm
    R = Ranks()
    rank, world = R.rank, R.world
    fast = args.kernel == "fast"
    scaling = WORKLOADS[args.workload][1]
    fb_all = make_workload(args.workload, world)
    fb, _ = shard_of(fb_all, rank, world, fast, args.shard_by) if scaling != "replicas" else (fb_all, None)
    eng = phmm.PairHMM(device=R.local, fast=fast, fast64=fast)
    peak_ops, _ = eng.microbench(0)            # FFMA lane-ops/s measured on this device, now
    peak = 2 * peak_ops / 1e12
    kind = cpu_kind() if (rank == 0 and not args.no_cpu) else None

    m = measure(R, eng, fb, args.steps, args.warmup, e2e=not args.no_e2e, clocks=True)
    per_rank = R.gather(m["ms"])
    dev_ms = max(per_rank)
    cells_all, pairs_all = R.sum(fb.n_cells), R.sum(fb.n_pairs)
    line = None
    if not args.no_e2e:
        e2e_ms, e2e_pin_ms = R.max(m["e2e_ms"]), R.max(m["e2e_pinned_ms"])
        h2d_all, d2h_all = R.sum(m["h2d"]), R.sum(m["d2h"])
        e2e_ok = R.sum(0.0 if m["e2e_ok"] else 1.0) == 0
    if rank == 0:
        st = m["st"]
        ach = FLOP_PER_CELL * (fb.n_cells - int(st.get("n_spec_cells", 0))) / (m["ms_fp32"] * 1e-3) / 1e12
        clocks = m["clocks"]
        mhz = (clocks or {}).get("sm_mhz") or 1965.0
        traffic, traffic_src = ncu_traffic() if (fast and args.workload == "c2") else (None, None)
        line = {
            "metric": METRIC, "value": cells_all / (dev_ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms, "higher_is_better": True,
            "scaling": scaling if scaling != "replicas" else "weak", "vs_baseline": None,
            "dtype": "f32+f64", "data": "synthetic",
            "config": config_block(args.workload, fb_all, world),
            "kernel": args.kernel, "clocks": clocks,
            "imbalance": {"max_over_mean_rank_ms": max(per_rank) / (sum(per_rank) / len(per_rank)),
                          "rank_ms": [round(x, 3) for x in per_rank], "balanced_by": args.shard_by},
            "gpu_launches": int(st["n_launches"]) * args.steps,
            "roofline": {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                         "peak_nominal": 148 * 128 * 2 * mhz * 1e6 / 1e12, "frac_of_nominal": ach / (148 * 128 * 2 * mhz * 1e6 / 1e12),
                         "traffic": traffic, "traffic_source": traffic_src,
                         "hbm": {"algorithmic_bytes_per_step": int(m.get("h2d", 0) + 8 * fb.n_pairs),
                                 "achieved_gbs": (m.get("h2d", 0) + 8 * fb.n_pairs) / (m["ms"] * 1e-3) / 1e9,
                                 "peak_gbs": hbm_peak(),
                                 "note": "inputs as uploaded + 8 B per result; `traffic` = dram__bytes_read.sum + "
                                         "dram__bytes_write.sum over the FP32 launches of one step from the ncu --set full "
                                         "capture named in traffic_source (C2, one GPU, a constant of that capture, not of "
                                         "this run); both are ~1e-3 of what HBM moves in the step time"},
                         "kernel": ("phmm_forward_hp_kernel<KH, LANES>" if fast else "phmm_forward_kernel<float, KH, LANES>")
                                   + " (all FP32 class instantiations of the step together, rank 0)",
                         "how": "12 flop/cell x cells / CUDA-event time of the FP32 launches; peak = 2 x FFMA lane-ops/s "
                                "measured by phmm_microbench on this device in this run (MEASURED_PEAKS.json has no FP32 "
                                "entry); peak_nominal = 148 SMs x 128 lanes x 2 flop x the SM clock sampled in this run",
                         # FP32 lane-operations issued per cell / FFMA lane-ops peak: 6.25 packed instructions per
                         # cell pair in the fast kernels (scaled recurrence, 16-row tiles), 12 unfused in the exact ones
                         "fma_pipe_slot_frac": ach / peak * 2 * ((6.25 if fast else 12.0) / 12.0),
                         "ms_fp32": m["ms_fp32"], "ms_fp64": m["ms_fp64"], "tile_fill": tile_fill(fb, fast)},
            "rescue": {"pairs": int(m["rescued"].sum()), "frac": float(m["rescued"].mean()) if m["rescued"].size else 0.0,
                       "cells": m["resc_cells"], "guard_band_pairs": int(st["n_guard"]),
                       "fp64_first": {"pairs": int(st.get("n_spec_pairs", 0)), "cells": int(st.get("n_spec_cells", 0)),
                                      "passed_after_all": int(st.get("n_spec_direct", 0)),
                                      "note": "reads the planner predicts to fail the FP32 pass against their whole batch (k-mer "
                                              "anchored score, phmm_cost.cpp) skip it: FP64 first, and the rescaled FP64 sum tells "
                                              "what the FP32 pass would have decided; their cells are left out of roofline.achieved"}},
        }
        if not args.no_e2e:
            line["e2e"] = {"value": cells_all / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d_all),
                           "d2h_bytes_per_step": int(d2h_all), "ms_per_step": e2e_ms, "matches_resident_run": e2e_ok,
                           "pinned": {"value": cells_all / (e2e_pin_ms * 1e-3) / 1e9, "ms_per_step": e2e_pin_ms,
                                      "note": "same call, caller's pools and results array page-locked once with "
                                              "phmm_host_register"},
                           "includes": "host planning + H2D from pageable host buffers + kernels + D2H of the 8 B/pair "
                                       "results into the caller's buffer; chunks of batches pipelined (plan + upload of one "
                                       "under the kernels of the others)"}
        if kind:
            if world == 1:
                line["cpu_baseline"], ref = cpu_baseline(fb, kind)
                line["parity"] = (parity_block(ref, (m["out"], m["rescued"]), args.kernel) if kind == "reference"
                                  else {"ok": None, "note": "oracle/_ref not built"})
                if args.cpu_variants and kind == "reference":
                    line["cpu_baseline"].update(cpu_variants(fb))
            else:
                line["parity"] = check_parity(fb, m, args.kernel, 4.0, kind)
    del m
    if subs:
        blocks = {}
        for name in subs:
            blocks[name if WORKLOADS[name][1] != "strong" else name + "_strong"] = sub_config(R, eng, name, fast, kind, peak)
        if rank == 0:
            line["configs"] = blocks
    eng.close()
    if with_sw and rank == 0:
        # SURVEY.md 8f row 3 (GKL Smith-Waterman with backtrack, libsw_cuda.so): one GPU, not sharded
        line["sw"] = sw_block(R, args.no_cpu)
    if with_gatk and rank == 0:
        # SURVEY.md 8f row 1 (GATK native-bindings-shaped batch API): one GPU, host buffers in and out
        line["gatk"] = gatk_block(R, args.no_cpu)
    R.barrier()
    R.close()
    if rank == 0:
        print(json.dumps(line))
        bad = [k for k, b in [("headline", line)] + list(line.get("configs", {}).items())
               if (b.get("parity") or {}).get("ok") is False]
        if bad:
            sys.stderr.write("bench.py: PARITY BROKEN in %s\n" % ", ".join(bad))
            sys.exit(3)


if __name__ == "__main__":
    main()
