"""phmm_forward_region_noalloc (include/phmm_cuda.h): the allocation-free entry point, whose kernels' scratch lives in a
static __device__ array.  No counterpart in the reference: it is checked against the library's existing entry points
(which the other GPU tests hold to the oracle).  Collected last on purpose."""
import numpy as np
import pytest

from genomicsbench_palisade_b200 import phmm, synth
from genomicsbench_palisade_b200.batches import FlatBatches


def _region(seed, n_reads, n_haps, rlen, hlen):
    fb = synth.generic(synth.SEED0 + seed, 1, (n_reads, n_reads), (n_haps, n_haps), rlen, hlen, foreign_frac=0.25,
                       name="noalloc")
    assert fb.n_batches == 1
    return fb


def _packed(fb: FlatBatches):
    """Pools with the reads (haplotypes) back to back and the n + 1 int32 offset tables the entry point takes."""
    def pack(pools, off, ln):
        idx = np.concatenate([np.arange(o, o + n) for o, n in zip(off, ln)]) if len(ln) else np.zeros(0, np.int64)
        noff = np.zeros(len(ln) + 1, np.int32)
        noff[1:] = np.cumsum(ln)
        return [np.ascontiguousarray(p[idx]) for p in pools], noff
    (rs, q, i, d, c), roff = pack((fb.rs, fb.q, fb.i, fb.d, fb.c), fb.read_off, fb.read_len)
    (hap,), hoff = pack((fb.hap,), fb.hap_off, fb.hap_len)
    return rs, q, i, d, c, roff, hap, hoff


def _existing(eng, fb):
    w = eng.workload(fb)
    w.run()
    out, raw32, raw64, resc = w.results(raw=True)
    w.close()
    return out, resc


def _noalloc(torch, fb, max_read_len=None):
    rs, q, i, d, c, roff, hap, hoff = _packed(fb)
    dev = [torch.from_numpy(a).cuda() for a in (rs, q, i, d, c, roff, hap, hoff)]
    n_reads, n_haps = len(fb.read_len), len(fb.hap_len)
    out = torch.full((n_reads * n_haps,), 12345.0, dtype=torch.float64, device="cuda")
    resc = torch.full((n_reads * n_haps,), 7, dtype=torch.uint8, device="cuda")
    mrl = int(fb.read_len.max()) if max_read_len is None else max_read_len
    p = [t.data_ptr() for t in dev]
    phmm.forward_region_noalloc(0, None, p[0], p[1], p[2], p[3], p[4], p[5], n_reads, mrl, p[6], p[7], n_haps,
                                int(hoff[-1]), out.data_ptr(), resc.data_ptr())
    torch.cuda.synchronize()
    return out.cpu().numpy(), resc.cpu().numpy()


@pytest.mark.gpu
def test_noalloc_entry_point_matches_the_existing_one_on_calls_in_a_row():
    import torch
    assert torch.cuda.is_available()
    regions = [_region(301, 96, 24, (40, 255), (100, 473)),      # several runs of haplotypes, rescued pairs
               _region(302, 7, 3, (20, 90), (30, 120)),           # smaller: finds the first call's contents in the scratch
               _region(303, 255, 128, (100, 151), (150, 400)),    # 32 640 pairs: next to the maximum
               _region(301, 96, 24, (40, 255), (100, 473))]       # the first one again
    eng = phmm.PairHMM(device=0, fast=False)
    n_rescued = 0
    for fb in regions:
        want, want_resc = _existing(eng, fb)
        got, got_resc = _noalloc(torch, fb)
        assert np.array_equal(got_resc, want_resc), "FP64-rescue set differs from phmm_workload_run's"
        r = want_resc.astype(bool)
        # the bars of tests/test_gpu_parity.py (BASELINE.json: 1e-4 FP32 path, 1e-9 rescued pairs) ...
        assert np.all(np.abs(got[~r] - want[~r]) <= 1e-4 * np.abs(want[~r]))
        assert np.all(np.abs(got[r] - want[r]) <= 1e-9 * np.abs(want[r]))
        # ... and, both being the bit-exact kernels, the same bits
        assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
        n_rescued += int(r.sum())
    eng.close()
    assert n_rescued > 0, "the regions were meant to exercise the FP64 rescue list"


@pytest.mark.gpu
def test_noalloc_tables_that_contradict_the_stated_sizes_give_nan():
    import torch
    fb = _region(304, 12, 5, (50, 120), (80, 200))
    got, _ = _noalloc(torch, fb, max_read_len=int(fb.read_len.max()) - 1)
    assert np.all(np.isnan(got))
    got, resc = _noalloc(torch, fb)          # and the next call is not affected
    eng = phmm.PairHMM(device=0, fast=False)
    want, want_resc = _existing(eng, fb)
    eng.close()
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64)) and np.array_equal(resc, want_resc)


def test_noalloc_rejects_sizes_beyond_the_maximum():
    """Argument checks come before any CUDA call: runs without a GPU.  (Pointers are never dereferenced.)"""
    ok = dict(n_reads=8, max_read_len=100, n_haps=4, hap_bytes=1000)
    for key, value, needle in (("n_reads", phmm.NOALLOC_MAX_READS + 1, "PHMM_NOALLOC_MAX_READS"),
                               ("n_haps", phmm.NOALLOC_MAX_HAPS + 1, "PHMM_NOALLOC_MAX_HAPS"),
                               ("max_read_len", phmm.NOALLOC_MAX_READ_LEN + 1, "PHMM_NOALLOC_MAX_READ_LEN"),
                               ("hap_bytes", phmm.NOALLOC_MAX_HAP_BYTES + 1, "PHMM_NOALLOC_MAX_HAP_BYTES")):
        a = dict(ok, **{key: value})
        with pytest.raises(phmm.PhmmError) as e:
            phmm.forward_region_noalloc(0, None, 256, 256, 256, 256, 256, 256, a["n_reads"], a["max_read_len"], 256, 256,
                                        a["n_haps"], a["hap_bytes"], 256)
        assert e.value.code == 3 and needle in str(e.value)
    with pytest.raises(phmm.PhmmError) as e:
        phmm.forward_region_noalloc(0, None, 256, 256, 256, 256, 256, 256, 1024, 100, 256, 256, 128, 1000, 256)
    assert e.value.code == 3 and "PHMM_NOALLOC_MAX_PAIRS" in str(e.value)
    with pytest.raises(phmm.PhmmError) as e:
        phmm.forward_region_noalloc(0, None, 256, 256, 256, 256, 256, None, 8, 100, 256, 256, 4, 1000, 256)
    assert e.value.code == 1


@pytest.mark.gpu
def test_noalloc_entry_point_under_cuda_graph_capture():
    """The reason the entry point exists: it can be captured (no allocation, no synchronisation, nothing read back) and
    the graph replayed on new contents of the same buffers."""
    import torch
    fbs = [_region(305, 40, 9, (60, 200), (100, 300)), _region(306, 40, 9, (60, 200), (100, 300))]
    eng = phmm.PairHMM(device=0, fast=False)
    want = [_existing(eng, fb) for fb in fbs]
    eng.close()
    packed = [_packed(fb) for fb in fbs]
    # fixed-size device buffers (the graph bakes the pointers and the host-side sizes in): sized for the larger region
    sizes = [max(len(p[k]) for p in packed) for k in range(8)]
    dev = [torch.zeros(n + 1, dtype=torch.from_numpy(packed[0][k]).dtype, device="cuda") for k, n in enumerate(sizes)]
    n_reads, n_haps = 40, 9
    out = torch.zeros(n_reads * n_haps, dtype=torch.float64, device="cuda")
    resc = torch.zeros(n_reads * n_haps, dtype=torch.uint8, device="cuda")
    p = [t.data_ptr() for t in dev]
    mrl, hap_cap = 200, sizes[6]

    def load(k):
        for t, a in zip(dev, packed[k]):
            t[:len(a)].copy_(torch.from_numpy(a))

    def call(stream):
        phmm.forward_region_noalloc(0, stream, p[0], p[1], p[2], p[3], p[4], p[5], n_reads, mrl, p[6], p[7], n_haps,
                                    hap_cap, out.data_ptr(), resc.data_ptr())

    load(0)
    call(None)                      # the first call on a device sets it up: outside capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        call(torch.cuda.current_stream().cuda_stream)
    for k in (1, 0, 1):
        load(k)
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy().view(np.uint64), want[k][0].view(np.uint64))
        assert np.array_equal(resc.cpu().numpy(), want[k][1])
