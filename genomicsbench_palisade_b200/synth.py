"""Seeded synthetic phmm workloads in the shapes BASELINE.json names.

The reference's datasets (`input-datasets/phmm/small/5m.in`, `large/large.in`) are a separate
download that is not available offline, so every test and benchmark here runs on synthetic
batches with the dataset shape facts the reference records
(`benchmarks/phmm/PairHMMUnitTest.cpp:1-2,9-10,30`; pair cap `PairHMMUnitTest_orig.cpp:55`).

Generator: numpy `PCG64(seed)`; the haplotypes of a batch are variants (window jitter + a few
SNPs) of one uniform-ACGT base sequence with 0.1 % N, as the candidate haplotypes of one active
region are; a read is a substring of one haplotype of its batch at a uniform offset, each base then replaced by a
uniform ACGT base with probability `s`; q = Phred ~ U[qlo, qhi]; insertion/deletion GOP = 45
with 5 % of positions U[20, 45]; GCP = 10.  A fraction of "foreign" reads gets a much higher
substitution rate so that the FP64-rescue set is not empty.
"""
from __future__ import annotations

import numpy as np

from .batches import FlatBatches

ACGT = np.frombuffer(b"ACGT", np.uint8)
SEED0 = 20261104


def _rand_bases(rng, n, n_frac=0.001):
    b = ACGT[rng.integers(0, 4, n)]
    if n_frac > 0 and n:
        b = np.where(rng.random(n) < n_frac, np.uint8(ord("N")), b)
    return b.astype(np.uint8)


def _related_haps(rng, bnh, base_len, n_snps=3, jitter=8):
    """Haplotypes of one batch are variants of one base sequence (as candidate haplotypes of an
    active region are): a window of the batch's base sequence shifted/resized by up to
    `jitter` bases, plus up to `n_snps` substitutions.  Returns (hap_len, hap_pool)."""
    bnh = np.asarray(bnh, np.int64)
    base_len = np.asarray(base_len, np.int64)
    nh = int(bnh.sum())
    ext = base_len + 2 * jitter
    base_off = np.concatenate(([0], np.cumsum(ext)))[:-1]
    base_pool = _rand_bases(rng, int(ext.sum()))
    bid = np.repeat(np.arange(len(bnh)), bnh)
    start = rng.integers(0, jitter + 1, nh)
    hap_len = np.maximum(1, base_len[bid] + rng.integers(-jitter, jitter + 1, nh))
    hap_len = np.minimum(hap_len, ext[bid] - start)
    total = int(hap_len.sum())
    hoff = np.concatenate(([0], np.cumsum(hap_len)))[:-1]
    hid = np.repeat(np.arange(nh), hap_len)
    pos = np.arange(total) - hoff[hid]
    pool = base_pool[base_off[bid][hid] + start[hid] + pos].copy()
    for _ in range(n_snps):
        on = rng.random(nh) < 0.6
        p = hoff + (rng.random(nh) * hap_len).astype(np.int64)
        pool[p[on]] = ACGT[rng.integers(0, 4, int(on.sum()))]
    return hap_len, pool


def _gen_reads(rng, bnr, bnh, hap_len, hap_off, hap_pool, read_len,
               q_rng, subst, foreign_frac, foreign_q, foreign_subst, exact_offset0=False):
    bnr = np.asarray(bnr, np.int64)
    bnh = np.asarray(bnh, np.int64)
    nr = int(bnr.sum())
    read_len = np.asarray(read_len, np.int64)
    bhs = np.concatenate(([0], np.cumsum(bnh)))
    bid = np.repeat(np.arange(len(bnr)), bnr)
    h = bhs[bid] + np.minimum((rng.random(nr) * bnh[bid]).astype(np.int64), bnh[bid] - 1)
    H = hap_len[h].astype(np.int64)
    span = np.maximum(H - read_len + 1, 1)
    off = np.minimum((rng.random(nr) * span).astype(np.int64), span - 1)
    foreign = rng.random(nr) < foreign_frac
    s_read = np.where(foreign, foreign_subst, subst)
    qlo = np.where(foreign, foreign_q[0], q_rng[0])
    qhi = np.where(foreign, foreign_q[1], q_rng[1])

    total = int(read_len.sum())
    roff = np.concatenate(([0], np.cumsum(read_len)))[:-1]
    rid = np.repeat(np.arange(nr), read_len)
    pos = np.arange(total) - roff[rid]
    hp = off[rid] + pos
    valid = hp < H[rid]
    src = hap_off[h][rid] + np.where(valid, hp, 0)
    rnd = ACGT[rng.integers(0, 4, total)]
    bases = np.where(valid, hap_pool[src], rnd)
    sub = rng.random(total) < s_read[rid]
    bases = np.where(sub, ACGT[rng.integers(0, 4, total)], bases).astype(np.uint8)
    q = (qlo[rid] + (rng.random(total) * (qhi[rid] - qlo[rid] + 1)).astype(np.int64)).astype(np.uint8)
    gop_var = rng.random(total) < 0.05
    ins = np.where(gop_var, rng.integers(20, 46, total), 45).astype(np.uint8)
    gop_var = rng.random(total) < 0.05
    dele = np.where(gop_var, rng.integers(20, 46, total), 45).astype(np.uint8)
    gcp = np.full(total, 10, np.uint8)
    return read_len.astype(np.int32), roff, bases, q, ins, dele, gcp


def _assemble(rng, bnr, bnh, hap_len, hap_pool, read_len, meta, **kw):
    hap_len = np.asarray(hap_len, np.int64)
    hap_off = np.concatenate(([0], np.cumsum(hap_len)))[:-1]
    rl, roff, rs, q, i, d, c = _gen_reads(rng, bnr, bnh, hap_len, hap_off, hap_pool, read_len, **kw)
    meta = dict(meta, gen=dict(kw))
    return FlatBatches(bnr, bnh, rl, roff, rs, q, i, d, c, hap_len, hap_off, hap_pool, meta)


def same_shape(fb, content_seed, jitter=8, n_snps=3):
    """A set with exactly the shape of `fb` (batch sizes, every read and haplotype length, hence the
    same cell count) and new content: per batch a fresh base sequence, every haplotype a window of it
    plus a few substitutions, reads drawn from the new haplotypes with the quality / substitution
    parameters `fb` was generated with.  The parts of a weak-scaling run are made this way, so that
    every GPU gets the work the single GPU gets at N = 1."""
    rng = np.random.Generator(np.random.PCG64(content_seed))
    bnh = fb.batch_num_haps.astype(np.int64)
    hap_len = fb.hap_len.astype(np.int64)
    nh = len(hap_len)
    bid = np.repeat(np.arange(len(bnh)), bnh)
    ext = np.zeros(len(bnh), np.int64)
    np.maximum.at(ext, bid, hap_len)
    ext += 2 * jitter
    base_off = np.concatenate(([0], np.cumsum(ext)))[:-1]
    base_pool = _rand_bases(rng, int(ext.sum()))
    start = np.minimum(rng.integers(0, 2 * jitter + 1, nh), ext[bid] - hap_len)
    hoff = np.concatenate(([0], np.cumsum(hap_len)))[:-1]
    hid = np.repeat(np.arange(nh), hap_len)
    pos = np.arange(int(hap_len.sum())) - hoff[hid]
    pool = base_pool[base_off[bid][hid] + start[hid] + pos].copy()
    for _ in range(n_snps):
        on = rng.random(nh) < 0.6
        p = hoff + (rng.random(nh) * hap_len).astype(np.int64)
        pool[p[on]] = ACGT[rng.integers(0, 4, int(on.sum()))]
    meta = dict(fb.meta, content_seed=content_seed)
    gen = meta.pop("gen")
    return _assemble(rng, fb.batch_num_reads, fb.batch_num_haps, hap_len, pool, fb.read_len.astype(np.int64), meta, **gen)


def generic(seed, n_batches, reads_rng, haps_rng, rlen_rng, hlen_rng, q_rng=(10, 40), subst=0.01,
            foreign_frac=0.0, foreign_q=(20, 40), foreign_subst=0.25, max_pairs=50000, name="generic"):
    """Batches with uniform shape parameters; every batch is capped at `max_pairs` pairs by
    dropping reads (the reference asserts batch_size <= MAX_BATCH_SIZE, `_orig.cpp:280`)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    bnh = rng.integers(haps_rng[0], haps_rng[1] + 1, n_batches)
    bnr = rng.integers(reads_rng[0], reads_rng[1] + 1, n_batches)
    bnr = np.maximum(1, np.minimum(bnr, max_pairs // bnh))
    base_len = rng.integers(hlen_rng[0], hlen_rng[1] + 1, n_batches)
    hap_len, hap_pool = _related_haps(rng, bnh, base_len)
    read_len = rng.integers(rlen_rng[0], rlen_rng[1] + 1, int(bnr.sum()))
    return _assemble(rng, bnr, bnh, hap_len, hap_pool, read_len, {"workload": name, "seed": seed},
                     q_rng=q_rng, subst=subst, foreign_frac=foreign_frac, foreign_q=foreign_q,
                     foreign_subst=foreign_subst)


def c1_small(scale=1.0, seed=SEED0 + 1):
    """C1: `5m.in`-shaped: 550 batches, <=110 reads x <=37 haps, R 100-150, H 100-500."""
    return generic(seed, max(1, int(550 * scale)), (1, 110), (1, 37), (100, 150), (100, 500),
                   q_rng=(10, 40), subst=0.01, foreign_frac=0.05, name="C1-small")


def c2_large(min_pairs=2_000_000, seed=SEED0 + 2):
    """C2: `large.in`-shaped: reads 200-1193 x haps 16-128 per batch, capped at 50 000 pairs,
    R 100-250, H 150-473, batches added until >= min_pairs pairs."""
    rng = np.random.Generator(np.random.PCG64(seed))
    bnr, bnh, pairs = [], [], 0
    while pairs < min_pairs:
        nh = int(rng.integers(16, 129))
        nr = max(1, min(int(rng.integers(200, 1194)), 50000 // nh))
        bnr.append(nr)
        bnh.append(nh)
        pairs += nr * nh
    bnr, bnh = np.asarray(bnr), np.asarray(bnh)
    hap_len, hap_pool = _related_haps(rng, bnh, rng.integers(158, 466, len(bnh)))
    read_len = rng.integers(100, 251, int(bnr.sum()))
    return _assemble(rng, bnr, bnh, hap_len, hap_pool, read_len,
                     {"workload": "C2-large", "seed": seed, "min_pairs": min_pairs},
                     q_rng=(10, 40), subst=0.01, foreign_frac=0.02, foreign_q=(20, 40),
                     foreign_subst=0.25)


def c3_regions(n_regions=4096, seed=SEED0 + 3):
    """C3: HaplotypeCaller-style active regions: {4,8,16,32} candidate haplotypes that are
    SNP/indel variants of one 300-500 bp base haplotype, 64-256 reads of exactly 250 bp."""
    rng = np.random.Generator(np.random.PCG64(seed))
    bnh = rng.choice([4, 8, 16, 32], n_regions)
    bnr = rng.integers(64, 257, n_regions)
    haps, hap_len = [], []
    for g in range(n_regions):
        base = _rand_bases(rng, int(rng.integers(300, 501)))
        for k in range(int(bnh[g])):
            v = base.copy()
            if k:
                for _ in range(int(rng.integers(1, 4))):
                    p = int(rng.integers(0, len(v)))
                    kind = rng.integers(0, 3)
                    if kind == 0:
                        v[p] = ACGT[rng.integers(0, 4)]
                    elif kind == 1 and len(v) > 260:
                        v = np.delete(v, slice(p, p + int(rng.integers(1, 6))))
                    else:
                        v = np.insert(v, p, _rand_bases(rng, int(rng.integers(1, 6)), 0))
            haps.append(v)
            hap_len.append(len(v))
    hap_pool = np.concatenate(haps)
    read_len = np.full(int(bnr.sum()), 250)
    return _assemble(rng, bnr, bnh, np.asarray(hap_len), hap_pool, read_len,
                     {"workload": "C3-regions", "seed": seed, "n_regions": n_regions},
                     q_rng=(20, 40), subst=0.005, foreign_frac=0.0, foreign_q=(20, 40),
                     foreign_subst=0.25)


def c4_stress(n_pairs=200_000, seed=SEED0 + 4):
    """C4: low-quality / long-haplotype stress set maximising the FP64 rescue fraction:
    R 200-250, H 1000-5000, Q 6-15, 20 % substitutions (about half the pairs are rescued)."""
    nb = max(1, n_pairs // 400)
    return generic(seed, nb, (40, 40), (10, 10), (200, 250), (1000, 5000), q_rng=(6, 15),
                   subst=0.20, foreign_frac=0.0, name="C4-stress")


def c5_skewed(n_pairs=10_000_000, seed=SEED0 + 5):
    """C5: length-skewed sweep: R in {100,150,250} w.p. {.5,.3,.2}, H log-uniform in
    [100, 5000], Q 10-40, 2 % substitutions + 2 % foreign reads; batches of 100 reads x 20 haps."""
    rng = np.random.Generator(np.random.PCG64(seed))
    nb = max(1, n_pairs // 2000)
    bnr = np.full(nb, 100)
    bnh = np.full(nb, 20)
    base_len = np.exp(rng.uniform(np.log(108), np.log(4992), nb)).astype(np.int64)
    hap_len, hap_pool = _related_haps(rng, bnh, base_len)
    read_len = rng.choice([100, 150, 250], int(bnr.sum()), p=[0.5, 0.3, 0.2])
    return _assemble(rng, bnr, bnh, hap_len, hap_pool, read_len,
                     {"workload": "C5-skewed", "seed": seed, "n_pairs": n_pairs},
                     q_rng=(10, 40), subst=0.02, foreign_frac=0.02, foreign_q=(20, 40),
                     foreign_subst=0.25)


def skewed_rescue(min_pairs=2_000_000, seed=SEED0 + 6, dirty_every=3, dirty_frac=0.6):
    """C2-shaped batches of which every `dirty_every`-th holds mostly reads that fit none of its haplotypes (a
    contaminated or mis-assembled region): their pairs are computed twice, the second time in FP64, so equal
    cells are far from equal device time.  The scheduler test case for balancing by predicted time."""
    base = c2_large(min_pairs, seed=seed)
    rng = np.random.Generator(np.random.PCG64(seed + 1))
    brs = base.batch_read_start
    rs = base.rs.copy()
    for b in range(0, base.n_batches, dirty_every):
        for r in range(brs[b], brs[b + 1]):
            if rng.random() < dirty_frac:
                o, n = int(base.read_off[r]), int(base.read_len[r])
                sub = rng.random(n) < 0.25
                rs[o:o + n] = np.where(sub, ACGT[rng.integers(0, 4, n)], rs[o:o + n])
    return FlatBatches(base.batch_num_reads, base.batch_num_haps, base.read_len, base.read_off, rs, base.q, base.i,
                       base.d, base.c, base.hap_len, base.hap_off, base.hap, dict(base.meta, workload="skewed-rescue"))


def edge_cases(seed=SEED0 + 9, max_read_len=None):
    """Small adversarial batches: R > H, length-1 reads and haplotypes, N bases on both sides,
    bytes outside ACGTN (the reference maps them to 'A', `pairhmm_common.h:27,34-38`), reads
    whose length sits on every register-tile boundary, extreme quality values."""
    rng = np.random.Generator(np.random.PCG64(seed))
    batches = []

    def read(n, qlo=6, qhi=45):
        return (_rand_bases(rng, n, 0.05), rng.integers(qlo, qhi + 1, n).astype(np.uint8),
                rng.integers(0, 60, n).astype(np.uint8), rng.integers(0, 60, n).astype(np.uint8),
                rng.integers(0, 40, n).astype(np.uint8))

    # every read length 1..70 and around multiples of 32 against short/long haplotypes
    lens = list(range(1, 71)) + [95, 96, 97, 127, 128, 129, 159, 160, 161, 191, 192, 193,
                                 223, 224, 225, 254, 255, 256, 257, 300, 511, 512, 513, 700]
    if max_read_len is not None:
        lens = [n for n in lens if n <= max_read_len]
    reads = [read(n) for n in lens]
    haps = [_rand_bases(rng, n, 0.05) for n in (1, 2, 3, 31, 32, 33, 64, 100, 255, 256, 257, 600)]
    batches.append((reads, haps))
    # odd bytes: lower case, IUPAC codes, NUL-free binary junk
    odd = np.frombuffer(b"acgtnRYKMSWBDHVacgtXxZ*-.", np.uint8)
    r = list(read(25))
    r[0] = odd
    batches.append(([tuple(r), read(40)], [odd, np.frombuffer(b"ACGTNNNNACGT", np.uint8),
                                           np.frombuffer(b"N", np.uint8)]))
    # extreme qualities: 0, 127 and bytes with bit 7 set (kernel masks &127,
    # avx-pairhmm-template.h:110-112,125)
    n = 50
    rq = (_rand_bases(rng, n, 0), np.r_[np.zeros(10), np.full(10, 127), rng.integers(128, 256, 30)].astype(np.uint8),
          np.r_[np.zeros(25), np.full(25, 127)].astype(np.uint8),
          rng.integers(0, 256, n).astype(np.uint8), rng.integers(0, 256, n).astype(np.uint8))
    batches.append(([rq], [_rand_bases(rng, 80, 0), _rand_bases(rng, 20, 0)]))
    # identical read and haplotype (likelihood near its maximum) and a total mismatch
    h = _rand_bases(rng, 120, 0)
    same = (h.copy(), np.full(120, 40, np.uint8), np.full(120, 45, np.uint8),
            np.full(120, 45, np.uint8), np.full(120, 10, np.uint8))
    comp = np.frombuffer(bytes(h).translate(bytes.maketrans(b"ACGT", b"CATG")), np.uint8)
    diff = (comp, np.full(120, 40, np.uint8), np.full(120, 45, np.uint8),
            np.full(120, 45, np.uint8), np.full(120, 10, np.uint8))
    batches.append(([same, diff], [h]))
    from .batches import from_lists
    fb = from_lists(batches)
    fb.meta = {"workload": "edge-cases", "seed": seed}
    return fb
