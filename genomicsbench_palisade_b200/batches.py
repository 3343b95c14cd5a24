"""Flat structure-of-arrays form of phmm input batches (host side, numpy).

This is the in-memory image of what the reference driver parses from its text input
(`benchmarks/phmm/PairHMMUnitTest_orig.cpp:97-169`, `read_batch`): a sequence of batches,
each `num_reads` reads (bases + four quality strings) and `num_haps` haplotypes, evaluated
all-pairs in read-major order (`:284-297`, result index `r * num_haps + h`).  Instead of one
heap string per field (the reference `strndup`s every string) everything lives in six byte
pools plus offset tables, which is exactly the layout `include/phmm_cuda.h`'s
`phmm_flat_batches` hands to the CUDA library.

Quality pools hold the values the reference hands to the kernel, i.e. AFTER the driver's
`normalize` (`:88-92,120-123`: `max(min, byte - 33)`, min 6 for `q`, 0 for `i/d/c`).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class FlatBatches:
    batch_num_reads: np.ndarray   # int32[nb]
    batch_num_haps: np.ndarray    # int32[nb]
    read_len: np.ndarray          # int32[nr]   reads of batch 0, then batch 1, ...
    read_off: np.ndarray          # int64[nr]   offset into the five read pools
    rs: np.ndarray                # uint8 pools, parallel
    q: np.ndarray
    i: np.ndarray
    d: np.ndarray
    c: np.ndarray
    hap_len: np.ndarray           # int32[nh]
    hap_off: np.ndarray           # int64[nh]
    hap: np.ndarray               # uint8 pool
    meta: dict = field(default_factory=dict)

    def __post_init__(self):
        self.batch_num_reads = np.ascontiguousarray(self.batch_num_reads, dtype=np.int32)
        self.batch_num_haps = np.ascontiguousarray(self.batch_num_haps, dtype=np.int32)
        self.read_len = np.ascontiguousarray(self.read_len, dtype=np.int32)
        self.read_off = np.ascontiguousarray(self.read_off, dtype=np.int64)
        self.hap_len = np.ascontiguousarray(self.hap_len, dtype=np.int32)
        self.hap_off = np.ascontiguousarray(self.hap_off, dtype=np.int64)
        for name in ("rs", "q", "i", "d", "c", "hap"):
            setattr(self, name, np.ascontiguousarray(getattr(self, name), dtype=np.uint8))
        assert int(self.batch_num_reads.sum()) == len(self.read_len)
        assert int(self.batch_num_haps.sum()) == len(self.hap_len)

    # ---- derived tables -------------------------------------------------------------------
    @property
    def n_batches(self) -> int:
        return len(self.batch_num_reads)

    @property
    def batch_read_start(self) -> np.ndarray:
        return np.concatenate(([0], np.cumsum(self.batch_num_reads, dtype=np.int64)))

    @property
    def batch_hap_start(self) -> np.ndarray:
        return np.concatenate(([0], np.cumsum(self.batch_num_haps, dtype=np.int64)))

    @property
    def batch_pairs(self) -> np.ndarray:
        return self.batch_num_reads.astype(np.int64) * self.batch_num_haps.astype(np.int64)

    @property
    def batch_out_start(self) -> np.ndarray:
        return np.concatenate(([0], np.cumsum(self.batch_pairs)))

    @property
    def n_pairs(self) -> int:
        return int(self.batch_pairs.sum())

    @property
    def batch_cells(self) -> np.ndarray:
        """Exact cell count per batch = sum(read_len) * sum(hap_len) (`_orig.cpp:166`)."""
        rs, hs = self.batch_read_start, self.batch_hap_start
        rl = np.concatenate(([0], np.cumsum(self.read_len, dtype=np.int64)))
        hl = np.concatenate(([0], np.cumsum(self.hap_len, dtype=np.int64)))
        return (rl[rs[1:]] - rl[rs[:-1]]) * (hl[hs[1:]] - hl[hs[:-1]])

    @property
    def n_cells(self) -> int:
        return int(self.batch_cells.sum())

    def input_bytes(self) -> int:
        return int(5 * self.read_len.astype(np.int64).sum() + self.hap_len.astype(np.int64).sum())

    # ---- the reference's pair list ----------------------------------------------------------
    def pairs(self):
        """Expand to the testcase list the reference driver builds (`_orig.cpp:284-297`).

        Returns (rslen, haplen, rs_off, hap_off), one entry per pair, batch-major then
        read-major then haplotype -- the order of the reference's `results` arrays."""
        rsl, hpl, rso, hpo = [], [], [], []
        brs, bhs = self.batch_read_start, self.batch_hap_start
        for b in range(self.n_batches):
            nr, nh = int(self.batch_num_reads[b]), int(self.batch_num_haps[b])
            r = np.arange(brs[b], brs[b] + nr)
            h = np.arange(bhs[b], bhs[b] + nh)
            rsl.append(np.repeat(self.read_len[r], nh))
            rso.append(np.repeat(self.read_off[r], nh))
            hpl.append(np.tile(self.hap_len[h], nr))
            hpo.append(np.tile(self.hap_off[h], nr))
        cat = lambda xs, dt: (np.concatenate(xs).astype(dt) if xs else np.zeros(0, dt))
        return (cat(rsl, np.int32), cat(hpl, np.int32), cat(rso, np.int64), cat(hpo, np.int64))

    def pair_cells(self) -> np.ndarray:
        rsl, hpl, _, _ = self.pairs()
        return rsl.astype(np.int64) * hpl.astype(np.int64)

    # ---- slicing (multi-GPU sharding keeps whole batches) -------------------------------------
    def select_batches(self, idx, compact=False) -> "FlatBatches":
        idx = np.asarray(idx, dtype=np.int64)
        brs, bhs = self.batch_read_start, self.batch_hap_start
        r_idx = (np.concatenate([np.arange(brs[b], brs[b + 1]) for b in idx])
                 if len(idx) else np.zeros(0, np.int64))
        h_idx = (np.concatenate([np.arange(bhs[b], bhs[b + 1]) for b in idx])
                 if len(idx) else np.zeros(0, np.int64))
        if not compact:
            return FlatBatches(self.batch_num_reads[idx], self.batch_num_haps[idx],
                               self.read_len[r_idx], self.read_off[r_idx],
                               self.rs, self.q, self.i, self.d, self.c,
                               self.hap_len[h_idx], self.hap_off[h_idx], self.hap, dict(self.meta))
        # own pools holding only the selected reads / haplotypes (what a rank of a sharded run keeps)
        def gather(lens, offs, pools):
            lens = lens.astype(np.int64)
            new_off = np.concatenate(([0], np.cumsum(lens)))[:-1] if len(lens) else np.zeros(0, np.int64)
            src = (np.repeat(offs - new_off, lens) + np.arange(int(lens.sum()))) if len(lens) else np.zeros(0, np.int64)
            return new_off, [np.ascontiguousarray(p[src]) for p in pools]
        r_off, (rs, q, i, d, c) = gather(self.read_len[r_idx], self.read_off[r_idx], (self.rs, self.q, self.i, self.d, self.c))
        h_off, (hap,) = gather(self.hap_len[h_idx], self.hap_off[h_idx], (self.hap,))
        return FlatBatches(self.batch_num_reads[idx], self.batch_num_haps[idx], self.read_len[r_idx], r_off,
                           rs, q, i, d, c, self.hap_len[h_idx], h_off, hap, dict(self.meta))

    def select_reads(self, keep):
        """Drop reads where `keep` is False.  Returns (batches, pair_index) with pair_index[k] the
        position of the new pair k in this object's pair order."""
        keep = np.asarray(keep, bool)
        brs, bos = self.batch_read_start, self.batch_out_start
        bnr, idx = [], []
        for b in range(self.n_batches):
            nh = int(self.batch_num_haps[b])
            k = np.nonzero(keep[brs[b]:brs[b + 1]])[0]
            bnr.append(len(k))
            idx.append((bos[b] + k[:, None] * nh + np.arange(nh)[None, :]).ravel())
        fb = FlatBatches(bnr, self.batch_num_haps, self.read_len[keep], self.read_off[keep],
                         self.rs, self.q, self.i, self.d, self.c, self.hap_len, self.hap_off, self.hap,
                         dict(self.meta))
        return fb, (np.concatenate(idx).astype(np.int64) if idx else np.zeros(0, np.int64))

    # ---- the reference's text format ----------------------------------------------------------
    def to_text(self) -> bytes:
        """Serialise in the reference's input format (`_orig.cpp:97-169`): qualities are
        written Phred+33 so that the driver's `normalize` gives back the pool values (a `q`
        below the floor of 6 cannot be represented and is rejected)."""
        out = []
        brs, bhs = self.batch_read_start, self.batch_hap_start
        for b in range(self.n_batches):
            out.append(b"%d %d\n" % (self.batch_num_reads[b], self.batch_num_haps[b]))
            for r in range(brs[b], brs[b + 1]):
                o, n = int(self.read_off[r]), int(self.read_len[r])
                if n and int(self.q[o:o + n].min()) < 6:
                    raise ValueError("q below the driver's floor of 6 cannot round-trip")
                flds = [self.rs[o:o + n].tobytes()]
                for pool in (self.q, self.i, self.d, self.c):
                    flds.append((pool[o:o + n] + 33).astype(np.uint8).tobytes())
                out.append(b" ".join(flds) + b"\n")
            for h in range(bhs[b], bhs[b + 1]):
                o, n = int(self.hap_off[h]), int(self.hap_len[h])
                out.append(self.hap[o:o + n].tobytes() + b"\n")
        return b"".join(out)

    @staticmethod
    def from_text(data: bytes) -> "FlatBatches":
        """Python mirror of `read_testfile`/`read_batch` (`_orig.cpp:97-213`), used to check the
        C++ parser.  Tokens are whitespace separated, as `operator>>` reads them."""
        tok = data.split()
        p = 0
        bnr, bnh, rl, hl = [], [], [], []
        pools = {k: [] for k in ("rs", "q", "i", "d", "c", "hap")}
        while p < len(tok):
            nr, nh = int(tok[p]), int(tok[p + 1])
            p += 2
            bnr.append(nr)
            bnh.append(nh)
            for _ in range(nr):
                bases, q, i, d, c = tok[p:p + 5]
                p += 5
                n = len(bases)
                if not (len(q) == len(i) == len(d) == len(c) == n):
                    raise ValueError("read fields differ in length")
                rl.append(n)
                pools["rs"].append(np.frombuffer(bases, np.uint8))
                for key, s, floor in (("q", q, 6), ("i", i, 0), ("d", d, 0), ("c", c, 0)):
                    v = np.frombuffer(s, np.uint8).astype(np.int32) - 33
                    pools[key].append(np.maximum(v, floor).astype(np.uint8))
            for _ in range(nh):
                hl.append(len(tok[p]))
                pools["hap"].append(np.frombuffer(tok[p], np.uint8))
                p += 1
        cat = lambda xs: (np.concatenate(xs) if xs else np.zeros(0, np.uint8))
        rl = np.asarray(rl, np.int32)
        hl = np.asarray(hl, np.int32)
        roff = np.concatenate(([0], np.cumsum(rl, dtype=np.int64)))[:-1]
        hoff = np.concatenate(([0], np.cumsum(hl, dtype=np.int64)))[:-1]
        return FlatBatches(bnr, bnh, rl, roff, cat(pools["rs"]), cat(pools["q"]), cat(pools["i"]),
                           cat(pools["d"]), cat(pools["c"]), hl, hoff, cat(pools["hap"]))


def from_lists(batches) -> FlatBatches:
    """Build from [(reads, haps)], reads = [(bases, q, i, d, c)] as bytes / uint8 arrays
    (quality values already normalised), haps = [bases]."""
    bnr, bnh, rl, hl = [], [], [], []
    pools = {k: [] for k in ("rs", "q", "i", "d", "c", "hap")}
    as_u8 = lambda x: (np.frombuffer(x, np.uint8) if isinstance(x, (bytes, bytearray))
                       else np.asarray(x, np.uint8))
    for reads, haps in batches:
        bnr.append(len(reads))
        bnh.append(len(haps))
        for rd in reads:
            arrs = [as_u8(x) for x in rd]
            n = len(arrs[0])
            assert all(len(a) == n for a in arrs)
            rl.append(n)
            for key, a in zip(("rs", "q", "i", "d", "c"), arrs):
                pools[key].append(a)
        for hp in haps:
            a = as_u8(hp)
            hl.append(len(a))
            pools["hap"].append(a)
    cat = lambda xs: (np.concatenate(xs) if xs else np.zeros(0, np.uint8))
    rl = np.asarray(rl, np.int32)
    hl = np.asarray(hl, np.int32)
    roff = np.concatenate(([0], np.cumsum(rl, dtype=np.int64)))[:-1]
    hoff = np.concatenate(([0], np.cumsum(hl, dtype=np.int64)))[:-1]
    return FlatBatches(bnr, bnh, rl, roff, cat(pools["rs"]), cat(pools["q"]), cat(pools["i"]),
                       cat(pools["d"]), cat(pools["c"]), hl, hoff, cat(pools["hap"]))
