"""Flat structure-of-arrays form of Smith-Waterman inputs and seeded synthetic generators.

A pair is (ref, alt) as the reference's C entry takes it (`align(char *ref, char *alt, ...)`,
tools/GKL/src/main/native/smithwaterman/IntelSmithWatermanCSource.cpp:45; JNI
`alignNative(byte[] ref, byte[] alt, ...)`, IntelSmithWaterman.cc:52-75): `ref` indexes the rows of
the DP matrix (seq1, len1), `alt` the columns (seq2, len2) (PairWiseSW.h:423-434).  Sequences live in
two byte pools with offset tables -- the layout `include/sw_cuda.h`'s `sw_flat_pairs` hands to
the CUDA library.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

MAX_SEQ_LEN = 1024        # smithwaterman_common.h:50


@dataclass
class SwPairs:
    ref_len: np.ndarray     # int32[n]
    ref_off: np.ndarray     # int64[n]
    ref_pool: np.ndarray    # uint8
    alt_len: np.ndarray
    alt_off: np.ndarray
    alt_pool: np.ndarray

    def __post_init__(self):
        self.ref_len = np.ascontiguousarray(self.ref_len, np.int32)
        self.alt_len = np.ascontiguousarray(self.alt_len, np.int32)
        self.ref_off = np.ascontiguousarray(self.ref_off, np.int64)
        self.alt_off = np.ascontiguousarray(self.alt_off, np.int64)
        self.ref_pool = np.ascontiguousarray(self.ref_pool, np.uint8)
        self.alt_pool = np.ascontiguousarray(self.alt_pool, np.uint8)
        assert len(self.ref_len) == len(self.alt_len) == len(self.ref_off) == len(self.alt_off)

    @property
    def n_pairs(self) -> int:
        return len(self.ref_len)

    @property
    def n_cells(self) -> int:
        return int((self.ref_len.astype(np.int64) * self.alt_len).sum())

    def pair(self, p):
        r = self.ref_pool[self.ref_off[p]:self.ref_off[p] + self.ref_len[p]].tobytes()
        a = self.alt_pool[self.alt_off[p]:self.alt_off[p] + self.alt_len[p]].tobytes()
        return r, a

    def select(self, idx) -> "SwPairs":
        idx = np.asarray(idx, np.int64)
        return SwPairs(self.ref_len[idx], self.ref_off[idx], self.ref_pool, self.alt_len[idx],
                       self.alt_off[idx], self.alt_pool)


def from_lists(pairs) -> SwPairs:
    """[(ref bytes, alt bytes)] -> SwPairs; identical ref objects share pool bytes."""
    rl, ro, al, ao, rp, ap = [], [], [], [], [], []
    seen, rpos, apos = {}, 0, 0
    for r, a in pairs:
        if r in seen:
            off = seen[r]
        else:
            off = seen[r] = rpos
            rp.append(np.frombuffer(r, np.uint8))
            rpos += len(r)
        rl.append(len(r)); ro.append(off)
        al.append(len(a)); ao.append(apos)
        ap.append(np.frombuffer(a, np.uint8))
        apos += len(a)
    cat = lambda xs: np.concatenate(xs) if xs else np.zeros(0, np.uint8)
    return SwPairs(rl, ro, cat(rp), al, ao, cat(ap))


def _mutate(rng, s, snp, indel, max_indel):
    """SNPs and short indels, the way candidate haplotypes differ from the reference window."""
    out, k = [], 0
    n = len(s)
    while k < n:
        u = rng.random()
        if u < indel / 2:                                   # deletion from ref
            k += int(rng.integers(1, max_indel + 1))
            continue
        if u < indel:                                       # insertion into alt
            out.append(rng.integers(0, 4, int(rng.integers(1, max_indel + 1))).astype(np.uint8))
        b = s[k]
        if rng.random() < snp:
            b = (b + rng.integers(1, 4)) & 3
        out.append(np.array([b], np.uint8))
        k += 1
    return np.concatenate(out) if out else np.zeros(1, np.uint8)


def synth(seed: int, n_pairs: int, ref_rng=(100, 600), alts_per_ref=(1, 16), snp=0.01, indel=0.004,
          max_indel=12, trim=0.3, read_like=0.0, alphabet=b"ACGT") -> SwPairs:
    """HaplotypeCaller-shaped inputs: one reference window per active region, aligned against its
    candidate haplotypes (SNPs / indels of the window, ends sometimes trimmed or extended) and,
    with probability `read_like`, against short read-like substrings (dangling-end recovery)."""
    rng = np.random.default_rng(seed)
    lut = np.frombuffer(alphabet, np.uint8)
    pairs = []
    while len(pairs) < n_pairs:
        R = int(rng.integers(ref_rng[0], ref_rng[1] + 1))
        ref = rng.integers(0, 4, R).astype(np.uint8)
        refb = lut[ref].tobytes()
        for _ in range(int(rng.integers(alts_per_ref[0], alts_per_ref[1] + 1))):
            if len(pairs) >= n_pairs:
                break
            if rng.random() < read_like:
                L = int(rng.integers(max(1, min(20, R)), max(2, min(R, 150)) + 1))
                o = int(rng.integers(0, R - L + 1))
                alt = _mutate(rng, ref[o:o + L], snp * 2, indel, max_indel)
            else:
                alt = _mutate(rng, ref, snp, indel, max_indel)
                if rng.random() < trim:
                    a, b = int(rng.integers(0, 20)), int(rng.integers(0, 20))
                    alt = alt[a:len(alt) - b] if len(alt) > a + b + 1 else alt
                if rng.random() < trim / 2:
                    alt = np.concatenate((rng.integers(0, 4, int(rng.integers(1, 15))).astype(np.uint8), alt))
            alt = alt[:MAX_SEQ_LEN]
            pairs.append((refb, lut[alt].tobytes()))
    return from_lists(pairs)


def edge_cases(seed: int = 7) -> SwPairs:
    """Length-1 sequences, identical / disjoint sequences, homopolymers (maximal score ties),
    len1 >> len2 and len2 >> len1, the maximum length, lower-case and N bytes (plain byte equality,
    PairWiseSW.h:24-26)."""
    rng = np.random.default_rng(seed)
    rnd = lambda n: bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), n))
    ps = [(b"A", b"A"), (b"A", b"C"), (b"ACGT", b"A"), (b"A", b"ACGT"), (b"ACGT", b"ACGT"),
          (b"AAAAAAAAAA", b"AAAA"), (b"AAAA", b"AAAAAAAAAA"), (b"AAAAAAAA", b"CCCCCCCC"),
          (b"ACGTACGTACGTACGT", b"ACGTACGTTTACGTACGT"), (b"ACGTNNACGT", b"ACGTnnACGT"),
          (b"TTTTACGTACGTTTTT", b"ACGTACGT"), (b"ACGTACGT", b"GGGGACGTACGTGGGG")]
    # the reference itself is only defined up to 1023 rows: with len1 = MAX_SEQ_LEN it writes
    # E[MAX_SEQ_LEN - ilo] = E[-1] (PairWiseSW.h:207 with ilo = nrow + 1) and corrupts its heap
    big = rnd(MAX_SEQ_LEN)
    top = big[:MAX_SEQ_LEN - 1]
    ps += [(top, top), (top, rnd(MAX_SEQ_LEN)), (top, big[100:400]), (big[100:400], big),
           (big[:513], big[3:516]), (big[:257], big[:256]), (big[:256], big[:257]), (big[:33], big[:31])]
    for n1, n2 in ((255, 256), (256, 255), (257, 300), (31, 700), (700, 31), (1023, 1), (1, 1024)):
        ps.append((rnd(n1), rnd(n2)))
    return from_lists(ps)


def full_length_cases(seed: int = 8) -> SwPairs:
    """len1 = MAX_SEQ_LEN rows: beyond what the reference survives (see edge_cases), defined here."""
    rng = np.random.default_rng(seed)
    rnd = lambda n: bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), n))
    big = rnd(MAX_SEQ_LEN)
    return from_lists([(big, big), (big, rnd(MAX_SEQ_LEN)), (big, big[5:900]), (big, b"A")])
