/*
 * TEST INFRASTRUCTURE -- plain-C restatement of the GKL Smith-Waterman with backtrack
 * (SURVEY.md 8f row 3).  Only tests/, tools/sw_bench.py's cpu_baseline leg and
 * __graft_entry__.smoke() may use it; the product never links or calls it.
 *
 * Follows tools/GKL/src/main/native/smithwaterman/PairWiseSW.h (paths relative to
 * /root/reference): smithWatermanBackTrack :40-263, getCIGAR :265-412, runSWOnePairBT_* :415-450,
 * constants smithwaterman_common.h:21-55.  The reference sweeps anti-diagonals with AVX vectors
 * over three rolling arrays; the cell recurrence, the boundary values, the order in which the
 * last-row / last-column candidates are visited and the traceback are restated here on full
 * matrices in scalar code.  Pinned against the reference's own AVX2 / AVX-512 objects by
 * tests/test_sw.py (oracle/_ref/libsw_ref.so, live) and tests/golden/sw_*.npz (generated from it).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

enum { OP_MATCH = 0, OP_INSERT = 1, OP_DELETE = 2, INSERT_EXT = 4, DELETE_EXT = 8,       /* :21-25 */
       SOFTCLIP = 9, INDEL = 10, LEADING_INDEL = 11, IGNORE = 12 };                      /* :26-29 */
#define MATRIX_MIN_CUTOFF (-100000000)                                                   /* :52 */
#define LOW_INIT_VALUE (INT32_MIN / 2)                                                   /* :53 */

static int iabs(int x) { return x < 0 ? -x : x; }

/* Returns the alignment offset; fills cigar (NUL-terminated, at most cap - 1 characters; returns
 * -1 - needed in *cigar_len when it does not fit) and the int32 score (the reference truncates it
 * to int16 on the way out, PairWiseSW.h:441). */
int sw_oracle_align(const uint8_t *seq1, int nrow, const uint8_t *seq2, int ncol, int match, int mismatch,
                    int open, int extend, int strategy, char *cigar, int cap, int *cigar_len, int *score_out)
{
    const size_t W = (size_t)ncol + 1;
    int32_t *H = (int32_t *)malloc((size_t)(nrow + 1) * W * sizeof(int32_t));
    int32_t *E = (int32_t *)malloc(W * sizeof(int32_t) * 2);       /* E of the current row, F per column */
    int32_t *F = E + W;
    uint8_t *bt = (uint8_t *)calloc((size_t)(nrow + 1) * W, 1);
    const int lead = (strategy == INDEL || strategy == LEADING_INDEL);                   /* :196-205 */
    H[0] = 0;                                                                            /* :86 */
    for (int j = 1; j <= ncol; ++j) H[j] = lead ? open + (j - 1) * extend : 0;
    for (int j = 0; j <= ncol; ++j) F[j] = LOW_INIT_VALUE;                               /* :77-80,206 */
    for (int i = 1; i <= nrow; ++i) {
        H[i * W] = lead ? open + (i - 1) * extend : 0;
        int32_t e = LOW_INIT_VALUE;                                                      /* :81-84,207 */
        for (int j = 1; j <= ncol; ++j) {
            /* MAIN_CODE :3-37; E = gap along the row (INSERT: j--), F = gap along the column */
            const int32_t ext_h = e + extend, open_h = H[i * W + j - 1] + open;
            const int32_t e11 = open_h > ext_h ? open_h : ext_h;
            int flags = (open_h > ext_h) ? 0 : INSERT_EXT;
            const int32_t ext_v = F[j] + extend, open_v = H[(i - 1) * W + j] + open;
            const int32_t f11 = ext_v > open_v ? ext_v : open_v;
            flags |= (open_v > ext_v) ? 0 : DELETE_EXT;
            const int32_t m11 = H[(i - 1) * W + j - 1] + (seq1[i - 1] == seq2[j - 1] ? match : mismatch);
            int32_t h11 = m11 > MATRIX_MIN_CUTOFF ? m11 : MATRIX_MIN_CUTOFF;
            int b = (e11 > h11) ? OP_INSERT : OP_MATCH;
            if (e11 > h11) h11 = e11;
            if (f11 > h11) { b = OP_DELETE; h11 = f11; }
            e = e11; F[j] = f11;
            H[i * W + j] = h11;
            bt[i * W + j] = (uint8_t)(b | flags);
        }
    }
    /* last-row and last-column candidates in anti-diagonal order, row candidate first (:209-236) */
    int32_t maxScore = INT32_MIN;
    int max_i = 0, max_j = 0;
    for (int d = 1; d <= nrow + ncol; ++d) {
        if (d >= nrow + 1) {
            const int i = nrow, j = d - nrow;
            const int32_t s = H[i * W + j];
            if (strategy == SOFTCLIP || strategy == IGNORE)
                if (maxScore < s || (maxScore == s && iabs(i - j) < iabs(max_i - max_j))) { maxScore = s; max_i = i; max_j = j; }
        }
        if (d >= ncol + 1) {
            const int i = d - ncol, j = ncol;
            const int32_t s = H[i * W + j];
            if (maxScore < s || (maxScore == s && (max_j == ncol || iabs(i - j) <= iabs(max_i - max_j)))) { maxScore = s; max_i = i; max_j = j; }
        }
    }
    *score_out = maxScore;                                                               /* :257 (the INDEL branch :254 is overwritten) */

    /* getCIGAR :265-412 */
    int16_t *el = (int16_t *)malloc(4 * (size_t)(nrow + ncol + 4) * sizeof(int16_t));
    int n = 0, i, j, state = 0, offset;
    if (strategy == INDEL) { i = nrow; j = ncol; }
    else if (strategy == LEADING_INDEL) { i = max_i; j = ncol; }
    else { i = max_i; j = max_j; }
    if (j < ncol) { el[2 * n] = SOFTCLIP; el[2 * n + 1] = (int16_t)(ncol - j); ++n; }
    while (i > 0 && j > 0) {
        const int btr = bt[i * W + j];
        if (state == INSERT_EXT) { --j; el[2 * n - 1]++; state = btr & INSERT_EXT; }
        else if (state == DELETE_EXT) { --i; el[2 * n - 1]++; state = btr & DELETE_EXT; }
        else switch (btr & 3) {
            case OP_MATCH:  --i; --j; el[2 * n] = OP_MATCH; el[2 * n + 1] = 1; state = 0; ++n; break;
            case OP_INSERT: --j; el[2 * n] = OP_INSERT; el[2 * n + 1] = 1; state = btr & INSERT_EXT; ++n; break;
            case OP_DELETE: --i; el[2 * n] = OP_DELETE; el[2 * n + 1] = 1; state = btr & DELETE_EXT; ++n; break;
        }
    }
    if (strategy == SOFTCLIP) {
        if (j > 0) { el[2 * n] = SOFTCLIP; el[2 * n + 1] = (int16_t)j; ++n; }
        offset = i;
    } else if (strategy == IGNORE) {
        if (j > 0) { el[2 * n] = el[2 * (n - 1)]; el[2 * n + 1] = (int16_t)j; ++n; }
        offset = i - j;
    } else {
        if (i > 0) { el[2 * n] = OP_DELETE; el[2 * n + 1] = (int16_t)i; ++n; }
        else if (j > 0) { el[2 * n] = OP_INSERT; el[2 * n + 1] = (int16_t)j; ++n; }
        offset = 0;
    }
    int m = 0;                                                                           /* merge equal neighbours :372-388 */
    for (int k = 1; k < n; ++k) {
        if (el[2 * k] == el[2 * m]) el[2 * m + 1] += el[2 * k + 1];
        else { ++m; el[2 * m] = el[2 * k]; el[2 * m + 1] = el[2 * k + 1]; }
    }
    size_t need = 0;
    char tmp[16];
    for (int k = m; k >= 0; --k) need += (size_t)snprintf(tmp, sizeof tmp, "%d", el[2 * k + 1]) + 1;
    if ((int)need + 1 > cap) {
        *cigar_len = -1 - (int)need;
    } else {
        size_t at = 0;
        for (int k = m; k >= 0; --k) {                                                   /* :390-410 */
            at += (size_t)sprintf(cigar + at, "%d", el[2 * k + 1]);
            const int op = el[2 * k];
            cigar[at++] = op == OP_MATCH ? 'M' : op == OP_INSERT ? 'I' : op == OP_DELETE ? 'D' : op == SOFTCLIP ? 'S' : 'R';
        }
        cigar[at] = 0;
        *cigar_len = (int)at;
    }
    free(el); free(bt); free(E); free(H);
    return offset;
}

/* Batch form over flat pools, OpenMP over pairs (the timed CPU baseline of tools/sw_bench.py when
 * the reference library is absent). */
void sw_oracle_run(int64_t n, const uint8_t *ref_pool, const int64_t *ref_off, const int32_t *ref_len,
                   const uint8_t *alt_pool, const int64_t *alt_off, const int32_t *alt_len, int match, int mismatch,
                   int open, int extend, int strategy, int32_t *score, int32_t *offset, char *cigar, int stride,
                   int32_t *cigar_len, int nthreads)
{
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads > 0 ? nthreads : 1)
    for (int64_t p = 0; p < n; ++p) {
        int sc = 0, cl = 0;
        offset[p] = sw_oracle_align(ref_pool + ref_off[p], ref_len[p], alt_pool + alt_off[p], alt_len[p], match,
                                    mismatch, open, extend, strategy, cigar + p * (int64_t)stride, stride, &cl, &sc);
        score[p] = sc;
        cigar_len[p] = cl;
    }
}
