// Predicted device time per batch, for the host-side length-balanced scheduler (phmm_partition_lpt).
//
// The reference balances nothing: its driver hands whole batches to OpenMP threads and `num_cells`
// (benchmarks/phmm/PairHMMUnitTest_orig.cpp:166) is only the numerator of its GCUPS figure.  Cells alone
// mispredict GPU time in two ways this file models:
//   * a read runs in the register tile of its kernel class (rows per thread x lanes, phmm_plan.cpp
//     choose_class): the device pays for the padded tile rows plus a per-step overhead per lane (shuffles,
//     emission-table loads, loop control), not for the read's own rows;
//   * a pair whose FP32 sum falls below MIN_ACCEPTED is computed a second time in FP64
//     (IntelPairHmmCSource.cpp:73-75) at less than half the FP32 cell rate.  Whether that happens is predicted
//     per read from a k-mer anchored, quality-weighted mismatch score against the first haplotype of the batch
//     (the haplotypes of a batch are variants of one another): no dynamic programming on the host.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "phmm_host.h"

namespace phmm {

CostModel default_cost_model()
{
    CostModel m;
    if (const char *s = getenv("PHMM_COST_MODEL")) {
        double v[5];
        const int n = sscanf(s, "%lf,%lf,%lf,%lf,%lf", &v[0], &v[1], &v[2], &v[3], &v[4]);
        if (n >= 1) m.step_rows_fast = v[0];
        if (n >= 2) m.step_rows_exact = v[1];
        if (n >= 3) m.step_rows_f64 = v[2];
        if (n >= 4) m.f64_row_cost = v[3];
        if (n >= 5) m.exact_row_cost = v[4];
    }
    return m;
}

namespace {

const int K = 10;                                  // anchor length
const uint32_t KMASK = (1u << (2 * K)) - 1;

// Direct-address table of the k-mers of one haplotype; entries are validated by a stamp, so that moving on
// to the next batch costs nothing.
struct KmerIndex {
    std::vector<int32_t> pos, stamp;
    int32_t cur = 0;
    KmerIndex() : pos(KMASK + 1, 0), stamp(KMASK + 1, 0) {}
};

inline int base2(uint8_t b)
{
    switch (b) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'T': return 3;
    default:  return -1;                           // N and everything else interrupts a k-mer
    }
}

void index_hap(KmerIndex &ix, const uint8_t *hap, int len)
{
    ++ix.cur;
    uint32_t key = 0;
    int run = 0;
    for (int i = 0; i < len; ++i) {
        const int c = base2(hap[i]);
        if (c < 0) { run = 0; continue; }
        key = ((key << 2) | (uint32_t)c) & KMASK;
        if (++run >= K) { ix.pos[key] = i - K + 1; ix.stamp[key] = ix.cur; }
    }
}

struct LogTables {
    float lm[128], lx[128];                        // log10(1 - eps_q), log10(eps_q / 3)
    LogTables()
    {
        for (int q = 0; q < 128; ++q) {
            const double eps = std::pow(10.0, -q / 10.0);
            lm[q] = (float)std::log10(std::max(1e-30, 1.0 - eps));
            lx[q] = (float)std::log10(eps / 3.0);
        }
    }
};

// log10 of the likelihood the best gap-free placements of the read on the haplotype give, up to the 1 / haplen
// prior: the sum over read positions of the match or mismatch emission, a position counting as matched when it
// matches on any of the (up to three) diagonals its k-mers voted for.
float anchored_score(const KmerIndex &ix, const LogTables &lt, const uint8_t *hap, int hlen, const uint8_t *rs,
                     const uint8_t *q, const uint8_t *gcp, int rlen)
{
    struct Vote { int d, n; } votes[8];
    int nv = 0;
    uint32_t key = 0;
    int run = 0;
    for (int r = 0; r < rlen; ++r) {
        const int c = base2(rs[r]);
        if (c < 0) { run = 0; continue; }
        key = ((key << 2) | (uint32_t)c) & KMASK;
        if (++run < K || ((r + 1) & 1)) continue;                       // every second k-mer
        if (ix.stamp[key] != ix.cur) continue;
        const int d = ix.pos[key] - (r - K + 1);
        int j = 0;
        while (j < nv && votes[j].d != d) ++j;
        if (j < nv) ++votes[j].n;
        else if (nv < 8) votes[nv++] = Vote{d, 1};
    }
    if (nv > 3) {
        std::partial_sort(votes, votes + 3, votes + nv, [](const Vote &a, const Vote &b) { return a.n > b.n; });
        nv = 3;
    }
    float s = 0.0f;
    if (nv == 0) {
        // no anchor: an unrelated read matches a quarter of its positions by chance
        for (int r = 0; r < rlen; ++r) s += 0.25f * lt.lm[q[r] & 127] + 0.75f * lt.lx[q[r] & 127];
        return s;
    }
    for (int r = 0; r < rlen; ++r) {
        bool m = rs[r] == 'N', inside = false;
        for (int j = 0; j < nv && !m; ++j) {
            const int c = r + votes[j].d;
            if (c < 0 || c >= hlen) continue;
            inside = true;
            m = hap[c] == rs[r] || hap[c] == 'N';
        }
        // a read base hanging over either end of the haplotype is an insertion: one gap continuation
        // (pXX = ph2pr[gcp], avx-pairhmm-template.h:117), not a mismatch
        s += (m || rs[r] == 'N') ? lt.lm[q[r] & 127] : inside ? lt.lx[q[r] & 127] : -0.1f * (float)(gcp[r] & 127);
    }
    return s;
}

}  // namespace

int batch_costs(const phmm_flat_batches &fb, int fast, int precision, const CostModel &m, int64_t *cost,
                uint8_t *read_rescued)
{
    if (fb.n_batches < 0 || (fb.n_batches && (!fb.batch_num_reads || !fb.batch_num_haps || !cost))) return PHMM_ERR_ARG;
    static const LogTables lt;
    thread_local KmerIndex ix;
    PlanOptions opt;
    opt.precision = precision;
    opt.hp = fast ? 1 : 0;
    opt.latency = 0;
    double w_of_len[512];
    for (double &v : w_of_len) v = -1.0;
    // tile weight of a read: device time per haplotype column, in units of one FP32 row step of the fast kernels
    auto weight32 = [&](int len) -> double {
        if (len < 512 && w_of_len[len] >= 0) return w_of_len[len];
        KClass kc;
        choose_class(len, opt, &kc);
        double w;
        if (kc.longr) {
            const int rows = kc.kh * kc.lanes, passes = (len + rows) / rows;
            w = passes * kc.lanes * (kc.kh + m.step_rows_exact) * m.exact_row_cost;
        } else if (precision == 2) {
            w = kc.lanes * (kc.kh + m.step_rows_f64) * m.f64_row_cost;
        } else if (kc.hp) {
            w = kc.lanes * (kc.kh + m.step_rows_fast) * (kc.lanes == 8 ? m.lanes8_cost : kc.lanes == 32 ? m.lanes32_cost : 1.0);
        } else {
            w = kc.lanes * (kc.kh + m.step_rows_exact) * m.exact_row_cost * (fast ? 0.8 : 1.0);
        }
        if (len < 512) w_of_len[len] = w;
        return w;
    };
    auto weight64 = [&](int len) -> double {
        const int kh = (len + 1 + 31) / 32;
        if (kh > MAX_KH_SCALAR) return ((len + 256) / 256) * 32 * (LONG_KH + m.step_rows_f64) * m.f64_row_cost;
        return 32 * (kh + m.step_rows_f64) * m.f64_row_cost;
    };
    int64_t r0 = 0, h0 = 0;
    for (int64_t b = 0; b < fb.n_batches; ++b) {
        const int nr = fb.batch_num_reads[b], nh = fb.batch_num_haps[b];
        if (nr < 0 || nh < 0 || r0 + nr > fb.n_reads || h0 + nh > fb.n_haps) return PHMM_ERR_ARG;
        double cols = 0;
        for (int k = 0; k < nh; ++k) cols += fb.hap_len[h0 + k] + SEP_COLS;
        double c32 = 0, c64 = 0;
        const bool predict = precision == 0 && nh > 0 && nr > 0 && fb.hap_len[h0] > 0;
        if (predict) index_hap(ix, fb.hap + fb.hap_off[h0], fb.hap_len[h0]);
        for (int k = 0; k < nr; ++k) {
            const int len = fb.read_len[r0 + k];
            if (len <= 0) return PHMM_ERR_EMPTY_SEQ;
            c32 += weight32(len);
            if (!predict) continue;
            const int64_t ro = fb.read_off[r0 + k];
            const float s = anchored_score(ix, lt, fb.hap + fb.hap_off[h0], fb.hap_len[h0], fb.rs + ro, fb.q + ro, fb.c + ro, len);
            // FP32 sum = likelihood * 2^120 with prior 1 / haplen; rescued when below MIN_ACCEPTED = 1e-28
            double rc = 0;
            int n_resc = 0;
            for (int j = 0; j < nh; ++j) {
                const int hl = fb.hap_len[h0 + j];
                if (s - std::log10((double)hl) + 36.1236 < -28.0) { rc += hl + SEP_COLS + 32; ++n_resc; }
            }
            c64 += rc * weight64(len);
            if (read_rescued) read_rescued[r0 + k] = (uint8_t)(2 * n_resc >= nh ? 1 : 0);
        }
        if (!predict && read_rescued) memset(read_rescued + r0, 0, (size_t)nr);
        cost[b] = (int64_t)std::llround(c32 * cols + c64);
        r0 += nr;
        h0 += nh;
    }
    return PHMM_OK;
}

}  // namespace phmm

extern "C" int phmm_batch_costs(const phmm_flat_batches *fb, int fp32_fast, int precision, int64_t *cost,
                                uint8_t *read_rescued)
{
    if (!fb || precision < 0 || precision > 2) return PHMM_ERR_ARG;
    return phmm::batch_costs(*fb, fp32_fast, precision, phmm::default_cost_model(), cost, read_rescued);
}
