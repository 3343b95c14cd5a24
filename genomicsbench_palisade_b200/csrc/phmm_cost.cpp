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

// Hashed table of the k-mers of one haplotype (64 K entries of stamp | position: 512 KB, so it stays in the
// core's L2); a hit is verified against the haplotype's bases, entries are validated by a stamp, so that moving on
// to the next batch costs nothing.  A later k-mer that hashes to the same entry replaces the earlier one.
const int HBITS = 16;
struct KmerIndex {
    std::vector<uint64_t> e;
    uint32_t cur = 0;
    KmerIndex() : e((size_t)1 << HBITS, 0) {}
};

struct Base2Table {
    int8_t t[256];
    Base2Table() { memset(t, -1, sizeof t); t['A'] = 0; t['C'] = 1; t['G'] = 2; t['T'] = 3; }
};
const Base2Table g_base2;
inline int base2(uint8_t b) { return g_base2.t[b]; }      // N and everything else (-1) interrupts a k-mer

inline uint32_t slot_of(uint32_t key) { return (key * 2654435761u) >> (32 - HBITS); }

void index_hap(KmerIndex &ix, const uint8_t *hap, int len)
{
    ++ix.cur;
    uint32_t key = 0;
    int run = 0;
    for (int i = 0; i < len; ++i) {
        const int c = base2(hap[i]);
        if (c < 0) { run = 0; continue; }
        key = ((key << 2) | (uint32_t)c) & KMASK;
        if (++run >= K) ix.e[slot_of(key)] = ((uint64_t)ix.cur << 32) | (uint32_t)(i - K + 1);
    }
}

// position in the haplotype of the k-mer that starts at `kmer` (K bases, key `key`), or -1
inline int lookup(const KmerIndex &ix, uint32_t key, const uint8_t *hap, const uint8_t *kmer)
{
    const uint64_t v = ix.e[slot_of(key)];
    if ((uint32_t)(v >> 32) != ix.cur) return -1;
    const int pos = (int)(uint32_t)v;
    return memcmp(hap + pos, kmer, K) == 0 ? pos : -1;
}

// Six k-mers spread over the read; four of them on (nearly) one diagonal of the haplotype: the read sits on it with few
// mismatches and is nowhere near failing the FP32 pass.  This is what nine reads in ten look like, and it costs a
// twentieth of anchored_score.
bool sits_on_a_diagonal(const KmerIndex &ix, const uint8_t *hap, const uint8_t *rs, int rlen)
{
    const int NS = 6;
    if (rlen < 4 * K) return false;
    int d[NS], nd = 0;
    for (int i = 0; i < NS; ++i) {
        const int p = (int)((int64_t)i * (rlen - K) / (NS - 1));
        uint32_t key = 0;
        bool ok = true;
        for (int j = 0; j < K && ok; ++j) {
            const int c = base2(rs[p + j]);
            ok = c >= 0;
            key = (key << 2) | (uint32_t)(c & 3);
        }
        if (!ok) continue;
        const int pos = lookup(ix, key & KMASK, hap, rs + p);
        if (pos >= 0) d[nd++] = pos - p;
        if (i == 3 && nd == 4 && d[1] >= d[0] - 48 && d[1] <= d[0] + 48 && d[2] >= d[0] - 48 && d[2] <= d[0] + 48 &&
            d[3] >= d[0] - 48 && d[3] <= d[0] + 48)
            return true;                                   // the common case: the first four agree
    }
    // (diagonals within a few dozen columns of each other: an indel between two anchors shifts the later ones)
    for (int i = 0; i + 3 < nd; ++i) {
        int same = 1;
        for (int j = i + 1; j < nd; ++j) same += (d[j] >= d[i] - 48 && d[j] <= d[i] + 48);
        if (same >= 4) return true;
    }
    return false;
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
        const int pos = lookup(ix, key, hap, rs + (r - K + 1));
        if (pos < 0) continue;
        const int d = pos - (r - K + 1);
        int j = 0;
        while (j < nv && votes[j].d != d) ++j;
        if (j < nv) ++votes[j].n;
        else if (nv < 8) votes[nv++] = Vote{d, 1};
    }
    if (nv > 3) {
        std::partial_sort(votes, votes + 3, votes + nv, [](const Vote &a, const Vote &b) { return a.n > b.n; });
        nv = 3;
    }
    // a diagonal that a single k-mer voted for is a chance hit (an unrelated read collects one or two of those):
    // scoring the read along it costs a pass over the read per diagonal and changes nothing
    {
        int kept = 0;
        for (int j = 0; j < nv; ++j)
            if (votes[j].n >= 2) votes[kept++] = votes[j];
        nv = kept;
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

bool spec_enabled()
{
    static const bool on = !(getenv("PHMM_SPEC64") && atoi(getenv("PHMM_SPEC64")) == 0);
    return on;
}

// Whether a call is planned with FP64-first reads (PlanOptions::spec = -1: from the input).  The predictor costs the
// planner 0.12 us per read that sits on a diagonal and ~2 us per read that does not -- more than the rest of the plan
// (0.085 us per read) -- so it runs only where it can pay: when a read is at least ~0.7 us of device work (2.5e6
// cells: C1-shaped input, whose end-to-end rate is bound by planning, stays as it is) and when a sample of the reads
// (about 600 of them, spread over every third batch) has at least 1 % of them failing the FP32 pass by a factor of ten.
// Decided once per host-buffer call, like the latency plan: a pair's result does not depend on the cut into chunks.
int plan_spec(const phmm_flat_batches &fb, const PlanOptions &opt)
{
    if (opt.spec >= 0) return opt.spec;
    if (!spec_enabled()) return 0;
    if (!opt.hp || opt.precision != 0 || opt.latency == 1 || opt.force_kh > 0 || fb.n_reads <= 0 || fb.n_batches <= 0) return 0;
    if (!fb.batch_num_reads || !fb.batch_num_haps || !fb.read_len || !fb.hap_len || !fb.read_off || !fb.hap_off) return 0;
    double cells = 0;
    {
        int64_t r0 = 0, h0 = 0;
        for (int64_t b = 0; b < fb.n_batches; ++b) {
            const int nr = fb.batch_num_reads[b], nh = fb.batch_num_haps[b];
            if (nr < 0 || nh < 0 || r0 + nr > fb.n_reads || h0 + nh > fb.n_haps) return 0;
            double sr = 0, sh = 0;
            for (int k = 0; k < nr; ++k) sr += fb.read_len[r0 + k];
            for (int k = 0; k < nh; ++k) sh += fb.hap_len[h0 + k];
            cells += sr * sh;
            r0 += nr; h0 += nh;
        }
    }
    if (cells < 2.5e6 * (double)fb.n_reads) return 0;
    static const LogTables lt;
    thread_local KmerIndex ix;
    // about 600 reads, spread over the batches
    const int64_t bstride = fb.n_batches >= 12 ? 3 : 1;
    const int rstride = (int)std::max<int64_t>(4, fb.n_reads / (bstride * 600));
    int64_t seen = 0, sure = 0, r0 = 0, h0 = 0;
    for (int64_t b = 0; b < fb.n_batches; ++b) {
        const int nr = fb.batch_num_reads[b], nh = fb.batch_num_haps[b];
        if (b % bstride == 0 && nr > 0 && nh > 0 && fb.hap_len[h0] > 0) {
            const uint8_t *hap = fb.hap + fb.hap_off[h0];
            const int hlen = fb.hap_len[h0];
            index_hap(ix, hap, hlen);
            for (int k = (int)(b % rstride); k < nr; k += rstride) {
                const int len = fb.read_len[r0 + k];
                if (len <= 0) return 0;
                const int64_t ro = fb.read_off[r0 + k];
                ++seen;
                if (sits_on_a_diagonal(ix, hap, fb.rs + ro, len)) continue;
                const float s = anchored_score(ix, lt, hap, hlen, fb.rs + ro, fb.q + ro, fb.c + ro, len);
                // batch_costs' test with the 1 / haplen prior left out: stricter than for any haplotype
                if (s + 36.1236 + 28.0 < -1.0) ++sure;
            }
        }
        r0 += nr; h0 += nh;
    }
    return seen > 0 && sure * 100 >= seen ? 1 : 0;
}

int batch_costs(const phmm_flat_batches &fb, int fast, int precision, const CostModel &m, int64_t *cost,
                uint8_t *read_rescued, int spec_plan)
{
    if (fb.n_batches < 0 || (fb.n_batches && (!fb.batch_num_reads || !fb.batch_num_haps || !cost))) return PHMM_ERR_ARG;
    static const LogTables lt;
    thread_local KmerIndex ix;
    std::vector<double> log_hl;
    PlanOptions opt;
    opt.precision = precision;
    opt.hp = fast ? 1 : 0;
    opt.latency = 0;
    // whether reads named "sure" cost no FP32 pass: as the library would plan this very input (the scheduler hands the
    // whole workload over), unless the caller knows (the planner, which asks per slice of a chunk)
    bool spec = fast && precision == 0 && spec_enabled();
    if (spec && spec_plan < 0) {
        PlanOptions po;
        po.precision = 0; po.hp = 1; po.spec = -1;
        po.latency = plan_latency(fb, po);
        spec = plan_spec(fb, po) > 0;
    } else if (spec) {
        spec = spec_plan > 0;
    }
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
        if (predict) {
            index_hap(ix, fb.hap + fb.hap_off[h0], fb.hap_len[h0]);
            log_hl.resize((size_t)nh);
            for (int j = 0; j < nh; ++j) log_hl[(size_t)j] = std::log10((double)std::max(fb.hap_len[h0 + j], 1));
        }
        for (int k = 0; k < nr; ++k) {
            const int len = fb.read_len[r0 + k];
            if (len <= 0) return PHMM_ERR_EMPTY_SEQ;
            c32 += weight32(len);
            if (!predict) continue;
            const int64_t ro = fb.read_off[r0 + k];
            if (sits_on_a_diagonal(ix, fb.hap + fb.hap_off[h0], fb.rs + ro, len)) {
                if (read_rescued) read_rescued[r0 + k] = 0;
                continue;
            }
            const float s = anchored_score(ix, lt, fb.hap + fb.hap_off[h0], fb.hap_len[h0], fb.rs + ro, fb.q + ro, fb.c + ro, len);
            // FP32 sum = likelihood * 2^120 with prior 1 / haplen; rescued when below MIN_ACCEPTED = 1e-28
            double rc = 0;
            int n_resc = 0, n_sure = 0;
            for (int j = 0; j < nh; ++j) {
                const int hl = fb.hap_len[h0 + j];
                const double margin = s - log_hl[(size_t)j] + 36.1236 + 28.0;
                if (margin < 0) { rc += hl + SEP_COLS + 32; ++n_resc; }
                if (margin < -1.0) ++n_sure;
            }
            c64 += rc * weight64(len);
            // bit 0: most pairs of the read predicted rescued; bit 1: all of them, by at least a factor of ten -- such a
            // read is planned FP64-first (PlanOptions::spec) and costs no FP32 pass
            const bool sure = n_sure == nh && len + 1 <= 32 * MAX_KH_SCALAR;
            if (sure && spec) c32 -= weight32(len);
            if (read_rescued) read_rescued[r0 + k] = (uint8_t)((2 * n_resc >= nh ? 1 : 0) | (sure ? 2 : 0));
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
