// Host-side planner: flat batches -> haplotype stream, read table, per-class warp tasks.
//
// The reference driver turns a batch into num_reads x num_haps `testcase`s that point at the
// same strings over and over (PairHMMUnitTest_orig.cpp:284-297) and the library then redoes the
// per-read setup for every haplotype and the per-haplotype mask precomputation for every read
// (avx-pairhmm-template.h:256,262).  Here a batch is kept in its factored form: every read is
// uploaded once, every haplotype is coded once into the "haplotype stream", and a task pairs
// one read (or 2-4 short reads) with a run of consecutive haplotypes of its batch.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "phmm_host.h"

namespace phmm {

bool choose_class(int rslen, const PlanOptions &opt, KClass *out)
{
    const int vl = opt.precision == 2 ? 1 : 2;
    if (opt.force_kh > 0 && opt.force_lanes > 0) {
        KClass kc{opt.force_kh, opt.precision == 2 ? 32 : opt.force_lanes, vl, 0};
        if (kc.capacity() >= rslen) { *out = kc; return true; }
    }
    // Fewest padded rows wins; the register-file-bound cell update costs the same per row in
    // every class.  16 lanes (two read pairs per warp, half the wavefront fill) for rows that fit,
    // 32 lanes above; FP64 always 32 lanes.
    const int rows = rslen + 1;                // + the dummy row that carries DP row 0
    if (vl == 2 && rows <= 16 * MAX_KH) { *out = KClass{(rows + 15) / 16, 16, vl, 0}; return true; }
    if (rows <= 32 * MAX_KH) { *out = KClass{(rows + 31) / 32, 32, vl, 0}; return true; }
    *out = KClass{LONG_KH, 32, vl, 1};         // longer than every single-pass class: multi-pass
    return true;
}

static int scalar_class_of(int rslen)
{
    const int kh = (rslen + 1 + 31) / 32;
    return kh <= MAX_KH_SCALAR ? kh : 0;
}

int build_plan(const phmm_flat_batches &fb, const PlanOptions &opt, Plan *plan, std::string *err)
{
    char msg[256];
    auto fail = [&](int code, const char *m) { if (err) *err = m; return code; };
    if (fb.n_batches < 0 || fb.n_reads < 0 || fb.n_haps < 0) return fail(PHMM_ERR_ARG, "negative count");
    if (fb.n_batches && (!fb.batch_num_reads || !fb.batch_num_haps)) return fail(PHMM_ERR_ARG, "NULL batch table");
    if (fb.n_reads && (!fb.read_len || !fb.read_off || !fb.rs || !fb.q || !fb.i || !fb.d || !fb.c))
        return fail(PHMM_ERR_ARG, "NULL read table or pool");
    if (fb.n_haps && (!fb.hap_len || !fb.hap_off || !fb.hap)) return fail(PHMM_ERR_ARG, "NULL haplotype table or pool");

    Plan &P = *plan;
    P = Plan();
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) P.resc_cap[k] = 0;

    // ---- validate tables, count ---------------------------------------------------------------
    int64_t nr = 0, nh = 0;
    for (int64_t b = 0; b < fb.n_batches; ++b) {
        if (fb.batch_num_reads[b] < 0 || fb.batch_num_haps[b] < 0) return fail(PHMM_ERR_ARG, "negative batch size");
        nr += fb.batch_num_reads[b];
        nh += fb.batch_num_haps[b];
    }
    if (nr != fb.n_reads || nh != fb.n_haps) return fail(PHMM_ERR_ARG, "batch tables do not add up to n_reads / n_haps");
    if (nh + 1 >= INT32_MAX) return fail(PHMM_ERR_ARG, "too many haplotypes");
    int64_t lo = INT64_MAX, hi = 0;
    for (int64_t r = 0; r < fb.n_reads; ++r) {
        const int len = fb.read_len[r];
        if (len <= 0) return fail(PHMM_ERR_EMPTY_SEQ, "read of length <= 0 (undefined in the reference)");
        if (len > PHMM_MAX_SEQ_LEN) return fail(PHMM_ERR_TOO_LONG, "read longer than PHMM_MAX_SEQ_LEN");
        if (fb.read_off[r] < 0) return fail(PHMM_ERR_ARG, "negative read offset");
        lo = std::min(lo, fb.read_off[r]);
        hi = std::max(hi, fb.read_off[r] + len);
        P.max_rslen = std::max(P.max_rslen, len);
    }
    if (fb.n_reads == 0) lo = 0;
    P.pool_lo = lo;
    P.pool_hi = hi;
    for (int64_t h = 0; h < fb.n_haps; ++h) {
        const int len = fb.hap_len[h];
        if (len <= 0) return fail(PHMM_ERR_EMPTY_SEQ, "haplotype of length <= 0 (division by haplen in the reference)");
        if (len > PHMM_MAX_SEQ_LEN) return fail(PHMM_ERR_TOO_LONG, "haplotype longer than PHMM_MAX_SEQ_LEN");
        if (fb.hap_off[h] < 0) return fail(PHMM_ERR_ARG, "negative haplotype offset");
        P.max_haplen = std::max(P.max_haplen, len);
    }

    // ---- haplotype stream: one code byte per column, SEP_COLS zero columns after each haplotype,
    //      every batch's first column 16-byte aligned (cp.async granularity) ---------------------
    std::vector<int64_t> hap_col(fb.n_haps + 1, 0);
    {
        int64_t total = 0, h = 0;
        for (int64_t b = 0; b < fb.n_batches; ++b) {
            total = (total + 15) & ~(int64_t)15;
            for (int k = 0; k < fb.batch_num_haps[b]; ++k, ++h) {
                hap_col[h] = total;
                total += fb.hap_len[h] + SEP_COLS;
            }
        }
        hap_col[fb.n_haps] = total;
        P.stream.assign((size_t)total + STREAM_SLACK, 0);
        P.hap_len.assign((size_t)fb.n_haps + 1, 1);
        for (h = 0; h < fb.n_haps; ++h) {
            const uint8_t *src = fb.hap + fb.hap_off[h];
            uint8_t *dst = P.stream.data() + hap_col[h];
            const int len = fb.hap_len[h];
            for (int k = 0; k < len; ++k) dst[k] = base_code(src[k]);
            P.hap_len[h] = len;
        }
    }

    // ---- reads, classes, tasks -----------------------------------------------------------------
    P.reads.resize((size_t)fb.n_reads);
    struct ClassKey { int kh, lanes, vl; };
    auto class_index = [&](const KClass &kc) -> int {
        for (size_t k = 0; k < P.classes.size(); ++k)
            if (P.classes[k].kc.kh == kc.kh && P.classes[k].kc.lanes == kc.lanes && P.classes[k].kc.vl == kc.vl &&
                P.classes[k].kc.longr == kc.longr)
                return (int)k;
        ClassPlan cp;
        cp.kc = kc;
        P.classes.push_back(cp);
        return (int)P.classes.size() - 1;
    };

    // adaptive run length: aim at >= 3 tasks per resident warp, but keep runs long enough that
    // the wavefront fill (<= 63 steps) stays small
    int seg_cols = opt.seg_cols > 0 ? opt.seg_cols : 4096;
    std::vector<int> read_class((size_t)fb.n_reads);
    {
        int64_t r = 0;
        for (int64_t b = 0; b < fb.n_batches; ++b)
            for (int k = 0; k < fb.batch_num_reads[b]; ++k, ++r) {
                KClass kc;
                if (!choose_class(fb.read_len[r], opt, &kc)) {
                    snprintf(msg, sizeof msg, "read %lld of length %d exceeds the largest single-pass kernel class",
                             (long long)r, fb.read_len[r]);
                    return fail(PHMM_ERR_TOO_LONG, msg);
                }
                read_class[r] = class_index(kc);
            }
    }
    auto count_tasks = [&](int cols) -> int64_t {
        int64_t n = 0, r0 = 0, h0 = 0;
        std::vector<int64_t> per_class(P.classes.size());
        for (int64_t b = 0; b < fb.n_batches; ++b) {
            const int bnr = fb.batch_num_reads[b], bnh = fb.batch_num_haps[b];
            int64_t runs = 0, acc = 0;
            for (int k = 0, n_in = 0; k < bnh; ++k) {
                acc += fb.hap_len[h0 + k] + SEP_COLS;
                if (acc >= cols || ++n_in >= MAX_RUN_HAPS || k == bnh - 1) { ++runs; acc = 0; n_in = 0; }
            }
            std::fill(per_class.begin(), per_class.end(), 0);
            for (int k = 0; k < bnr; ++k) per_class[read_class[r0 + k]]++;
            for (size_t c = 0; c < P.classes.size(); ++c) {
                const int g = P.classes[c].kc.group();
                n += runs * ((per_class[c] + g - 1) / g);
            }
            r0 += bnr;
            h0 += bnh;
        }
        return n;
    };
    if (opt.seg_cols <= 0)
        while (seg_cols > 512 && count_tasks(seg_cols) < 3 * (int64_t)opt.resident_warps) seg_cols /= 2;

    int64_t r0 = 0, h0 = 0, out0 = 0;
    std::vector<std::vector<int32_t>> by_class(P.classes.size());
    for (int64_t b = 0; b < fb.n_batches; ++b) {
        const int bnr = fb.batch_num_reads[b], bnh = fb.batch_num_haps[b];
        int64_t sum_h = 0;
        for (int k = 0; k < bnh; ++k) sum_h += fb.hap_len[h0 + k];
        for (auto &v : by_class) v.clear();
        for (int k = 0; k < bnr; ++k) {
            const int64_t r = r0 + k;
            ReadRec &rr = P.reads[r];
            rr.off = fb.read_off[r] - P.pool_lo;
            rr.len = fb.read_len[r];
            rr.batch = (int32_t)b;
            rr.out_base = out0 + (int64_t)k * bnh;
            by_class[read_class[r]].push_back((int32_t)r);
            P.n_cells += (int64_t)rr.len * sum_h;
            P.resc_cap[scalar_class_of(rr.len)] += bnh;
        }
        // a class's leftover reads (fewer than one full task) move up to the next larger class of
        // the batch, where they fill slots instead of leaving half a warp idle
        {
            std::vector<int> order;
            for (size_t c = 0; c < P.classes.size(); ++c)
                if (!by_class[c].empty()) order.push_back((int)c);
            std::sort(order.begin(), order.end(), [&](int x, int y) {
                const KClass &kx = P.classes[x].kc, &ky = P.classes[y].kc;
                if (kx.longr != ky.longr) return kx.longr < ky.longr;
                return kx.capacity() < ky.capacity();
            });
            for (size_t oi = 0; oi + 1 < order.size(); ++oi) {
                std::vector<int32_t> &src = by_class[order[oi]], &dst = by_class[order[oi + 1]];
                const size_t rem = src.size() % (size_t)P.classes[order[oi]].kc.group();
                for (size_t j = 0; j < rem; ++j) { dst.push_back(src.back()); src.pop_back(); }
            }
        }
        if (bnh > 0)
            for (size_t c = 0; c < P.classes.size(); ++c)
                for (int32_t r : by_class[c]) {
                    P.classes[c].cells += (int64_t)P.reads[r].len * sum_h;
                    P.classes[c].pairs += bnh;
                }
        // runs of consecutive haplotypes
        int k = 0;
        while (k < bnh) {
            int k1 = k;
            int64_t acc = 0;
            while (k1 < bnh && acc < seg_cols && k1 - k < MAX_RUN_HAPS) { acc += fb.hap_len[h0 + k1] + SEP_COLS; ++k1; }
            for (size_t c = 0; c < P.classes.size(); ++c) {
                const std::vector<int32_t> &rl = by_class[c];
                const int g = P.classes[c].kc.group();
                for (size_t i = 0; i < rl.size(); i += g) {
                    Task t;
                    t.stream_off = hap_col[h0 + k];
                    t.ncols = (int32_t)acc;
                    t.hap0 = (int32_t)(h0 + k);
                    t.out0 = k;
                    for (int j = 0; j < MAX_GROUP; ++j)
                        t.read[j] = (j < g && i + j < rl.size()) ? rl[i + j] : -1;
                    t.nh = k1 - k;
                    P.classes[c].tasks.push_back(t);
                    if (P.classes[c].kc.longr) P.long_max_cols = std::max(P.long_max_cols, t.ncols);
                }
            }
            k = k1;
        }
        P.n_pairs += (int64_t)bnr * bnh;
        out0 += (int64_t)bnr * bnh;
        r0 += bnr;
        h0 += bnh;
    }
    // longest first: the device-side queue then behaves like LPT list scheduling
    for (auto &cp : P.classes)
        std::stable_sort(cp.tasks.begin(), cp.tasks.end(),
                         [](const Task &a, const Task &b) { return a.ncols > b.ncols; });
    // drop classes that ended up empty (possible when a batch has no haplotypes)
    P.classes.erase(std::remove_if(P.classes.begin(), P.classes.end(),
                                   [](const ClassPlan &c) { return c.tasks.empty(); }),
                    P.classes.end());
    return PHMM_OK;
}

}  // namespace phmm

// Longest-processing-time-first partition (include/phmm_cuda.h).
extern "C" int phmm_partition_lpt(const int64_t *cost, int64_t n, int parts, int32_t *owner)
{
    if (n < 0 || parts <= 0 || (n && (!cost || !owner))) return PHMM_ERR_ARG;
    std::vector<int64_t> order((size_t)n);
    for (int64_t k = 0; k < n; ++k) order[k] = k;
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return cost[a] > cost[b]; });
    std::vector<int64_t> load((size_t)parts, 0);
    for (int64_t k = 0; k < n; ++k) {
        int best = 0;
        for (int p = 1; p < parts; ++p)
            if (load[p] < load[best]) best = p;
        owner[order[k]] = best;
        load[best] += cost[order[k]];
    }
    return PHMM_OK;
}
