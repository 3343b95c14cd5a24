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
#include <thread>

#include "phmm_host.h"

namespace phmm {

bool choose_class(int rslen, const PlanOptions &opt, KClass *out)
{
    const int vl = opt.precision == 2 ? 1 : 2;
    const bool hp = opt.hp && opt.precision != 2;
    const int rows = rslen + 1;                // + the dummy row that carries DP row 0
    if (opt.force_kh > 0 && opt.force_lanes > 0) {
        KClass kc{opt.force_kh, opt.precision == 2 ? 32 : opt.force_lanes, vl, 0, hp ? 1 : 0};
        const bool ok = hp ? (kc.kh <= MAX_KH_HP && (kc.lanes == 32 || kc.lanes == 16 || kc.lanes == 8))
                           : (kc.kh <= MAX_KH && (kc.lanes == 32 || (kc.lanes == 16 && vl == 2)));
        if (ok && kc.capacity() >= rslen) { *out = kc; return true; }
    }
    if (hp) {
        // Haplotype-pair kernels: the tallest register tile that fits wins (fewest lanes: least
        // wavefront fill, shuffles and sums amortised over the most rows).  A small call (one active region)
        // leaves most of the device idle and waits for its longest task: there the shortest tile wins --
        // 32 lanes per read, half the rows per step of the 16-lane class and twice the tasks.
        for (int lanes = opt.latency == 1 ? 32 : 8; lanes <= 32; lanes *= 2)
            if (rows <= lanes * MAX_KH_HP) { *out = KClass{(rows + lanes - 1) / lanes, lanes, 2, 0, 1}; return true; }
        *out = KClass{LONG_KH, 32, 2, 1, 0};
        return true;
    }
    // Fewest padded rows wins; the register-file-bound cell update costs the same per row in
    // every class.  16 lanes (two read pairs per warp, half the wavefront fill) for rows that fit,
    // 32 lanes above; FP64 always 32 lanes.
    if (vl == 2 && rows <= 16 * MAX_KH) { *out = KClass{(rows + 15) / 16, 16, vl, 0, 0}; return true; }
    if (rows <= 32 * MAX_KH) { *out = KClass{(rows + 31) / 32, 32, vl, 0, 0}; return true; }
    *out = KClass{LONG_KH, 32, vl, 1, 0};      // longer than every single-pass class: multi-pass
    return true;
}

static int scalar_class_of(int rslen)
{
    const int kh = (rslen + 1 + 31) / 32;
    return kh <= MAX_KH_SCALAR ? kh : 0;
}

// Whether a call is planned for latency (PlanOptions::latency = -1: by its size).  A host-buffer call decides once
// for all of its chunks, so that a pair's result does not depend on how the call was cut.
int plan_latency(const phmm_flat_batches &fb, const PlanOptions &opt)
{
    if (opt.latency >= 0) return opt.latency;
    // tasks if every read took one warp per pair of haplotypes (~50 us each): below eight per resident warp
    // the call ends before the throughput plan (long runs, tall tiles, few kernels at once) has filled the
    // device -- measured on 16 active regions per call: 0.55 ms of FP32 kernels against 0.25.  Not when the
    // tasks are few but long (kilobase haplotypes): above ~1 ms of device work the throughput plan's shorter
    // runs fill the device as well and its tall tiles run the cells faster.
    const int64_t limit = 8 * (int64_t)opt.resident_warps;
    int64_t t = 0, r = 0, h = 0;
    double cells = 0;
    for (int64_t b = 0; b < fb.n_batches && t < limit && cells < 3e9; ++b) {
        const int nr = std::max(fb.batch_num_reads[b], 0), nh = std::max(fb.batch_num_haps[b], 0);
        t += (int64_t)nr * ((nh + 1) / 2);
        if (r + nr <= fb.n_reads && h + nh <= fb.n_haps && fb.read_len && fb.hap_len) {
            double sr = 0, sh = 0;
            for (int k = 0; k < nr; ++k) sr += fb.read_len[r + k];
            for (int k = 0; k < nh; ++k) sh += fb.hap_len[h + k];
            cells += sr * sh;
        }
        r += nr; h += nh;
    }
    return (t < limit && cells < 3e9) ? 1 : 0;
}

// One-wave calls (one active region, a few): one kernel class.  When the tasks of a latency-planned call fit the
// device in one wave, the call lasts as long as its longest task whichever classes the shorter reads run in -- but
// every class is a launch of its own in each of the phases (FP32, guard band, FP64 rescue), with its fork and join
// events, and a single-region call is mostly launch overhead.  All reads then take the 32-lane class of the longest
// read, and all rescue / guard work goes to that read's scalar class.  Returns that class's height, 0 = no.  Like the
// latency decision it is taken once per host-buffer call, so a pair's result does not depend on the cut into chunks.
int plan_one_class(const phmm_flat_batches &fb, const PlanOptions &opt)
{
    if (opt.one_class >= 0) return opt.one_class;
    if (opt.latency != 1 || opt.force_kh > 0 || fb.n_reads <= 0 || !fb.read_len || !fb.batch_num_reads || !fb.batch_num_haps) return 0;
    int max_rslen = 0;
    for (int64_t r = 0; r < fb.n_reads; ++r) max_rslen = std::max(max_rslen, fb.read_len[r]);
    if (max_rslen < 1 || max_rslen + 1 > 32 * MAX_KH_SCALAR) return 0;
    const bool hp = opt.hp && opt.precision != 2;
    const int g = (hp || opt.precision == 2) ? 1 : 2;          // reads per 32-lane task
    int64_t t = 0;
    for (int64_t b = 0; b < fb.n_batches; ++b) {
        const int64_t nr = std::max(fb.batch_num_reads[b], 0), nh = std::max(fb.batch_num_haps[b], 0);
        t += ((nr + g - 1) / g) * (hp ? (nh + 1) / 2 : nh);     // one haplotype per run
        if (t > opt.resident_warps / 2) return 0;
    }
    return (max_rslen + 1 + 31) / 32;
}

int build_plan(const phmm_flat_batches &fb, const PlanOptions &opt_in, Plan *plan, std::string *err)
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
    PlanOptions opt = opt_in;
    if (opt.latency < 0) opt.latency = plan_latency(fb, opt);
    P.latency = opt.latency;

    // ---- validate tables, count ---------------------------------------------------------------
    int64_t nr = 0, nh = 0;
    for (int64_t b = 0; b < fb.n_batches; ++b) {
        if (fb.batch_num_reads[b] < 0 || fb.batch_num_haps[b] < 0) return fail(PHMM_ERR_ARG, "negative batch size");
        nr += fb.batch_num_reads[b];
        nh += fb.batch_num_haps[b];
    }
    if (nr != fb.n_reads || nh != fb.n_haps) return fail(PHMM_ERR_ARG, "batch tables do not add up to n_reads / n_haps");
    if (nh + 1 >= INT32_MAX) return fail(PHMM_ERR_ARG, "too many haplotypes");
    bool sorted = true;
    for (int64_t r = 0; r < fb.n_reads; ++r) {
        const int len = fb.read_len[r];
        if (len <= 0) return fail(PHMM_ERR_EMPTY_SEQ, "read of length <= 0 (undefined in the reference)");
        if (len > PHMM_MAX_SEQ_LEN) return fail(PHMM_ERR_TOO_LONG, "read longer than PHMM_MAX_SEQ_LEN");
        if (fb.read_off[r] < 0) return fail(PHMM_ERR_ARG, "negative read offset");
        if (r && fb.read_off[r] < fb.read_off[r - 1]) sorted = false;
        P.max_rslen = std::max(P.max_rslen, len);
    }
    if (opt.one_class < 0) opt.one_class = plan_one_class(fb, opt);
    if (opt.spec < 0) opt.spec = plan_spec(fb, opt);
    if (opt.one_class > 0 && opt.force_kh <= 0) {
        opt.force_kh = opt.one_class;
        opt.force_lanes = 32;
        P.resc_force = opt.one_class;
    }
    // ---- pool segments: merge the reads' byte ranges (gaps up to 64 bytes are carried along) so that
    //      only referenced data is uploaded, whatever else the caller's pools hold ------------------
    std::vector<int64_t> seg_of_read((size_t)fb.n_reads);
    {
        std::vector<int64_t> order;
        if (!sorted) {
            order.resize((size_t)fb.n_reads);
            for (int64_t r = 0; r < fb.n_reads; ++r) order[r] = r;
            std::sort(order.begin(), order.end(), [&](int64_t x, int64_t y) { return fb.read_off[x] < fb.read_off[y]; });
        }
        int64_t dst = 0;
        for (int64_t k = 0; k < fb.n_reads; ++k) {
            const int64_t r = sorted ? k : order[k];
            const int64_t off = fb.read_off[r], end = off + fb.read_len[r];
            if (P.segments.empty() || off > P.segments.back().src + P.segments.back().len + 64) {
                if (!P.segments.empty()) dst += (P.segments.back().len + 15) & ~(int64_t)15;
                P.segments.push_back(Plan::Segment{off, 0, dst});
            }
            Plan::Segment &sg = P.segments.back();
            sg.len = std::max(sg.len, end - sg.src);
            seg_of_read[r] = (int64_t)P.segments.size() - 1;
        }
        if (!P.segments.empty()) dst += P.segments.back().len;
        P.pool_len = dst;
    }
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
        // The code bytes themselves are produced on the device (phmm_code_stream_kernel, phmm_api.cu): the host only
        // says where every haplotype's raw bases go.  (Coding them here, byte-wise, was 40 % of the planning time of
        // small-batch input, and planning is what bounds the end-to-end rate there.)
        P.stream_len = total + STREAM_SLACK;
        P.hap_len.assign((size_t)fb.n_haps + 1, 1);
        if (P.stream_len <= HOST_CODED_STREAM_MAX) {
            // a small call (an active region or a few): coding a few kilobytes here is cheaper than a memset and a
            // kernel launch, and the stream then goes up with everything else in the one staged copy
            P.stream.assign((size_t)P.stream_len, 0);
            for (h = 0; h < fb.n_haps; ++h) {
                const uint8_t *src = fb.hap + fb.hap_off[h];
                uint8_t *dstb = P.stream.data() + hap_col[h];
                for (int k = 0; k < fb.hap_len[h]; ++k) dstb[k] = base_code(src[k]);
                P.hap_len[h] = fb.hap_len[h];
            }
        } else {
        P.hap_src.resize((size_t)fb.n_haps);
        bool hsorted = true;
        for (h = 0; h < fb.n_haps; ++h) {
            P.hap_len[h] = fb.hap_len[h];
            if (h && fb.hap_off[h] < fb.hap_off[h - 1]) hsorted = false;
        }
        // referenced pieces of the caller's haplotype pool, packed as the read pools are
        std::vector<int64_t> horder;
        if (!hsorted) {
            horder.resize((size_t)fb.n_haps);
            for (h = 0; h < fb.n_haps; ++h) horder[h] = h;
            std::sort(horder.begin(), horder.end(), [&](int64_t x, int64_t y) { return fb.hap_off[x] < fb.hap_off[y]; });
        }
        std::vector<int64_t> seg_of_hap((size_t)fb.n_haps);
        int64_t dst = 0;
        for (int64_t k = 0; k < fb.n_haps; ++k) {
            const int64_t hh = hsorted ? k : horder[k];
            const int64_t off = fb.hap_off[hh], end = off + fb.hap_len[hh];
            if (P.hap_segments.empty() || off > P.hap_segments.back().src + P.hap_segments.back().len + 64) {
                if (!P.hap_segments.empty()) dst += (P.hap_segments.back().len + 15) & ~(int64_t)15;
                P.hap_segments.push_back(Plan::Segment{off, 0, dst});
            }
            Plan::Segment &sg = P.hap_segments.back();
            sg.len = std::max(sg.len, end - sg.src);
            seg_of_hap[hh] = (int64_t)P.hap_segments.size() - 1;
        }
        if (!P.hap_segments.empty()) dst += P.hap_segments.back().len;
        P.hap_pool_len = dst;
        for (h = 0; h < fb.n_haps; ++h) {
            const Plan::Segment &sg = P.hap_segments[seg_of_hap[h]];
            P.hap_src[h] = HapSrc{fb.hap_off[h] - sg.src + sg.dst, hap_col[h], fb.hap_len[h], 0};
        }
        }
    }

    // ---- reads, classes -----------------------------------------------------------------------
    P.reads.resize((size_t)fb.n_reads);
    auto class_index = [&](const KClass &kc) -> int {
        for (size_t k = 0; k < P.classes.size(); ++k) {
            const KClass &o = P.classes[k].kc;
            if (o.kh == kc.kh && o.lanes == kc.lanes && o.vl == kc.vl && o.longr == kc.longr && o.hp == kc.hp) return (int)k;
        }
        ClassPlan cp;
        cp.kc = kc;
        P.classes.push_back(cp);
        return (int)P.classes.size() - 1;
    };
    std::vector<int> read_class((size_t)fb.n_reads);
    {
        int class_of_len[512];                 // memo for the common lengths
        for (int &v : class_of_len) v = -1;
        for (int64_t r = 0; r < fb.n_reads; ++r) {
            const int len = fb.read_len[r];
            if (len < 512 && class_of_len[len] >= 0) { read_class[r] = class_of_len[len]; continue; }
            KClass kc;
            if (!choose_class(len, opt, &kc)) {
                snprintf(msg, sizeof msg, "read %lld of length %d exceeds the largest kernel class", (long long)r, len);
                return fail(PHMM_ERR_TOO_LONG, msg);
            }
            read_class[r] = class_index(kc);
            if (len < 512) class_of_len[len] = read_class[r];
        }
    }
    const size_t ncls = P.classes.size();

    // ---- FP64-first reads (PlanOptions::spec) -----------------------------------------------------------
    // A read that the k-mer predictor (phmm_cost.cpp) expects to fail the FP32 pass against every haplotype of its
    // batch, by a factor of ten at least, skips that pass: its tasks (runs of up to 4 haplotypes / 2048 columns, as
    // the fast kernels hand rescued pairs over) are laid into the rescue lists here, and the FP64 kernel checks on the
    // rescaled sum what the FP32 pass would have decided (phmm_kernels.cuh: emit_result).
    const bool spec = opt.spec > 0 && opt.hp && opt.precision == 0 && opt.latency != 1 && opt.force_kh <= 0 && fb.n_reads > 0;
    std::vector<uint8_t> pred;
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) P.spec_cap[k] = 0;
    if (spec) {
        pred.assign((size_t)fb.n_reads, 0);
        std::vector<int64_t> cost_unused((size_t)std::max<int64_t>(fb.n_batches, 1));
        // the prediction is the longest part of planning a large chunk (2 us per read that sits on no diagonal), and
        // the chunk's upload waits for its plan: slices of batches on up to four threads
        const unsigned hw = std::thread::hardware_concurrency();
        const int nthr = (fb.n_reads >= 4000 && fb.n_batches >= 4) ? (hw >= 16 ? 4 : hw >= 8 ? 2 : 1) : 1;
        std::vector<int64_t> cut_b(1, 0), cut_r(1, 0), cut_h(1, 0);
        {
            int64_t r = 0, h = 0;
            for (int64_t b = 0; b < fb.n_batches; ++b) {
                r += fb.batch_num_reads[b];
                h += fb.batch_num_haps[b];
                if (b + 1 == fb.n_batches || r >= (int64_t)cut_b.size() * fb.n_reads / nthr) {
                    cut_b.push_back(b + 1); cut_r.push_back(r); cut_h.push_back(h);
                }
            }
        }
        std::vector<int> rcs(cut_b.size() - 1, PHMM_OK);
        auto slice = [&](size_t j) {
            phmm_flat_batches v = fb;
            v.n_batches = cut_b[j + 1] - cut_b[j];
            v.batch_num_reads = fb.batch_num_reads + cut_b[j];
            v.batch_num_haps = fb.batch_num_haps + cut_b[j];
            v.n_reads = cut_r[j + 1] - cut_r[j];
            v.read_len = fb.read_len + cut_r[j];
            v.read_off = fb.read_off + cut_r[j];
            v.n_haps = cut_h[j + 1] - cut_h[j];
            v.hap_len = fb.hap_len + cut_h[j];
            v.hap_off = fb.hap_off + cut_h[j];
            rcs[j] = batch_costs(v, 1, 0, default_cost_model(), cost_unused.data() + cut_b[j], pred.data() + cut_r[j], 1);
        };
        std::vector<std::thread> helpers;
        for (size_t j = 1; j + 1 < cut_b.size(); ++j) helpers.emplace_back(slice, j);
        slice(0);
        for (std::thread &t : helpers) t.join();
        for (int rc1 : rcs) if (rc1 != PHMM_OK) { pred.assign((size_t)fb.n_reads, 0); break; }
    }
    struct SpecRun { int k0, k1; int32_t cols; };
    std::vector<SpecRun> spec_runs;

    // ---- per batch: reads by class --------------------------------------------------------------
    // rl[b * ncls + c] = reads of batch b handled by class c.  A class's leftover reads (fewer
    // than one full task) move up to the next larger class of the batch, where they fill slots
    // instead of leaving part of a warp idle.
    struct BatchInfo { int64_t r0, h0, out0, cols; };
    std::vector<BatchInfo> binfo((size_t)fb.n_batches);
    // flat storage: the lists of batch b, class c are rl_idx[rl_off[b * ncls + c] .. rl_off[b * ncls + c + 1])
    // (one vector per batch and class made thousands of small allocations on 'small'-shaped input)
    std::vector<int32_t> rl_idx;
    std::vector<int64_t> rl_off((size_t)fb.n_batches * ncls + 1, 0);
    rl_idx.reserve((size_t)fb.n_reads);
    std::vector<std::vector<int32_t>> scratch(ncls);     // the lists of the batch at hand; capacity is kept
    struct ListView {
        const int32_t *p; size_t n;
        size_t size() const { return n; }
        bool empty() const { return n == 0; }
        int32_t operator[](size_t k) const { return p[k]; }
    };
    auto rl = [&](size_t k) { return ListView{rl_idx.data() + rl_off[k], (size_t)(rl_off[k + 1] - rl_off[k])}; };
    std::vector<int> order(ncls);
    for (size_t c = 0; c < ncls; ++c) order[c] = (int)c;
    std::sort(order.begin(), order.end(), [&](int x, int y) {
        const KClass &kx = P.classes[x].kc, &ky = P.classes[y].kc;
        if (kx.longr != ky.longr) return kx.longr < ky.longr;
        return kx.capacity() < ky.capacity();
    });
    {
        int64_t r0 = 0, h0 = 0, out0 = 0;
        for (int64_t b = 0; b < fb.n_batches; ++b) {
            const int bnr = fb.batch_num_reads[b], bnh = fb.batch_num_haps[b];
            int64_t sum_h = 0;
            for (int k = 0; k < bnh; ++k) sum_h += fb.hap_len[h0 + k];
            binfo[b] = BatchInfo{r0, h0, out0, sum_h + (int64_t)bnh * SEP_COLS};
            for (auto &v : scratch) v.clear();
            spec_runs.clear();
            std::vector<int32_t> *by_class = scratch.data();
            for (int k = 0; k < bnr; ++k) {
                const int64_t r = r0 + k;
                ReadRec &rr = P.reads[r];
                rr.off = fb.read_off[r] - P.segments[seg_of_read[r]].src + P.segments[seg_of_read[r]].dst;
                rr.len = fb.read_len[r];
                rr.batch = (int32_t)b;
                rr.out_base = out0 + (int64_t)k * bnh;
                P.n_cells += (int64_t)rr.len * sum_h;
                const int sc = P.resc_force > 0 ? P.resc_force : scalar_class_of(rr.len);
                P.resc_cap[sc] += bnh;
                if (spec && (pred[(size_t)r] & 2) && sc > 0 && bnh > 0) {
                    if (spec_runs.empty())
                        for (int j = 0; j < bnh;) {
                            int j1 = j;
                            int32_t cols = 0;
                            while (j1 < bnh && (j1 == j || (j1 - j < 4 && cols + fb.hap_len[h0 + j1] + SEP_COLS <= 2048))) {
                                cols += fb.hap_len[h0 + j1] + SEP_COLS;
                                ++j1;
                            }
                            spec_runs.push_back(SpecRun{j, j1, cols});
                            j = j1;
                        }
                    for (const SpecRun &sr : spec_runs) {
                        Task t;
                        t.stream_off = hap_col[h0 + sr.k0];
                        t.ncols = sr.cols;
                        t.hap0 = (int32_t)(h0 + sr.k0);
                        t.out0 = sr.k0;
                        t.read[0] = (int32_t)r; t.read[1] = SPEC_TASK; t.read[2] = -1; t.read[3] = -1;
                        t.nh = sr.k1 - sr.k0;
                        P.spec_tasks[sc].push_back(t);
                    }
                    P.resc_cap[sc] += (int64_t)spec_runs.size();      // the list also holds what the kernels append
                    P.spec_cap[sc] += bnh;
                    P.spec_pairs += bnh;
                    P.spec_cells += (int64_t)rr.len * sum_h;
                    continue;
                }
                by_class[read_class[r]].push_back((int32_t)r);
            }
            int prev = -1;
            for (size_t oi = 0; oi < ncls; ++oi) {
                const int c = order[oi];
                if (by_class[c].empty()) continue;
                if (prev >= 0 && P.classes[prev].kc.longr == P.classes[c].kc.longr) {
                    std::vector<int32_t> &src = by_class[prev], &dst = by_class[c];
                    const size_t rem = src.size() % (size_t)P.classes[prev].kc.group();
                    for (size_t j = 0; j < rem; ++j) { dst.push_back(src.back()); src.pop_back(); }
                }
                prev = c;
            }
            if (bnh > 0)
                for (size_t c = 0; c < ncls; ++c)
                    for (int32_t r : by_class[c]) {
                        P.classes[c].cells += (int64_t)P.reads[r].len * sum_h;
                        P.classes[c].pairs += bnh;
                    }
            for (size_t c = 0; c < ncls; ++c) {
                rl_idx.insert(rl_idx.end(), by_class[c].begin(), by_class[c].end());
                rl_off[(size_t)b * ncls + c + 1] = (int64_t)rl_idx.size();
            }
            P.n_pairs += (int64_t)bnr * bnh;
            out0 += (int64_t)bnr * bnh;
            r0 += bnr;
            h0 += bnh;
        }
    }

    // ---- runs of consecutive haplotypes -----------------------------------------------------------
    // A batch's haplotypes are cut into n contiguous runs of about `cols` columns each (n even for
    // the haplotype-pair classes, which take two runs per task), balanced so that the two runs of a
    // pair end within half a haplotype of each other.
    auto runs_of = [&](int64_t b, int cols, bool pair) -> int {
        const int bnh = fb.batch_num_haps[b];
        if (bnh == 0) return 0;
        int64_t n = (binfo[b].cols + cols - 1) / cols;
        n = std::max<int64_t>(n, (bnh + MAX_RUN_HAPS - 1) / MAX_RUN_HAPS);
        if (pair) n = (n + 1) & ~(int64_t)1;
        return (int)std::min<int64_t>(n, bnh);
    };
    auto count_tasks = [&](int cols) -> int64_t {
        int64_t n = 0;
        for (int64_t b = 0; b < fb.n_batches; ++b)
            for (size_t c = 0; c < ncls; ++c) {
                const KClass &kc = P.classes[c].kc;
                const int64_t nr = (int64_t)rl((size_t)b * ncls + c).size();
                if (!nr) continue;
                const int runs = runs_of(b, cols, kc.hp != 0);
                n += (kc.hp ? (runs + 1) / 2 : runs) * ((nr + kc.group() - 1) / kc.group());
            }
        return n;
    };
    // adaptive run length: aim at >= 3 tasks per resident warp, but keep runs long enough that
    // the wavefront fill (<= 31 steps) stays small
    int seg_cols = opt.seg_cols > 0 ? opt.seg_cols : 4096;
    if (opt.seg_cols <= 0) {
        while (seg_cols > 512 && count_tasks(seg_cols) < 3 * (int64_t)opt.resident_warps) seg_cols /= 2;
        // a few active regions per call: the longest run is the latency of the call, so the fill no longer
        // matters -- down to one haplotype per run
        while (opt.latency == 1 && seg_cols > 64 && count_tasks(seg_cols) < 3 * (int64_t)opt.resident_warps) seg_cols /= 2;
    }

    struct Run { int64_t b; int k0, k1; int32_t cols; };
    std::vector<Run> runs[2];                  // [0] single runs (classic classes), [1] run pairs: 2j, 2j + 1
    bool any_kind[2] = {false, false};
    for (size_t c = 0; c < ncls; ++c) any_kind[P.classes[c].kc.hp ? 1 : 0] = true;
    std::vector<int> cuts;
    for (int kind = 0; kind < 2; ++kind) {
        if (!any_kind[kind]) continue;
        for (int64_t b = 0; b < fb.n_batches; ++b) {
            const int bnh = fb.batch_num_haps[b];
            if (bnh == 0) continue;
            bool used = false;
            for (size_t c = 0; c < ncls && !used; ++c)
                used = (P.classes[c].kc.hp ? 1 : 0) == kind && !rl((size_t)b * ncls + c).empty();
            if (!used) continue;
            const int64_t h0 = binfo[b].h0, total = binfo[b].cols;
            int n = runs_of(b, seg_cols, kind == 1);
            for (;;) {
                cuts.assign(1, 0);
                int64_t acc = 0;
                int k = 0;
                bool ok = true;
                for (int j = 1; j <= n; ++j) {
                    const int64_t target = total * j / n;
                    const int kmax = bnh - (n - j);              // leave one haplotype per remaining run
                    int k1 = k;
                    while (k1 < kmax && (k1 == k || acc + (fb.hap_len[h0 + k1] + SEP_COLS) / 2 < target)) {
                        acc += fb.hap_len[h0 + k1] + SEP_COLS;
                        ++k1;
                    }
                    if (j == n) while (k1 < bnh) { acc += fb.hap_len[h0 + k1] + SEP_COLS; ++k1; }
                    if (k1 - k > MAX_RUN_HAPS) ok = false;
                    cuts.push_back(k1);
                    k = k1;
                }
                if (ok || n >= bnh) break;
                n = std::min(bnh, n + (kind == 1 ? 2 : 1));
            }
            const size_t first = runs[kind].size();
            for (size_t j = 0; j + 1 < cuts.size(); ++j) {
                int64_t w = 0;
                for (int k = cuts[j]; k < cuts[j + 1]; ++k) w += fb.hap_len[h0 + k] + SEP_COLS;
                runs[kind].push_back(Run{b, cuts[j], cuts[j + 1], (int32_t)w});
            }
            if (kind == 1) {
                // A task of the haplotype-pair kernels lasts as long as the longer of its two runs.  Runs are whole
                // haplotypes, so their lengths differ (by kilobases when the haplotypes are that long): pair the
                // runs of a batch in order of length, not of position -- the two runs of a task need not be adjacent
                // (C4: FP32 phase 37.5 -> 30.9 ms; C5 -6 %).
                std::stable_sort(runs[1].begin() + (std::ptrdiff_t)first, runs[1].end(),
                                 [](const Run &x, const Run &y) { return x.cols > y.cols; });
                if ((cuts.size() - 1) % 2 == 1) runs[1].push_back(Run{b, 0, 0, 0});   // idle second half (pairs with the shortest)
            }
        }
    }

    // ---- tasks, longest first: the device-side queue then behaves like LPT list scheduling ---------
    {
        // exact task counts per class, so that the task vectors are allocated once
        std::vector<size_t> n_of(ncls, 0);
        for (int kind = 0; kind < 2; ++kind)
            for (size_t j = 0; j < runs[kind].size(); j += (kind ? 2 : 1)) {
                const int64_t b = runs[kind][j].b;
                for (size_t c = 0; c < ncls; ++c) {
                    const KClass &kc = P.classes[c].kc;
                    if ((kc.hp ? 1 : 0) != kind) continue;
                    const size_t nr = rl((size_t)b * ncls + c).size();
                    n_of[c] += (nr + kc.group() - 1) / kc.group();
                }
            }
        for (size_t c = 0; c < ncls; ++c) {
            if (P.classes[c].kc.hp) P.classes[c].htasks.reserve(n_of[c]);
            else P.classes[c].tasks.reserve(n_of[c]);
        }
    }
    {
        std::vector<int32_t> idx(runs[0].size());
        for (size_t j = 0; j < idx.size(); ++j) idx[j] = (int32_t)j;
        std::stable_sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) { return runs[0][x].cols > runs[0][y].cols; });
        for (int32_t j : idx) {
            const Run &rn = runs[0][j];
            const BatchInfo &bi = binfo[rn.b];
            for (size_t c = 0; c < ncls; ++c) {
                const KClass &kc = P.classes[c].kc;
                if (kc.hp) continue;
                const ListView v = rl((size_t)rn.b * ncls + c);
                const int g = kc.group();
                for (size_t i = 0; i < v.size(); i += g) {
                    Task t;
                    t.stream_off = hap_col[bi.h0 + rn.k0];
                    t.ncols = rn.cols;
                    t.hap0 = (int32_t)(bi.h0 + rn.k0);
                    t.out0 = rn.k0;
                    for (int q = 0; q < MAX_GROUP; ++q) t.read[q] = (q < g && i + q < v.size()) ? v[i + q] : -1;
                    t.nh = rn.k1 - rn.k0;
                    P.classes[c].tasks.push_back(t);
                    if (kc.longr) P.long_max_cols = std::max(P.long_max_cols, t.ncols);
                }
            }
        }
    }
    {
        const size_t npairs = runs[1].size() / 2;
        std::vector<int32_t> idx(npairs);
        for (size_t j = 0; j < npairs; ++j) idx[j] = (int32_t)j;
        auto cost = [&](int32_t x) { return std::max(runs[1][2 * x].cols, runs[1][2 * x + 1].cols); };
        std::stable_sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) { return cost(x) > cost(y); });
        for (int32_t j : idx) {
            const Run *rn = &runs[1][2 * (size_t)j];
            const BatchInfo &bi = binfo[rn[0].b];
            for (size_t c = 0; c < ncls; ++c) {
                const KClass &kc = P.classes[c].kc;
                if (!kc.hp) continue;
                const ListView v = rl((size_t)rn[0].b * ncls + c);
                const int g = kc.group();
                for (size_t i = 0; i < v.size(); i += g) {
                    HTask t;
                    memset(&t, 0, sizeof t);
                    for (int h = 0; h < 2; ++h) {
                        const bool idle = rn[h].k1 == rn[h].k0;
                        t.stream_off[h] = hap_col[bi.h0 + (idle ? rn[0].k0 : rn[h].k0)];
                        t.ncols[h] = rn[h].cols;
                        t.hap0[h] = (int32_t)(bi.h0 + (idle ? rn[0].k0 : rn[h].k0));
                        t.out0[h] = idle ? rn[0].k0 : rn[h].k0;
                        t.nh[h] = (int16_t)(rn[h].k1 - rn[h].k0);
                    }
                    for (int q = 0; q < MAX_HP_GROUP; ++q) t.read[q] = (q < g && i + q < v.size()) ? v[i + q] : -1;
                    P.classes[c].htasks.push_back(t);
                }
            }
        }
    }
    // drop classes that ended up empty (possible when a batch has no haplotypes)
    P.classes.erase(std::remove_if(P.classes.begin(), P.classes.end(),
                                   [](const ClassPlan &c) { return c.ntasks() == 0; }),
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
