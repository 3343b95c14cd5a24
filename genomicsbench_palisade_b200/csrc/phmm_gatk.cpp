// GATK native-bindings shaped batch API on top of the context.
//
// Mirrors IntelPairHmm.initNative / computeLikelihoodsNative / doneNative
// (tools/GKL/src/main/native/pairhmm/IntelPairHmm.cc:32-96,103-156,164-168): one call = all reads x
// all haplotypes of an active region, results read-major (JavaData.h:44-68), FP32 with FP64 rescue
// or -- with use_double -- FP64 for every pair (:123-124).  Host-only translation unit: the work
// is done by phmm_compute_flat.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "phmm_host.h"

struct phmm_gatk {
    phmm_ctx *ctx = nullptr;
    int use_double = 0, max_threads = 1;
    std::string error;
    // scratch reused across calls (GATK calls once per active region, tens of thousands of times)
    phmm::ParsedBatches b;
    int64_t read_bytes = 0;          // extent of the read pools in use by the current call
    std::vector<double> out;
    // large calls (many regions): groups of regions staged by helper threads while the previous groups compute
    // A slot's pools are page-locked once an allocation has survived a call (GATK calls thousands of times with
    // regions of similar size): the uploads of a group are then DMA the calling thread only queues (pageable copies
    // cost it ~1 ms per 6 MB group, as long as the group's kernels), and a slot is reused only after its group retired.
    static const int N_STAGE = 6;
    struct PoolPin { void *ptr = nullptr; size_t bytes = 0; bool registered = false; int64_t alloc_call = 0; };
    struct StageSlot { phmm::ParsedBatches b; PoolPin pin[6]; };
    StageSlot stage[N_STAGE];
    int64_t n_calls = 0;
    int device = 0;
};

namespace {
thread_local std::string g_gatk_create_error;

void clear(phmm::ParsedBatches &b)
{
    b.batch_num_reads.clear(); b.batch_num_haps.clear();
    b.read_len.clear(); b.read_off.clear(); b.hap_len.clear(); b.hap_off.clear();
    // the byte pools keep their size: they are overwritten up to the new extent, and growing a vector
    // value-initialises (memsets) the new part -- a second pass over tens of megabytes per call
}

// Validates the regions and lays out the tables; the bytes are copied afterwards, one host thread per pool.
int layout_regions(phmm::ParsedBatches &b, std::string *error, const phmm_gatk_region *regions, int64_t n_regions,
                   int64_t *npairs, int64_t *read_bytes, phmm_gatk::PoolPin *pins = nullptr, int64_t call = 0)
{
    int64_t rbytes = 0, hbytes = 0;
    *npairs = 0;
    for (int64_t k = 0; k < n_regions; ++k) {
        const phmm_gatk_region &rg = regions[k];
        if (rg.n_reads < 0 || rg.n_haps < 0 || (rg.n_reads && !rg.reads) || (rg.n_haps && !rg.haps)) {
            *error = "negative count or NULL array";
            return PHMM_ERR_ARG;
        }
        b.batch_num_reads.push_back(rg.n_reads);
        b.batch_num_haps.push_back(rg.n_haps);
        for (int32_t r = 0; r < rg.n_reads; ++r) {
            const phmm_gatk_read &rd = rg.reads[r];
            if (rd.length <= 0) { *error = "readBases of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (rd.length > PHMM_MAX_SEQ_LEN) { *error = "read longer than PHMM_MAX_SEQ_LEN"; return PHMM_ERR_TOO_LONG; }
            if (!rd.readBases || !rd.readQuals || !rd.insertionGOP || !rd.deletionGOP || !rd.overallGCP) {
                *error = "NULL field in ReadDataHolder";
                return PHMM_ERR_ARG;
            }
            b.read_len.push_back(rd.length);
            b.read_off.push_back(rbytes);
            rbytes += rd.length;
        }
        for (int32_t h = 0; h < rg.n_haps; ++h) {
            const phmm_gatk_hap &hp = rg.haps[h];
            if (hp.length <= 0) { *error = "haplotypeBases of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (hp.length > PHMM_MAX_SEQ_LEN) { *error = "haplotype longer than PHMM_MAX_SEQ_LEN"; return PHMM_ERR_TOO_LONG; }
            if (!hp.haplotypeBases) { *error = "NULL haplotypeBases"; return PHMM_ERR_ARG; }
            b.hap_len.push_back(hp.length);
            b.hap_off.push_back(hbytes);
            hbytes += hp.length;
        }
        const int64_t n = (int64_t)rg.n_reads * rg.n_haps;
        if (n && !rg.likelihoods) { *error = "NULL likelihoods"; return PHMM_ERR_ARG; }
        *npairs += n;
    }
    std::vector<uint8_t> *pools[6] = {&b.rs, &b.q, &b.i, &b.d, &b.c, &b.hap};
    for (int k = 0; k < 6; ++k) {
        std::vector<uint8_t> *v = pools[k];
        const size_t need = (size_t)(k < 5 ? rbytes : hbytes);
        if (v->size() >= need) continue;
        // a page-locked allocation is released before the vector lets go of it
        if (pins && pins[k].registered) { cudaHostUnregister(pins[k].ptr); pins[k].registered = false; }
        v->resize(need + need / 4);
        if (pins) { pins[k].ptr = v->data(); pins[k].bytes = v->size(); pins[k].alloc_call = call; }
    }
    *read_bytes = rbytes;
    return PHMM_OK;
}

void copy_pool(phmm::ParsedBatches &b, const phmm_gatk_region *regions, int64_t n_regions, int which)
{
    uint8_t *dst = which == 0 ? b.rs.data() : which == 1 ? b.q.data() : which == 2 ? b.i.data()
                 : which == 3 ? b.d.data() : which == 4 ? b.c.data() : b.hap.data();
    for (int64_t k = 0; k < n_regions; ++k) {
        const phmm_gatk_region &rg = regions[k];
        if (which == 5) {
            for (int32_t h = 0; h < rg.n_haps; ++h) { memcpy(dst, rg.haps[h].haplotypeBases, (size_t)rg.haps[h].length); dst += rg.haps[h].length; }
            continue;
        }
        for (int32_t r = 0; r < rg.n_reads; ++r) {
            const phmm_gatk_read &rd = rg.reads[r];
            const uint8_t *src = which == 0 ? rd.readBases : which == 1 ? rd.readQuals : which == 2 ? rd.insertionGOP
                               : which == 3 ? rd.deletionGOP : rd.overallGCP;
            memcpy(dst, src, (size_t)rd.length);
            dst += rd.length;
        }
    }
}

// Large calls (hundreds of regions): laying the holders out into pools is memory-bound host work of the same order
// as the kernels (62 MB for 1024 regions: 3 ms on six threads, against ~9 ms of kernels).  The regions are cut into
// groups by cells (a small first group, so the device starts early), six helper threads (one per pool; the first
// also validates and lays out the tables) stage group k + 1 .. k + 2 while the calling thread plans, uploads and
// launches group k through the asynchronous entry (phmm_submit_flat, up to three groups in flight), and a group's
// results are handed to its regions when it retires.  A seventh thread plans a group as soon as it is staged.  A
// staging buffer is free again when its group has retired (its uploads may be asynchronous DMA out of page-locked
// pools, see phmm_gatk::PoolPin).  Returns -1 when the call is too small to be worth cutting.
int compute_regions_pipelined(phmm_gatk *g, const phmm_gatk_region *regions, int64_t n_regions)
{
    static const bool trace = getenv("PHMM_TRACE") != nullptr;
    static const bool off = getenv("PHMM_GATK_PIPELINE") && atoi(getenv("PHMM_GATK_PIPELINE")) == 0;
    if (off) return -1;
    std::vector<int64_t> gstart;
    int64_t total_pairs = 0;
    {
        double acc = 0, total = 0, target = 1e9;
        std::vector<double> cells((size_t)n_regions);
        for (int64_t k = 0; k < n_regions; ++k) {
            const phmm_gatk_region &rg = regions[k];
            // malformed input is reported by the one-shot path, which validates everything
            if (rg.n_reads < 0 || rg.n_haps < 0 || (rg.n_reads && !rg.reads) || (rg.n_haps && !rg.haps)) return -1;
            double sr = 0, sh = 0;
            for (int32_t r = 0; r < rg.n_reads; ++r) sr += std::max(rg.reads[r].length, 0);
            for (int32_t h = 0; h < rg.n_haps; ++h) sh += std::max(rg.haps[h].length, 0);
            cells[(size_t)k] = sr * sh;
            total += sr * sh;
            total_pairs += (int64_t)rg.n_reads * rg.n_haps;
        }
        if (total < 6e9) return -1;
        gstart.push_back(0);
        for (int64_t k = 0; k < n_regions; ++k) {
            acc += cells[(size_t)k];
            if (acc >= target && k + 1 < n_regions) { gstart.push_back(k + 1); acc = 0; target = 3e9; }
        }
        gstart.push_back(n_regions);
    }
    const int64_t ngroups = (int64_t)gstart.size() - 1;
    if (ngroups < 3) return -1;
    struct timespec ts0, ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts0);
    g->out.resize((size_t)total_pairs);

    struct Group {
        int64_t npairs = 0, read_bytes = 0, ticket = -1, out_off = 0;
        int rc = PHMM_OK, done = 0;
        bool laid = false, planned = false;
        std::string err;
        phmm::Plan plan;
    };
    std::vector<Group> gs((size_t)ngroups);
    std::mutex mu;
    std::condition_variable cv;
    int64_t n_retired = 0;          // groups whose results are back: slot k % NS of group k is free once k - NS has retired
    bool stop = false;
    static const bool pin_off = getenv("PHMM_GATK_PIN") && atoi(getenv("PHMM_GATK_PIN")) == 0;
    const int64_t call = ++g->n_calls;
    const int NS = phmm_gatk::N_STAGE;
    auto worker = [&](int w) {
        if (w == 0) cudaSetDevice(g->device);
        for (int64_t k = 0; k < ngroups; ++k) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&]() { return stop || k < n_retired + NS; });
                if (stop) return;
            }
            phmm::ParsedBatches &b = g->stage[k % NS].b;
            const phmm_gatk_region *rg = regions + gstart[(size_t)k];
            const int64_t n = gstart[(size_t)k + 1] - gstart[(size_t)k];
            Group &G = gs[(size_t)k];
            if (w == 0) {
                clear(b);
                phmm_gatk::PoolPin *pins = g->stage[k % NS].pin;
                const int rc = layout_regions(b, &G.err, rg, n, &G.npairs, &G.read_bytes, pins, call);
                if (rc == PHMM_OK && !pin_off)
                    for (int q = 0; q < 6; ++q)
                        if (!pins[q].registered && pins[q].ptr && pins[q].alloc_call < call) {
                            // the allocation served an earlier call unchanged: page-lock it
                            if (cudaHostRegister(pins[q].ptr, pins[q].bytes, cudaHostRegisterPortable) == cudaSuccess) pins[q].registered = true;
                            else cudaGetLastError();
                        }
                std::lock_guard<std::mutex> lk(mu);
                G.rc = rc; G.laid = true;
                if (rc != PHMM_OK) stop = true;
                cv.notify_all();
                if (rc != PHMM_OK) return;
            } else {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&]() { return stop || G.laid; });
                if (stop) return;
            }
            copy_pool(b, rg, n, w);
            std::lock_guard<std::mutex> lk(mu);
            G.done += 1;
            cv.notify_all();
        }
    };
    // a seventh thread plans a group as soon as its six pools are in place (planning a group of small regions takes a
    // third of its kernels' time: left to the calling thread, next to the uploads and launches, it bounded the call)
    const phmm::PlanOptions popt0 = phmm::context_plan_options(g->ctx);
    auto planner = [&]() {
        for (int64_t k = 0; k < ngroups; ++k) {
            Group &G = gs[(size_t)k];
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&]() { return stop || G.done == 6; });
                if (G.done != 6) return;
            }
            int rc = PHMM_OK;
            if (G.npairs > 0) {
                phmm::ParsedBatches &b = g->stage[k % NS].b;
                b.finish();
                phmm::PlanOptions popt = popt0;
                popt.latency = phmm::plan_latency(b.view, popt);
                popt.one_class = phmm::plan_one_class(b.view, popt);
                popt.spec = phmm::plan_spec(b.view, popt);
                rc = phmm::build_plan(b.view, popt, &G.plan, &G.err);
            }
            std::lock_guard<std::mutex> lk(mu);
            if (rc != PHMM_OK) { G.rc = rc; stop = true; }
            G.planned = true;
            cv.notify_all();
            if (rc != PHMM_OK) return;
        }
    };
    std::thread th[7];
    for (int w = 0; w < 6; ++w) th[w] = std::thread(worker, w);
    th[6] = std::thread(planner);
    auto finish_threads = [&]() {
        { std::lock_guard<std::mutex> lk(mu); stop = true; }
        cv.notify_all();
        for (int w = 0; w < 7; ++w) if (th[w].joinable()) th[w].join();
    };
    auto retire = [&](int64_t k) -> int {
        Group &G = gs[(size_t)k];
        if (G.ticket < 0) return PHMM_OK;
        const int rc = phmm_wait(g->ctx, G.ticket, nullptr);
        if (rc != PHMM_OK) { g->error = phmm_last_error(g->ctx); return rc; }
        const double *src = g->out.data() + G.out_off;
        for (int64_t r = gstart[(size_t)k]; r < gstart[(size_t)k + 1]; ++r) {
            const int64_t n = (int64_t)regions[r].n_reads * regions[r].n_haps;
            if (n) memcpy(regions[r].likelihoods, src, (size_t)n * sizeof(double));
            src += n;
        }
        return PHMM_OK;
    };
    int rc = PHMM_OK;
    int64_t out_off = 0, last_ticket_group = -1, retired = 0;
    auto now_ms = [&]() {
        struct timespec t;
        clock_gettime(CLOCK_MONOTONIC, &t);
        return (t.tv_sec - ts0.tv_sec) * 1e3 + (t.tv_nsec - ts0.tv_nsec) * 1e-6;
    };
    for (int64_t k = 0; k < ngroups && rc == PHMM_OK; ++k) {
        Group &G = gs[(size_t)k];
        const double t_wait0 = trace ? now_ms() : 0;
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&]() { return stop || G.planned; });
            if (!G.planned || G.rc != PHMM_OK) {   // a group failed validation or planning (this one or one staged ahead)
                for (const Group &x : gs) if (x.rc != PHMM_OK) { rc = x.rc; g->error = x.err; break; }
                if (rc == PHMM_OK) { rc = PHMM_ERR_ARG; g->error = "staging stopped"; }
                break;
            }
        }
        // at most three groups in flight: the slot this one takes has been retired before
        const double t_ready = trace ? now_ms() : 0;
        while (rc == PHMM_OK && k - retired >= 3) {
            rc = retire(retired++);
            { std::lock_guard<std::mutex> lk(mu); n_retired = retired; }
            cv.notify_all();
        }
        if (rc != PHMM_OK) break;
        const double t_retired = trace ? now_ms() : 0;
        G.out_off = out_off;
        out_off += G.npairs;
        if (G.npairs > 0) {
            phmm::ParsedBatches &b = g->stage[k % NS].b;
            rc = phmm::submit_planned(g->ctx, &b.view, &G.plan, g->out.data() + G.out_off, &G.ticket);
            if (rc != PHMM_OK) { g->error = phmm_last_error(g->ctx); G.ticket = -1; break; }
            last_ticket_group = k;
        }
        if (trace)
            fprintf(stderr, "phmm trace: gatk group %lld/%lld regions %lld pairs %lld: waited %.2f ms for its staging + plan, %.2f ms "
                    "retiring, submit %.2f ms, t=%.2f\n", (long long)k, (long long)ngroups,
                    (long long)(gstart[(size_t)k + 1] - gstart[(size_t)k]), (long long)G.npairs, t_ready - t_wait0,
                    t_retired - t_ready, now_ms() - t_retired, now_ms());
    }
    finish_threads();
    if (rc == PHMM_OK) {
        while (rc == PHMM_OK && retired < ngroups) rc = retire(retired++);
    }
    if (rc != PHMM_OK && last_ticket_group >= 0) {
        // leave the context drained: wait for what is still in flight, results are dropped
        std::string keep = g->error;
        phmm_wait(g->ctx, gs[(size_t)last_ticket_group].ticket, nullptr);
        g->error = keep;
    }
    if (trace) {
        clock_gettime(CLOCK_MONOTONIC, &ts1);
        fprintf(stderr, "phmm trace: gatk pipelined call: %lld regions in %lld groups, %.2f ms\n", (long long)n_regions,
                (long long)ngroups, (ts1.tv_sec - ts0.tv_sec) * 1e3 + (ts1.tv_nsec - ts0.tv_nsec) * 1e-6);
    }
    return rc;
}
}  // namespace

extern "C" const char *phmm_gatk_last_error(const phmm_gatk *g)
{
    return g ? g->error.c_str() : g_gatk_create_error.c_str();
}

extern "C" phmm_ctx *phmm_gatk_context(phmm_gatk *g) { return g ? g->ctx : nullptr; }

extern "C" int phmm_gatk_init(int device, int use_double, int max_threads, int use_fpga, phmm_gatk **out)
{
    if (!out) return PHMM_ERR_ARG;
    *out = nullptr;
    if (use_fpga) {
        // the reference only consults the FPGA when shacc_pairhmm::calculate succeeds, which the
        // shipped stub never does (shacc_pairhmm.h); there is nothing to mirror
        g_gatk_create_error = "use_fpga is not supported";
        return PHMM_ERR_ARG;
    }
    phmm_ctx *ctx = nullptr;
    int rc = phmm_create(device, &ctx);
    if (rc != PHMM_OK) { g_gatk_create_error = phmm_last_error(nullptr); return rc; }
    // computeLikelihoodsNative's policy: FP32 then FP64 below MIN_ACCEPTED, or FP64 only
    phmm_set_option(ctx, PHMM_OPT_PRECISION, use_double ? 2 : 0);
    // Bit-exact kernels by default: a variant caller expects a pair's likelihood not to depend on which other
    // regions share the call, and the low bits of the fast FP32 kernels depend on the kernel class a read runs in,
    // which the planner picks by the size of the call.  PHMM_GATK_FAST=1, or phmm_set_option(phmm_gatk_context(g),
    // PHMM_OPT_FP32_KERNEL / PHMM_OPT_FP64_KERNEL, 1), opts into the fast ones (same rescue set, |rel err| ~1e-6).
    const char *fast_env = getenv("PHMM_GATK_FAST");
    const int fast = (fast_env && atoi(fast_env) != 0) ? 1 : 0;
    phmm_set_option(ctx, PHMM_OPT_FP32_KERNEL, fast);
    phmm_set_option(ctx, PHMM_OPT_FP64_KERNEL, fast);
    phmm_gatk *g = new phmm_gatk();
    g->ctx = ctx;
    g->device = device;
    g->use_double = use_double ? 1 : 0;
    g->max_threads = max_threads;
    *out = g;
    return PHMM_OK;
}

extern "C" int phmm_gatk_compute_regions(phmm_gatk *g, const phmm_gatk_region *regions, int64_t n_regions)
{
    if (!g || n_regions < 0 || (n_regions && !regions)) return PHMM_ERR_ARG;
    static const bool trace = getenv("PHMM_TRACE") != nullptr;
    struct timespec ts0, ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts0);
    clear(g->b);
    int64_t npairs = 0;
    if (n_regions >= 32) {
        int rcp = compute_regions_pipelined(g, regions, n_regions);
        if (rcp >= 0) return rcp;              // < 0: the call is too small to be worth cutting
    }
    int rc0 = layout_regions(g->b, &g->error, regions, n_regions, &npairs, &g->read_bytes);
    if (rc0 != PHMM_OK) return rc0;
    if (g->read_bytes > (1 << 20)) {
        // many regions: the staging copy is memory-bound host work, one thread per pool
        std::thread th[5];
        for (int w = 0; w < 5; ++w) th[w] = std::thread(copy_pool, std::ref(g->b), regions, n_regions, w + 1);
        copy_pool(g->b, regions, n_regions, 0);
        for (int w = 0; w < 5; ++w) th[w].join();
    } else {
        for (int w = 0; w < 6; ++w) copy_pool(g->b, regions, n_regions, w);
    }
    if (npairs == 0) return PHMM_OK;
    g->b.finish();
    if (trace) {
        clock_gettime(CLOCK_MONOTONIC, &ts1);
        fprintf(stderr, "phmm trace: gatk staging of %lld regions, %zu read bytes x 5: %.2f ms\n", (long long)n_regions,
                (size_t)g->read_bytes, (ts1.tv_sec - ts0.tv_sec) * 1e3 + (ts1.tv_nsec - ts0.tv_nsec) * 1e-6);
    }
    double *dst = g->out.data();
    if (n_regions == 1) {
        dst = regions[0].likelihoods;          // straight into the caller's array
    } else {
        g->out.resize((size_t)npairs);
        dst = g->out.data();
    }
    int rc = phmm_compute_flat(g->ctx, &g->b.view, dst, nullptr);
    if (rc != PHMM_OK) { g->error = phmm_last_error(g->ctx); return rc; }
    if (n_regions > 1) {
        int64_t p = 0;
        for (int64_t k = 0; k < n_regions; ++k) {
            const int64_t n = (int64_t)regions[k].n_reads * regions[k].n_haps;
            if (n) memcpy(regions[k].likelihoods, dst + p, (size_t)n * sizeof(double));
            p += n;
        }
    }
    return PHMM_OK;
}

extern "C" int phmm_gatk_compute_likelihoods(phmm_gatk *g, const phmm_gatk_read *reads, int32_t n_reads,
                                             const phmm_gatk_hap *haps, int32_t n_haps, double *likelihoods)
{
    phmm_gatk_region rg;
    rg.reads = reads; rg.haps = haps; rg.likelihoods = likelihoods;
    rg.n_reads = n_reads; rg.n_haps = n_haps;
    return phmm_gatk_compute_regions(g, &rg, 1);
}

extern "C" int phmm_gatk_done(phmm_gatk *g)
{
    if (!g) return PHMM_OK;
    for (phmm_gatk::StageSlot &sl : g->stage)
        for (phmm_gatk::PoolPin &pp : sl.pin)
            if (pp.registered) { cudaHostUnregister(pp.ptr); pp.registered = false; }
    phmm_destroy(g->ctx);
    delete g;
    return PHMM_OK;
}
