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
int layout_regions(phmm_gatk *g, const phmm_gatk_region *regions, int64_t n_regions, int64_t *npairs)
{
    phmm::ParsedBatches &b = g->b;
    int64_t rbytes = 0, hbytes = 0;
    *npairs = 0;
    for (int64_t k = 0; k < n_regions; ++k) {
        const phmm_gatk_region &rg = regions[k];
        if (rg.n_reads < 0 || rg.n_haps < 0 || (rg.n_reads && !rg.reads) || (rg.n_haps && !rg.haps)) {
            g->error = "negative count or NULL array";
            return PHMM_ERR_ARG;
        }
        b.batch_num_reads.push_back(rg.n_reads);
        b.batch_num_haps.push_back(rg.n_haps);
        for (int32_t r = 0; r < rg.n_reads; ++r) {
            const phmm_gatk_read &rd = rg.reads[r];
            if (rd.length <= 0) { g->error = "readBases of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (rd.length > PHMM_MAX_SEQ_LEN) { g->error = "read longer than PHMM_MAX_SEQ_LEN"; return PHMM_ERR_TOO_LONG; }
            if (!rd.readBases || !rd.readQuals || !rd.insertionGOP || !rd.deletionGOP || !rd.overallGCP) {
                g->error = "NULL field in ReadDataHolder";
                return PHMM_ERR_ARG;
            }
            b.read_len.push_back(rd.length);
            b.read_off.push_back(rbytes);
            rbytes += rd.length;
        }
        for (int32_t h = 0; h < rg.n_haps; ++h) {
            const phmm_gatk_hap &hp = rg.haps[h];
            if (hp.length <= 0) { g->error = "haplotypeBases of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (hp.length > PHMM_MAX_SEQ_LEN) { g->error = "haplotype longer than PHMM_MAX_SEQ_LEN"; return PHMM_ERR_TOO_LONG; }
            if (!hp.haplotypeBases) { g->error = "NULL haplotypeBases"; return PHMM_ERR_ARG; }
            b.hap_len.push_back(hp.length);
            b.hap_off.push_back(hbytes);
            hbytes += hp.length;
        }
        const int64_t n = (int64_t)rg.n_reads * rg.n_haps;
        if (n && !rg.likelihoods) { g->error = "NULL likelihoods"; return PHMM_ERR_ARG; }
        *npairs += n;
    }
    for (std::vector<uint8_t> *v : {&b.rs, &b.q, &b.i, &b.d, &b.c})
        if (v->size() < (size_t)rbytes) v->resize((size_t)rbytes + (size_t)rbytes / 4);
    if (b.hap.size() < (size_t)hbytes) b.hap.resize((size_t)hbytes + (size_t)hbytes / 4);
    g->read_bytes = rbytes;
    return PHMM_OK;
}

void copy_pool(phmm_gatk *g, const phmm_gatk_region *regions, int64_t n_regions, int which)
{
    phmm::ParsedBatches &b = g->b;
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
    int rc0 = layout_regions(g, regions, n_regions, &npairs);
    if (rc0 != PHMM_OK) return rc0;
    if (g->read_bytes > (1 << 20)) {
        // many regions: the staging copy is memory-bound host work, one thread per pool
        std::thread th[5];
        for (int w = 0; w < 5; ++w) th[w] = std::thread(copy_pool, g, regions, n_regions, w + 1);
        copy_pool(g, regions, n_regions, 0);
        for (int w = 0; w < 5; ++w) th[w].join();
    } else {
        for (int w = 0; w < 6; ++w) copy_pool(g, regions, n_regions, w);
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
    phmm_destroy(g->ctx);
    delete g;
    return PHMM_OK;
}
