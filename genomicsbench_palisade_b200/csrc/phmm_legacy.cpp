// The four C++-linkage entry points (and one data symbol) of libgkl_pairhmm_c, so that the
// reference driver object links against libphmm_cuda.so unchanged.
//
//   void initPairHMM()                                  IntelPairHmmCSource.cpp:29-51
//   void computelikelihoodsboth(testcase*, double*, int)                    :61-85
//   void computelikelihoodsfloat(testcase*, float*)                         :89-99
//   void computelikelihoodsdouble(testcase*, double*)                       :103-114
//   uint8_t ConvertChar::conversionTable[255]           pairhmm_common.cc:3 (the driver calls the
//                                                       inline ConvertChar::init(), _orig.cpp:240)
// (paths under /root/reference/tools/GKL/src/main/native/pairhmm/).  The driver re-declares the
// functions itself (benchmarks/phmm/PairHMMUnitTest_orig.cpp:84-86), there is no header.
//
// Semantics kept: blocking, out[k] in input order, callable concurrently from the driver's
// OpenMP threads (the calls of all threads are coalesced by one dispatcher, see Coalescer below;
// PHMM_LEGACY_COALESCE=0: each thread gets its own context/stream on first use).  Not kept: the fork's
// per-pair printf (:80) and the "AVX512 supported!" banner.  The reference has no error channel;
// failures abort() with a message.
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <stdint.h>
#include <string>
#include <thread>
#include <vector>

#include "../../include/phmm_cuda.h"

// Same layout and name as the reference's struct so that the mangled names match.
typedef struct {
    int rslen, haplen;
    const char *q, *i, *d, *c;
    const char *hap, *rs;
} testcase;

class ConvertChar {
  public:
    static uint8_t conversionTable[255];
};
uint8_t ConvertChar::conversionTable[255];

namespace {

int legacy_device()
{
    const char *s = getenv("PHMM_DEVICE");
    return s ? atoi(s) : 0;
}

struct ThreadCtx {
    phmm_ctx *both = nullptr, *f32 = nullptr, *f64 = nullptr;
    ~ThreadCtx() { phmm_destroy(both); phmm_destroy(f32); phmm_destroy(f64); }
};
thread_local ThreadCtx g_tls;

phmm_ctx *get_ctx(phmm_ctx **slot, int precision)
{
    if (!*slot) {
        int rc = phmm_create(legacy_device(), slot);
        if (rc != PHMM_OK) {
            fprintf(stderr, "libphmm_cuda: cannot create context: %s\n", phmm_last_error(nullptr));
            abort();
        }
        phmm_set_option(*slot, PHMM_OPT_PRECISION, precision);
        if (const char *s = getenv("PHMM_FAST")) {
            phmm_set_option(*slot, PHMM_OPT_FP32_KERNEL, atoi(s));
            phmm_set_option(*slot, PHMM_OPT_FP64_KERNEL, atoi(s));
        }
    }
    return *slot;
}

void check(phmm_ctx *ctx, int rc, const char *what)
{
    if (rc != PHMM_OK) {
        fprintf(stderr, "libphmm_cuda: %s failed (%d): %s\n", what, rc, phmm_last_error(ctx));
        abort();
    }
}

// ---- coalescing front end ---------------------------------------------------------------------------
// The benchmark driver calls computelikelihoodsboth once per batch from every OpenMP thread
// (PairHMMUnitTest_orig.cpp:271-299): at most 50 000 pairs per call, a few hundred on 'small'-shaped input, far
// too little to fill a B200 and mostly call latency.  All calls of the process therefore go through one
// dispatcher thread with one context: whatever requests are pending when a pipeline slot is free are merged into
// one planned chunk (phmm_submit_pairs), up to four chunks are in flight, and each caller sleeps until the chunk
// holding its batch has been retired.  Under light load a call is submitted at once; under load the chunks grow
// by themselves, because requests pile up while the dispatcher waits for a slot.  PHMM_LEGACY_COALESCE=0 goes back
// to one context per calling thread.
class Coalescer {
  public:
    struct Request { const phmm_flat_batches *flat; double *out; int n; bool done = false; int rc = PHMM_OK; };

    static Coalescer &get()
    {
        static Coalescer c;
        return c;
    }
    static bool enabled()
    {
        static const bool on = !(getenv("PHMM_LEGACY_COALESCE") && atoi(getenv("PHMM_LEGACY_COALESCE")) == 0);
        return on;
    }

    void run(testcase *tc, double *out, int n)
    {
        // the per-pair work (folding the testcase slab back into reads x haplotypes, copying the strings) is done
        // here, by the calling thread: sixteen callers do it in parallel, the dispatcher only merges and launches
        phmm_parsed *flat = nullptr;
        char err[160];
        const int frc = phmm_flatten_pairs(reinterpret_cast<const phmm_testcase *>(tc), n, &flat, err, sizeof err);
        if (frc != PHMM_OK) {
            fprintf(stderr, "libphmm_cuda: computelikelihoodsboth failed (%d): %s\n", frc, err);
            abort();
        }
        struct Free { phmm_parsed *p; ~Free() { phmm_parsed_free(p); } } free_flat{flat};
        Request r{phmm_parsed_batches(flat), out, n};
        std::unique_lock<std::mutex> lk(mu_);
        if (!started_) { worker_ = std::thread([this]() { loop(); }); started_ = true; }
        pending_.push_back(&r);
        cv_.notify_all();
        done_cv_.wait(lk, [&]() { return r.done; });
        if (r.rc != PHMM_OK) {
            fprintf(stderr, "libphmm_cuda: computelikelihoodsboth failed (%d): %s\n", r.rc, error_.c_str());
            abort();
        }
    }

    ~Coalescer()
    {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
        cv_.notify_all();
        if (worker_.joinable()) worker_.join();
        phmm_destroy(ctx_);
    }

  private:
    struct Flight { int64_t ticket; std::vector<Request *> reqs; };
    static const int MAX_FLIGHTS = 4;                  // = pipeline slots of a context
    static const int64_t MAX_PAIRS = 1500000;          // per chunk (device memory of a slot)

    void finish(std::vector<Request *> &reqs, int rc)
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (rc != PHMM_OK) error_ = phmm_last_error(ctx_);
        for (Request *r : reqs) { r->rc = rc; r->done = true; }
        done_cv_.notify_all();
    }

    void loop()
    {
        ctx_ = nullptr;
        if (phmm_create(legacy_device(), &ctx_) != PHMM_OK) {
            fprintf(stderr, "libphmm_cuda: cannot create context: %s\n", phmm_last_error(nullptr));
            abort();
        }
        if (const char *s = getenv("PHMM_FAST")) {
            phmm_set_option(ctx_, PHMM_OPT_FP32_KERNEL, atoi(s));
            phmm_set_option(ctx_, PHMM_OPT_FP64_KERNEL, atoi(s));
        }
        std::deque<Flight> flights;
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            cv_.wait(lk, [&]() { return stop_ || !pending_.empty() || !flights.empty(); });
            if (pending_.empty() && flights.empty()) { if (stop_) return; continue; }
            if (!pending_.empty() && (int)flights.size() < MAX_FLIGHTS) {
                Flight f;
                int64_t pairs = 0;
                while (!pending_.empty() && (f.reqs.empty() || pairs + pending_.front()->n <= MAX_PAIRS)) {
                    pairs += pending_.front()->n;
                    f.reqs.push_back(pending_.front());
                    pending_.pop_front();
                }
                lk.unlock();
                std::vector<phmm_flat_request> rq;
                for (Request *r : f.reqs) rq.push_back({r->flat, r->out});
                const int rc = phmm_submit_flat(ctx_, rq.data(), (int)rq.size(), &f.ticket);
                if (rc != PHMM_OK) finish(f.reqs, rc);
                lk.lock();
                if (rc == PHMM_OK) flights.push_back(std::move(f));
                continue;
            }
            // nothing new to launch, or every slot is taken: retire the oldest chunk and wake its callers
            Flight f = std::move(flights.front());
            flights.pop_front();
            lk.unlock();
            const int rc = phmm_wait(ctx_, f.ticket, nullptr);
            finish(f.reqs, rc);
            lk.lock();
        }
    }

    std::mutex mu_;
    std::condition_variable cv_, done_cv_;
    std::deque<Request *> pending_;
    std::thread worker_;
    phmm_ctx *ctx_ = nullptr;
    std::string error_;
    bool started_ = false, stop_ = false;
};

}  // namespace

void initPairHMM()
{
    // table generation and device binding happen on first use (per thread, or in the dispatcher); fail early here
    if (phmm_device_count() <= 0) {
        fprintf(stderr, "libphmm_cuda: no CUDA device (there is no CPU fallback)\n");
        abort();
    }
    if (!Coalescer::enabled()) get_ctx(&g_tls.both, 0);
    ConvertChar::conversionTable['A'] = 0;
    ConvertChar::conversionTable['C'] = 1;
    ConvertChar::conversionTable['T'] = 2;
    ConvertChar::conversionTable['G'] = 3;
    ConvertChar::conversionTable['N'] = 4;
}

void computelikelihoodsboth(testcase *testcases, double *expected_results, int batch_size)
{
    static_assert(sizeof(testcase) == sizeof(phmm_testcase), "testcase layout");
    if (batch_size <= 0) return;
    if (Coalescer::enabled()) { Coalescer::get().run(testcases, expected_results, batch_size); return; }
    phmm_ctx *ctx = get_ctx(&g_tls.both, 0);
    check(ctx, phmm_compute_pairs(ctx, reinterpret_cast<const phmm_testcase *>(testcases), expected_results, batch_size),
          "computelikelihoodsboth");
}

void computelikelihoodsfloat(testcase *testcases, float *expected_results)
{
    phmm_ctx *ctx = get_ctx(&g_tls.f32, 1);
    double r = 0;
    check(ctx, phmm_compute_pairs(ctx, reinterpret_cast<const phmm_testcase *>(testcases), &r, 1), "computelikelihoodsfloat");
    *expected_results = (float)r;
}

void computelikelihoodsdouble(testcase *testcases, double *expected_results)
{
    phmm_ctx *ctx = get_ctx(&g_tls.f64, 2);
    check(ctx, phmm_compute_pairs(ctx, reinterpret_cast<const phmm_testcase *>(testcases), expected_results, 1),
          "computelikelihoodsdouble");
}
