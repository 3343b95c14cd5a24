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
// OpenMP threads (each thread gets its own context/stream on first use).  Not kept: the fork's
// per-pair printf (:80) and the "AVX512 supported!" banner.  The reference has no error channel;
// failures abort() with a message.
#include <cstdio>
#include <cstdlib>
#include <stdint.h>

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
        if (const char *s = getenv("PHMM_FAST")) phmm_set_option(*slot, PHMM_OPT_FP32_KERNEL, atoi(s));
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

}  // namespace

void initPairHMM()
{
    // table generation and device binding happen on first use per thread; fail early here
    get_ctx(&g_tls.both, 0);
    ConvertChar::conversionTable['A'] = 0;
    ConvertChar::conversionTable['C'] = 1;
    ConvertChar::conversionTable['T'] = 2;
    ConvertChar::conversionTable['G'] = 3;
    ConvertChar::conversionTable['N'] = 4;
}

void computelikelihoodsboth(testcase *testcases, double *expected_results, int batch_size)
{
    phmm_ctx *ctx = get_ctx(&g_tls.both, 0);
    static_assert(sizeof(testcase) == sizeof(phmm_testcase), "testcase layout");
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
