/*
 * ref_harness.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Thin extern "C" shim around the UNMODIFIED reference pair-HMM sources, which
 * oracle/build_ref.sh compiles in place from /root/reference into
 * oracle/_ref/libphmm_ref.so.  It only #includes reference headers (via -I); no
 * reference source is copied into this repository.
 *
 * What it exposes, with the same flat pair-list ABI as phmm_oracle.c:
 *   phmm_ref_run      the policy loop of computelikelihoodsboth
 *                     (tools/GKL/src/main/native/pairhmm/IntelPairHmmCSource.cpp:69-82)
 *                     re-stated around the reference's exported kernel pointers
 *                     compute_fp_avx512s/d (avx512_impl.cc:7-8) and compute_fp_avxs/d
 *                     (avx_impl.cc:4-5).  It does not go through computelikelihoodsboth
 *                     itself because the fork put a per-pair printf inside that loop (:80).
 *   phmm_ref_tables   the reference's Context<float>/Context<double> tables (Context.h).
 *   phmm_ref_kat      computelikelihoodsfloat/double (IntelPairHmmCSource.cpp:89-114)
 *                     on one pair, i.e. the real C interface, used for the known-answer test.
 *   phmm_ref_isa      1 if the reference's own cpuid dispatch would pick AVX-512.
 */
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <xmmintrin.h>
#include <pmmintrin.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <avx.h>
#include "pairhmm_common.h"
#include "avx_impl.h"
#include "avx512_impl.h"
#include "Context.h"

void initPairHMM();
void computelikelihoodsfloat(testcase *testcases, float *expected_result);
void computelikelihoodsdouble(testcase *testcases, double *expected_result);

extern Context<float> g_ctxf;      /* IntelPairHmmCSource.cpp:19-20 */
extern Context<double> g_ctxd;

static int g_inited = 0;
static void ensure_init()
{
    if (!g_inited) {
        /* initPairHMM prints one line ("AVX512 supported!"); keep stdout clean for callers */
        FILE *saved = stdout;
        FILE *nul = fopen("/dev/null", "w");
        if (nul) stdout = nul;
        initPairHMM();
        if (nul) { fflush(nul); stdout = saved; fclose(nul); }
        g_inited = 1;
    }
}

extern "C" {

int phmm_ref_isa(void) { return (is_avx512_supported() ? 1 : 0); }

void phmm_ref_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d,
                     float *init_f, float *log10_init_f, double *init_d, double *log10_init_d)
{
    ensure_init();
    const int MM = ((MAX_QUAL + 1) * (MAX_QUAL + 2)) >> 1;
    if (ph2pr_f) memcpy(ph2pr_f, ContextBase<float>::ph2pr, 128 * sizeof(float));
    if (ph2pr_d) memcpy(ph2pr_d, ContextBase<double>::ph2pr, 128 * sizeof(double));
    if (mm_f) memcpy(mm_f, ContextBase<float>::matchToMatchProb, MM * sizeof(float));
    if (mm_d) memcpy(mm_d, ContextBase<double>::matchToMatchProb, MM * sizeof(double));
    if (init_f) *init_f = ContextBase<float>::INITIAL_CONSTANT;
    if (log10_init_f) *log10_init_f = ContextBase<float>::LOG10_INITIAL_CONSTANT;
    if (init_d) *init_d = ContextBase<double>::INITIAL_CONSTANT;
    if (log10_init_d) *log10_init_d = ContextBase<double>::LOG10_INITIAL_CONSTANT;
}

/* isa: 0 = the reference's own dispatch, 1 = force AVX-256 pointers, 2 = force AVX-512 pointers.
 * ftz: 0 = as the benchmark ships (IntelPairHmmCSource.cpp never touches MXCSR),
 *      1 = flush-to-zero as the JNI entry sets it (IntelPairHmm.cc:66-70). */
void phmm_ref_run(int64_t n, const int32_t *rslen, const int32_t *haplen,
                  const int64_t *rs_off, const int64_t *hap_off,
                  const uint8_t *rs, const uint8_t *q, const uint8_t *i, const uint8_t *d,
                  const uint8_t *c, const uint8_t *hap,
                  double *out, float *raw32, double *raw64, uint8_t *rescued,
                  int mode, int nthreads, int isa, int ftz)
{
    ensure_init();
    float (*kf)(testcase *);
    double (*kd)(testcase *);
    bool use512 = (isa == 2) || (isa == 0 && is_avx512_supported());
    if (use512) { kf = compute_fp_avx512s; kd = compute_fp_avx512d; }
    else        { kf = compute_fp_avxs;    kd = compute_fp_avxd; }
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
    {
        unsigned int saved_ftz = _MM_GET_FLUSH_ZERO_MODE();
        unsigned int saved_daz = _MM_GET_DENORMALS_ZERO_MODE();
        if (ftz) { _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON); }
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t p = 0; p < n; p++) {
            testcase tc;
            tc.rslen = rslen[p]; tc.haplen = haplen[p];
            tc.rs = (const char *)rs + rs_off[p];
            tc.q = (const char *)q + rs_off[p];
            tc.i = (const char *)i + rs_off[p];
            tc.d = (const char *)d + rs_off[p];
            tc.c = (const char *)c + rs_off[p];
            tc.hap = (const char *)hap + hap_off[p];
            double result;
            float f = 0.f;
            double dv = 0.0;
            int resc = 0;
            if (mode != 2) f = kf(&tc);
            if (mode == 2 || (mode == 0 && f < MIN_ACCEPTED)) {
                dv = kd(&tc);
                result = log10(dv) - g_ctxd.LOG10_INITIAL_CONSTANT;
                resc = (mode == 0);
            } else {
                result = (double)(log10f(f) - g_ctxf.LOG10_INITIAL_CONSTANT);
            }
            out[p] = result;
            if (raw32) raw32[p] = f;
            if (raw64) raw64[p] = dv;
            if (rescued) rescued[p] = (uint8_t)resc;
        }
        _MM_SET_FLUSH_ZERO_MODE(saved_ftz);
        _MM_SET_DENORMALS_ZERO_MODE(saved_daz);
    }
}

/* The reference's real single-pair C entry points, for the known-answer test. */
void phmm_ref_kat(int rslen, int haplen, const char *rs, const char *q, const char *i,
                  const char *d, const char *c, const char *hap, float *out_f, double *out_d)
{
    ensure_init();
    testcase tc;
    tc.rslen = rslen; tc.haplen = haplen;
    tc.rs = rs; tc.q = q; tc.i = i; tc.d = d; tc.c = c; tc.hap = hap;
    computelikelihoodsfloat(&tc, out_f);
    computelikelihoodsdouble(&tc, out_d);
}

} /* extern "C" */
