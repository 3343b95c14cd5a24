/*
 * phmm_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C scalar restatement of the GenomicsBench `phmm` hot path (GKL pair-HMM
 * forward algorithm, FP32 with FP64 rescue).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product (libphmm_cuda.so) never links, loads or calls anything in oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement against
 *   (1) the reference's only in-repo known-answer test
 *       (tools/GKL/src/test/java/com/intel/gkl/pairhmm/PairHmmUnitTest.java:42-49,55),
 *   (2) golden vectors produced by the reference's own AVX/AVX-512 code compiled
 *       from /root/reference (oracle/build_ref.sh -> oracle/_ref/libphmm_ref.so;
 *       generator: tests/golden/make_golden.py), bit-exact on the raw FP32/FP64 sums,
 *   (3) when oracle/_ref is present, the live reference library on random pairs.
 *
 * All paths below are relative to /root/reference/tools/GKL/src/main/native/pairhmm/.
 * Must be compiled with -ffp-contract=off (the reference's canonical build has no FMA
 * in the kernel).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAX_QUAL 254                       /* Context.h:7 */
#define MM_SIZE (((MAX_QUAL + 1) * (MAX_QUAL + 2)) >> 1)   /* Context.h:24 */
#define JAC_TOL 8.0                        /* Context.h:8 */
#define JAC_STEP 0.0001                    /* Context.h:9 */
#define JAC_INV_STEP (1.0 / JAC_STEP)      /* Context.h:10 */
#define JAC_SIZE 80001                     /* Context.h:11: (int)(8.0 / 0.0001) + 1 */
#define MIN_ACCEPTED 1e-28f                /* pairhmm_common.h:16 */

static float  g_jac_f[JAC_SIZE];
static double g_jac_d[JAC_SIZE];
static float  g_mm_f[MM_SIZE];
static double g_mm_d[MM_SIZE];
static float  g_ph2pr_f[128];
static double g_ph2pr_d[128];
static float  g_init_f, g_log10_init_f;
static double g_init_d, g_log10_init_d;
static uint8_t g_conv[256];
static int g_ready = 0;

/* Context.h:63-65 fastRound, in the table's own type */
static int fast_round_f(float d)  { return (d > 0.0f) ? (int)(d + 0.5f) : (int)(d - 0.5f); }
static int fast_round_d(double d) { return (d > 0.0)  ? (int)(d + 0.5)  : (int)(d - 0.5); }

/* Context.h:67-90 approximateLog10SumLog10 -- evaluated in type T (float table != rounded double table) */
static float approx_log10_sum_f(float small, float big)
{
    if (small > big) { float t = big; big = small; small = t; }
    float diff = big - small;
    if (diff >= (float)JAC_TOL) return big;
    int ind = fast_round_f((float)(diff * ((float)JAC_INV_STEP)));
    return big + g_jac_f[ind];
}
static double approx_log10_sum_d(double small, double big)
{
    if (small > big) { double t = big; big = small; small = t; }
    double diff = big - small;
    if (diff >= (double)JAC_TOL) return big;
    int ind = fast_round_d((double)(diff * ((double)JAC_INV_STEP)));
    return big + g_jac_d[ind];
}

void phmm_oracle_init(void)
{
    if (g_ready) return;
    /* Context.h:42-47 initializeJacobianLogTable */
    for (int k = 0; k < JAC_SIZE; k++) {
        double v = log10(1.0 + pow(10.0, -((double)k) * JAC_STEP));
        g_jac_f[k] = (float)v;
        g_jac_d[k] = v;
    }
    /* Context.h:50-61 initializeMatchToMatchProb */
    double LN10 = log(10);
    double INV_LN10 = 1.0 / LN10;
    for (int i = 0, offset = 0; i <= MAX_QUAL; offset += ++i)
        for (int j = 0; j <= i; j++) {
            double s_f = approx_log10_sum_f(-0.1 * i, -0.1 * j);
            double s_d = approx_log10_sum_d(-0.1 * i, -0.1 * j);
            double l_f = log1p(-fmin(1.0, pow(10, s_f))) * INV_LN10;
            double l_d = log1p(-fmin(1.0, pow(10, s_d))) * INV_LN10;
            g_mm_f[offset + j] = (float)(pow(10, l_f));
            g_mm_d[offset + j] = (double)(pow(10, l_d));
        }
    /* Context.h:105-112 (double), :145-152 (float) */
    for (int x = 0; x < 128; x++) {
        g_ph2pr_d[x] = pow(10.0, -((double)x) / 10.0);
        g_ph2pr_f[x] = powf(10.f, -((float)x) / 10.f);
    }
    g_init_d = ldexp(1.0, 1020.0);
    g_log10_init_d = log10(g_init_d);
    g_init_f = ldexpf(1.f, 120.f);
    g_log10_init_f = log10f(g_init_f);
    /* pairhmm_common.h:30-39 ConvertChar::init; every other byte stays 0 ('A') */
    memset(g_conv, 0, sizeof g_conv);
    g_conv['A'] = 0; g_conv['C'] = 1; g_conv['T'] = 2; g_conv['G'] = 3; g_conv['N'] = 4;
    g_ready = 1;
}

/* Context.h:120-132 / :160-172 set_mm_prob (quals are &127 so the MAX_QUAL branch is unreachable) */
static int mm_index(int ins, int del)
{
    int mn = del, mx = ins;
    if (ins <= del) { mn = ins; mx = del; }
    return ((mx * (mx + 1)) >> 1) + mn;
}

void phmm_oracle_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d,
                        float *init_f, float *log10_init_f, double *init_d, double *log10_init_d)
{
    phmm_oracle_init();
    if (ph2pr_f) memcpy(ph2pr_f, g_ph2pr_f, sizeof g_ph2pr_f);
    if (ph2pr_d) memcpy(ph2pr_d, g_ph2pr_d, sizeof g_ph2pr_d);
    if (mm_f) memcpy(mm_f, g_mm_f, sizeof g_mm_f);
    if (mm_d) memcpy(mm_d, g_mm_d, sizeof g_mm_d);
    if (init_f) *init_f = g_init_f;
    if (log10_init_f) *log10_init_f = g_log10_init_f;
    if (init_d) *init_d = g_init_d;
    if (log10_init_d) *log10_init_d = g_log10_init_d;
}

/*
 * One pair, type T.  Restates compute_full_prob_<T> (avx-pairhmm-template.h:210-346):
 *   row constants   initializeVectors :83-128, stripeINITIALIZATION :151-158
 *   match rule      precompute_masks_ :3-35 (hap N sets all rows; read N row is all ones)
 *   cell update     computeMXY :183-198 -- association ((M*a + X*b) + Y*b) * e, no FMA
 *   boundary        Y[0][c] = INITIAL_CONSTANT/haplen (:86,:162), everything else 0
 *   result          sumM and sumX accumulated separately in ascending column order, then
 *                   added once (:328,:331,:342-343)
 */
#define DEFINE_PAIR(NAME, T, PH2PR, MMTAB, INITC, ONE, THREE)                                   \
T NAME(int R, int H, const uint8_t *rs, const uint8_t *q, const uint8_t *ii,                    \
       const uint8_t *dd, const uint8_t *cc, const uint8_t *hap)                                \
{                                                                                               \
    T *buf = (T *)malloc(sizeof(T) * 6 * (size_t)(H + 1));                                      \
    T *Mp = buf, *Xp = Mp + (H + 1), *Yp = Xp + (H + 1);                                        \
    T *Mc = Yp + (H + 1), *Xc = Mc + (H + 1), *Yc = Xc + (H + 1);                               \
    T init_Y = INITC / (T)H;                                                                    \
    for (int c = 0; c <= H; c++) { Mp[c] = 0; Xp[c] = 0; Yp[c] = init_Y; }                      \
    T sumM = 0, sumX = 0;                                                                       \
    for (int r = 1; r <= R; r++) {                                                              \
        int _i = ii[r - 1] & 127, _d = dd[r - 1] & 127, _c = cc[r - 1] & 127;                   \
        int _q = q[r - 1] & 127;                                                                \
        T pMM = MMTAB[mm_index(_i, _d)];                                                        \
        T pGAPM = ONE - PH2PR[_c];                                                              \
        T pMX = PH2PR[_i], pXX = PH2PR[_c], pMY = PH2PR[_d], pYY = PH2PR[_c];                   \
        T eps = PH2PR[_q];                                                                      \
        T em = ONE - eps;                                                                       \
        T ex = eps / THREE;                                                                     \
        int rc = g_conv[rs[r - 1]];                                                             \
        Mc[0] = 0; Xc[0] = 0; Yc[0] = 0;                                                        \
        for (int c = 1; c <= H; c++) {                                                          \
            int hc = g_conv[hap[c - 1]];                                                        \
            T e = (rc == hc || rc == 4 || hc == 4) ? em : ex;                                   \
            T t1 = Mp[c - 1] * pMM;                                                             \
            T t2 = Xp[c - 1] * pGAPM;                                                           \
            T t3 = Yp[c - 1] * pGAPM;                                                           \
            T s12 = t1 + t2;                                                                    \
            T s = s12 + t3;                                                                     \
            Mc[c] = s * e;                                                                      \
            T u1 = Mp[c] * pMX;                                                                 \
            T u2 = Xp[c] * pXX;                                                                 \
            Xc[c] = u1 + u2;                                                                    \
            T v1 = Mc[c - 1] * pMY;                                                             \
            T v2 = Yc[c - 1] * pYY;                                                             \
            Yc[c] = v1 + v2;                                                                    \
        }                                                                                       \
        T *t;                                                                                   \
        t = Mp; Mp = Mc; Mc = t;  t = Xp; Xp = Xc; Xc = t;  t = Yp; Yp = Yc; Yc = t;            \
    }                                                                                           \
    for (int c = 1; c <= H; c++) { sumM = sumM + Mp[c]; sumX = sumX + Xp[c]; }                  \
    T res = sumM + sumX;                                                                        \
    free(buf);                                                                                  \
    return res;                                                                                 \
}

DEFINE_PAIR(phmm_oracle_pair_f32, float,  g_ph2pr_f, g_mm_f, g_init_f, 1.0f, 3.0f)
DEFINE_PAIR(phmm_oracle_pair_f64, double, g_ph2pr_d, g_mm_d, g_init_d, 1.0,  3.0)

/*
 * Policy loop of computelikelihoodsboth (IntelPairHmmCSource.cpp:61-85) without the
 * fork's per-pair printf (:80): FP32 first, FP64 rescue when the FP32 sum is below
 * MIN_ACCEPTED (:73), log10 finalisation in float on the FP32 path (:78) and in double
 * on the rescued path (:75).
 *   mode 0 = both (the benchmark's path), 1 = FP32 only (:89-99), 2 = FP64 only (:103-114)
 * Pair p reads its read at rs_off[p] in the five parallel read pools and its haplotype at
 * hap_off[p] in the hap pool.  raw32/raw64/rescued may be NULL.
 */
void phmm_oracle_run(int64_t n, const int32_t *rslen, const int32_t *haplen,
                     const int64_t *rs_off, const int64_t *hap_off,
                     const uint8_t *rs, const uint8_t *q, const uint8_t *i, const uint8_t *d,
                     const uint8_t *c, const uint8_t *hap,
                     double *out, float *raw32, double *raw64, uint8_t *rescued,
                     int mode, int nthreads)
{
    phmm_oracle_init();
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
#endif
    for (int64_t p = 0; p < n; p++) {
        int R = rslen[p], H = haplen[p];
        int64_t ro = rs_off[p], ho = hap_off[p];
        double result;
        float f = 0.f;
        double dv = 0.0;
        int resc = 0;
        if (mode != 2)
            f = phmm_oracle_pair_f32(R, H, rs + ro, q + ro, i + ro, d + ro, c + ro, hap + ho);
        if (mode == 2 || (mode == 0 && f < MIN_ACCEPTED)) {
            dv = phmm_oracle_pair_f64(R, H, rs + ro, q + ro, i + ro, d + ro, c + ro, hap + ho);
            result = log10(dv) - g_log10_init_d;
            resc = (mode == 0);
        } else {
            result = (double)(log10f(f) - g_log10_init_f);
        }
        out[p] = result;
        if (raw32) raw32[p] = f;
        if (raw64) raw64[p] = dv;
        if (rescued) rescued[p] = (uint8_t)resc;
    }
}
