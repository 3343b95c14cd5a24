/*
 * sw_cuda.h -- C ABI of libsw_cuda.so, the B200 (sm_100a) Smith-Waterman aligner with backtrack
 * (the step before pair-HMM in HaplotypeCaller).
 *
 * Reference behind this boundary (paths relative to /root/reference):
 *   tools/GKL/src/main/native/smithwaterman/PairWiseSW.h          the algorithm (:40-450)
 *   tools/GKL/src/main/native/smithwaterman/IntelSmithWatermanCSource.{h,cpp}   C entry: initNative(), align()
 *   tools/GKL/src/main/native/smithwaterman/IntelSmithWaterman.cc  JNI entry: alignNative (:52-75)
 * Results are bit-exact: the same score, alignment offset and CIGAR string as the reference's
 * AVX2 / AVX-512 code, for all four overhang strategies.
 *
 * There is no CPU fallback: every compute entry point needs a context on an sm_100 device.
 */
#ifndef SW_CUDA_H
#define SW_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SW_ABI_VERSION 1

#define SW_OK            0
#define SW_ERR_ARG       1   /* NULL pointer, negative count, unknown strategy, parameter out of range */
#define SW_ERR_EMPTY_SEQ 2   /* a sequence of length 0 (the reference reads uninitialised state)        */
#define SW_ERR_TOO_LONG  3   /* longer than SW_MAX_SEQ_LEN                                              */
#define SW_ERR_CUDA      4
#define SW_ERR_NOMEM     5

/* MAX_SEQ_LEN of the reference (smithwaterman_common.h:50): its work arrays are sized by it.
 * (The reference itself corrupts its heap at len1 == 1024 -- E[MAX_SEQ_LEN - ilo] with
 * ilo = nrow + 1, PairWiseSW.h:207; this library is defined up to and including 1024.) */
#define SW_MAX_SEQ_LEN 1024
#define SW_MAX_ABS_PARAM 100000   /* |match|, |mismatch|, |open|, |extend|: keeps every score inside int32 */

/* Overhang strategies: the integers IntelSmithWaterman.java:106-121 passes down
 * (smithwaterman_common.h:26-29). */
#define SW_SOFTCLIP      9
#define SW_INDEL         10
#define SW_LEADING_INDEL 11
#define SW_IGNORE        12

/* SWParameters of the GATK native bindings, in the order align() takes them (:45). */
typedef struct {
    int32_t match, mismatch, open, extend;
} sw_params;

/* Pairs as two byte pools plus offset tables; pair p is ref = rows of the DP matrix (seq1/len1 in
 * PairWiseSW.h:423-434), alt = columns.  Several pairs may point at the same ref bytes (one
 * reference window against all candidate haplotypes of an active region): they are uploaded once. */
typedef struct {
    int64_t n_pairs;
    const int32_t *ref_len;      /* [n_pairs] */
    const int64_t *ref_off;      /* [n_pairs] into ref_pool */
    const uint8_t *ref_pool;
    const int32_t *alt_len;
    const int64_t *alt_off;
    const uint8_t *alt_pool;
} sw_flat_pairs;

typedef struct {
    int64_t n_pairs;
    int64_t n_cells;             /* sum len1 * len2 */
    int64_t n_overflow;          /* pairs whose CIGAR did not fit cigar_stride (cigar_len < 0) */
    int64_t h2d_bytes, d2h_bytes;
    int64_t bt_bytes;            /* backtrack words written to HBM by the last run */
    int32_t n_launches, n_classes;
    float ms_kernels;            /* CUDA-event time of the kernels of the last run */
    float ms_total;              /* host-buffer entry points: upload + kernels + download */
} sw_stats;

typedef struct sw_ctx sw_ctx;
typedef struct sw_workload sw_workload;

int sw_abi_version(void);
int sw_create(int device, sw_ctx **ctx);            /* replaces initNative() (cpuid dispatch, :22-40) */
int sw_destroy(sw_ctx *ctx);
const char *sw_last_error(const sw_ctx *ctx);       /* ctx may be NULL: last sw_create failure */

/* ---- host-buffer entry points (blocking) --------------------------------------------------------
 * score[p]    : the int32 alignment score (the reference hands it back through an int16_t*,
 *               PairWiseSW.h:441; compare modulo 2^16 against it)
 * offset[p]   : alignment offset, the return value of runSWOnePairBT / alignNative
 * cigar       : n_pairs slots of cigar_stride bytes; slot p receives the NUL-terminated CIGAR
 * cigar_len[p]: its length, or -1 - needed when it does not fit the slot (nothing is written;
 *               call again for those pairs with a larger stride)
 * Any of score / offset may be NULL. */
int sw_align_flat(sw_ctx *ctx, const sw_flat_pairs *pairs, const sw_params *params, int strategy,
                  int32_t *score, int32_t *offset, char *cigar, int32_t cigar_stride,
                  int32_t *cigar_len, sw_stats *stats);
/* One pair: the argument list of alignNative / align() (IntelSmithWaterman.cc:52,
 * IntelSmithWatermanCSource.cpp:45) with explicit lengths and an explicit CIGAR capacity. */
int sw_align_one(sw_ctx *ctx, const uint8_t *ref, int32_t ref_len, const uint8_t *alt, int32_t alt_len,
                 const sw_params *params, int strategy, char *cigar, int32_t cigar_cap,
                 int32_t *score, int32_t *offset);

/* ---- device-resident workloads (upload once, run many times: kernel-only timing) ----------------- */
int sw_workload_create(sw_ctx *ctx, const sw_flat_pairs *pairs, int32_t cigar_stride, sw_workload **w);
int sw_workload_run(sw_ctx *ctx, sw_workload *w, const sw_params *params, int strategy, sw_stats *stats);
int sw_workload_results(sw_ctx *ctx, sw_workload *w, int32_t *score, int32_t *offset, char *cigar,
                        int32_t *cigar_len);
int sw_workload_destroy(sw_workload *w);

/* ---- the reference's own C entry points ------------------------------------------------------------
 * libsw_cuda.so also defines, with C linkage as IntelSmithWatermanCSource.h:1-10 declares them,
 *     void initNative();
 *     void align(char *ref, char *alt, char *cigar, int match, int mismatch, int open, int extend,
 *                int strategy, int *score, int *offset);
 * ref / alt are NUL-terminated (the reference strlen()s them, :51-52); cigar must hold the result
 * (the reference sprintf()s without a bound).  Differences: *score IS written (the reference passes
 * &score of its own parameter to the kernel, :54, so the caller's variable is never touched), and
 * errors abort() with a message because the signature has no way to report one.  Device
 * $SW_DEVICE (default 0). */
void initNative(void);
void align(char *ref, char *alt, char *cigar, int match, int mismatch, int open, int extend,
           int strategy, int *score, int *offset);

#ifdef __cplusplus
}
#endif
#endif /* SW_CUDA_H */
