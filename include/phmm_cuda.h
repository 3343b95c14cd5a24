/*
 * phmm_cuda.h -- C ABI of libphmm_cuda.so, the B200 (sm_100a) pair-HMM forward engine.
 *
 * Drop-in boundary for the GenomicsBench `phmm` hot path.  Reference paths below are relative
 * to /root/reference.  The reference library behind this boundary is libgkl_pairhmm_c
 * (tools/GKL/src/main/native/pairhmm/IntelPairHmmCSource.cpp), which the benchmark driver
 * links with -lgkl_pairhmm_c (benchmarks/phmm/Makefile:32-34).
 *
 * Three layers, all in this one library:
 *   1. the reference's own four C++-linkage symbols (section "legacy symbols" at the end), so
 *      the unmodified driver object links against libphmm_cuda.so instead of libgkl_pairhmm_c;
 *   2. plain `extern "C"` entry points taking HOST buffers (blocking, results written when they
 *      return) -- what a cgo / JNI / ctypes binding would call;
 *   3. a device-resident workload API (plan + upload once, run many times) used for kernel-only
 *      timing and by the rewritten driver's pipeline.
 *
 * There is no CPU fallback: every compute entry point returns PHMM_ERR_CUDA when no sm_100
 * device is usable.
 */
#ifndef PHMM_CUDA_H
#define PHMM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PHMM_ABI_VERSION 5

/* ---- status codes (the reference defines none: it asserts / exit(0)s) ---- */
#define PHMM_OK              0
#define PHMM_ERR_ARG         1   /* NULL pointer, negative count, inconsistent table          */
#define PHMM_ERR_EMPTY_SEQ   2   /* rslen == 0 or haplen == 0: undefined in the reference      */
                                 /* (division by haplen, avx-pairhmm-template.h:86)            */
#define PHMM_ERR_TOO_LONG    3   /* read or haplotype longer than PHMM_MAX_SEQ_LEN             */
#define PHMM_ERR_CUDA        4   /* no usable device / CUDA runtime error (see last_error)     */
#define PHMM_ERR_NOMEM       5
#define PHMM_ERR_PARSE       6   /* malformed text input                                       */
#define PHMM_ERR_IO          7

/* Declared limits of the reference: MAX_READ_LENGTH / MAX_HAP_LENGTH = 32K
 * (benchmarks/phmm/PairHMMUnitTest.h:35-36). */
#define PHMM_MAX_SEQ_LEN 32768

/* ---- the reference's per-pair descriptor ------------------------------------------------
 * Layout-identical to `testcase` (benchmarks/phmm/pairhmm_common.h:20-24,
 * tools/GKL/src/main/native/pairhmm/pairhmm_common.h:20-24): q/i/d/c are Phred values already
 * normalised by the driver (PairHMMUnitTest_orig.cpp:88-92,120-123), rs/hap are ASCII bases,
 * none of them NUL-terminated beyond rslen/haplen. */
typedef struct {
    int rslen, haplen;
    const char *q, *i, *d, *c;
    const char *hap, *rs;
} phmm_testcase;

/* ---- the reference driver's batch containers ----------------------------------------------
 * Field-for-field the plaintext part of shacc_pairhmm::Read / Haplotype / Batch
 * (tools/GKL/src/main/native/pairhmm/shacc_pairhmm.h:12-35; benchmarks/phmm/shacc_pairhmm.h:14-46
 * adds HE ciphertext members that are out of scope).  A batch is evaluated all-pairs and
 * results[r * num_haps + h] receives the log10 likelihood (PairHMMUnitTest_orig.cpp:284-299). */
typedef struct {
    int length;
    const char *bases, *q, *i, *d, *c;
} phmm_read;

typedef struct {
    int length;
    const char *bases;
} phmm_hap;

typedef struct {
    int num_reads, num_haps;
    const phmm_read *reads;
    const phmm_hap *haps;
    double *results;              /* num_reads * num_haps doubles, caller-owned */
} phmm_batch;

/* ---- flat structure-of-arrays batches -------------------------------------------------------
 * What the rewritten driver's parser produces directly (no per-string heap allocation) and
 * what a binding for a GC'd language can build without pinning thousands of small objects:
 * six byte pools plus offset tables.  Reads of batch 0 come first, then batch 1, ...; the same
 * for haplotypes.  Read r occupies [read_off[r], read_off[r] + read_len[r]) in EACH of the five
 * read pools.  Results are batch-major, then r * num_haps + h, exactly the order in which the
 * reference fills its per-batch `results` arrays. */
typedef struct {
    int64_t n_batches;
    const int32_t *batch_num_reads;     /* [n_batches] */
    const int32_t *batch_num_haps;      /* [n_batches] */
    int64_t n_reads;
    const int32_t *read_len;            /* [n_reads] */
    const int64_t *read_off;            /* [n_reads] */
    const uint8_t *rs, *q, *i, *d, *c;  /* read pools */
    int64_t n_haps;
    const int32_t *hap_len;             /* [n_haps] */
    const int64_t *hap_off;             /* [n_haps] */
    const uint8_t *hap;                 /* haplotype pool */
} phmm_flat_batches;

/* ---- options ------------------------------------------------------------------------------ */
#define PHMM_OPT_PRECISION   1   /* 0 both: FP32 + FP64 rescue (computelikelihoodsboth, default)
                                    1 FP32 only (computelikelihoodsfloat, :89-99)
                                    2 FP64 only (computelikelihoodsdouble, :103-114)          */
#define PHMM_OPT_FP32_KERNEL 2   /* 0 bit-exact: unfused mul/add in the reference's association
                                      order, raw FP32 sums identical to the AVX code
                                    1 fast: FMA-contracted; pairs whose sum lands in a guard band
                                      around MIN_ACCEPTED are re-decided by the bit-exact kernel,
                                      so the FP64-rescue set stays identical                    */
#define PHMM_OPT_FP64_KERNEL 3   /* 0 bit-exact (unfused), 1 FMA-contracted (|rel err| ~1e-13)  */
#define PHMM_OPT_SEG_COLS    4   /* target haplotype-stream columns per task (load balancing)    */
#define PHMM_OPT_FORCE_CLASS 6   /* benchmarking: pin the kernel class, value = kh << 8 | lanes
                                    (0 = choose per read length); also env PHMM_FORCE_CLASS=kh,lanes */

#define PHMM_OPT_CHUNK_CELLS 7   /* host-buffer entry points: cells per pipelined chunk of batches
                                    (0 = default: a fifth of the work, within [4e9, 4e10]); also env PHMM_CHUNK_CELLS */

#define PHMM_OPT_PLAN        8   /* how tasks are cut: 0 by the size of the call (default), 1 for throughput (tall
                                    register tiles, long runs of haplotypes, four kernels at once), 2 for the latency
                                    of a small call such as one active region (32 lanes per read, one haplotype per
                                    run, every class kernel at once); results of the bit-exact kernels do not depend
                                    on it, the low bits of the fast FP32 kernels do */

typedef struct phmm_ctx phmm_ctx;
typedef struct phmm_workload phmm_workload;

/* Statistics of one run (all counts are for the last call on that workload). */
typedef struct {
    int64_t n_pairs;
    int64_t n_cells;            /* sum rslen * haplen, rescued pairs counted once              */
    int64_t n_rescued;          /* pairs re-run in FP64 (FP32 sum < MIN_ACCEPTED = 1e-28f)     */
    int64_t n_rescued_cells;
    int64_t n_guard;            /* fast kernel only: pairs re-decided by the bit-exact kernel  */
    int64_t n_tasks;            /* warp tasks over all kernel classes                          */
    int64_t h2d_bytes, d2h_bytes;
    int32_t n_launches;         /* kernels launched by the last run                            */
    int32_t n_classes;          /* distinct (rows-per-lane, lanes-per-read) kernel classes     */
    float   ms_fp32;            /* CUDA-event time of the FP32 kernels of the last run         */
    float   ms_fp64;            /* ... of the FP64 rescue kernels                              */
    float   ms_total;           /* first launch to last kernel end                            */
    float   ms_h2d, ms_d2h;     /* host-buffer entry points only                               */
    /* ABI 5.  FP64-first reads (fast FP32 kernels, both precisions): reads the planner predicts to fail the FP32
     * pass against their whole batch skip it; the FP64 kernel checks on the rescaled sum what that pass would have
     * decided.  n_spec_direct: pairs of such reads that pass after all (finalised from the FP64 sum in float). */
    int64_t n_spec_pairs, n_spec_cells, n_spec_direct;
} phmm_stats;

/* ---- lifetime ------------------------------------------------------------------------------
 * phmm_create replaces initPairHMM (IntelPairHmmCSource.cpp:29-51): instead of cpuid dispatch it
 * binds a CUDA device, generates the Context<float>/Context<double> probability tables
 * (Context.h:42-61,101-114,141-154) on the host with the host libm and uploads them. */
int phmm_create(int device, phmm_ctx **ctx);
int phmm_destroy(phmm_ctx *ctx);
int phmm_set_option(phmm_ctx *ctx, int option, int64_t value);
int64_t phmm_get_option(const phmm_ctx *ctx, int option);
const char *phmm_last_error(const phmm_ctx *ctx);   /* ctx may be NULL: last create() failure */
int phmm_abi_version(void);
int phmm_device_count(void);

/* ---- host-buffer entry points (blocking) --------------------------------------------------- */
/* computelikelihoodsboth(testcase*, double*, int) (IntelPairHmmCSource.cpp:61-85): out[k] is the
 * log10 likelihood of pair k, order preserved.  Pairs sharing rs/hap pointers (the driver's
 * all-pairs slabs) are de-duplicated before upload. */
int phmm_compute_pairs(phmm_ctx *ctx, const phmm_testcase *pairs, double *out, int64_t n);
/* The reference driver's inner loop over batches (PairHMMUnitTest_orig.cpp:271-300) in one call. */
int phmm_compute_batches(phmm_ctx *ctx, const phmm_batch *batches, int64_t n_batches);
int phmm_compute_flat(phmm_ctx *ctx, const phmm_flat_batches *fb, double *out, phmm_stats *stats);

/* Asynchronous form of phmm_compute_pairs for a dispatcher that serves several callers: the requests' testcase
 * arrays (each what one computelikelihoodsboth call receives, IntelPairHmmCSource.cpp:61) are merged into ONE planned
 * chunk, whose uploads and kernels are queued in a free pipeline slot; the call returns without waiting (the pageable
 * inputs have been staged by then and may be reused).  phmm_wait(ticket, stats) blocks until that submission -- and every
 * earlier one -- has finished and copies each request's results into its own `out`.  Up to four submissions are in
 * flight per context; one thread at a time per context.  libphmm_cuda's own legacy symbols use this to coalesce the
 * concurrent per-batch calls of the driver's OpenMP threads (PairHMMUnitTest_orig.cpp:271-299). */
typedef struct {
    const phmm_testcase *pairs;
    double *out;
    int64_t n;
} phmm_pairs_request;
int phmm_submit_pairs(phmm_ctx *ctx, const phmm_pairs_request *reqs, int n_reqs, int64_t *ticket);
/* stats may be NULL; otherwise the counts and kernel times of every submission this call retires (and of those an
 * earlier phmm_submit_* had to retire to free a slot) are ADDED to it: zero it before the first use. */
int phmm_wait(phmm_ctx *ctx, int64_t ticket, phmm_stats *stats);
/* The same with requests that are flat batches already (what a caller thread gets from phmm_flatten_pairs, doing the
 * per-pair work of a legacy call itself and in parallel with the other callers): tables are appended, pools copied
 * back to back.  out receives the request's results in the order of phmm_compute_flat. */
typedef struct phmm_parsed phmm_parsed;
typedef struct {
    const phmm_flat_batches *batches;
    double *out;
} phmm_flat_request;
int phmm_submit_flat(phmm_ctx *ctx, const phmm_flat_request *reqs, int n_reqs, int64_t *ticket);
/* testcase[] -> flat batches (de-duplicating the strings the pairs share), owned by the returned handle: read with
 * phmm_parsed_batches, release with phmm_parsed_free.  No device needed. */
int phmm_flatten_pairs(const phmm_testcase *pairs, int64_t n, phmm_parsed **out, char *err, size_t err_len);

/* Optional: page-lock caller-owned buffers that live across many calls (the pools of phmm_flat_batches, the results
 * array), so that the library's copies run as asynchronous DMA at full PCIe / NVLink-C2C rate instead of being staged
 * through the driver's bounce buffer (cudaHostRegister behind a plain-pointer interface: a cgo / JNI / ctypes caller
 * has no CUDA headers).  The reference has no counterpart -- its buffers never leave the host
 * (PairHMMUnitTest_orig.cpp:246 allocates them with _mm_malloc).  Unregister before freeing the buffer. */
int phmm_host_register(phmm_ctx *ctx, void *ptr, size_t bytes);
int phmm_host_unregister(phmm_ctx *ctx, void *ptr);

/* ---- device-resident workloads -------------------------------------------------------------- */
/* Plan (kernel classes, haplotype-stream segments, LPT-ordered warp tasks) and upload. */
int phmm_workload_create(phmm_ctx *ctx, const phmm_flat_batches *fb, phmm_workload **w);
/* Launch the FP32 kernels and the FP64 rescue kernels on the context's stream and wait. */
int phmm_workload_run(phmm_ctx *ctx, phmm_workload *w, phmm_stats *stats);
/* Copy results back: out = log10 likelihoods; the others may be NULL.  raw32 = FP32 sums scaled
 * by 2^120 (what compute_full_prob_<float> returns), raw64 = FP64 sums scaled by 2^1020 for
 * rescued pairs (0 elsewhere), rescued = 1 where the FP64 path produced `out`. */
int phmm_workload_results(phmm_ctx *ctx, phmm_workload *w, double *out, float *raw32,
                          double *raw64, uint8_t *rescued);
int phmm_workload_destroy(phmm_workload *w);

/* ---- host-side scheduler ---------------------------------------------------------------------
 * Longest-processing-time-first assignment of n items (batches) with the given costs (cells) to
 * `parts` devices; owner[k] in [0, parts).  Deterministic, so every rank computes the same map. */
int phmm_partition_lpt(const int64_t *cost, int64_t n, int parts, int32_t *owner);
/* Predicted device time of every batch (arbitrary unit), the cost phmm_partition_lpt should be given.  The
 * reference has no counterpart: its only per-batch figure is Batch::num_cells
 * (benchmarks/phmm/PairHMMUnitTest_orig.cpp:166), which ignores that the device pays for the register tile of
 * a read's kernel class (padded rows + a per-step overhead per lane) and that a pair below MIN_ACCEPTED is
 * computed twice, the second time in FP64 (IntelPairHmmCSource.cpp:73-75).  fp32_fast / precision as
 * PHMM_OPT_FP32_KERNEL / PHMM_OPT_PRECISION.  The FP64 share is predicted per read on the host, without dynamic
 * programming, from a k-mer anchored quality-weighted mismatch score against the first haplotype of the batch;
 * read_rescued (may be NULL, [n_reads]) receives 1 where most pairs of the read are expected to be rescued. */
int phmm_batch_costs(const phmm_flat_batches *fb, int fp32_fast, int precision, int64_t *cost,
                     uint8_t *read_rescued);

/* ---- text input (the reference's on-disk format) ---------------------------------------------
 * Parses what read_testfile / read_batch read (PairHMMUnitTest_orig.cpp:97-213) into flat
 * batches owned by the returned handle; qualities are normalised as the driver does
 * (`normalize`, :88-92: q floor 6, i/d/c floor 0, after subtracting 33). */
int phmm_parse_file(const char *path, phmm_parsed **out, char *err, size_t err_len);
int phmm_parse_buffer(const char *data, size_t len, phmm_parsed **out, char *err, size_t err_len);
const phmm_flat_batches *phmm_parsed_batches(const phmm_parsed *p);
void phmm_parsed_free(phmm_parsed *p);

/* ---- binary cache of parsed batches ----------------------------------------
 * At GPU speed the text parse dominates the wall time of the reference driver's untimed phase
 * (read_testfile, PairHMMUnitTest_orig.cpp:186-213).  phmm_cache_write stores flat batches as the
 * library consumes them: a 128-byte header, then the offset tables and the six byte pools as
 * 64-byte aligned little-endian sections, quality pools already normalised.  phmm_cache_open
 * maps the file (no copy, no parse; pages fault in as the planner / uploads touch them) and
 * checks magic, version, section bounds and a 64-bit FNV-1a checksum of every section; the
 * handle is released with phmm_parsed_free like a parsed text file. */
int phmm_cache_write(const phmm_flat_batches *fb, const char *path, char *err, size_t err_len);
int phmm_cache_open(const char *path, phmm_parsed **out, char *err, size_t err_len);

/* ---- GATK native-bindings shaped batch API -------------------------------
 * The upstream caller of this kernel: IntelPairHmm.initNative / computeLikelihoodsNative /
 * doneNative (tools/GKL/src/main/native/pairhmm/IntelPairHmm.cc:32-96,103-156,164-168; Java side
 * tools/GKL/src/main/java/com/intel/gkl/pairhmm/IntelPairHmm.java:112-117).  One call evaluates
 * all reads x all haplotypes of an active region; likelihoods[r * n_haps + h] is the log10
 * likelihood (JavaData.h:44-68 builds the testcases read-major).  Field names are those of GATK's
 * ReadDataHolder / HaplotypeDataHolder (JavaData.h:15-20); quality bytes are raw Phred values
 * (no +33), masked & 127 by the kernel as in avx-pairhmm-template.h:110-112,125.
 *   use_double  : the reference skips the FP32 pass and computes every pair in FP64
 *                 (IntelPairHmm.cc:123-124: result_float = 0.0f < MIN_ACCEPTED).
 *   max_threads : host threads of the reference's OpenMP loop; accepted and ignored.
 *   use_fpga    : the reference's FPGA hook (shacc_pairhmm::calculate, always false); must be 0.
 * Kernels: the bit-exact ones by default (raw FP32 / FP64 sums identical to the reference's AVX code, so a pair's
 * likelihood does not depend on which other regions share the call).  The fast kernels are an opt-in
 * (environment PHMM_GATK_FAST=1, or phmm_set_option(phmm_gatk_context(g), PHMM_OPT_FP32_KERNEL, 1)): same
 * FP64-rescue set, FP32-path likelihoods within ~1e-6 relative, but their low bits depend on the kernel class a
 * read runs in, which the planner picks by the size of the call (one region: 32 lanes per read).
 * log10: the returned value goes through the device's log10f / log10 (documented maximum error 2 ulp / 1 ulp)
 * where the reference calls glibc's; "bit-exact" in this library refers to the raw sums, the returned log10
 * likelihoods agree within those ulps (tests hold them to 1e-4 / 1e-9 relative, BASELINE.json's bars).
 * FTZ: initNative turns on x86 flush-to-zero (:66-70).  The kernels here honour denormals, which
 * changes a result only when an FP32 intermediate falls below 2^-126 (the pair is then far below
 * MIN_ACCEPTED = 1e-28f or the lost term is < 2^-50 of the sum) or an FP64 intermediate below
 * 2^-1022 (likelihood < 1e-614); tests/test_oracle.py checks the reference with FTZ on against
 * FTZ off on the test sets. */
typedef struct {
    const uint8_t *readBases, *readQuals, *insertionGOP, *deletionGOP, *overallGCP;
    int32_t length;
} phmm_gatk_read;

typedef struct {
    const uint8_t *haplotypeBases;
    int32_t length;
} phmm_gatk_hap;

typedef struct {                        /* one active region = one computeLikelihoodsNative call */
    const phmm_gatk_read *reads;
    const phmm_gatk_hap *haps;
    double *likelihoods;                /* n_reads * n_haps, caller-owned */
    int32_t n_reads, n_haps;
} phmm_gatk_region;

typedef struct phmm_gatk phmm_gatk;
int phmm_gatk_init(int device, int use_double, int max_threads, int use_fpga, phmm_gatk **out);
int phmm_gatk_compute_likelihoods(phmm_gatk *g, const phmm_gatk_read *reads, int32_t n_reads,
                                  const phmm_gatk_hap *haps, int32_t n_haps, double *likelihoods);
/* Many regions in one call: the regions are planned and launched together, which is what makes
 * a GPU worthwhile when a region holds only tens of reads. */
int phmm_gatk_compute_regions(phmm_gatk *g, const phmm_gatk_region *regions, int64_t n_regions);
int phmm_gatk_done(phmm_gatk *g);
const char *phmm_gatk_last_error(const phmm_gatk *g);
phmm_ctx *phmm_gatk_context(phmm_gatk *g);    /* the underlying context (options, statistics) */

/* ---- GATK regression-vector loader ------------------------------------------
 * Parses the line format of GATK's pairhmm-testdata.txt as the reference's dataFileTest reads it
 * (tools/GKL/src/test/java/com/intel/gkl/pairhmm/PairHmmUnitTest.java:150-197): one pair per line,
 *     <hap-bases> <read-bases> <read-quals> <ins-quals> <del-quals> <gcp-quals> <expected>
 * lines starting with '#' are skipped, quality strings are Phred+33 ASCII and are converted as
 * the test does (normalize(): byte - 33, read quals floored at 6; :174-177,:268-278).  Each line becomes a
 * 1 x 1 batch; expected[k] receives the expected log10 likelihood of line k.
 * phmm_testdata_expected returns n (= number of pairs) and copies min(n, cap) values. */
int phmm_testdata_parse_buffer(const char *data, size_t len, phmm_parsed **out, char *err, size_t err_len);
int phmm_testdata_parse_file(const char *path, phmm_parsed **out, char *err, size_t err_len);
int64_t phmm_testdata_expected(const phmm_parsed *p, double *expected, int64_t cap);

/* ---- measurement helpers ----------------------------------------------------------------------
 * FP32 pipe microbenchmarks on the context's device, for the roofline denominator
 * ("builder must measure an FFMA-only microbenchmark on the box").
 * which: 0 FFMA, 1 FFMA2 (packed), 2 FMUL+FADD mix, 3 FMUL2+FADD mix, 4 DFMA.
 * Returns lane-operations per second (an FMA counts as one operation). */
int phmm_microbench(phmm_ctx *ctx, int which, double *lane_ops_per_s, float *ms);

/* ---- test hooks (no device needed) ------------------------------------------------------------
 * phmm_get_tables: the host-generated probability tables the library uploads (Context.h); returns
 * the number of matchToMatchProb entries (8256).  consts_f/d = {INITIAL_CONSTANT, LOG10_INITIAL_CONSTANT}.
 * phmm_plan_summary: runs the planner only (precision | 0x100 = plan as PHMM_OPT_FP32_KERNEL 1 does:
 * haplotype-pair tasks, two runs of haplotypes per read).  info[0..6] = pairs, cells, classes, stream bytes,
 * max rslen, max haplen, tasks, tile cells (what the FP32 kernels compute: kh x lanes rows per read, separator
 * columns included); info[8 + 4c ..] = {kh, lanes, vl, tasks} and info[72 + 2c ..] = {cells, tile cells} of class c
 * (c < 16); precision | 0x200 adds FP64-first reads to that plan (| 0x400: only if the library would, by its per-call rule): info[104] = their pairs, info[105] = their cells (info
 * holds 106 values);
 * pair_cover[p] is incremented once for every task that produces pair p. */
int phmm_get_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d, float *consts_f,
                    double *consts_d);
int phmm_plan_summary(const phmm_flat_batches *fb, int precision, int seg_cols,
                      int64_t *info, uint8_t *pair_cover);

/* ---- allocation-free entry point ----------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * One region = all n_reads x n_haps pairs, computelikelihoodsboth policy (FP32, pairs below MIN_ACCEPTED
 * again in FP64), through the library's bit-exact kernels.  The call allocates nothing, on the device or
 * on the host, frees nothing and reads nothing back: the kernels' scratch (probability tables, read
 * records, coded haplotype stream, task lists, queue heads, FP64-rescue list, raw sums) lives in one
 * file-scope static __device__ array of PHMM_NOALLOC_SCRATCH_BYTES that every GPU that loads the library
 * holds until exit (csrc/phmm_noalloc.cu); planning runs on the device.  Inputs and outputs are the
 * CALLER'S DEVICE buffers on `device`:
 *   d_rs, d_q, d_i, d_d, d_c   read pools (bases; Phred qualities as the driver normalises them)
 *   d_read_off[n_reads + 1]    read r = pool bytes [d_read_off[r], d_read_off[r + 1]), 1..max_read_len long
 *   d_hap, d_hap_off[n_haps+1] haplotype h = d_hap bytes [d_hap_off[h], d_hap_off[h + 1]), ascending;
 *                              hap_bytes >= d_hap_off[n_haps] - d_hap_off[0]
 *   d_out[n_reads * n_haps]    log10 likelihoods, read-major (r * n_haps + h)
 *   d_rescued                  same shape, 1 = pair went through FP64; may be NULL
 * The work is queued on `cuda_stream` (a cudaStream_t; NULL = the default stream) and the call returns
 * without waiting.  The first call on a device also sets that device up (cudaGetSymbolAddress of the array,
 * table upload into it, kernel attributes -- still no allocation) and must not happen under capture.
 * Sizes beyond the maxima below are refused with PHMM_ERR_TOO_LONG and a message in `err` (may be NULL).
 * Tables on the device that contradict the stated sizes (a read longer than max_read_len, haplotypes
 * beyond hap_bytes, offsets not ascending) cannot be reported without a read-back: every result of such a
 * call is NaN.  All calls on a GPU share the array, which keeps the last call's contents (what must start
 * at zero is cleared by every call): calls that OVERLAP on one GPU race -- serialise them (one stream, or
 * events). */
#define PHMM_NOALLOC_MAX_READS      1024
#define PHMM_NOALLOC_MAX_HAPS       128
#define PHMM_NOALLOC_MAX_PAIRS      32768      /* n_reads * n_haps */
#define PHMM_NOALLOC_MAX_READ_LEN   255
#define PHMM_NOALLOC_MAX_HAP_BYTES  65536      /* sum of the haplotype lengths */
#define PHMM_NOALLOC_SCRATCH_BYTES  (3u << 20) /* the static device array, per GPU */
int phmm_forward_region_noalloc(int device, void *cuda_stream,
                                const uint8_t *d_rs, const uint8_t *d_q, const uint8_t *d_i,
                                const uint8_t *d_d, const uint8_t *d_c, const int32_t *d_read_off,
                                int32_t n_reads, int32_t max_read_len,
                                const uint8_t *d_hap, const int32_t *d_hap_off, int32_t n_haps,
                                int32_t hap_bytes, double *d_out, uint8_t *d_rescued,
                                char *err, size_t err_cap);

#ifdef __cplusplus
} /* extern "C" */
#endif

/* ---- legacy symbols ------------------------------------------------------------------------
 * libphmm_cuda.so also defines, with C++ linkage and no header (the driver re-declares them,
 * PairHMMUnitTest_orig.cpp:84-86), exactly the symbols libgkl_pairhmm_c exports:
 *     void initPairHMM();                                           _Z11initPairHMMv
 *     void computelikelihoodsboth(testcase*, double*, int);         _Z22computelikelihoodsbothP8testcasePdi
 *     void computelikelihoodsfloat(testcase*, float*);              _Z23computelikelihoodsfloatP8testcasePf
 *     void computelikelihoodsdouble(testcase*, double*);            _Z24computelikelihoodsdoubleP8testcasePd
 *     uint8_t ConvertChar::conversionTable[255];                    _ZN11ConvertChar15conversionTableE
 * (IntelPairHmmCSource.cpp:29,61,89,103; pairhmm_common.cc:3).  They bind device
 * $PHMM_DEVICE (default 0), are thread-safe, and abort() with a message on error because the
 * reference signatures have no way to report one. */

#endif /* PHMM_CUDA_H */
