// Internal host/device shared types of libphmm_cuda (not part of the C ABI).
#ifndef PHMM_TYPES_H
#define PHMM_TYPES_H

#include <stdint.h>

namespace phmm {

// Haplotype-stream code per base: the row index of the per-warp emission table.  The match rule is
// the one precompute_masks_ builds (avx-pairhmm-template.h:3-35) on top of ConvertChar
// (pairhmm_common.h:30-43; every byte that is not ACGTN converts to 0 = 'A'): read base matches
// haplotype base when they are equal or either is N.  0 marks separator / padding columns, whose
// emission is 0 for every row.
static const uint8_t CODE_SEP = 0, CODE_A = 1, CODE_C = 2, CODE_T = 3, CODE_G = 4, CODE_N = 5;
static const int N_CODES = 6;

#ifdef __CUDACC__
__host__ __device__
#endif
static inline uint8_t base_code(uint8_t b)
{
    switch (b) {
    case 'C': return CODE_C;
    case 'T': return CODE_T;
    case 'G': return CODE_G;
    case 'N': return CODE_N;
    default:  return CODE_A;
    }
}

// One read of the workload.  `off` indexes the five device read pools.
struct ReadRec {
    int64_t off;
    int64_t out_base;    // result index of (this read, haplotype 0 of its batch)
    int32_t len;
    int32_t batch;
};

// Where the raw bases of one haplotype sit in the uploaded haplotype pool and where its code bytes go in the
// haplotype stream (phmm_code_stream_kernel).
struct HapSrc {
    int64_t src;         // offset into the device copy of the referenced haplotype-pool pieces
    int64_t dst;         // first column in the haplotype stream
    int32_t len;
    int32_t pad_;
};

// One unit of work for one warp: up to MAX_GROUP reads streamed against the same run of
// consecutive haplotypes of their batch.  FP32 kernels hold two reads per lane group (one per
// f32x2 half): read[2 * group + half]; FP64 kernels one: read[group].  The FP32 kernels also emit
// these (one read, one haplotype) as FP64-rescue / guard-band work.
static const int MAX_GROUP = 4;
static const int32_t SPEC_TASK = -2;          // Task::read[1] of an FP64-first task (FP64 kernels read read[0] only)
struct Task {
    int64_t stream_off;  // first column of the run in the haplotype stream
    int32_t ncols;       // columns incl. SEP_COLS separator columns after every haplotype
    int32_t hap0;        // index of the run's first haplotype in the haplotype tables
    int32_t out0;        // added to ReadRec::out_base: index of that haplotype within its batch
    int32_t read[MAX_GROUP];   // ReadRec indices, -1 = slot idle
    int32_t nh;          // haplotypes in the run (<= MAX_RUN_HAPS)
};

// Unit of work of the fast FP32 "haplotype-pair" kernels: each lane group holds ONE read and
// streams it against TWO runs of haplotypes at once, one in each half of the packed f32x2
// registers.  Both halves then share every per-row constant, which FFMA2/FMUL2 take as a single
// 32-bit broadcast operand (Rn.F32) instead of a register pair.
static const int MAX_HP_GROUP = 4;
struct HTask {
    int64_t stream_off[2];   // first column of each run in the haplotype stream
    int32_t ncols[2];        // columns incl. separators; 0 = half idle
    int32_t hap0[2];         // first haplotype of each run (valid index even when idle)
    int32_t out0[2];         // index of that haplotype within its batch
    int16_t nh[2];           // haplotypes per run (<= MAX_RUN_HAPS)
    int32_t read[MAX_HP_GROUP];   // ReadRec index per lane group, -1 = idle
    int32_t pad_;
};

// Kernel class = how reads are laid over a warp.
//   kh    rows per thread (register tile height)
//   lanes threads per read (32 or 16); a warp works on 32/lanes lane groups at once
//   vl    reads per lane group: 2 = FP32, one read in each half of Blackwell's packed f32x2
//         registers (FMUL2/FFMA2); 1 = FP64
// capacity = kh * lanes - 1 read rows (row 0 of the first lane is always a dummy row that
// re-creates the DP's row-0 boundary, so no per-step select is needed there).
//   longr 1 = multi-pass kernel for reads longer than any single-pass class: the read is cut
//         into blocks of kh * lanes rows, processed one block per pass over the same columns;
//         the bottom row of a block travels to the next pass through an HBM scratch row
//         (the reference does the same one level down with its shiftOut arrays,
//         avx-pairhmm-template.h:224,265-297).
//   hp    1 = haplotype-pair kernel (fast FP32 only): one read per lane group, vl counts the
//         haplotype runs instead; lanes 32, 16 or 8, kh <= MAX_KH_HP
struct KClass {
    int kh, lanes, vl, longr, hp;
    int capacity() const { return longr ? 0x7fffffff : kh * lanes - 1; }
    int group() const { return hp ? 32 / lanes : (32 / lanes) * vl; }      // reads per task
};

static const int SEP_COLS = 4;                // separator columns after each haplotype (= columns per loop iteration)
static const int MAX_RUN_HAPS = 62;           // haplotypes per run: per-warp table of INITIAL_CONSTANT / haplen
static const int STREAM_CHUNK = 512;          // haplotype-stream bytes staged per cp.async round
static const int STREAM_RING = 2048;          // per-warp shared-memory ring (4 chunks)
static const int STREAM_SLACK = 2 * STREAM_CHUNK + 128;   // readable bytes past the last column
static const int WARPS_PER_BLOCK = 1;          // one warp per block: a persistent block frees its SM slot as soon as its
                                               // warp runs out of tasks, so the next kernel's blocks fill the tail
static const int MAX_KH = 8;                  // single-pass classes: kh <= 8 (255 rows at 32 lanes)
static const int MAX_KH_HP = 16;              // haplotype-pair classes: kh <= 16 (511 rows at 32 lanes)
static const int LONG_KH = 8;                 // register tile of the multi-pass kernels
static const int MAX_KH_SCALAR = MAX_KH;      // FP64 rescue classes: 32 lanes x kh rows

// dynamic shared memory per warp: emission table [kh][N_CODES][32 lanes] of 8 bytes, the stream
// ring, and the per-run haplotype table (INITIAL_CONSTANT / haplen as 8 bytes + length)
static inline int smem_per_warp(int kh) { return kh * N_CODES * 256 + STREAM_RING + (MAX_RUN_HAPS + 2) * 12; }
// haplotype-pair kernels: emission table [N_CODES][kh / 4][32 lanes] of 16 bytes, two stream rings, two
// run tables (float INITIAL_CONSTANT / haplen + int length)
// + the pending FP64-rescue run of every (lane group, haplotype run): 8 x 24 bytes
static inline int smem_per_warp_hp(int kh) { return ((kh + 3) / 4) * N_CODES * 512 + 2 * STREAM_RING + 2 * (MAX_RUN_HAPS + 2) * 8 + 8 * 24; }

template <typename T>
struct KernelParams {
    const uint8_t *rs, *q, *i, *d, *c;   // read pools
    const ReadRec *reads;
    const uint8_t *stream;               // one-hot coded haplotype columns + separators
    const int32_t *hap_len;              // per haplotype, stream order
    const Task *tasks;
    const HTask *htasks;                 // haplotype-pair kernels
    const int32_t *ntasks_ptr;           // task count lives on the device (rescue lists)
    int32_t *counter;                    // dynamic task queue head
    const T *ph2pr;                      // [128]
    const T *mm;                         // matchToMatchProb prefix [8256] (quals are & 127)
    T init_const;                        // 2^120 or 2^1020
    T log10_init;                        // log10f(2^120) / log10(2^1020), from the host libm
    double *out;                         // log10 likelihoods
    float *raw32;                        // FP32 sums
    double *raw64;                       // FP64 sums (rescued pairs)
    uint8_t *rescued;
    Task *resc;                          // FP32 kernel: rescue list, FP64 kernel: unused
    int32_t *resc_count;                 // [MAX_KH_SCALAR + 1] one list per scalar kernel class (tasks)
    int32_t *resc_pairs;                 // pairs in those tasks (a task may hold a run of rescued haplotypes)
    int32_t resc_off[MAX_KH_SCALAR + 1]; // start of each class's list inside resc / guard
    int32_t mode;                        // PHMM_OPT_PRECISION
    int32_t resc_force;                  // > 0: every rescue / guard task goes to this scalar class's list (one-wave calls,
                                         // Plan::resc_force); 0: by read length
    Task *guard;                         // guard-band list (fast kernel only)
    int32_t *guard_count;
    // FP64-first ("speculative") tasks, FP64 kernels only: a read the planner predicts to fail the FP32 pass against
    // every haplotype of its batch skips that pass; its tasks (read[1] == SPEC_TASK) sit in the rescue lists from the
    // start.  The FP64 sum, rescaled to the FP32 recurrence's 2^120, tells what the FP32 pass would have decided:
    // clearly below MIN_ACCEPTED -> rescued, as predicted; clearly above -> the FP32 result is that value to ~1e-5,
    // finalised here in float; within the band around the threshold -> spec_list, where the bit-exact FP32 kernel
    // decides (mode 3).  See phmm_kernels.cuh: emit_result.
    Task *spec_list;
    int32_t *spec_count;                 // [MAX_KH_SCALAR + 1] tasks per class in spec_list, then [+ 9] pairs rescued as predicted,
                                         // [+ 10] pairs finalised in float
    int32_t spec_off[MAX_KH_SCALAR + 1];
    float log10_init32;                  // log10f(2^120)
    float spec_band;                     // > 0: relative half-width of the band is max(this, (R + H) * 4.768e-6)
    T *scratch;                          // multi-pass kernels: per-warp boundary rows
    int64_t scratch_stride;              // elements of T per warp (2 buffers x 3 arrays x columns)
};

}  // namespace phmm

#endif
