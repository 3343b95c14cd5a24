// Internal host/device shared types of libphmm_cuda (not part of the C ABI).
#ifndef PHMM_TYPES_H
#define PHMM_TYPES_H

#include <stdint.h>

namespace phmm {

// Match code per base, one-hot so that "read base matches haplotype base" is (rb & hb) != 0 and
// N on either side matches everything -- the rule precompute_masks_ builds
// (avx-pairhmm-template.h:3-35) on top of ConvertChar (pairhmm_common.h:30-43; every byte that
// is not ACGTN converts to 0 = 'A').  0 never matches: it marks separator / padding columns.
static const uint8_t CODE_A = 1, CODE_C = 2, CODE_T = 4, CODE_G = 8, CODE_N = 15;

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

// One unit of work for one warp: up to MAX_GROUP reads (one per lane group) streamed against
// the same run of haplotype-stream columns.  The FP32 kernel also emits these (one read, one
// haplotype) as FP64-rescue / guard-band work.
static const int MAX_GROUP = 4;
struct Task {
    int64_t stream_off;  // first column of the run in the haplotype stream
    int32_t ncols;       // columns incl. one separator after every haplotype
    int32_t hap0;        // index of the run's first haplotype in the haplotype tables
    int32_t out0;        // added to ReadRec::out_base: index of that haplotype within its batch
    int32_t read[MAX_GROUP];   // ReadRec indices, -1 = lane group idle
    int32_t pad;
};

// Kernel class = how a read is laid over a warp.
//   kh    rows per virtual lane (register tile height)
//   lanes threads per read (32, 16 or 8); a warp works on 32/lanes reads at once
//   vl    virtual lanes per thread: 2 = FP32 packed f32x2 (Blackwell FMUL2/FFMA2), 1 = scalar
// capacity = vl * kh * lanes - 1 read rows (row 0 of the first virtual lane is always a dummy
// row that re-creates the DP's row-0 boundary, so no per-step select is needed there).
//   longr 1 = multi-pass kernel for reads longer than any single-pass class: the read is cut
//         into blocks of vl * kh * lanes rows, processed one block per pass over the same columns;
//         the bottom row of a block travels to the next pass through an HBM scratch row
//         (the reference does the same one level down with its shiftOut arrays,
//         avx-pairhmm-template.h:224,265-297).
struct KClass {
    int kh, lanes, vl, longr;
    int capacity() const { return longr ? 0x7fffffff : vl * kh * lanes - 1; }
    int group() const { return 32 / lanes; }
};

static const int STREAM_CHUNK = 512;          // haplotype-stream bytes staged per cp.async round
static const int STREAM_RING = 2048;          // per-warp shared-memory ring (4 chunks)
static const int STREAM_SLACK = 2 * STREAM_CHUNK + 128;   // readable bytes past the last column
static const int WARPS_PER_BLOCK = 4;
static const int LONG_KH_F32 = 4, LONG_KH_F32S = 8, LONG_KH_F64 = 4;   // register tiles of the multi-pass kernels
static const int MAX_KH_SCALAR = 8;           // scalar single-pass classes: 32 lanes x kh rows, kh <= 8

template <typename T>
struct KernelParams {
    const uint8_t *rs, *q, *i, *d, *c;   // read pools
    const ReadRec *reads;
    const uint8_t *stream;               // one-hot coded haplotype columns + separators
    const int32_t *hap_len;              // per haplotype, stream order
    const Task *tasks;
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
    int32_t *resc_count;                 // [MAX_KH_SCALAR + 1] one list per scalar kernel class
    int32_t resc_off[MAX_KH_SCALAR + 1]; // start of each class's list inside resc / guard
    int32_t mode;                        // PHMM_OPT_PRECISION
    float guard_lo, guard_hi;            // fast kernel: sums in [lo, hi) go to the exact kernel
    Task *guard;                         // guard-band list (fast kernel only)
    int32_t *guard_count;
    T *scratch;                          // multi-pass kernels: per-warp boundary rows
    int64_t scratch_stride;              // elements of T per warp (2 buffers x 3 arrays x columns)
};

}  // namespace phmm

#endif
