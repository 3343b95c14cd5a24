// Pair-HMM forward kernels for sm_100a (B200).
//
// What is computed: compute_full_prob_<float|double> of the reference
// (tools/GKL/src/main/native/pairhmm/avx-pairhmm-template.h:210-346) followed by the FP32 ->
// FP64 rescue policy and log10 finalisation of computelikelihoodsboth
// (tools/GKL/src/main/native/pairhmm/IntelPairHmmCSource.cpp:61-85).
//
// How (not a port of the AVX striping):
//   * One lane group (32 or 16 threads) streams its reads against a whole run of haplotypes of
//     their batch, laid end to end in HBM as code bytes with SEP_COLS separator columns after each
//     haplotype.  The anti-diagonal wavefront therefore fills once per run instead of once per
//     pair, and the reads' per-row constants stay in registers for the run.
//   * A thread owns KH consecutive read rows.  FP32: of TWO reads at once, one in each half of
//     Blackwell's packed f32x2 registers, so every multiply of a cell update issues as FMUL2 (and
//     everything as FFMA2 in the fast kernel) and no value ever crosses between the halves.
//     FP64: of one read.
//   * Lane l works on column (step - l).  The bottom row of a thread moves to the next lane
//     through __shfl_up_sync.
//   * The emission factor of a cell, (match ? 1 - eps : eps / 3), is not selected per cell: it is
//     read from a per-warp shared-memory table e[row][code][lane] (one 64-bit LDS per row and
//     step, conflict-free), indexed by the haplotype code of the lane's current column.  Code 0
//     (separator) maps to 0 for every row, so M and X clear themselves in separator columns.
//   * Read rows are right-aligned: unused rows at the top are "dummy rows" whose constants make
//     them reproduce the DP's row-0 boundary (M = X = 0, Y = INITIAL_CONSTANT / haplen) exactly,
//     so the first real row needs no special case and the result is always the last row of the
//     last lane.
//   * The loop body handles SEP_COLS = 4 columns: one 32-bit word of haplotype codes per thread
//     and iteration.  A lane passes a separator run in four consecutive steps, exactly one of
//     which is the last step of an iteration: there it resets Y to the next haplotype's column-0
//     boundary (and the last lane emits the finished pair).  The values it carries through the
//     rest of the separator run stay at the boundary because M = X = 0 there.
//   * Haplotype columns are staged per warp into a shared-memory ring with 128-bit cp.async
//     copies one 512-column chunk ahead of use.
//   * Bit-exact kernel: __fmul_rn/__fadd_rn semantics, the reference's association order
//     ((M*a + X*b) + Y*b) * e (:188), sumM and sumX accumulated separately in column order
//     (:328,:331,:342).  ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even though both
//     carry .rn, so the bit-exact kernel pairs FMUL2 with scalar FADD, which it never fuses.
#ifndef PHMM_KERNELS_CUH
#define PHMM_KERNELS_CUH
#include <cuda_runtime.h>
#include <stdint.h>

#include "phmm_types.h"

namespace phmm {

#define FULL_MASK 0xffffffffu
#define MIN_ACCEPTED_F 1e-28f   // pairhmm_common.h:16

// ------------------------------------------------------------------------------------------
// Arithmetic back-ends: V = register type holding VL values of T.
// ------------------------------------------------------------------------------------------
template <typename T, bool FAST>
struct Ops;

// FP32, two reads packed in f32x2.
template <bool FAST>
struct Ops<float, FAST> {
    typedef float2 V;
    static constexpr int VL = 2;
    static __device__ __forceinline__ V splat(float x) { return make_float2(x, x); }
    static __device__ __forceinline__ float get(const V &a, int c) { return c ? a.y : a.x; }
    static __device__ __forceinline__ void set(V &a, int c, float x) { if (c) a.y = x; else a.x = x; }
    static __device__ __forceinline__ V mul(V a, V b) { return __fmul2_rn(a, b); }
    static __device__ __forceinline__ V add(V a, V b)
    {
        if (FAST) return __fadd2_rn(a, b);
        // scalar adds: ptxas would fuse FMUL2 + FADD2 into FFMA2
        return make_float2(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y));
    }
    static __device__ __forceinline__ V fma(V a, V b, V c) { return __ffma2_rn(a, b, c); }
    static __device__ __forceinline__ V shfl_up(V v, int width)
    {
        return make_float2(__shfl_up_sync(FULL_MASK, v.x, 1, width), __shfl_up_sync(FULL_MASK, v.y, 1, width));
    }
    static __device__ __forceinline__ V lds(uint32_t addr)
    {
        V v;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
        return v;
    }
};

// FP64 (rescue path), one read.
template <bool FAST>
struct Ops<double, FAST> {
    typedef double V;
    static constexpr int VL = 1;
    static __device__ __forceinline__ V splat(double x) { return x; }
    static __device__ __forceinline__ double get(const V &a, int) { return a; }
    static __device__ __forceinline__ void set(V &a, int, double x) { a = x; }
    static __device__ __forceinline__ V mul(V a, V b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ V add(V a, V b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ V fma(V a, V b, V c) { return __fma_rn(a, b, c); }
    static __device__ __forceinline__ V shfl_up(V v, int width) { return __shfl_up_sync(FULL_MASK, v, 1, width); }
    static __device__ __forceinline__ V lds(uint32_t addr)
    {
        V v;
        asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
        return v;
    }
};

template <typename T> struct Scalar;
template <> struct Scalar<float> {
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
};
template <> struct Scalar<double> {
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
};

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.wait_all;\n" ::: "memory");
}

// FP64 / guard kernel class index of a read (list the FP32 kernel appends rescue work to):
// 32 lanes, kh rows per lane, capacity 32*kh - 1.  0 = longer than any single-pass class.
__host__ __device__ __forceinline__ int rescue_class(int rslen)
{
    int kh = (rslen + 1 + 31) / 32;
    return kh <= MAX_KH_SCALAR ? kh : 0;
}

// the list a kernel appends rescue / guard work of a read to
template <typename T>
__device__ __forceinline__ int list_class(const KernelParams<T> &P, int rslen)
{
    return P.resc_force > 0 ? P.resc_force : rescue_class(rslen);
}

// ------------------------------------------------------------------------------------------
// Result of one (read, haplotype): policy of computelikelihoodsboth (IntelPairHmmCSource.cpp:69-82).
// ------------------------------------------------------------------------------------------
// Decision for one (read, haplotype) FP32 sum: 0 = final (log10 written), 1 = FP64 rescue, 2 = re-decide with the
// bit-exact FP32 kernel (fast kernels only).
template <bool FAST>
__device__ __forceinline__ int classify_result(const KernelParams<float> &P, float res, int64_t oidx, int hap_len, int rslen)
{
    P.raw32[oidx] = res;
    bool redo = false;
    if (FAST) {
        const float rel = fmaxf(0.0078125f, (float)(rslen + hap_len) * 2.384e-6f);      // see emit_result
        redo = !(res <= 3.402823466e+38f) ||
               (P.mode == 0 && res >= MIN_ACCEPTED_F * (1.0f - rel) && res < MIN_ACCEPTED_F * (1.0f + rel));
    }
    if (redo) return 2;
    if (P.mode == 0 && res < MIN_ACCEPTED_F) return 1;                                    // :73-75
    P.out[oidx] = (double)__fsub_rn(log10f(res), P.log10_init);                           // :78 (and :97)
    P.rescued[oidx] = 0;
    return 0;
}

// A run of consecutive rescued haplotypes of one read, collected by the last lane of a lane group while it walks a
// run of haplotypes: the FP64 kernel then fills its wavefront and sets the read up once per run instead of once per
// pair (C2-shaped input: most rescued pairs belong to reads that fit none of their batch's haplotypes).
struct RescRun {
    int64_t stream_off;
    int32_t ncols, hap0, out0, nh;
};
static const int RESC_RUN_HAPS = 4;        // FP64 tasks stay short enough for the queue to balance them
static const int RESC_RUN_COLS = 2048;

__device__ __forceinline__ void resc_flush(const KernelParams<float> &P, RescRun &pr, int rid, int rslen)
{
    Task nt;
    nt.stream_off = pr.stream_off;
    nt.ncols = pr.ncols;
    nt.hap0 = pr.hap0;
    nt.out0 = pr.out0;
    nt.read[0] = rid; nt.read[1] = -1; nt.read[2] = -1; nt.read[3] = -1;
    nt.nh = pr.nh;
    const int ci = list_class(P, rslen);
    const int k = atomicAdd(&P.resc_count[ci], 1);
    P.resc[P.resc_off[ci] + k] = nt;
    atomicAdd(P.resc_pairs, pr.nh);
    pr.nh = 0;
}

template <bool FAST>
__device__ __forceinline__ void emit_result(const KernelParams<float> &P, float res, int64_t oidx,
                                            int64_t hap_stream_off, int hap_len, int hap_idx,
                                            int out_rel, int rid, int rslen)
{
    P.raw32[oidx] = res;
    // Fast kernel: a sum inside the guard band around MIN_ACCEPTED, or a non-finite one (overflow
    // of the scaled recurrence), is re-decided by the bit-exact FP32 kernel.
    // The band is a function of the pair alone (|contracted - unfused| <= ~5 (R + H) ulp relative on an
    // all-positive recurrence; 8x that, never less than 2^-7), so that a pair's result does not depend on what
    // else is in the workload or on how a host-buffer call was cut into chunks.
    bool redo = false;
    if (FAST) {
        const float rel = fmaxf(0.0078125f, (float)(rslen + hap_len) * 2.384e-6f);
        redo = !(res <= 3.402823466e+38f) ||
               (P.mode == 0 && res >= MIN_ACCEPTED_F * (1.0f - rel) && res < MIN_ACCEPTED_F * (1.0f + rel));
    }
    if (P.mode == 3 && res < MIN_ACCEPTED_F) {
        // deciding an FP64-first pair whose rescaled FP64 sum fell inside the band: the FP32 pass fails, the FP64
        // result the pair already holds stands (KernelParams::spec_list)
        atomicAdd(P.resc_pairs, 1);
        return;
    }
    if (redo || (P.mode == 0 && res < MIN_ACCEPTED_F)) {
        // :73-75 -- below MIN_ACCEPTED the pair is re-run in FP64 (by the FP64 kernel of the
        // read's class).
        Task nt;
        nt.stream_off = hap_stream_off;
        nt.ncols = hap_len + SEP_COLS;
        nt.hap0 = hap_idx;
        nt.out0 = out_rel;
        nt.read[0] = rid; nt.read[1] = -1; nt.read[2] = -1; nt.read[3] = -1;
        nt.nh = 1;
        const int ci = list_class(P, rslen);
        if (redo) {
            int k = atomicAdd(&P.guard_count[ci], 1);
            P.guard[P.resc_off[ci] + k] = nt;
        } else {
            int k = atomicAdd(&P.resc_count[ci], 1);
            P.resc[P.resc_off[ci] + k] = nt;
            atomicAdd(P.resc_pairs, 1);
        }
        return;
    }
    // :78 (and :97) -- log10 in float, subtraction in float, then widened
    P.out[oidx] = (double)__fsub_rn(log10f(res), P.log10_init);
    P.rescued[oidx] = 0;
}

template <bool FAST>
__device__ __forceinline__ void emit_result(const KernelParams<double> &P, double res, int64_t oidx,
                                            int64_t hap_stream_off, int hap_len, int hap_idx,
                                            int out_rel, int rid, int rslen, bool spec = false)
{
    if (spec && P.mode == 0) {
        // FP64-first pair: what would the FP32 pass have summed?  The same recurrence from 2^120 instead of 2^1020;
        // its rounding error is <= ~5 (R + H) ulp relative on an all-positive recurrence, underflow only lowers it by
        // < 1e-33 absolute.  Band = 16 x that bound, never less than spec_band (2^-6).
        const double est = res * 0x1p-900;
        const double rel = fmax((double)P.spec_band, (double)(rslen + hap_len) * 4.768e-6);
        const double thr = (double)MIN_ACCEPTED_F;
        if (!(est < thr * (1.0 - rel))) {
            if (est >= thr * (1.0 + rel) && est <= 3.0e38) {
                // not rescued after all: the FP32 path's result (:78), from the rescaled sum
                const float s32 = (float)est;
                P.raw32[oidx] = s32;
                P.out[oidx] = (double)__fsub_rn(log10f(s32), P.log10_init32);
                P.rescued[oidx] = 0;
                atomicAdd(P.spec_count + MAX_KH_SCALAR + 2, 1);
                return;
            }
            // inside the band (or not finite): the FP64 result below stands unless the bit-exact FP32 kernel, which
            // runs on spec_list afterwards, says the pair passes
            Task nt;
            nt.stream_off = hap_stream_off;
            nt.ncols = hap_len + SEP_COLS;
            nt.hap0 = hap_idx;
            nt.out0 = out_rel;
            nt.read[0] = rid; nt.read[1] = -1; nt.read[2] = -1; nt.read[3] = -1;
            nt.nh = 1;
            const int ci = list_class(P, rslen);
            const int k = atomicAdd(&P.spec_count[ci], 1);
            P.spec_list[P.spec_off[ci] + k] = nt;
            // (a non-finite sum of the fast FP64 kernel also takes the branch below to the bit-exact FP64 re-run, which
            // wl_finish launches before the FP32 decision pass)
        } else {
            atomicAdd(P.spec_count + MAX_KH_SCALAR + 1, 1);
        }
        P.raw32[oidx] = (float)est;
    }
    // Fast FP64 kernel (scaled recurrence): X / sigma can overflow for exotic gap-continuation
    // penalties; a non-finite sum is re-run by the bit-exact FP64 kernel of the read's class.
    if (FAST && P.guard && !(res <= 1.7976931348623157e+308)) {
        Task nt;
        nt.stream_off = hap_stream_off;
        nt.ncols = hap_len + SEP_COLS;
        nt.hap0 = hap_idx;
        nt.out0 = out_rel;
        nt.read[0] = rid; nt.read[1] = -1; nt.read[2] = -1; nt.read[3] = -1;
        nt.nh = 1;
        const int ci = list_class(P, rslen);
        const int k = atomicAdd(&P.guard_count[ci], 1);
        P.guard[P.resc_off[ci] + k] = nt;
        return;
    }
    // :75 (and :109)
    P.raw64[oidx] = res;
    P.out[oidx] = __dsub_rn(log10(res), P.log10_init);
    P.rescued[oidx] = (P.mode == 0) ? 1 : 0;
}

template <bool FAST>
__device__ __forceinline__ void emit_result_spec(const KernelParams<double> &P, double res, int64_t oidx, int64_t hap_stream_off,
                                                 int hap_len, int hap_idx, int out_rel, int rid, int rslen, bool spec)
{
    emit_result<FAST>(P, res, oidx, hap_stream_off, hap_len, hap_idx, out_rel, rid, rslen, spec);
}
template <bool FAST>
__device__ __forceinline__ void emit_result_spec(const KernelParams<float> &P, float res, int64_t oidx, int64_t hap_stream_off,
                                                 int hap_len, int hap_idx, int out_rel, int rid, int rslen, bool)
{
    emit_result<FAST>(P, res, oidx, hap_stream_off, hap_len, hap_idx, out_rel, rid, rslen);
}

// ------------------------------------------------------------------------------------------
// The forward kernel.  Persistent warps pull tasks from a device-side queue.
// ------------------------------------------------------------------------------------------
template <typename T, int KH, int LANES, bool FAST, bool LONG, int MINB>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MINB)
phmm_forward_kernel(const KernelParams<T> P)
{
    typedef Ops<T, FAST> O;
    typedef typename O::V V;
    typedef Scalar<T> S;
    constexpr int VL = O::VL;
    // Scaled recurrence of the fast FP32 single-pass kernels (6 + NTOP / KH packed FP instructions
    // per cell instead of 8; see "scaled recurrence" in the cell update below).
    constexpr bool SCALED = FAST && !LONG;      // FP32 and FP64: a non-finite sum goes to the bit-exact kernel
    constexpr int NTOP = (KH + 3) / 4;                 // groups of chained rows per thread
    constexpr int GRP = (KH + NTOP - 1) / NTOP;        // rows per group (<= 4): group tops at p % GRP == 0
    constexpr int ROWS = KH * LANES;          // read rows per pass
    constexpr int ETAB = KH * N_CODES * 256;  // bytes
    constexpr int WSM = ETAB + STREAM_RING + (MAX_RUN_HAPS + 2) * 12;
    constexpr uint32_t WMASK = STREAM_RING / 4 - 1;
    static_assert(sizeof(V) == 8, "emission table entries are 8 bytes");

    extern __shared__ __align__(16) uint8_t smem_all[];
    const int lane = threadIdx.x & 31;
    uint8_t *etab = smem_all + (threadIdx.x >> 5) * WSM;
    uint8_t *ring = etab + ETAB;
    const uint32_t *ring32 = reinterpret_cast<const uint32_t *>(ring);
    uint8_t *iytab = ring + STREAM_RING;                                   // 8-byte slots holding a T
    int32_t *hltab = reinterpret_cast<int32_t *>(iytab + (MAX_RUN_HAPS + 2) * 8);
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(etab) + lane * 8;
    const int gl = lane & (LANES - 1);        // lane within its group
    const int grp = lane / LANES;
    const bool last_lane = (gl == LANES - 1);
    const int ntasks = *P.ntasks_ptr;
    const T zero = (T)0;
    const uint32_t sh = ((4 - (gl & 3)) & 3) * 8;

    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(P.counter, 1);
        t = __shfl_sync(FULL_MASK, t, 0);
        if (t >= ntasks) break;

        const Task *tk = P.tasks + t;
        const int64_t stream_off = tk->stream_off;
        const int ncols = tk->ncols;
        const int hap0 = tk->hap0;
        const int out0 = tk->out0;
        const int nh = tk->nh;
        int rid[VL], R[VL];
        int64_t roff[VL], out_base[VL];
        int rmax = 0;
#pragma unroll
        for (int c = 0; c < VL; ++c) {
            rid[c] = tk->read[grp * VL + c];
            R[c] = 0; roff[c] = 0; out_base[c] = 0;
            if (rid[c] >= 0) {
                const ReadRec rr = P.reads[rid[c]];
                R[c] = rr.len; roff[c] = rr.off; out_base[c] = rr.out_base;
            }
            rmax = max(rmax, R[c]);
        }
        // per-run haplotype table: column-0 boundary INITIAL_CONSTANT / haplen (:86) and length
        __syncwarp();
        for (int j = lane; j <= nh; j += 32) {
            const int hl = P.hap_len[hap0 + j];
            *reinterpret_cast<T *>(iytab + 8 * j) = S::div(P.init_const, (T)hl);
            hltab[j] = hl;
        }

        // Multi-pass (LONG) kernels cut the reads into blocks of ROWS rows; block `pass` reads the
        // previous block's bottom row from spill_in and leaves its own in spill_out
        // (index 0 = column-0 boundary of the first haplotype, index c + 1 = column c).
        int npass = 1;
        V *spill_in = nullptr, *spill_out = nullptr;
        int64_t spill_cols = 0;
        if (LONG) {
            npass = (__reduce_max_sync(FULL_MASK, rmax) + ROWS) / ROWS;
            const int64_t wslot = (int64_t)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
            spill_cols = P.scratch_stride / (6 * VL);
            spill_in = reinterpret_cast<V *>(P.scratch + wslot * P.scratch_stride);
            spill_out = spill_in + 3 * spill_cols;
        }
      for (int pass = 0; pass < npass; ++pass) {
        const bool final_pass = (pass == npass - 1);
        // ---- per-row constants (initializeVectors :83-128, stripeINITIALIZATION :151-158) and
        //      the emission table ----------------------------------------------------------------
        V pMM[KH], pG[KH], pMX[KH], pXX[KH], pMY[KH];
        V pYY0;                       // pYY of row 0 (see dummy rows below)
        uint32_t dmask[VL];
#pragma unroll
        for (int c = 0; c < VL; ++c) dmask[c] = 0;
        __syncwarp();                 // every lane is done with the previous table
#pragma unroll
        for (int p = 0; p < KH; ++p) {
            T em[VL], ex[VL];
            int rcode[VL];
#pragma unroll
            for (int c = 0; c < VL; ++c) {
                const int nd = npass * ROWS - R[c];                       // dummy rows on top; >= 1
                const int r = pass * ROWS + gl * KH + p - nd;             // 0-based read position
                T mm = zero, g = zero, mx = zero, xx = zero, my = zero, yy = zero;
                em[c] = zero; ex[c] = zero; rcode[c] = -1;
                if (r >= 0) {
                    const int _q = P.q[roff[c] + r] & 127, _i = P.i[roff[c] + r] & 127;
                    const int _d = P.d[roff[c] + r] & 127, _c = P.c[roff[c] + r] & 127;
                    const int mxq = max(_i, _d), mnq = min(_i, _d);
                    mm = P.mm[((mxq * (mxq + 1)) >> 1) + mnq];        // set_mm_prob, Context.h:120-132
                    const T pc = P.ph2pr[_c];
                    g = S::sub((T)1, pc);                             // :115
                    mx = P.ph2pr[_i]; my = P.ph2pr[_d]; xx = pc; yy = pc;
                    const T eps = P.ph2pr[_q];
                    em[c] = S::sub((T)1, eps);                        // :156
                    ex[c] = S::div(eps, (T)3);                        // :158 true division
                    const uint8_t b = P.rs[roff[c] + r];
                    rcode[c] = (b == 'C') ? CODE_C : (b == 'T') ? CODE_T : (b == 'G') ? CODE_G
                             : (b == 'N') ? CODE_N : CODE_A;
                } else {
                    // Dummy row: stays at (M, X, Y) = (0, 0, INITIAL_CONSTANT/haplen), the row-0
                    // boundary (:86,:162): emission 0 kills M, pMX = 0 and a zero "up" X keep X = 0,
                    // pMY = 0 and pYY = 1 carry Y along the row exactly.  pXX doubles as pYY in
                    // real rows; only row 0 of a thread can see a non-zero "up" X (lane 0 receives
                    // its own registers from the shuffle), so it gets pXX = 0 and its own pYY.
                    // (Scaled recurrence: pXX only multiplies Y there, the "up" X of row 0 is cut
                    // through K2 instead, and a dummy row's pGAPM of 1 is what the row above folds
                    // into its Y.)
                    xx = (p == 0 && !SCALED) ? zero : (T)1;
                    yy = (T)1;
                    if (SCALED) g = (T)1;
                    dmask[c] |= 1u << p;
                }
                O::set(pMM[p], c, mm); O::set(pG[p], c, g); O::set(pMX[p], c, mx);
                O::set(pXX[p], c, xx); O::set(pMY[p], c, my);
                if (p == 0) O::set(pYY0, c, yy);
            }
            // e[p][code]: match <=> equal or either is N (precompute_masks_ :3-35); separator -> 0
#pragma unroll
            for (int code = 0; code < N_CODES; ++code) {
                V ev;
#pragma unroll
                for (int c = 0; c < VL; ++c) {
                    const bool match = (code == rcode[c]) || (code == CODE_N) || (rcode[c] == CODE_N);
                    O::set(ev, c, (code == CODE_SEP || rcode[c] < 0) ? zero : (match ? em[c] : ex[c]));
                }
                *reinterpret_cast<V *>(etab + (p * N_CODES + code) * 256 + lane * 8) = ev;
            }
        }

        // ---- scaled recurrence: fold the constants ---------------------------------------------
        // State kept per row r (gn = pGAPM of row r + 1, 1 below the last row; sigma_r = 1 on a
        // group top, else pXX_r * sigma_{r-1}):
        //     Xs = X / sigma_r            Ys = Y * gn            M as is
        // so that  Z  = gn * (X + Y)                     = fma(Xs, gn * sigma_r, Ys)
        //          M' = (M_d * pMM + Z_d) * e            (Z_d of the row above, previous column)
        //          Ys'= M_l * (pMY * gn) + Ys_l * pYY
        //          Xs'= M_u * (pMX / sigma_r) + Xs_u                       chained row: one FMA
        //          Xs'= M_u * pMX + Xs_u * (pXX * sigma_{r-1})             group top
        // Registers: pMM = pMM, pG = gn * sigma, pMX = pMX / sigma, pMY = pMY * gn, pXX = pYY,
        // K2 = pXX * sigma_{r-1} of the group tops.  The last row's X sum is rescaled by `fin`.
        // sigma >= pXX^3 (1e-3 at the usual GCP of Q10): an overflow of X / sigma makes the pair's
        // sum non-finite, and emit_result hands such pairs to the bit-exact kernel.
        V K2[NTOP];
        V fin = O::splat((T)1);
        if (SCALED) {
            V gdn, sig[KH], sup;
#pragma unroll
            for (int c = 0; c < VL; ++c) {
                const T v = __shfl_down_sync(FULL_MASK, O::get(pG[0], c), 1, LANES);
                O::set(gdn, c, last_lane ? (T)1 : v);
            }
#pragma unroll
            for (int p = 0; p < KH; ++p)
#pragma unroll
                for (int c = 0; c < VL; ++c)
                    O::set(sig[p], c, (p % GRP == 0) ? (T)1 : O::get(pXX[p], c) * O::get(sig[p - (p % GRP ? 1 : 0)], c));
            sup = O::shfl_up(sig[KH - 1], LANES);
#pragma unroll
            for (int p = 0; p < KH; ++p)
#pragma unroll
                for (int c = 0; c < VL; ++c) {
                    const T gn = (p < KH - 1) ? O::get(pG[p < KH - 1 ? p + 1 : p], c) : O::get(gdn, c);
                    const T s = O::get(sig[p], c);
                    if (p % GRP == 0) {
                        const T sprev = p ? O::get(sig[p ? p - 1 : 0], c) : O::get(sup, c);
                        T k2 = O::get(pXX[p], c) * sprev;
                        if (p == 0 && (dmask[c] & 1u)) k2 = zero;   // nothing above row 0 of the tile
                        O::set(K2[p / GRP], c, k2);
                    }
                    O::set(pMX[p], c, S::div(O::get(pMX[p], c), s));
                    O::set(pMY[p], c, O::get(pMY[p], c) * gn);
                    O::set(pG[p], c, gn * s);
                }
            fin = sig[KH - 1];
        }

        // ---- state: column-0 boundary -------------------------------------------------------
        int jh = 0;
        V M[KH], X[KH], Y[KH];
        __syncwarp();
        {
            const T iy = *reinterpret_cast<const T *>(iytab);
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                M[p] = O::splat(zero); X[p] = O::splat(zero);
#pragma unroll
                for (int c = 0; c < VL; ++c)
                    O::set(Y[p], c, ((dmask[c] >> p) & 1) ? (SCALED ? iy * O::get(pG[p], c) : iy) : zero);
            }
        }
        V U1m = O::splat(zero), U1x = O::splat(zero), U1y = O::splat(zero);
        V U2m, U2x, U2y;
        V sumM = O::splat(zero), sumX = O::splat(zero);
        int hapstart = 0;
        V sp_m = O::splat(zero), sp_x = O::splat(zero), sp_y = O::splat(zero);

        // ---- haplotype-stream ring -------------------------------------------------------------
        const int a = (int)(stream_off & 15);
        const uint8_t *sbase = P.stream + (stream_off - a);
        const int need = a + ncols + LANES + 16;    // ring bytes that must hold stream data
        const int s_begin = a & ~3;                 // s = step + a; lane gl is at column s - a - gl
        const int s_end = (a + ncols + LANES - 1 + 3) & ~3;
        if (lane < 4) *reinterpret_cast<uint4 *>(ring + STREAM_RING - 64 + 16 * lane) = make_uint4(0, 0, 0, 0);
        cp_async16(ring + 16 * lane, sbase + 16 * lane);
        if (STREAM_CHUNK < need) cp_async16(ring + STREAM_CHUNK + 16 * lane, sbase + STREAM_CHUNK + 16 * lane);

        if (LONG) {
            if (!final_pass && last_lane) {          // boundary entry for the block below
                spill_out[0] = M[KH - 1]; spill_out[spill_cols] = X[KH - 1]; spill_out[2 * spill_cols] = Y[KH - 1];
            }
            if (pass > 0 && gl == 0) {
                U1m = __ldcg(spill_in); U1x = __ldcg(spill_in + spill_cols); U1y = __ldcg(spill_in + 2 * spill_cols);
                const int idx = min(max(s_begin - a + 1, 0), ncols);
                sp_m = __ldcg(spill_in + idx); sp_x = __ldcg(spill_in + spill_cols + idx);
                sp_y = __ldcg(spill_in + 2 * spill_cols + idx);
            }
        }

        uint32_t wi = (uint32_t)((s_begin >> 2) - ((gl + 3) >> 2));
        uint32_t wprev = 0;
        for (int cb = 0, ch = 0; cb < s_end; cb += STREAM_CHUNK, ++ch) {
            cp_async_wait_all();
            __syncwarp();
            if (ch == 0) {
                if (lane < a) ring[lane] = 0;       // columns -1 .. -a of an unaligned run
                __syncwarp();
                wprev = ring32[wi & WMASK];
            }
            if ((ch + 2) * STREAM_CHUNK < need)
                cp_async16(ring + (((ch + 2) * STREAM_CHUNK) & (STREAM_RING - 1)) + 16 * lane,
                           sbase + (ch + 2) * STREAM_CHUNK + 16 * lane);
            const int s_hi = min(cb + STREAM_CHUNK, s_end);
            for (int s0 = max(cb, s_begin); s0 < s_hi; s0 += 4) {
                ++wi;
                const uint32_t wnext = ring32[wi & WMASK];
                const uint32_t codes = __funnelshift_r(wprev, wnext, sh);   // codes of columns s0..s0+3 - a - gl
                wprev = wnext;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    // -- emission factors of this column: table row (code << 8) + p * N_CODES * 256
                    const uint32_t eaddr = ebase + __byte_perm(codes, 0, 0x4404 | (k << 4));
                    V e[KH];
#pragma unroll
                    for (int p = 0; p < KH; ++p) e[p] = O::lds(eaddr + p * N_CODES * 256);
                    // -- neighbours: bottom row of the upstream lane, one and two steps ago
                    //    (scaled recurrence: U*y carries Z = gn * (X + Y) of that row)
                    U2m = U1m; U2x = U1x; U2y = U1y;
                    if (SCALED) {
                        const V zb = O::fma(X[KH - 1], pG[KH - 1], Y[KH - 1]);
                        U1m = O::shfl_up(M[KH - 1], LANES); U1x = O::shfl_up(X[KH - 1], LANES);
                        U1y = O::shfl_up(zb, LANES);
                    } else {
                        U1m = O::shfl_up(M[KH - 1], LANES); U1x = O::shfl_up(X[KH - 1], LANES);
                        U1y = O::shfl_up(Y[KH - 1], LANES);
                    }
                    if (LONG && pass > 0 && gl == 0) {      // first row of a later block: "up" is the
                        U1m = sp_m; U1x = sp_x; U1y = sp_y; // block above
                        const int idx = min(max(s0 + k - a + 2, 0), ncols);
                        sp_m = __ldcg(spill_in + idx); sp_x = __ldcg(spill_in + spill_cols + idx);
                        sp_y = __ldcg(spill_in + 2 * spill_cols + idx);
                    }
                    // -- cell updates (computeMXY :183-198)
                    V nM[KH], nX[KH], nY[KH];
                    if (SCALED) {
#pragma unroll
                        for (int p = 0; p < KH; ++p) {
                            const V dm = p ? M[p - 1] : U2m;
                            const V dz = p ? O::fma(X[p ? p - 1 : 0], pG[p ? p - 1 : 0], Y[p ? p - 1 : 0]) : U2y;
                            nM[p] = O::mul(O::fma(dm, pMM[p], dz), e[p]);
                            nY[p] = O::fma(M[p], pMY[p], O::mul(Y[p], pXX[p]));
                        }
#pragma unroll
                        for (int p = 0; p < KH; ++p) {
                            const V um = p ? nM[p ? p - 1 : 0] : U1m;
                            const V ux = p ? nX[p ? p - 1 : 0] : U1x;
                            if (p % GRP == 0) nX[p] = O::fma(um, pMX[p], O::mul(ux, K2[p / GRP]));
                            else              nX[p] = O::fma(um, pMX[p], ux);
                        }
                    } else {
#pragma unroll
                        for (int p = 0; p < KH; ++p) {
                            const V dm = p ? M[p - 1] : U2m;
                            const V dx = p ? X[p - 1] : U2x;
                            const V dy = p ? Y[p - 1] : U2y;
                            const V yy = p ? pXX[p] : pYY0;
                            if (FAST) {
                                nM[p] = O::mul(O::fma(dm, pMM[p], O::mul(O::add(dx, dy), pG[p])), e[p]);
                                nY[p] = O::fma(M[p], pMY[p], O::mul(Y[p], yy));
                            } else {
                                nM[p] = O::mul(O::add(O::add(O::mul(dm, pMM[p]), O::mul(dx, pG[p])),
                                                      O::mul(dy, pG[p])), e[p]);
                                nY[p] = O::add(O::mul(M[p], pMY[p]), O::mul(Y[p], yy));
                            }
                        }
#pragma unroll
                        for (int p = 0; p < KH; ++p) {
                            const V um = p ? nM[p - 1] : U1m;
                            const V ux = p ? nX[p - 1] : U1x;
                            if (FAST) nX[p] = O::fma(ux, pXX[p], O::mul(um, pMX[p]));
                            else      nX[p] = O::add(O::mul(um, pMX[p]), O::mul(ux, pXX[p]));
                        }
                    }
#pragma unroll
                    for (int p = 0; p < KH; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; }
                    // -- last-row sums (:328,:331): two separate sequential sums in column order
                    //    (separator columns add exact zeros)
                    sumM = O::add(sumM, M[KH - 1]);
                    sumX = O::add(sumX, X[KH - 1]);
                    if (LONG && !final_pass && last_lane) {
                        const int col = s0 + k - a - gl;
                        if (col >= 0 && col < ncols) {
                            spill_out[col + 1] = M[KH - 1];
                            spill_out[spill_cols + col + 1] = X[KH - 1];
                            spill_out[2 * spill_cols + col + 1] = Y[KH - 1];
                        }
                    }
                }
                // -- separator run: the step that ends an iteration resets the lane to the next
                //    haplotype's column-0 boundary; the last lane emits the finished pair first
                const bool z = ((codes >> 24) == 0) && (s0 + 3 - a - gl >= 0) && (jh < nh);
                if (__any_sync(FULL_MASK, z)) {
                    if (z) {
                        if (last_lane && final_pass) {
                            const int hl = hltab[jh];
#pragma unroll
                            for (int c = 0; c < VL; ++c)
                                if (rid[c] >= 0) {
                                    const T res = SCALED ? O::get(sumM, c) + O::get(sumX, c) * O::get(fin, c)
                                                         : S::add(O::get(sumM, c), O::get(sumX, c));   // :342
                                    if (VL == 1) emit_result_spec<FAST>(P, res, out_base[c] + out0 + jh, stream_off + hapstart, hl,
                                                                        hap0 + jh, out0 + jh, rid[c], R[c], tk->read[1] == SPEC_TASK);
                                    else emit_result<FAST>(P, res, out_base[c] + out0 + jh, stream_off + hapstart, hl,
                                                           hap0 + jh, out0 + jh, rid[c], R[c]);
                                }
                            hapstart += hl + SEP_COLS;
                        }
                        sumM = O::splat(zero); sumX = O::splat(zero);
                        ++jh;
                        const T iy = *reinterpret_cast<const T *>(iytab + 8 * jh);
#pragma unroll
                        for (int p = 0; p < KH; ++p)
#pragma unroll
                            for (int c = 0; c < VL; ++c)
                                O::set(Y[p], c, ((dmask[c] >> p) & 1) ? (SCALED ? iy * O::get(pG[p], c) : iy) : zero);
                    }
                }
            }
        }
        cp_async_wait_all();
        if (LONG) {
            __syncwarp();
            V *tmp = spill_in; spill_in = spill_out; spill_out = tmp;
        }
      }   // pass
    }
}

// Minimum resident groups of four warps per SM requested from ptxas (caps registers per thread).
// FP64 with KH <= 7 fits 12 warps per SM (168 registers) without spills; KH = 8 would spill and stays at 8
// warps (measured on C4: 43.0 -> 42.25 ms of FP64 kernels with KH <= 7 at 12 warps, 42.8 with KH = 8 too).
template <typename T, int KH>
struct MinBlocks {
    static constexpr int K = KH * (sizeof(T) / 4);
    static constexpr int value = K <= 2 ? 6 : K <= 4 ? 4 : K <= 6 ? 3 : (sizeof(T) == 8 && KH <= 7) ? 3 : 2;
};

typedef void (*KernelF32)(const KernelParams<float>);
typedef void (*KernelF64)(const KernelParams<double>);

template <int KH, int LANES, bool FAST, bool LONG = false>
static KernelF32 f32_kernel() { return phmm_forward_kernel<float, KH, LANES, FAST, LONG, MinBlocks<float, KH>::value * (4 / WARPS_PER_BLOCK)>; }
template <int KH, bool FAST, bool LONG = false>
static KernelF64 f64_kernel() { return phmm_forward_kernel<double, KH, 32, FAST, LONG, MinBlocks<double, KH>::value * (4 / WARPS_PER_BLOCK)>; }

}  // namespace phmm
#endif
