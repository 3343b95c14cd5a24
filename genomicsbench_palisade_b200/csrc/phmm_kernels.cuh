// Pair-HMM forward kernels for sm_100a (B200).
//
// What is computed: compute_full_prob_<float|double> of the reference
// (tools/GKL/src/main/native/pairhmm/avx-pairhmm-template.h:210-346) followed by the FP32 ->
// FP64 rescue policy and log10 finalisation of computelikelihoodsboth
// (tools/GKL/src/main/native/pairhmm/IntelPairHmmCSource.cpp:61-85).
//
// How (not a port of the AVX striping):
//   * One warp streams ONE read (or 2/4 short reads, one per lane group) against a whole run of
//     haplotypes of its batch, laid end to end in HBM as one-hot codes with a separator column
//     after each haplotype.  The anti-diagonal wavefront therefore fills once per run instead
//     of once per pair, and the read's seven per-row constants stay in registers for the run.
//   * A thread owns VL virtual lanes of KH consecutive read rows each.  For FP32, VL = 2 and the
//     two virtual lanes live in the two halves of Blackwell's packed f32x2 registers: the 8
//     multiplies of a cell update issue as FMUL2 (bit-exact kernel) or everything as FFMA2
//     (fast kernel), halving the issue slots the FMA pipe needs.  The second virtual lane runs
//     one column behind the first, so its "up" neighbour is the first lane's previous result
//     and only one shuffle per state array crosses threads per step.
//   * Virtual lane v works on column (step - v).  The bottom row of each virtual lane moves to
//     the next one through __shfl_up_sync (between threads) or a register (within a thread).
//   * Read rows are right-aligned: unused rows at the top are "dummy rows" whose constants make
//     them reproduce the DP's row-0 boundary (M = X = 0, Y = INITIAL_CONSTANT / haplen) exactly,
//     so the first real row needs no special case and the result is always the last row of the
//     last virtual lane.
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
template <typename T, int VL, bool FAST>
struct Ops;

// FP32, two virtual lanes packed in f32x2.
template <bool FAST>
struct Ops<float, 2, FAST> {
    typedef float2 V;
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
};

// Scalar back-ends (FP64 rescue, FP32 guard-band re-decision, non-packed FP32 variant).
template <bool FAST>
struct Ops<float, 1, FAST> {
    typedef float V;
    static __device__ __forceinline__ V splat(float x) { return x; }
    static __device__ __forceinline__ float get(const V &a, int) { return a; }
    static __device__ __forceinline__ void set(V &a, int, float x) { a = x; }
    static __device__ __forceinline__ V mul(V a, V b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ V add(V a, V b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ V fma(V a, V b, V c) { return __fmaf_rn(a, b, c); }
};

template <bool FAST>
struct Ops<double, 1, FAST> {
    typedef double V;
    static __device__ __forceinline__ V splat(double x) { return x; }
    static __device__ __forceinline__ double get(const V &a, int) { return a; }
    static __device__ __forceinline__ void set(V &a, int, double x) { a = x; }
    static __device__ __forceinline__ V mul(V a, V b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ V add(V a, V b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ V fma(V a, V b, V c) { return __fma_rn(a, b, c); }
};

template <typename T> struct Scalar;
template <> struct Scalar<float> {
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float shfl_up(float v, int width) { return __shfl_up_sync(FULL_MASK, v, 1, width); }
};
template <> struct Scalar<double> {
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double shfl_up(double v, int width) { return __shfl_up_sync(FULL_MASK, v, 1, width); }
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

// FP64 kernel class index of a read (list the FP32 kernel appends rescue work to):
// scalar kernel, 32 lanes, kh rows per lane, capacity 32*kh - 1.  0 = longer than any
// single-pass class.
__host__ __device__ __forceinline__ int rescue_class(int rslen)
{
    int kh = (rslen + 1 + 31) / 32;
    return kh <= MAX_KH_SCALAR ? kh : 0;
}

// ------------------------------------------------------------------------------------------
// Result of one (read, haplotype): policy of computelikelihoodsboth (IntelPairHmmCSource.cpp:69-82).
// ------------------------------------------------------------------------------------------
template <bool FAST>
__device__ __forceinline__ void emit_result(const KernelParams<float> &P, float res, int64_t oidx,
                                            int64_t hap_stream_off, int hap_len, int hap_idx,
                                            int out_rel, int rid, int rslen)
{
    P.raw32[oidx] = res;
    if (P.mode == 0) {
        const bool guard = FAST && (res >= P.guard_lo) && (res < P.guard_hi);
        if (guard || res < MIN_ACCEPTED_F) {
            // :73-75 -- below MIN_ACCEPTED the pair is re-run in FP64 (by the FP64 kernel of the
            // read's scalar class); the fast kernel hands near-threshold sums to the bit-exact
            // FP32 kernel first so that the rescue decision is the reference's.
            Task nt;
            nt.stream_off = hap_stream_off;
            nt.ncols = hap_len + 1;
            nt.hap0 = hap_idx;
            nt.out0 = out_rel;
            nt.read[0] = rid; nt.read[1] = -1; nt.read[2] = -1; nt.read[3] = -1;
            nt.pad = 0;
            const int ci = rescue_class(rslen);
            if (guard) {
                int k = atomicAdd(&P.guard_count[ci], 1);
                P.guard[P.resc_off[ci] + k] = nt;
            } else {
                int k = atomicAdd(&P.resc_count[ci], 1);
                P.resc[P.resc_off[ci] + k] = nt;
            }
            return;
        }
    }
    // :78 (and :97) -- log10 in float, subtraction in float, then widened
    P.out[oidx] = (double)__fsub_rn(log10f(res), P.log10_init);
    P.rescued[oidx] = 0;
}

template <bool FAST>
__device__ __forceinline__ void emit_result(const KernelParams<double> &P, double res, int64_t oidx,
                                            int64_t, int, int, int, int, int)
{
    // :75 (and :109)
    P.raw64[oidx] = res;
    P.out[oidx] = __dsub_rn(log10(res), P.log10_init);
    P.rescued[oidx] = (P.mode == 0) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------
// The forward kernel.  Persistent warps pull tasks from a device-side queue.
// ------------------------------------------------------------------------------------------
template <typename T, int VL, int KH, int LANES, bool FAST, bool LONG, int MINB>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MINB)
phmm_forward_kernel(const KernelParams<T> P)
{
    typedef Ops<T, VL, FAST> O;
    typedef typename O::V V;
    typedef Scalar<T> S;
    constexpr int NV = VL * LANES;            // virtual lanes per read
    constexpr uint32_t RMASK = STREAM_RING - 1;

    __shared__ __align__(16) uint8_t ring_all[WARPS_PER_BLOCK][STREAM_RING];
    const int lane = threadIdx.x & 31;
    uint8_t *ring = ring_all[threadIdx.x >> 5];
    const int gl = lane & (LANES - 1);        // lane within its read's group
    const int grp = lane / LANES;
    const bool last_lane = (gl == LANES - 1);
    const int ntasks = *P.ntasks_ptr;
    const T zero = (T)0;

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
        const int rid = tk->read[grp];
        const bool active = rid >= 0;
        int R = 0;
        int64_t roff = 0, out_base = 0;
        if (active) {
            const ReadRec rr = P.reads[rid];
            R = rr.len; roff = rr.off; out_base = rr.out_base;
        }

        // Multi-pass (LONG) kernels cut the read into blocks of NV * KH rows; block `pass` reads
        // the previous block's bottom row from spill_in and leaves its own in spill_out.
        const int npass = LONG ? (R + NV * KH) / (NV * KH) : 1;
        T *spill_in = nullptr, *spill_out = nullptr;
        int64_t spill_cols = 0;
        if (LONG) {
            const int64_t wslot = (int64_t)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
            spill_cols = P.scratch_stride / 6;
            spill_in = P.scratch + wslot * P.scratch_stride;
            spill_out = spill_in + 3 * spill_cols;
        }
      for (int pass = 0; pass < npass; ++pass) {
        const bool final_pass = (pass == npass - 1);
        // ---- per-row constants (initializeVectors :83-128, stripeINITIALIZATION :151-158) ----
        V pMM[KH], pG[KH], pMX[KH], pXX[KH], pMY[KH], em[KH], ex[KH];
        V pYY0;                       // pYY of row 0 of each virtual lane (see dummy rows below)
        uint32_t rbits[VL], dmask[VL];
        const int nd = npass * NV * KH - R;   // dummy rows on top of block 0; always >= 1
#pragma unroll
        for (int c = 0; c < VL; ++c) {
            rbits[c] = 0; dmask[c] = 0;
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const int r = pass * NV * KH + (gl * VL + c) * KH + p - nd;    // 0-based read position
                T mm = zero, g = zero, mx = zero, xx = zero, my = zero, e1 = zero, e3 = zero, yy = zero;
                if (r >= 0) {
                    const int _q = P.q[roff + r] & 127, _i = P.i[roff + r] & 127;
                    const int _d = P.d[roff + r] & 127, _c = P.c[roff + r] & 127;
                    const int mxq = max(_i, _d), mnq = min(_i, _d);
                    mm = P.mm[((mxq * (mxq + 1)) >> 1) + mnq];        // set_mm_prob, Context.h:120-132
                    const T pc = P.ph2pr[_c];
                    g = S::sub((T)1, pc);                             // :115
                    mx = P.ph2pr[_i]; my = P.ph2pr[_d]; xx = pc; yy = pc;
                    const T eps = P.ph2pr[_q];
                    e1 = S::sub((T)1, eps);                           // :156
                    e3 = S::div(eps, (T)3);                           // :158 true division
                    uint8_t b = P.rs[roff + r];
                    uint32_t code = (b == 'C') ? CODE_C : (b == 'T') ? CODE_T : (b == 'G') ? CODE_G
                                  : (b == 'N') ? CODE_N : CODE_A;
                    rbits[c] |= code << (4 * p);
                } else {
                    // Dummy row: stays at (M, X, Y) = (0, 0, INITIAL_CONSTANT/haplen), the row-0
                    // boundary (:86,:162): e = 0 kills M, pMX = 0 and a zero "up" X keep X = 0,
                    // pMY = 0 and pYY = 1 carry Y along the row exactly.  pXX doubles as pYY in
                    // real rows; only row 0 of a virtual lane can see a non-zero "up" X (lane 0
                    // receives its own registers from the shuffle), so it gets pXX = 0 and its
                    // own pYY register.
                    xx = (p == 0) ? zero : (T)1;
                    yy = (T)1;
                    dmask[c] |= 1u << p;
                }
                O::set(pMM[p], c, mm); O::set(pG[p], c, g); O::set(pMX[p], c, mx);
                O::set(pXX[p], c, xx); O::set(pMY[p], c, my); O::set(em[p], c, e1);
                O::set(ex[p], c, e3);
                if (p == 0) O::set(pYY0, c, yy);
            }
        }

        // ---- state: column-0 boundary -------------------------------------------------------
        int jh[VL];
        T initY[VL];
        V M[KH], X[KH], Y[KH];
        {
            const T iy = S::div(P.init_const, (T)P.hap_len[hap0]);    // :86
#pragma unroll
            for (int c = 0; c < VL; ++c) { jh[c] = 0; initY[c] = iy; }
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                M[p] = O::splat(zero); X[p] = O::splat(zero);
#pragma unroll
                for (int c = 0; c < VL; ++c) O::set(Y[p], c, ((dmask[c] >> p) & 1) ? iy : zero);
            }
        }
        V U1m = O::splat(zero), U1x = O::splat(zero), U1y = O::splat(zero);
        // a later block whose block above is nothing but dummy rows sits directly under DP row 0:
        // its diagonal neighbour at the first column is Y[0][0] = INITIAL_CONSTANT / haplen
        if (LONG && pass > 0 && gl == 0 && nd >= pass * NV * KH) O::set(U1y, 0, initY[0]);
        V U2m, U2x, U2y;
        T sumM = zero, sumX = zero;
        int hapstart = 0;

        // ---- haplotype-stream ring -------------------------------------------------------------
        const int a = (int)(stream_off & 15);
        const uint8_t *sbase = P.stream + (stream_off - a);
        const int need = a + ncols;                 // ring bytes that must hold real columns
        const int s_end = a + ncols + NV - 1;       // loop runs s' = a .. s_end-1 (s' = step + a)
        __syncwarp();
        if (lane < 4) *reinterpret_cast<uint4 *>(ring + STREAM_RING - 64 + 16 * lane) = make_uint4(0, 0, 0, 0);
        cp_async16(ring + 16 * lane, sbase + 16 * lane);
        if (STREAM_CHUNK < need) cp_async16(ring + STREAM_CHUNK + 16 * lane, sbase + STREAM_CHUNK + 16 * lane);

        const int vpos = gl * VL;                   // first virtual lane of this thread
        uint32_t hb_next = 0, hrep_prev = 0;
        T sp_m = zero, sp_x = zero, sp_y = zero;
        if (LONG && pass > 0 && gl == 0) {
            sp_m = __ldcg(spill_in); sp_x = __ldcg(spill_in + spill_cols); sp_y = __ldcg(spill_in + 2 * spill_cols);
        }

        for (int cb = 0, ch = 0; cb < s_end; cb += STREAM_CHUNK, ++ch) {
            cp_async_wait_all();
            __syncwarp();
            if (ch == 0) {
                if (lane < a) ring[lane] = 0;       // columns -1 .. -a of an unaligned run
                __syncwarp();
                hb_next = ring[(a - vpos) & RMASK];
            }
            if ((ch + 2) * STREAM_CHUNK < need)
                cp_async16(ring + (((ch + 2) * STREAM_CHUNK) & RMASK) + 16 * lane,
                           sbase + (ch + 2) * STREAM_CHUNK + 16 * lane);
            const int s_hi = min(cb + STREAM_CHUNK, s_end);
#pragma unroll 2
            for (int s = max(cb, a); s < s_hi; ++s) {
                // -- neighbours: bottom row of the upstream virtual lane, one and two steps ago
                U2m = U1m; U2x = U1x; U2y = U1y;
                {
                    const T rm = S::shfl_up(O::get(M[KH - 1], VL - 1), LANES);
                    const T rx = S::shfl_up(O::get(X[KH - 1], VL - 1), LANES);
                    const T ry = S::shfl_up(O::get(Y[KH - 1], VL - 1), LANES);
                    if (VL == 2) {
                        O::set(U1m, 1, O::get(M[KH - 1], 0)); O::set(U1x, 1, O::get(X[KH - 1], 0));
                        O::set(U1y, 1, O::get(Y[KH - 1], 0));
                    }
                    if (LONG && pass > 0 && gl == 0) {     // first row of a later block: "up" is the
                        O::set(U1m, 0, sp_m); O::set(U1x, 0, sp_x); O::set(U1y, 0, sp_y);   // block above
                        const int ncol = s + 1 - a;
                        if (ncol < ncols) {
                            sp_m = __ldcg(spill_in + ncol); sp_x = __ldcg(spill_in + spill_cols + ncol);
                            sp_y = __ldcg(spill_in + 2 * spill_cols + ncol);
                        }
                    } else {
                        O::set(U1m, 0, rm); O::set(U1x, 0, rx); O::set(U1y, 0, ry);
                    }
                }
                // -- haplotype code of this step's column (prefetched) and of the next one
                const uint32_t hb = hb_next;
                hb_next = ring[(s + 1 - vpos) & RMASK];
                uint32_t hbits[VL];
                hbits[0] = __byte_perm(hb, 0, 0);           // code in all 8 nibbles
                if (VL == 2) hbits[1] = hrep_prev;          // second virtual lane: one column behind
                hrep_prev = hbits[0];

                // -- cell updates (computeMXY :183-198)
                V nM[KH], nX[KH], nY[KH];
#pragma unroll
                for (int p = 0; p < KH; ++p) {
                    const V dm = p ? M[p - 1] : U2m;
                    const V dx = p ? X[p - 1] : U2x;
                    const V dy = p ? Y[p - 1] : U2y;
                    V e;
#pragma unroll
                    for (int c = 0; c < VL; ++c) {
                        const bool match = (rbits[c] & hbits[c] & (0xFu << (4 * p))) != 0;
                        O::set(e, c, match ? O::get(em[p], c) : O::get(ex[p], c));
                    }
                    const V yy = p ? pXX[p] : pYY0;
                    if (FAST) {
                        nM[p] = O::mul(O::fma(dy, pG[p], O::fma(dx, pG[p], O::mul(dm, pMM[p]))), e);
                        nY[p] = O::fma(M[p], pMY[p], O::mul(Y[p], yy));
                    } else {
                        nM[p] = O::mul(O::add(O::add(O::mul(dm, pMM[p]), O::mul(dx, pG[p])),
                                              O::mul(dy, pG[p])), e);
                        nY[p] = O::add(O::mul(M[p], pMY[p]), O::mul(Y[p], yy));
                    }
                }
#pragma unroll
                for (int p = 0; p < KH; ++p) {
                    const V um = p ? nM[p - 1] : U1m;
                    const V ux = p ? nX[p - 1] : U1x;
                    if (FAST) nX[p] = O::fma(um, pMX[p], O::mul(ux, pXX[p]));
                    else      nX[p] = O::add(O::mul(um, pMX[p]), O::mul(ux, pXX[p]));
                }
#pragma unroll
                for (int p = 0; p < KH; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; }

                // -- separator / padding columns (code 0): emit, then reset to the column-0 boundary
                bool z = (hbits[0] == 0);
                if (VL == 2) z = z || (hbits[1] == 0);
                if (__any_sync(FULL_MASK, z)) {
#pragma unroll
                    for (int c = 0; c < VL; ++c) {
                        if (hbits[c] == 0) {
                            const int col = s - a - (vpos + c);
                            if (col >= 0 && col < ncols) {          // separator after haplotype jh[c]
                                if (c == VL - 1) {
                                    if (last_lane && active && final_pass) {
                                        const T res = S::add(sumM, sumX);   // :342
                                        emit_result<FAST>(P, res, out_base + out0 + jh[c],
                                                          stream_off + hapstart, col - hapstart,
                                                          hap0 + jh[c], out0 + jh[c], rid, R);
                                    }
                                    sumM = zero; sumX = zero; hapstart = col + 1;
                                }
                                jh[c]++;
                                initY[c] = S::div(P.init_const, (T)P.hap_len[hap0 + jh[c]]);
                            }
#pragma unroll
                            for (int p = 0; p < KH; ++p) {
                                O::set(M[p], c, zero); O::set(X[p], c, zero);
                                O::set(Y[p], c, ((dmask[c] >> p) & 1) ? initY[c] : zero);
                            }
                        }
                    }
                }
                // -- last-row sums (:328,:331): two separate sequential sums in column order
                sumM = S::add(sumM, O::get(M[KH - 1], VL - 1));
                sumX = S::add(sumX, O::get(X[KH - 1], VL - 1));
                if (LONG && !final_pass && last_lane) {
                    const int col = s - a - (NV - 1);
                    if (col >= 0 && col < ncols) {
                        spill_out[col] = O::get(M[KH - 1], VL - 1);
                        spill_out[spill_cols + col] = O::get(X[KH - 1], VL - 1);
                        spill_out[2 * spill_cols + col] = O::get(Y[KH - 1], VL - 1);
                    }
                }
            }
        }
        if (LONG) {
            __syncwarp();
            T *tmp = spill_in; spill_in = spill_out; spill_out = tmp;
        }
      }   // pass
    }
}

// Minimum resident blocks per SM requested from ptxas (caps registers per thread):
// rows per thread K = VL * KH need about 10 K (+25) 32-bit registers in FP32, twice that in FP64.
template <typename T, int VL, int KH>
struct MinBlocks {
    static constexpr int K = VL * KH * (sizeof(T) / 4);
    static constexpr int value = K <= 2 ? 8 : K <= 4 ? 6 : K <= 6 ? 5 : K <= 8 ? 4 : K <= 12 ? 3 : 2;
};

typedef void (*KernelF32)(const KernelParams<float>);
typedef void (*KernelF64)(const KernelParams<double>);

template <int VL, int KH, int LANES, bool FAST, bool LONG = false>
static KernelF32 f32_kernel() { return phmm_forward_kernel<float, VL, KH, LANES, FAST, LONG, MinBlocks<float, VL, KH>::value>; }
template <int KH, bool FAST, bool LONG = false>
static KernelF64 f64_kernel() { return phmm_forward_kernel<double, 1, KH, 32, FAST, LONG, MinBlocks<double, 1, KH>::value>; }

}  // namespace phmm
#endif
