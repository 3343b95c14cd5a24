// NOT PART OF THE BUILD: an experiment of round 2, kept for the record (profiles/r2_f64_rescue_register_port.md).
// It passed the GPU suite when wired in as lookup_f64_f1 and was not faster than phmm_forward_kernel<double, ...>.
//
// FP64 pair-HMM forward kernel, "column-block wavefront" (sm_100a).
//
// Same computation as phmm_forward_kernel<double, KH, 32, FAST = true> (phmm_kernels.cuh; reference:
// compute_full_prob_<double>, tools/GKL/src/main/native/pairhmm/avx-pairhmm-template.h:210-346, called for the
// pairs below MIN_ACCEPTED, IntelPairHmmCSource.cpp:73-75), laid out for what bounds FP64 on B200: a DFMA whose
// three operands are distinct registers reads six 32-bit words, and the register file delivers two words per
// cycle per scheduler, so it costs three issue cycles where the FP64 pipe needs two
// (profiles/r2_ubench6_fp64_issue.log; the column-at-a-time kernel sits at 2/3 of the DFMA peak for exactly this
// reason).  An operand that repeats in the same slot of the next instruction comes from the operand-reuse cache
// instead.  So:
//
//   * a lane owns KH consecutive read rows and advances FOUR columns per step, row by row: the four DFMAs of a
//     row that share a per-row constant (pMM, pMX / sigma, pYY, gn * sigma, pMY * gn) are adjacent and read it
//     once;
//   * lane l works on column block (step - l): the wavefront is skewed by four columns per lane.  The bottom row
//     of a lane reaches the next one as TWO values per column instead of three: the upstream lane applies the
//     downstream lane's row-0 constants itself (A = M * pMM' + Z for the diagonal term, B = M * pMX' + X * K2'
//     for the vertical one), which moves two DFMAs upstream and saves a third of the shuffles;
//   * emission factors come from a per-warp table e[code][row pair][lane] of double2: one 128-bit LDS per two
//     rows and column;
//   * scaled recurrence, dummy rows, separator columns, haplotype runs, rings: as in phmm_kernels.cuh.  A
//     non-finite sum is handed to the bit-exact FP64 kernel (emit_result<true>).
#ifndef PHMM_KERNELS_F64B_CUH
#define PHMM_KERNELS_F64B_CUH
#include "phmm_kernels.cuh"

namespace phmm {

__device__ __forceinline__ double2 f64b_lds128(uint32_t addr)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ double f64b_rot(double v, int src) { return __shfl_sync(FULL_MASK, v, src); }

static const int F64B_COLS = 4;      // columns per step = SEP_COLS

template <int KH, int MINB>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MINB)
phmm_forward_f64b_kernel(const KernelParams<double> P)
{
    constexpr int LANES = 32;
    constexpr int NTOP = (KH + 3) / 4;                 // groups of chained rows per thread
    constexpr int GRP = (KH + NTOP - 1) / NTOP;        // rows per group (<= 4)
    constexpr int ROWS = KH * LANES;
    constexpr int NP = (KH + 1) / 2;                   // row pairs per thread
    constexpr int EPAIR = 512;                         // bytes per (code, pair) table line: double2 per lane
    constexpr int ECODE = NP * EPAIR;
    constexpr int ETAB = N_CODES * ECODE;
    constexpr int WSM = ETAB + STREAM_RING + (MAX_RUN_HAPS + 2) * 12;
    constexpr uint32_t WMASK = STREAM_RING / 4 - 1;
    static_assert(F64B_COLS == SEP_COLS, "one separator run per column block");

    extern __shared__ __align__(16) uint8_t smem_all[];
    const int lane = threadIdx.x & 31;
    uint8_t *etab = smem_all + (threadIdx.x >> 5) * WSM;
    uint8_t *ring = etab + ETAB;
    const uint32_t *ring32 = reinterpret_cast<const uint32_t *>(ring);
    double *iytab = reinterpret_cast<double *>(ring + STREAM_RING);
    int32_t *hltab = reinterpret_cast<int32_t *>(iytab + (MAX_RUN_HAPS + 2));
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(etab) + lane * 16;
    const bool last_lane = (lane == LANES - 1);
    const int src_lane = (lane + LANES - 1) & (LANES - 1);     // rotation: lane 0 hears from the last lane
    const int ntasks = *P.ntasks_ptr;

    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(P.counter, 1);
        t = __shfl_sync(FULL_MASK, t, 0);
        if (t >= ntasks) break;

        const Task *tk = P.tasks + t;
        const int64_t stream_off = tk->stream_off;
        const int ncols = tk->ncols, hap0 = tk->hap0, out0 = tk->out0, nh = tk->nh;
        const int rid = tk->read[0];
        int R = 0;
        int64_t roff = 0, out_base = 0;
        if (rid >= 0) {
            const ReadRec rr = P.reads[rid];
            R = rr.len; roff = rr.off; out_base = rr.out_base;
        }
        __syncwarp();
        for (int j = lane; j <= nh; j += 32) {
            const int hl = P.hap_len[hap0 + j];
            iytab[j] = __ddiv_rn(P.init_const, (double)hl);     // :86
            hltab[j] = hl;
        }

        // ---- per-row constants (initializeVectors :83-128, stripeINITIALIZATION :151-158), emission table,
        //      folding of the scaled recurrence (phmm_kernels.cuh, "scaled recurrence") ---------------------
        double cA[KH], cP[KH], cK1[KH], cK3[KH], cYY[KH], cK2[NTOP];
        double fin = 1.0;
        uint32_t dmask = 0;
        __syncwarp();                 // every lane is done with the previous table
        {
            double g[KH], mx[KH], my[KH], sig[KH];
            const int nd = ROWS - R;                                  // dummy rows on top; >= 1
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const int r = lane * KH + p - nd;                     // 0-based read position
                double mm = 0.0, em = 0.0, ex = 0.0;
                int rcode = -1;
                g[p] = 1.0; mx[p] = 0.0; my[p] = 0.0; cYY[p] = 1.0;
                if (r >= 0) {
                    const int _q = P.q[roff + r] & 127, _i = P.i[roff + r] & 127;
                    const int _d = P.d[roff + r] & 127, _c = P.c[roff + r] & 127;
                    const int mxq = max(_i, _d), mnq = min(_i, _d);
                    mm = P.mm[((mxq * (mxq + 1)) >> 1) + mnq];        // set_mm_prob, Context.h:120-132
                    const double pc = P.ph2pr[_c];
                    g[p] = __dsub_rn(1.0, pc);                        // :115
                    mx[p] = P.ph2pr[_i]; my[p] = P.ph2pr[_d]; cYY[p] = pc;
                    const double eps = P.ph2pr[_q];
                    em = __dsub_rn(1.0, eps);                         // :156
                    ex = __ddiv_rn(eps, 3.0);                         // :158 true division
                    const uint8_t b = P.rs[roff + r];
                    rcode = (b == 'C') ? CODE_C : (b == 'T') ? CODE_T : (b == 'G') ? CODE_G
                          : (b == 'N') ? CODE_N : CODE_A;
                } else {
                    dmask |= 1u << p;      // dummy row: stays at (M, X, Y) = (0, 0, INITIAL_CONSTANT / haplen)
                }
                cA[p] = mm;
#pragma unroll
                for (int code = 0; code < N_CODES; ++code) {
                    const bool match = (code == rcode) || (code == CODE_N) || (rcode == CODE_N);
                    *reinterpret_cast<double *>(etab + code * ECODE + (p / 2) * EPAIR + lane * 16 + (p % 2) * 8) =
                        (code == CODE_SEP || rcode < 0) ? 0.0 : (match ? em : ex);
                }
            }
            if (KH & 1) {
#pragma unroll
                for (int code = 0; code < N_CODES; ++code)
                    *reinterpret_cast<double *>(etab + code * ECODE + (KH / 2) * EPAIR + lane * 16 + 8) = 0.0;
            }
            const double g_dn = __shfl_down_sync(FULL_MASK, g[0], 1);
#pragma unroll
            for (int p = 0; p < KH; ++p) sig[p] = (p % GRP == 0) ? 1.0 : cYY[p] * sig[p ? p - 1 : 0];
            const double s_up = __shfl_up_sync(FULL_MASK, sig[KH - 1], 1);
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const double gn = (p < KH - 1) ? g[p < KH - 1 ? p + 1 : p] : (last_lane ? 1.0 : g_dn);
                if (p % GRP == 0) {
                    double k2 = cYY[p] * (p ? sig[p ? p - 1 : 0] : s_up);
                    if (p == 0 && (dmask & 1u)) k2 = 0.0;             // nothing above row 0 of the tile
                    cK2[p / GRP] = k2;
                }
                cK1[p] = __ddiv_rn(mx[p], sig[p]);
                cK3[p] = my[p] * gn;
                cP[p] = gn * sig[p];
            }
            fin = sig[KH - 1];
        }
        // row-0 constants of the downstream lane (zero behind the last lane: lane 0 hears nothing from it)
        double nA = __shfl_down_sync(FULL_MASK, cA[0], 1), nK1 = __shfl_down_sync(FULL_MASK, cK1[0], 1);
        double nK2 = __shfl_down_sync(FULL_MASK, cK2[0], 1);
        if (last_lane) { nA = 0.0; nK1 = 0.0; nK2 = 0.0; }

        // ---- state: column-0 boundary -------------------------------------------------------
        int jh = 0, hapstart = 0;
        double M[KH], X[KH], Y[KH];
        __syncwarp();
        {
            const double iy = iytab[0];
#pragma unroll
            for (int p = 0; p < KH; ++p) { M[p] = 0.0; X[p] = 0.0; Y[p] = ((dmask >> p) & 1) ? iy * cP[p] : 0.0; }
        }
        double sumM = 0.0, sumX = 0.0;
        // what this lane hands down: per column of its last block A (diagonal term) and B (vertical term) of its
        // bottom row; before the first block: the column-0 boundary (M = X = 0, Z = Ys)
        double sA[F64B_COLS], sB[F64B_COLS];
#pragma unroll
        for (int k = 0; k < F64B_COLS; ++k) { sA[k] = Y[KH - 1]; sB[k] = 0.0; }
        double rAprev = 0.0;

        // ---- haplotype-stream ring -------------------------------------------------------------
        const int a = (int)(stream_off & 15);
        const uint8_t *sbase = P.stream + (stream_off - a);
        const int s_begin = a & ~3;                                   // lane l's block starts at ring byte s - 4 l
        const int s_end = ((a + ncols + 3) & ~3) + F64B_COLS * (LANES - 1);
        const int need = s_end + 16;                                  // ring bytes that must hold stream data
        if (lane < 8) *reinterpret_cast<uint4 *>(ring + STREAM_RING - 128 + 16 * lane) = make_uint4(0, 0, 0, 0);
        cp_async16(ring + 16 * lane, sbase + 16 * lane);
        if (STREAM_CHUNK < need) cp_async16(ring + STREAM_CHUNK + 16 * lane, sbase + STREAM_CHUNK + 16 * lane);

        uint32_t wi = (uint32_t)((s_begin >> 2) - lane);
        for (int cb = 0, ch = 0; cb < s_end; cb += STREAM_CHUNK, ++ch) {
            cp_async_wait_all();
            __syncwarp();
            if (ch == 0) {
                if (lane < a) ring[lane] = 0;       // columns -1 .. -a of an unaligned run
                __syncwarp();
            }
            if ((ch + 2) * STREAM_CHUNK < need)
                cp_async16(ring + (((ch + 2) * STREAM_CHUNK) & (STREAM_RING - 1)) + 16 * lane,
                           sbase + (ch + 2) * STREAM_CHUNK + 16 * lane);
            const int s_hi = min(cb + STREAM_CHUNK, s_end);
            for (int s0 = max(cb, s_begin); s0 < s_hi; s0 += F64B_COLS) {
                const uint32_t codes = ring32[wi & WMASK];            // codes of this lane's four columns
                ++wi;
                // -- from the upstream lane: its previous block is this lane's current one
                double dA[F64B_COLS], uB[F64B_COLS];
                dA[0] = rAprev;
#pragma unroll
                for (int k = 0; k < F64B_COLS; ++k) {
                    const double ra = f64b_rot(sA[k], src_lane);
                    uB[k] = f64b_rot(sB[k], src_lane);
                    if (k + 1 < F64B_COLS) dA[k + 1] = ra; else rAprev = ra;
                }
                // -- rows, top to bottom; in[]: the row above (diagonal M, Z at columns -1 .. 2; vertical M, X at 0 .. 3)
                double mD[F64B_COLS], zD[F64B_COLS], mU[F64B_COLS], xU[F64B_COLS];
                uint32_t ea[F64B_COLS];
#pragma unroll
                for (int k = 0; k < F64B_COLS; ++k) ea[k] = ebase + ((codes >> (8 * k)) & 0xffu) * ECODE;
                double2 ev[F64B_COLS];
#pragma unroll
                for (int p = 0; p < KH; ++p) {
                    if ((p & 1) == 0) {
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) ev[k] = f64b_lds128(ea[k] + (p / 2) * EPAIR);
                    }
                    const double zold = __fma_rn(X[p], cP[p], Y[p]);          // Z of this row at column -1
                    double tm[F64B_COLS], nM[F64B_COLS], ty[F64B_COLS], nY[F64B_COLS], nX[F64B_COLS];
                    if (p == 0) {
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) tm[k] = dA[k];
                    } else {
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) tm[k] = __fma_rn(mD[k], cA[p], zD[k]);
                    }
#pragma unroll
                    for (int k = 0; k < F64B_COLS; ++k) nM[k] = __dmul_rn(tm[k], (p & 1) ? ev[k].y : ev[k].x);
#pragma unroll
                    for (int k = 0; k < F64B_COLS; ++k) ty[k] = __dmul_rn(k ? nM[k ? k - 1 : 0] : M[p], cK3[p]);
                    nY[0] = __fma_rn(Y[p], cYY[p], ty[0]);
#pragma unroll
                    for (int k = 1; k < F64B_COLS; ++k) nY[k] = __fma_rn(nY[k - 1], cYY[p], ty[k]);
                    if (p == 0) {
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) nX[k] = uB[k];
                    } else if (p % GRP == 0) {
                        double xs[F64B_COLS];
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) xs[k] = __dmul_rn(xU[k], cK2[p / GRP]);
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) nX[k] = __fma_rn(mU[k], cK1[p], xs[k]);
                    } else {
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) nX[k] = __fma_rn(mU[k], cK1[p], xU[k]);
                    }
                    // the row below sees this row as its "row above"
                    mD[0] = M[p]; zD[0] = zold;
#pragma unroll
                    for (int k = 1; k < F64B_COLS; ++k) { mD[k] = nM[k - 1]; zD[k] = __fma_rn(nX[k - 1], cP[p], nY[k - 1]); }
#pragma unroll
                    for (int k = 0; k < F64B_COLS; ++k) { mU[k] = nM[k]; xU[k] = nX[k]; }
                    M[p] = nM[F64B_COLS - 1]; X[p] = nX[F64B_COLS - 1]; Y[p] = nY[F64B_COLS - 1];
                    if (p == KH - 1) {
                        // bottom row: last-row sums (:328,:331; separator columns add exact zeros) and the
                        // hand-down values, with the downstream lane's row-0 constants
                        const double z3 = __fma_rn(nX[F64B_COLS - 1], cP[p], nY[F64B_COLS - 1]);
#pragma unroll
                        for (int k = 0; k < F64B_COLS; ++k) {
                            sumM = __dadd_rn(sumM, nM[k]);
                            sumX = __dadd_rn(sumX, nX[k]);
                            const double zk = (k + 1 < F64B_COLS) ? zD[k + 1 < F64B_COLS ? k + 1 : k] : z3;
                            sA[k] = __fma_rn(nM[k], nA, zk);
                            sB[k] = __fma_rn(nM[k], nK1, __dmul_rn(nX[k], nK2));
                        }
                    }
                }
                // -- separator run: the block that ends on a separator column resets the lane to the next
                //    haplotype's column-0 boundary; the last lane emits the finished pair first
                const bool z = ((codes >> 24) == 0) && (s0 + 3 - a - F64B_COLS * lane >= 0) && (jh < nh);
                if (__any_sync(FULL_MASK, z)) {
                    if (z) {
                        if (last_lane) {
                            const int hl = hltab[jh];
                            if (rid >= 0) {
                                const double res = sumM + sumX * fin;                        // :342
                                emit_result<true>(P, res, out_base + out0 + jh, stream_off + hapstart, hl, hap0 + jh,
                                                  out0 + jh, rid, R);
                            }
                            hapstart += hl + SEP_COLS;
                        }
                        sumM = 0.0; sumX = 0.0;
                        ++jh;
                        const double iy = iytab[jh];
#pragma unroll
                        for (int p = 0; p < KH; ++p) Y[p] = ((dmask >> p) & 1) ? iy * cP[p] : 0.0;
                        // the hand-down value of the block's last column was made from the old Y: at the boundary
                        // M = X = 0, so A = Z = Ys
                        sA[F64B_COLS - 1] = Y[KH - 1];
                    }
                }
            }
        }
        cp_async_wait_all();
    }
}

// resident groups of four warps per SM requested from ptxas (caps registers per thread)
template <int KH>
struct MinBlocksF64B {
    static constexpr int value = KH <= 1 ? 6 : KH <= 2 ? 4 : KH <= 5 ? 3 : 2;
};

template <int KH>
static KernelF64 f64b_kernel() { return phmm_forward_f64b_kernel<KH, MinBlocksF64B<KH>::value * (4 / WARPS_PER_BLOCK)>; }

}  // namespace phmm
#endif
