// Fast FP32 pair-HMM forward kernel, "haplotype-pair" packing (sm_100a).
//
// Same computation as phmm_forward_kernel<float, ..., FAST = true> (phmm_kernels.cuh; reference:
// compute_full_prob_<float>, tools/GKL/src/main/native/pairhmm/avx-pairhmm-template.h:210-346, policy
// IntelPairHmmCSource.cpp:61-85), different use of Blackwell's packed f32x2 datapath:
//
//   * A lane group (32, 16 or 8 threads) holds ONE read, KH consecutive rows per thread, and streams
//     it against TWO runs of haplotypes of its batch at once: run 0 in the low half of every f32x2
//     state register, run 1 in the high half.  The two halves are independent DPs that only share
//     the step counter.
//   * Because both halves belong to the same read, every per-row constant is the same in both.
//     FFMA2 / FMUL2 take such an operand as one 32-bit register broadcast to both halves (SASS
//     "Rn.F32"), so a cell update reads 27 registers instead of 32.5 and the constants need half
//     the registers -- the register tile can be 16 rows high instead of 8.  The kernels are bound
//     by register-file operand bandwidth (profiles/README.md), so this is what buys speed.
//   * Scaled recurrence (see phmm_kernels.cuh, "scaled recurrence"): 6 + ceil(KH / 4) / KH packed FP
//     instructions per row and step.
//   * Emission factors: per-warp table e[code][row quad][lane] of float4; one 128-bit LDS per four
//     rows, step and haplotype run, multiplied in with two scalar FMULs per row.
//   * Everything else as in phmm_kernels.cuh: right-aligned reads with dummy rows on top, four
//     columns per loop iteration, separator columns with the reset on the step that ends an
//     iteration, cp.async-staged haplotype rings (one per run), FP64-rescue and guard lists.
#ifndef PHMM_KERNELS_HP_CUH
#define PHMM_KERNELS_HP_CUH
#include "phmm_kernels.cuh"

namespace phmm {

__device__ __forceinline__ float2 hp_bc(float c) { return make_float2(c, c); }
__device__ __forceinline__ float2 hp_shfl_up(float2 v, int width)
{
    return make_float2(__shfl_up_sync(FULL_MASK, v.x, 1, width), __shfl_up_sync(FULL_MASK, v.y, 1, width));
}
__device__ __forceinline__ float4 hp_lds128(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

template <int KH, int LANES, int MINB>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MINB)
phmm_forward_hp_kernel(const KernelParams<float> P)
{
    constexpr int NTOP = (KH + 3) / 4;                 // groups of chained rows per thread
    constexpr int GRP = (KH + NTOP - 1) / NTOP;        // rows per group (<= 4)
    constexpr int ROWS = KH * LANES;
    constexpr int NQ = (KH + 3) / 4;                   // row quads per thread
    constexpr int EQUAD = 512;                         // bytes per (code, quad) table line: float4 per lane
    constexpr int ECODE = NQ * EQUAD;                  // bytes per code
    constexpr int ETAB = N_CODES * ECODE;
    constexpr int RTAB = (MAX_RUN_HAPS + 2) * 8;       // per run: float iy[], int len[]
    constexpr int WSM = ETAB + 2 * STREAM_RING + 2 * RTAB + 8 * (int)sizeof(RescRun);
    static_assert(sizeof(RescRun) == 24, "smem_per_warp_hp");
    constexpr uint32_t WMASK = STREAM_RING / 4 - 1;

    extern __shared__ __align__(16) uint8_t smem_all[];
    const int lane = threadIdx.x & 31;
    uint8_t *etab = smem_all + (threadIdx.x >> 5) * WSM;
    uint8_t *ring[2] = {etab + ETAB, etab + ETAB + STREAM_RING};
    float *iytab[2];
    int32_t *hltab[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        iytab[h] = reinterpret_cast<float *>(etab + ETAB + 2 * STREAM_RING + h * RTAB);
        hltab[h] = reinterpret_cast<int32_t *>(iytab[h] + (MAX_RUN_HAPS + 2));
    }
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(etab) + lane * 16;
    const int gl = lane & (LANES - 1);
    const int grp = lane / LANES;
    // pending rescue runs of this lane group (touched by its last lane only, at haplotype ends)
    RescRun *pend = reinterpret_cast<RescRun *>(etab + ETAB + 2 * STREAM_RING + 2 * RTAB) + 2 * grp;
    const bool last_lane = (gl == LANES - 1);
    const int ntasks = *P.ntasks_ptr;
    const uint32_t sh = ((4 - (gl & 3)) & 3) * 8;

    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(P.counter, 1);
        t = __shfl_sync(FULL_MASK, t, 0);
        if (t >= ntasks) break;

        const HTask *tk = P.htasks + t;
        int64_t stream_off[2];
        int ncols[2], hap0[2], out0[2], nh[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            stream_off[h] = tk->stream_off[h]; ncols[h] = tk->ncols[h]; hap0[h] = tk->hap0[h];
            out0[h] = tk->out0[h]; nh[h] = tk->nh[h];
        }
        const int rid = tk->read[grp];
        int R = 0;
        int64_t roff = 0, out_base = 0;
        if (rid >= 0) {
            const ReadRec rr = P.reads[rid];
            R = rr.len; roff = rr.off; out_base = rr.out_base;
        }
        if (last_lane) { pend[0].nh = 0; pend[1].nh = 0; }
        // per-run haplotype tables: column-0 boundary INITIAL_CONSTANT / haplen (:86) and length
        __syncwarp();
#pragma unroll
        for (int h = 0; h < 2; ++h)
            for (int j = lane; j <= nh[h]; j += 32) {
                const int hl = P.hap_len[hap0[h] + j];
                iytab[h][j] = __fdiv_rn(P.init_const, (float)hl);
                hltab[h][j] = hl;
            }

        // ---- per-row constants (initializeVectors :83-128, stripeINITIALIZATION :151-158), the
        //      emission table, and the folding of the scaled recurrence ------------------------------
        float cA[KH], cP[KH], cK1[KH], cK3[KH], cYY[KH], cK2[NTOP];
        float fin = 1.0f;
        uint32_t dmask = 0;
        __syncwarp();                 // every lane is done with the previous table
        {
            float g[KH], mx[KH], my[KH], sig[KH];
            const int nd = ROWS - R;                                  // dummy rows on top; >= 1
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const int r = gl * KH + p - nd;                       // 0-based read position
                float mm = 0.0f, em = 0.0f, ex = 0.0f;
                int rcode = -1;
                g[p] = 1.0f; mx[p] = 0.0f; my[p] = 0.0f; cYY[p] = 1.0f;
                if (r >= 0) {
                    const int _q = P.q[roff + r] & 127, _i = P.i[roff + r] & 127;
                    const int _d = P.d[roff + r] & 127, _c = P.c[roff + r] & 127;
                    const int mxq = max(_i, _d), mnq = min(_i, _d);
                    mm = P.mm[((mxq * (mxq + 1)) >> 1) + mnq];        // set_mm_prob, Context.h:120-132
                    const float pc = P.ph2pr[_c];
                    g[p] = __fsub_rn(1.0f, pc);                       // :115
                    mx[p] = P.ph2pr[_i]; my[p] = P.ph2pr[_d]; cYY[p] = pc;
                    const float eps = P.ph2pr[_q];
                    em = __fsub_rn(1.0f, eps);                        // :156
                    ex = __fdiv_rn(eps, 3.0f);                        // :158 true division
                    const uint8_t b = P.rs[roff + r];
                    rcode = (b == 'C') ? CODE_C : (b == 'T') ? CODE_T : (b == 'G') ? CODE_G
                          : (b == 'N') ? CODE_N : CODE_A;
                } else {
                    // dummy row: stays at the row-0 boundary (M, X, Y) = (0, 0, INITIAL_CONSTANT / haplen)
                    dmask |= 1u << p;
                }
                cA[p] = mm;
                // e[p][code]: match <=> equal or either is N (precompute_masks_ :3-35); separator -> 0
#pragma unroll
                for (int code = 0; code < N_CODES; ++code) {
                    const bool match = (code == rcode) || (code == CODE_N) || (rcode == CODE_N);
                    *reinterpret_cast<float *>(etab + code * ECODE + (p / 4) * EQUAD + lane * 16 + (p % 4) * 4) =
                        (code == CODE_SEP || rcode < 0) ? 0.0f : (match ? em : ex);
                }
            }
            // sigma_r = 1 on a group top, else pXX_r * sigma_{r-1}; gn = pGAPM of the next row
            const float g_dn = __shfl_down_sync(FULL_MASK, g[0], 1, LANES);
#pragma unroll
            for (int p = 0; p < KH; ++p) sig[p] = (p % GRP == 0) ? 1.0f : cYY[p] * sig[p ? p - 1 : 0];
            const float s_up = __shfl_up_sync(FULL_MASK, sig[KH - 1], 1, LANES);
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const float gn = (p < KH - 1) ? g[p < KH - 1 ? p + 1 : p] : (last_lane ? 1.0f : g_dn);
                if (p % GRP == 0) {
                    float k2 = cYY[p] * (p ? sig[p ? p - 1 : 0] : s_up);
                    if (p == 0 && (dmask & 1u)) k2 = 0.0f;            // nothing above row 0 of the tile
                    cK2[p / GRP] = k2;
                }
                cK1[p] = __fdiv_rn(mx[p], sig[p]);
                cK3[p] = my[p] * gn;
                cP[p] = gn * sig[p];
            }
            fin = sig[KH - 1];
        }

        // ---- state: column-0 boundary -------------------------------------------------------
        int jh[2] = {0, 0}, hapstart[2] = {0, 0};
        float2 M[KH], X[KH], Y[KH];
        __syncwarp();
        {
            const float iy0 = iytab[0][0], iy1 = iytab[1][0];
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                M[p] = make_float2(0.0f, 0.0f); X[p] = M[p];
                Y[p] = ((dmask >> p) & 1) ? make_float2(iy0 * cP[p], iy1 * cP[p]) : make_float2(0.0f, 0.0f);
            }
        }
        float2 U1m = make_float2(0.0f, 0.0f), U1x = U1m, U1z = U1m, U2m, U2z;
        float2 sumM = U1m, sumX = U1m;

        // ---- haplotype-stream rings, one per run ---------------------------------------------------
        // ring byte j of run h holds column j - a[h]; both runs advance through ring position s
        int a[2], need[2], s_end_h[2];
        const uint8_t *sbase[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            a[h] = (int)(stream_off[h] & 15);
            sbase[h] = P.stream + (stream_off[h] - a[h]);
            need[h] = nh[h] ? a[h] + ncols[h] + LANES + 16 : 0;      // ring bytes that must hold stream data
            s_end_h[h] = nh[h] ? (a[h] + ncols[h] + LANES - 1 + 3) & ~3 : 0;
            if (lane < 4) *reinterpret_cast<uint4 *>(ring[h] + STREAM_RING - 64 + 16 * lane) = make_uint4(0, 0, 0, 0);
            if (0 < need[h]) cp_async16(ring[h] + 16 * lane, sbase[h] + 16 * lane);
            if (STREAM_CHUNK < need[h]) cp_async16(ring[h] + STREAM_CHUNK + 16 * lane, sbase[h] + STREAM_CHUNK + 16 * lane);
        }
        const int s_begin = (nh[1] ? min(a[0], a[1]) : a[0]) & ~3;
        const int s_end = max(s_end_h[0], s_end_h[1]);
        const uint32_t *ring32_0 = reinterpret_cast<const uint32_t *>(ring[0]);
        const uint32_t *ring32_1 = reinterpret_cast<const uint32_t *>(ring[1]);

        uint32_t wi = (uint32_t)((s_begin >> 2) - ((gl + 3) >> 2));
        uint32_t wprev0 = 0, wprev1 = 0;
        for (int cb = 0, ch = 0; cb < s_end; cb += STREAM_CHUNK, ++ch) {
            cp_async_wait_all();
            __syncwarp();
            if (ch == 0) {
                if (lane < a[0]) ring[0][lane] = 0;       // columns -1 .. -a of an unaligned run
                if (lane < a[1]) ring[1][lane] = 0;
                __syncwarp();
                wprev0 = ring32_0[wi & WMASK];
                wprev1 = ring32_1[wi & WMASK];
            }
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if ((ch + 2) * STREAM_CHUNK < need[h])
                    cp_async16(ring[h] + (((ch + 2) * STREAM_CHUNK) & (STREAM_RING - 1)) + 16 * lane,
                               sbase[h] + (ch + 2) * STREAM_CHUNK + 16 * lane);
            const int s_hi = min(cb + STREAM_CHUNK, s_end);
            for (int s0 = max(cb, s_begin); s0 < s_hi; s0 += 4) {
                ++wi;
                const uint32_t wnext0 = ring32_0[wi & WMASK], wnext1 = ring32_1[wi & WMASK];
                // codes of columns s0 .. s0 + 3 (- a - gl); a run that has ended sees separators only
                const uint32_t codes0 = (s0 < s_end_h[0]) ? __funnelshift_r(wprev0, wnext0, sh) : 0u;
                const uint32_t codes1 = (s0 < s_end_h[1]) ? __funnelshift_r(wprev1, wnext1, sh) : 0u;
                wprev0 = wnext0; wprev1 = wnext1;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    // -- emission factors of this column: one 128-bit LDS per row quad and run.  The
                    //    two runs' factors land in unrelated registers, so the emission multiply is
                    //    two scalar FMULs per row (same FP32-pipe time as one FMUL2) -- the kernel is
                    //    bound by the operand-read port, where four 32-bit LDS cost four cycles and
                    //    one 128-bit LDS costs one.
                    const uint32_t ea0 = ebase + ((codes0 >> (8 * k)) & 0xffu) * ECODE;
                    const uint32_t ea1 = ebase + ((codes1 >> (8 * k)) & 0xffu) * ECODE;
                    float e0[NQ * 4], e1[NQ * 4];
#pragma unroll
                    for (int q = 0; q < NQ; ++q) {
                        const float4 v0 = hp_lds128(ea0 + q * EQUAD), v1 = hp_lds128(ea1 + q * EQUAD);
                        e0[4 * q] = v0.x; e0[4 * q + 1] = v0.y; e0[4 * q + 2] = v0.z; e0[4 * q + 3] = v0.w;
                        e1[4 * q] = v1.x; e1[4 * q + 1] = v1.y; e1[4 * q + 2] = v1.z; e1[4 * q + 3] = v1.w;
                    }
                    // -- neighbours: bottom row of the upstream lane, one and two steps ago
                    U2m = U1m; U2z = U1z;
                    const float2 zb = __ffma2_rn(X[KH - 1], hp_bc(cP[KH - 1]), Y[KH - 1]);
                    U1m = hp_shfl_up(M[KH - 1], LANES); U1x = hp_shfl_up(X[KH - 1], LANES);
                    U1z = hp_shfl_up(zb, LANES);
                    // -- cell updates (computeMXY :183-198, scaled recurrence)
                    float2 nM[KH], nX[KH], nY[KH];
#pragma unroll
                    for (int p = 0; p < KH; ++p) {
                        const float2 dm = p ? M[p ? p - 1 : 0] : U2m;
                        const float2 dz = p ? __ffma2_rn(X[p ? p - 1 : 0], hp_bc(cP[p ? p - 1 : 0]), Y[p ? p - 1 : 0]) : U2z;
                        const float2 tm = __ffma2_rn(dm, hp_bc(cA[p]), dz);
                        nM[p] = make_float2(__fmul_rn(tm.x, e0[p]), __fmul_rn(tm.y, e1[p]));
                        nY[p] = __ffma2_rn(M[p], hp_bc(cK3[p]), __fmul2_rn(Y[p], hp_bc(cYY[p])));
                    }
#pragma unroll
                    for (int p = 0; p < KH; ++p) {
                        const float2 um = p ? nM[p ? p - 1 : 0] : U1m;
                        const float2 ux = p ? nX[p ? p - 1 : 0] : U1x;
                        if (p % GRP == 0) nX[p] = __ffma2_rn(um, hp_bc(cK1[p]), __fmul2_rn(ux, hp_bc(cK2[p / GRP])));
                        else              nX[p] = __ffma2_rn(um, hp_bc(cK1[p]), ux);
                    }
#pragma unroll
                    for (int p = 0; p < KH; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; }
                    // -- last-row sums (:328,:331); separator columns add exact zeros
                    sumM = __fadd2_rn(sumM, M[KH - 1]);
                    sumX = __fadd2_rn(sumX, X[KH - 1]);
                }
                // -- separator run: the step that ends an iteration resets that half of the lane to
                //    the next haplotype's column-0 boundary; the last lane emits the finished pair
                const bool z0 = ((codes0 >> 24) == 0) && (s0 + 3 - a[0] - gl >= 0) && (jh[0] < nh[0]);
                const bool z1 = ((codes1 >> 24) == 0) && (s0 + 3 - a[1] - gl >= 0) && (jh[1] < nh[1]);
                if (__any_sync(FULL_MASK, z0 || z1)) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (h ? z1 : z0) {
                            if (last_lane) {
                                const int hl = hltab[h][jh[h]];
                                if (rid >= 0) {
                                    const float res = h ? sumM.y + sumX.y * fin : sumM.x + sumX.x * fin;   // :342
                                    const int verdict = classify_result<true>(P, res, out_base + out0[h] + jh[h], hl, R);
                                    RescRun &pr = pend[h];
                                    if (verdict == 1) {
                                        // FP64 rescue: extend the run that ended on the previous haplotype, or start one
                                        // (reads of the multi-pass FP64 class keep one pair per task: its boundary-row
                                        // scratch is sized for one haplotype)
                                        if (pr.nh > 0 && pr.nh < RESC_RUN_HAPS && pr.ncols + hl + SEP_COLS <= RESC_RUN_COLS &&
                                            list_class(P, R) != 0) {
                                            pr.nh += 1;
                                            pr.ncols += hl + SEP_COLS;
                                        } else {
                                            if (pr.nh > 0) resc_flush(P, pr, rid, R);
                                            pr.stream_off = stream_off[h] + hapstart[h];
                                            pr.ncols = hl + SEP_COLS;
                                            pr.hap0 = hap0[h] + jh[h];
                                            pr.out0 = out0[h] + jh[h];
                                            pr.nh = 1;
                                        }
                                    } else {
                                        if (pr.nh > 0) resc_flush(P, pr, rid, R);
                                        if (verdict == 2) {      // guard band / non-finite: the bit-exact FP32 kernel decides
                                            Task nt;
                                            nt.stream_off = stream_off[h] + hapstart[h];
                                            nt.ncols = hl + SEP_COLS;
                                            nt.hap0 = hap0[h] + jh[h];
                                            nt.out0 = out0[h] + jh[h];
                                            nt.read[0] = rid; nt.read[1] = -1; nt.read[2] = -1; nt.read[3] = -1;
                                            nt.nh = 1;
                                            const int ci = list_class(P, R);
                                            const int k = atomicAdd(&P.guard_count[ci], 1);
                                            P.guard[P.resc_off[ci] + k] = nt;
                                        }
                                    }
                                }
                                hapstart[h] += hl + SEP_COLS;
                            }
                            ++jh[h];
                            const float iy = iytab[h][jh[h]];
                            if (h) { sumM.y = 0.0f; sumX.y = 0.0f; } else { sumM.x = 0.0f; sumX.x = 0.0f; }
#pragma unroll
                            for (int p = 0; p < KH; ++p) {
                                const float y = ((dmask >> p) & 1) ? iy * cP[p] : 0.0f;
                                if (h) Y[p].y = y; else Y[p].x = y;
                            }
                        }
                    }
                }
            }
        }
        cp_async_wait_all();
        if (last_lane && rid >= 0) {
            if (pend[0].nh > 0) resc_flush(P, pend[0], rid, R);
            if (pend[1].nh > 0) resc_flush(P, pend[1], rid, R);
        }
    }
}

// Minimum resident groups of four warps per SM requested from ptxas (caps registers per thread).
template <int KH>
struct MinBlocksHP {
    static constexpr int value = KH <= 2 ? 6 : KH <= 4 ? 5 : KH <= 6 ? 4 : KH <= 9 ? 3 : 2;
};

template <int KH, int LANES>
static KernelF32 f32_hp_kernel() { return phmm_forward_hp_kernel<KH, LANES, MinBlocksHP<KH>::value * (4 / WARPS_PER_BLOCK)>; }

}  // namespace phmm
#endif
