// Smith-Waterman with backtrack for sm_100a: one warp per (ref, alt) pair.
//
// Computes what smithWatermanBackTrack + getCIGAR compute
// (tools/GKL/src/main/native/smithwaterman/PairWiseSW.h:40-263, :265-412), bit for bit:
//
//   * Wavefront: lane l owns rows l*KH + 1 .. l*KH + KH in registers (H and E of the previous
//     column), works on column step - l, and hands the H and F of its bottom row to lane l + 1 with
//     __shfl_up_sync.  KH is a template parameter; a pair runs in the smallest class with
//     32 * KH >= len1 (KH <= 32 covers MAX_SEQ_LEN = 1024), so most lanes hold real rows.  The
//     reference sweeps anti-diagonals with 8/16-wide vectors through three memory arrays (:98-194);
//     here the diagonal state never leaves registers.
//   * max(a + b, c) is Blackwell's fused VIADDMNMX (__viaddmax_s32): three of them per cell.
//   * Backtrack: 4 bits per cell (MATCH/INSERT/DELETE + the two *_EXT flags, :21-25), eight cells of
//     a lane per 32-bit word.  The reference writes 16 bits per cell (:170).  Layout in the per-warp
//     scratch: tiles of eight steps, [tile][lane][step in tile][word], so that the words of one lane
//     over consecutive steps -- what the traceback walks -- are contiguous (the first layout,
//     [step][word][lane], made every traceback window touch 32 sectors for 128 useful bytes: DRAM reads as
//     large as the backtrack writes).  The DP collects a tile in shared memory (the lastrow buffer, idle
//     until the DP is done) and writes it out with full, coalesced 128-byte stores, lanes holding real
//     rows only.
//   * Last column scores go to shared memory.  The last row sits in a register slot that depends on
//     the pair (row nrow of lane (nrow - 1) / KH): picking it per cell would cost a compare and a
//     select on the ALU pipe, the pipe that bounds the kernel, so that lane stores all its KH row
//     values of the step with 128-bit stores into per-warp scratch in L2 and the scan reads the one
//     column it needs.  The best cell is chosen afterwards in the reference's visiting order
//     (:209-236): a warp-wide maximum, then the tie rule applied to the cells that reach it, in
//     anti-diagonal order.
//   * Traceback and CIGAR formatting (:265-412) by the warp on the backtrack words it just wrote,
//     through a 32-step window of words fetched with one warp-wide load.
#ifndef SW_KERNELS_CUH
#define SW_KERNELS_CUH
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>
#include <utility>

namespace sw {

// compile-time loop: the body sees its index as a constant expression (immediate operands of inline PTX)
template <int... Is, class F>
__device__ __forceinline__ void static_for_impl(std::integer_sequence<int, Is...>, F &&f)
{
    (f(std::integral_constant<int, Is>{}), ...);
}
template <int N, class F>
__device__ __forceinline__ void static_for(F &&f) { static_for_impl(std::make_integer_sequence<int, N>{}, f); }

static const int WARPS_PER_BLOCK = 4;
static const int MAX_LEN = 1024;
static const unsigned FULL = 0xffffffffu;

enum { OP_MATCH = 0, OP_INSERT = 1, OP_DELETE = 2, F_INSERT_EXT = 4, F_DELETE_EXT = 8,
       ST_SOFTCLIP = 9, ST_INDEL = 10, ST_LEADING_INDEL = 11, ST_IGNORE = 12 };
#define SW_MIN_CUTOFF (-100000000)              /* MATRIX_MIN_CUTOFF, smithwaterman_common.h:52 */
#define SW_LOW_INIT   (INT32_MIN / 2)           /* LOW_INIT_VALUE, :53 */

struct PairRec {            // 32 bytes
    int64_t ref_off, alt_off;   // into the device pools
    int32_t ref_len, alt_len;
    int32_t out;                // result index (the caller's pair number)
    int32_t pad;
};

struct Params {
    const uint8_t *ref_pool, *alt_pool;
    const PairRec *pairs;       // this class's tasks, longest first
    int32_t ntasks;
    int32_t *counter;           // queue head
    int32_t match, mismatch, open, extend, strategy;
    int32_t one;                // = 1, opaque to the compiler: see fma_add()
    uint32_t *bt;               // per-warp scratch: lr_words of last-row values, then the backtrack words
    int64_t bt_stride;          // words per warp (both parts)
    int32_t lr_words;           // multiple of 32
    int2 *elems;                // per-warp CIGAR element scratch
    int32_t elems_stride;
    int32_t *score, *offset, *cigar_len;
    char *cigar;
    int32_t cigar_stride;
};

// per-warp shared memory: last row (copied from the L2 scratch after the DP), last column, alt bytes
static const int BT_TILE = 8;                    // steps per backtrack tile
struct __align__(16) WarpSmem {
    int32_t lastrow[MAX_LEN + 1 + 32], lastcol[MAX_LEN + 1];   // lastrow doubles as the tile staging buffer: 32 x (8 NW + 1) words
    uint8_t alt[MAX_LEN + 16];
};

__device__ __forceinline__ int iabs(int x) { return x < 0 ? -x : x; }

// The DP loop saturates the integer ALU pipe (ncu: 92 % of cycles active) while the FMA pipe, which
// also executes IMAD, idles.  a * one + b with `one` a kernel parameter equal to 1 cannot be folded by
// ptxas into an IADD, so these additions issue as IMAD on the FMA pipe.
__device__ __forceinline__ int32_t fma_add(int32_t a, int32_t one, int32_t b)
{
    int32_t r;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(b));
    return r;
}

// Backtrack nibble of one cell (PairWiseSW.h:8-11,18-20,27-35) added into `word` at bit SHIFT: four
// compares on the ALU pipe, four predicated IMADs on the FMA pipe.
//   INSERT_EXT  <=> !(open_h > ext_h) <=> max(open_h, ext_h) == ext_h
//   DELETE_EXT  likewise for the vertical gap
//   DELETE      <=> f11 > max(h0, e11)            (overrides INSERT, :33)
//   INSERT      <=> e11 > h0 and not DELETE        (:29-31)
template <int SHIFT>
__device__ __forceinline__ void bt_flags(uint32_t &word, int32_t one, int32_t e11, int32_t ext_h, int32_t f11,
                                         int32_t ext_v, int32_t h0, int32_t h1)
{
    asm("{\n\t"
        ".reg .pred pa, pb, pc, pd;\n\t"
        "setp.eq.s32 pa, %1, %2;\n\t"
        "setp.eq.s32 pb, %3, %4;\n\t"
        "setp.gt.s32 pc, %3, %6;\n\t"
        "setp.gt.and.s32 pd, %1, %5, !pc;\n\t"
        "@pa mad.lo.u32 %0, %7, %8, %0;\n\t"
        "@pb mad.lo.u32 %0, %7, %9, %0;\n\t"
        "@pc mad.lo.u32 %0, %7, %10, %0;\n\t"
        "@pd mad.lo.u32 %0, %7, %11, %0;\n\t"
        "}"
        : "+r"(word)
        : "r"(e11), "r"(ext_h), "r"(f11), "r"(ext_v), "r"(h0), "r"(h1), "r"(one),
          "n"((uint32_t)F_INSERT_EXT << SHIFT), "n"((uint32_t)F_DELETE_EXT << SHIFT),
          "n"((uint32_t)OP_DELETE << SHIFT), "n"((uint32_t)OP_INSERT << SHIFT));
}

// Hd + (c1 == c2 ? match : mismatch) without the select and without the MATRIX_MIN_CUTOFF clamp
// (PairWiseSW.h:5-7): one compare on the ALU pipe, two IMADs on the FMA pipe.  Only for parameters
// under which the clamp can never bind (sw_api.cu: cutoff_can_bind).
__device__ __forceinline__ int32_t diag_score(int32_t hd, uint32_t c1, uint32_t c2, int32_t one, int32_t match, int32_t mismatch)
{
    int32_t r;
    asm("{\n\t"
        ".reg .pred pe;\n\t"
        "setp.eq.u32 pe, %2, %3;\n\t"
        "mad.lo.s32 %0, %1, %4, %6;\n\t"
        "@pe mad.lo.s32 %0, %1, %4, %5;\n\t"
        "}"
        : "=&r"(r)
        : "r"(hd), "r"(c1), "r"(c2), "r"(one), "r"(match), "r"(mismatch));
    return r;
}

// first backtrack word of (step, lane) in the tiled per-warp scratch
template <int NW>
__device__ __forceinline__ int64_t bt_word(int st, int ln)
{
    return (((int64_t)(st >> 3) * 32 + ln) * BT_TILE + (st & (BT_TILE - 1))) * NW;
}

// KH rows per lane; a pair runs in the smallest class with 32 * KH >= len1, so one stripe of the
// wavefront always covers the whole matrix (MAX_LEN = 32 * 32) and most lanes hold real rows.
// CUT: keep the MATRIX_MIN_CUTOFF clamp of the diagonal term (parameters large enough for it to bind).
// resident blocks per SM asked of ptxas (caps the registers at what it chose for the first version of this
// kernel; left alone it sometimes trades a block per SM for scheduling freedom the dependency chain of a
// cell cannot use -- measured 945 against 961 GCUPS)
template <int KH>
struct MinBlocks { static constexpr int value = KH <= 4 ? 8 : KH <= 8 ? 7 : KH <= 12 ? 6 : KH <= 20 ? 5 : KH <= 24 ? 4 : 3; };

template <int KH, bool CUT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MinBlocks<KH>::value)
sw_align_kernel(const Params P)
{
    static_assert(KH >= 1 && KH <= 32, "one stripe must cover MAX_LEN rows");
    constexpr int NW = (KH + 7) / 8;                     // backtrack words per lane and step
    constexpr int KH4 = (KH + 3) & ~3;                   // last-row scratch: KH4 values per column
    extern __shared__ __align__(16) uint8_t smem_raw[];
    WarpSmem &S = reinterpret_cast<WarpSmem *>(smem_raw)[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    int32_t *lr = reinterpret_cast<int32_t *>(P.bt + (int64_t)warp_global * P.bt_stride);
    uint32_t *bt = P.bt + (int64_t)warp_global * P.bt_stride + P.lr_words;
    const bool row_cands = (P.strategy == ST_SOFTCLIP || P.strategy == ST_IGNORE);
    int2 *elems = P.elems + (int64_t)warp_global * P.elems_stride;
    const bool lead = (P.strategy == ST_INDEL || P.strategy == ST_LEADING_INDEL);

    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(P.counter, 1);
        t = __shfl_sync(FULL, t, 0);
        if (t >= P.ntasks) break;
        const PairRec pr = P.pairs[t];
        const int nrow = pr.ref_len, ncol = pr.alt_len;
        const uint8_t *seq1 = P.ref_pool + pr.ref_off;
        __syncwarp();
        for (int k = lane; k < ncol; k += 32) S.alt[k] = P.alt_pool[pr.alt_off + k];
        __syncwarp();
        {
            const int i0 = lane * KH + 1;                // first row of this lane (1-based)
            const int lanes_used = (nrow + KH - 1) / KH;
            // column-0 boundary (:196-205): H(i, 0) = open + (i - 1) * extend or 0; E(i, 0) = LOW (:207)
            int32_t Hl[KH], E[KH];
            uint32_t c1[KH];
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const int i = i0 + p;
                Hl[p] = lead ? P.open + (i - 1) * P.extend : 0;
                E[p] = SW_LOW_INIT;
                c1[p] = (i <= nrow) ? (uint32_t)seq1[i - 1] : 0x100u;    // padding rows match nothing
            }
            // H(i0 - 1, 0): diagonal neighbour of the lane's first row at column 1
            int32_t Hd_next = (lane == 0) ? 0 : (lead ? P.open + (i0 - 2) * P.extend : 0);
            int32_t Hbot = 0, Fbot = SW_LOW_INIT;
            const bool lr_store = row_cands && lane == (nrow - 1) / KH;      // last-row candidates: SOFTCLIP / IGNORE only
            const int steps = ncol + lanes_used - 1;
            constexpr int LSTR = BT_TILE * NW + 1;                           // staging stride per lane (odd: conflict-free)
            uint32_t *stg = reinterpret_cast<uint32_t *>(S.lastrow);
            const int tile_words = lanes_used * BT_TILE * NW;                // lanes holding real rows only
            for (int st = 0; st < steps; ++st) {
                const int j = st - lane + 1;             // this lane's column (1-based)
                int32_t Ht = __shfl_up_sync(FULL, Hbot, 1);
                int32_t Ft = __shfl_up_sync(FULL, Fbot, 1);
                if (lane == 0) { Ht = lead ? P.open + (j - 1) * P.extend : 0; Ft = SW_LOW_INIT; }   // row 0 (:196-206)
                if (j >= 1 && j <= ncol && lane < lanes_used) {
                    const uint32_t c2 = S.alt[j - 1];
                    int32_t Hd = Hd_next;
                    Hd_next = Ht;
                    uint32_t word[NW];
#pragma unroll
                    for (int q = 0; q < NW; ++q) word[q] = 0;
                    static_for<KH>([&](auto pc) {
                        constexpr int p = decltype(pc)::value;
                        // MAIN_CODE (:3-37); max(a + b, c) is one VIADDMNMX
                        const int32_t ext_h = fma_add(E[p], P.one, P.extend);
                        const int32_t e11 = __viaddmax_s32(Hl[p], P.open, ext_h);
                        const int32_t ext_v = fma_add(Ft, P.one, P.extend);
                        const int32_t f11 = __viaddmax_s32(Ht, P.open, ext_v);
                        const int32_t h0 = CUT ? __viaddmax_s32(Hd, (c1[p] == c2) ? P.match : P.mismatch, SW_MIN_CUTOFF)
                                               : diag_score(Hd, c1[p], c2, P.one, P.match, P.mismatch);
                        const int32_t h1 = max(h0, e11);
                        const int32_t h11 = max(h1, f11);
                        bt_flags<4 * (p % 8)>(word[p / 8], P.one, e11, ext_h, f11, ext_v, h0, h1);
                        Hd = Hl[p];
                        Ht = h11; Ft = f11;
                        Hl[p] = h11; E[p] = e11;
                    });
                    Hbot = Ht; Fbot = Ft;
#pragma unroll
                    for (int q = 0; q < NW; ++q) stg[lane * LSTR + (st & (BT_TILE - 1)) * NW + q] = word[q];
                    if (lr_store) {
                        int4 *dst = reinterpret_cast<int4 *>(lr + (int64_t)j * KH4);
#pragma unroll
                        for (int q = 0; q < KH4 / 4; ++q)
                            dst[q] = make_int4(Hl[4 * q], Hl[min(4 * q + 1, KH - 1)], Hl[min(4 * q + 2, KH - 1)],
                                               Hl[min(4 * q + 3, KH - 1)]);    // slots >= KH: never read
                    }
                    if (j == ncol) {
#pragma unroll
                        for (int p = 0; p < KH; ++p)
                            if (i0 + p <= nrow) S.lastcol[i0 + p] = Hl[p];
                    }
                }
                if ((st & (BT_TILE - 1)) == BT_TILE - 1 || st == steps - 1) {
                    // tile complete: [lane][step][word] of the lanes in use, contiguous in the scratch
                    __syncwarp();
                    uint32_t *dst = bt + (int64_t)(st >> 3) * (32 * BT_TILE * NW) + lane;
                    if (NW == 3) {
                        for (int wd = lane; wd < tile_words; wd += 32) dst[wd - lane] = stg[(wd / (BT_TILE * NW)) * LSTR + wd % (BT_TILE * NW)];
                    } else {
                        // 32 output words per round = 32 / (8 NW) whole lanes: the staging address is linear in the round
                        constexpr int LPR = 32 / (BT_TILE * NW);                 // lanes per round: 4, 2, 1
                        const uint32_t *src = stg + (lane / (BT_TILE * NW)) * LSTR + lane % (BT_TILE * NW);
#pragma unroll
                        for (int it = 0; it < 32 / LPR; ++it)
                            if (it * LPR < lanes_used) dst[it * 32] = src[it * LPR * LSTR];
                    }
                    __syncwarp();
                }
            }
            __syncwarp();
        }

        // ---- best last-row / last-column cell in the reference's visiting order (:209-236) ----------
        // slot 2d = (nrow, d - nrow) for d >= nrow + 1 (SOFTCLIP / IGNORE only), slot 2d + 1 = (d - ncol, ncol)
        // for d >= ncol + 1, d = 1 .. nrow + ncol
        if (row_cands) {
            const int pr_row = (nrow - 1) % KH;          // register slot of row nrow in its lane
#pragma unroll 4
            for (int k = 1 + lane; k <= ncol; k += 32) S.lastrow[k] = __ldcg(lr + (int64_t)k * KH4 + pr_row);
            __syncwarp();
        }
        const int slot_lo = 2 * (min(nrow, ncol) + 1), slot_hi = 2 * (nrow + ncol) + 2;
        auto slot_score = [&](int slot, int32_t *sc) -> bool {
            const int d = slot >> 1;
            if (slot >= slot_hi) return false;
            if (slot & 1) { if (d < ncol + 1) return false; *sc = S.lastcol[d - ncol]; return true; }
            if (!row_cands || d < nrow + 1) return false;
            *sc = S.lastrow[d - nrow];
            return true;
        };
        int32_t best = INT32_MIN;
        for (int sl = slot_lo + lane; sl < slot_hi; sl += 32) {
            int32_t sc;
            if (slot_score(sl, &sc)) best = max(best, sc);
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) best = max(best, __shfl_xor_sync(FULL, best, o));
        int max_i = 0, max_j = 0;
        bool have = false;
        for (int sl0 = slot_lo; sl0 < slot_hi; sl0 += 32) {
            int32_t sc = 0;
            const bool hit = slot_score(sl0 + lane, &sc) && sc == best;
            unsigned m = __ballot_sync(FULL, hit);
            while (m) {
                const int sl = sl0 + __ffs(m) - 1;
                m &= m - 1;
                const int d = sl >> 1;
                int ci, cj;
                bool take;
                if (sl & 1) { ci = d - ncol; cj = ncol; take = !have || max_j == ncol || iabs(ci - cj) <= iabs(max_i - max_j); }
                else        { ci = nrow; cj = d - nrow; take = !have || iabs(ci - cj) < iabs(max_i - max_j); }
                if (take) { max_i = ci; max_j = cj; have = true; }
            }
        }

        // ---- traceback (getCIGAR :265-356), every lane in lock step on broadcast loads ---------------
        int i, j, n = 0;
        if (P.strategy == ST_INDEL) { i = nrow; j = ncol; }
        else if (P.strategy == ST_LEADING_INDEL) { i = max_i; j = ncol; }
        else { i = max_i; j = max_j; }
        int cur_op = -1, cur_len = 0;                    // the element being built (equal neighbours merge, :372-388)
        auto push = [&](int op, int len) {
            if (op == cur_op) { cur_len += len; return; }
            if (cur_op >= 0) { if (lane == 0) elems[n] = make_int2(cur_op, cur_len); ++n; }
            cur_op = op; cur_len = len;
        };
        if (j < ncol) push(ST_SOFTCLIP, ncol - j);
        int state = 0;
        // window: lane L holds the NW backtrack words of lane w_ln at step w_st - L, so a run of moves inside
        // one lane's rows costs one round trip to L2 (or HBM: the backtrack words of all resident warps exceed
        // L2) instead of one per move.  A second window is fetched ahead for where a diagonal run enters the
        // lane above (8 steps of slack for deletions on the way), so that round trip overlaps the moves.
        int w_ln = -1, w_st = 0, n_ln = -1, n_st = 0;
        uint32_t wv[NW], nv[NW];
#pragma unroll
        for (int q = 0; q < NW; ++q) wv[q] = nv[q] = 0;
        while (i > 0 && j > 0) {
            const int r = i - 1;
            const int ln = r / KH, p = r % KH, q = p / 8;
            const int st = j - 1 + ln;
            if (ln != w_ln || st > w_st || st < w_st - 31) {
                if (ln == n_ln && st <= n_st && st >= n_st - 31) {
                    w_st = n_st;
#pragma unroll
                    for (int k = 0; k < NW; ++k) wv[k] = nv[k];
                } else {
                    w_st = st;
                    const int s2 = st - lane;
#pragma unroll
                    for (int k = 0; k < NW; ++k) wv[k] = (s2 >= 0) ? __ldcg(bt + bt_word<NW>(s2, ln) + k) : 0u;
                }
                w_ln = ln;
                n_ln = ln - 1;
                if (n_ln >= 0) {
                    n_st = min(st - p - 2 + 8, ncol - 1 + n_ln);          // st of (i - p - 1, j - p - 1) in lane ln - 1, + slack
                    const int s2 = n_st - lane;
#pragma unroll
                    for (int k = 0; k < NW; ++k) nv[k] = (s2 >= 0) ? __ldcg(bt + bt_word<NW>(s2, n_ln) + k) : 0u;
                }
            }
            uint32_t mine = wv[0];
#pragma unroll
            for (int k = 1; k < NW; ++k) if (q == k) mine = wv[k];
            const uint32_t w = __shfl_sync(FULL, mine, w_st - st);
            const int btr = (int)((w >> (4 * (p % 8))) & 15u);
            if (state == F_INSERT_EXT) { --j; ++cur_len; state = btr & F_INSERT_EXT; }
            else if (state == F_DELETE_EXT) { --i; ++cur_len; state = btr & F_DELETE_EXT; }
            else {
                const int op = btr & 3;
                if (op == OP_MATCH) { --i; --j; push(OP_MATCH, 1); state = 0; }
                else if (op == OP_INSERT) { --j; push(OP_INSERT, 1); state = btr & F_INSERT_EXT; }
                else { --i; push(OP_DELETE, 1); state = btr & F_DELETE_EXT; }
            }
        }
        int off;
        if (P.strategy == ST_SOFTCLIP) { if (j > 0) push(ST_SOFTCLIP, j); off = i; }
        else if (P.strategy == ST_IGNORE) { if (j > 0) push(cur_op, j); off = i - j; }     // repeats the last op (:363)
        else { if (i > 0) push(OP_DELETE, i); else if (j > 0) push(OP_INSERT, j); off = 0; }
        if (cur_op >= 0) { if (lane == 0) elems[n] = make_int2(cur_op, cur_len); ++n; }
        __syncwarp();

        // ---- CIGAR string, last element first (:390-410) -----------------------------------------------
        if (lane == 0) {
            int need = 0;
            for (int k = 0; k < n; ++k) {
                const int len = elems[k].y;
                need += 1 + (len >= 1000 ? 4 : len >= 100 ? 3 : len >= 10 ? 2 : 1);
            }
            P.score[pr.out] = best;
            P.offset[pr.out] = off;
            if (need + 1 > P.cigar_stride) {
                P.cigar_len[pr.out] = -1 - need;
            } else {
                char *out = P.cigar + (int64_t)pr.out * P.cigar_stride;
                int at = 0;
                for (int k = n - 1; k >= 0; --k) {
                    const int2 e = elems[k];
                    int len = e.y;
                    const int nd = len >= 1000 ? 4 : len >= 100 ? 3 : len >= 10 ? 2 : 1;
                    for (int q = nd - 1; q >= 0; --q) { out[at + q] = (char)('0' + len % 10); len /= 10; }
                    at += nd;
                    out[at++] = e.x == OP_MATCH ? 'M' : e.x == OP_INSERT ? 'I' : e.x == OP_DELETE ? 'D'
                              : e.x == ST_SOFTCLIP ? 'S' : 'R';
                }
                out[at] = 0;
                P.cigar_len[pr.out] = at;
            }
        }
        __syncwarp();
    }
}

}  // namespace sw
#endif
