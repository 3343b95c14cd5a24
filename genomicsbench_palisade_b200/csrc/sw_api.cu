// C ABI of libsw_cuda: context, device-resident workloads, host-buffer entry points and the
// reference's own C entry points.  Contract and reference interfaces: include/sw_cuda.h.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/sw_cuda.h"
#include "sw_kernels.cuh"

using namespace sw;

namespace {
thread_local std::string g_create_error;
// Rows per lane of the kernel classes.  A pair runs in the smallest class with 32 * KH >= len1; the unused
// rows of the last lanes are issue slots spent on nothing, so up to 640 rows (where GATK's haplotypes and
// reference windows are) every KH has its own class; the last class covers SW_MAX_SEQ_LEN = 32 * 32.
#define SW_CLASS_LIST(X) X(0, 1) X(1, 2) X(2, 3) X(3, 4) X(4, 5) X(5, 6) X(6, 7) X(7, 8) X(8, 9) X(9, 10) X(10, 11) \
    X(11, 12) X(12, 13) X(13, 14) X(14, 15) X(15, 16) X(16, 17) X(17, 18) X(18, 19) X(19, 20) X(20, 24) X(21, 28) X(22, 32)
const int N_CLASSES = 23;
#define SW_X_KH(c, kh) kh,
const int CLASS_KH[N_CLASSES] = {SW_CLASS_LIST(SW_X_KH)};

typedef void (*Kernel)(const Params);
template <bool CUT>
Kernel kernel_of(int c)
{
    switch (c) {
#define SW_X_CASE(c, kh) case c: return sw_align_kernel<kh, CUT>;
        SW_CLASS_LIST(SW_X_CASE)
    default: return nullptr;
    }
}
Kernel kernel_for(int c, bool cut) { return cut ? kernel_of<true>(c) : kernel_of<false>(c); }

// The diagonal term max(H(i-1, j-1) + w, MATRIX_MIN_CUTOFF) (PairWiseSW.h:5-7) can only be clamped when some H
// falls below -1e8.  H(i, j) >= H(i-1, j-1) + min(match, mismatch), and the boundary values are 0 or
// open + (k - 1) * extend, so with MAX_LEN rows and columns this bound keeps every H above the cutoff and the
// kernels without the clamp give the same bits.
bool cutoff_can_bind(const sw_params *pm)
{
    const int64_t w = std::min<int64_t>(std::min(pm->match, pm->mismatch), 0);
    const int64_t lo = -(std::llabs((long long)pm->open) + (int64_t)(MAX_LEN - 1) * std::llabs((long long)pm->extend)) + (int64_t)MAX_LEN * w;
    return lo - 1 <= (int64_t)SW_MIN_CUTOFF;
}
}  // namespace

// Streams, events and the grow-only device arena one workload in flight needs.  Class kernels run side by
// side on the side streams: the tail of one persistent kernel is filled by the blocks of the next.
struct Slot {
    static const int N_SIDE = 4;
    cudaStream_t stream = nullptr;
    cudaStream_t side[N_SIDE] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[N_SIDE] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    uint8_t *arena = nullptr;               // host-buffer entry points: no cudaMalloc / cudaFree per call
    size_t arena_cap = 0;
};

struct sw_ctx {
    int device = 0, sm_count = 0;
    // slot 0 also serves the device-resident workload API; sw_align_flat pipelines chunks of pairs
    // through all three (planning + upload of one chunk under the kernels of the others)
    static const int N_SLOTS = 3;
    Slot slot[N_SLOTS];
    int blocks_per_sm[N_CLASSES] = {0};
    std::string error;
    std::recursive_mutex mu;           // sw_align_flat holds it across create / run / results (the arenas are shared)
};

struct sw_workload {
    sw_ctx *ctx = nullptr;
    // SW_CANARY=1: red zone (rounding slack + 256 bytes of 0xA5) behind every device buffer, verified by
    // sw_workload_results after the kernels; SW_CANARY=2 damages one on purpose (the test of the check)
    struct RedZone { const char *name; uint8_t *begin; size_t len; };
    std::vector<RedZone> redzones;
    int64_t n_pairs = 0, n_cells = 0, h2d_bytes = 0, bt_bytes = 0;
    int32_t cigar_stride = 0;
    int32_t class_ntasks[N_CLASSES] = {0};
    int64_t class_off[N_CLASSES] = {0};
    int64_t class_bt_stride[N_CLASSES] = {0};      // words per warp: last-row scratch + backtrack words
    int32_t class_lr_words[N_CLASSES] = {0};
    int class_grid[N_CLASSES] = {0};
    int64_t class_bt_off[N_CLASSES] = {0}, class_elems_off[N_CLASSES] = {0};   // own scratch: classes run concurrently
    int32_t elems_stride = 0;
    uint8_t *d_base = nullptr;           // one allocation
    uint8_t *d_ref = nullptr, *d_alt = nullptr;
    PairRec *d_pairs = nullptr;
    int32_t *d_score = nullptr, *d_offset = nullptr, *d_cigar_len = nullptr, *d_counters = nullptr;
    char *d_cigar = nullptr;
    uint32_t *d_bt = nullptr;
    int2 *d_elems = nullptr;
    int launches = 0;
    bool in_arena = false;               // buffers live in the slot's arena: nothing to free
    Slot *slot = nullptr;
};

#define SW_TRY(ctx, expr)                                                                      \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            (ctx)->error = std::string(#expr) + ": " + cudaGetErrorString(e__);                \
            return SW_ERR_CUDA;                                                                \
        }                                                                                      \
    } while (0)

extern "C" int sw_abi_version(void) { return SW_ABI_VERSION; }

extern "C" const char *sw_last_error(const sw_ctx *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

extern "C" int sw_create(int device, sw_ctx **out)
{
    if (!out) return SW_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                         " (libsw_cuda has no CPU fallback)";
        return SW_ERR_CUDA;
    }
    if (device < 0 || device >= n) { g_create_error = "device index out of range"; return SW_ERR_ARG; }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return SW_ERR_CUDA; }
    if (prop.major != 10) {
        char buf[160];
        snprintf(buf, sizeof buf, "device %d is sm_%d%d; libsw_cuda is built for sm_100a (B200) only", device, prop.major, prop.minor);
        g_create_error = buf;
        return SW_ERR_CUDA;
    }
    sw_ctx *ctx = new sw_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    auto bail = [&](const char *what, cudaError_t err) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
        sw_destroy(ctx);
        return SW_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    for (int j = 0; j < sw_ctx::N_SLOTS; ++j) {
        Slot &sl = ctx->slot[j];
        if ((e = cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
        for (int k = 0; k < 2; ++k)
            if ((e = cudaEventCreate(&sl.ev[k])) != cudaSuccess) return bail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&sl.ev_fork, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        for (int k = 0; k < Slot::N_SIDE; ++k) {
            if ((e = cudaStreamCreateWithFlags(&sl.side[k], cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
            if ((e = cudaEventCreateWithFlags(&sl.ev_join[k], cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        }
    }
    const int smem = WARPS_PER_BLOCK * (int)sizeof(WarpSmem);
    for (int c = 0; c < N_CLASSES; ++c) {
        Kernel k = kernel_for(c, false);
        for (Kernel kk : {k, kernel_for(c, true)})
            if ((e = cudaFuncSetAttribute((const void *)kk, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess)
                return bail("cudaFuncSetAttribute", e);
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void *)k, WARPS_PER_BLOCK * 32, smem) != cudaSuccess || nb < 1) nb = 1;
        ctx->blocks_per_sm[c] = nb;
    }
    *out = ctx;
    return SW_OK;
}

extern "C" int sw_destroy(sw_ctx *ctx)
{
    if (!ctx) return SW_OK;
    cudaSetDevice(ctx->device);
    for (int j = 0; j < sw_ctx::N_SLOTS; ++j) {
        Slot &sl = ctx->slot[j];
        for (int k = 0; k < 2; ++k) if (sl.ev[k]) cudaEventDestroy(sl.ev[k]);
        if (sl.ev_fork) cudaEventDestroy(sl.ev_fork);
        for (int k = 0; k < Slot::N_SIDE; ++k) {
            if (sl.ev_join[k]) cudaEventDestroy(sl.ev_join[k]);
            if (sl.side[k]) cudaStreamDestroy(sl.side[k]);
        }
        cudaFree(sl.arena);
        if (sl.stream) cudaStreamDestroy(sl.stream);
    }
    delete ctx;
    return SW_OK;
}

extern "C" int sw_workload_destroy(sw_workload *w)
{
    if (!w) return SW_OK;
    if (w->ctx && !w->in_arena) { cudaSetDevice(w->ctx->device); cudaFree(w->d_base); }
    delete w;
    return SW_OK;
}

static int check_params(sw_ctx *ctx, const sw_params *pm, int strategy)
{
    if (!pm) { ctx->error = "NULL parameters"; return SW_ERR_ARG; }
    if (strategy < SW_SOFTCLIP || strategy > SW_IGNORE) { ctx->error = "unknown overhang strategy (expected 9..12)"; return SW_ERR_ARG; }
    const int32_t v[4] = {pm->match, pm->mismatch, pm->open, pm->extend};
    for (int k = 0; k < 4; ++k)
        if (v[k] > SW_MAX_ABS_PARAM || v[k] < -SW_MAX_ABS_PARAM) { ctx->error = "scoring parameter out of range"; return SW_ERR_ARG; }
    return SW_OK;
}

static int wl_create(sw_ctx *ctx, const sw_flat_pairs *fp, int32_t cigar_stride, Slot *arena_slot, sw_workload **out);

extern "C" int sw_workload_create(sw_ctx *ctx, const sw_flat_pairs *fp, int32_t cigar_stride, sw_workload **out)
{
    if (!ctx) return SW_ERR_ARG;
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    return wl_create(ctx, fp, cigar_stride, nullptr, out);
}

// arena_slot: place the buffers in that slot's arena (which must be idle) and use its streams; NULL: own
// allocation, slot 0's streams.
static int wl_create(sw_ctx *ctx, const sw_flat_pairs *fp, int32_t cigar_stride, Slot *arena_slot, sw_workload **out)
{
    if (!ctx || !fp || !out) return SW_ERR_ARG;
    *out = nullptr;
    const int64_t n = fp->n_pairs;
    if (n < 0 || n > INT32_MAX || cigar_stride < 2) { ctx->error = "bad pair count or cigar stride"; return SW_ERR_ARG; }
    if (n && (!fp->ref_len || !fp->ref_off || !fp->ref_pool || !fp->alt_len || !fp->alt_off || !fp->alt_pool)) {
        ctx->error = "NULL array in sw_flat_pairs";
        return SW_ERR_ARG;
    }
    // validate; the referenced range of each pool is uploaded as one piece
    int64_t rlo = INT64_MAX, rhi = 0, alo = INT64_MAX, ahi = 0, cells = 0;
    for (int64_t p = 0; p < n; ++p) {
        const int32_t l1 = fp->ref_len[p], l2 = fp->alt_len[p];
        if (l1 <= 0 || l2 <= 0) { ctx->error = "sequence of length <= 0"; return SW_ERR_EMPTY_SEQ; }
        if (l1 > SW_MAX_SEQ_LEN || l2 > SW_MAX_SEQ_LEN) { ctx->error = "sequence longer than SW_MAX_SEQ_LEN"; return SW_ERR_TOO_LONG; }
        if (fp->ref_off[p] < 0 || fp->alt_off[p] < 0) { ctx->error = "negative pool offset"; return SW_ERR_ARG; }
        rlo = std::min(rlo, fp->ref_off[p]); rhi = std::max(rhi, fp->ref_off[p] + l1);
        alo = std::min(alo, fp->alt_off[p]); ahi = std::max(ahi, fp->alt_off[p] + l2);
        cells += (int64_t)l1 * l2;
    }
    if (n == 0) { rlo = rhi = alo = ahi = 0; }
    sw_workload *w = new sw_workload();
    w->ctx = ctx;
    w->n_pairs = n;
    w->n_cells = cells;
    w->cigar_stride = cigar_stride;
    w->slot = arena_slot ? arena_slot : &ctx->slot[0];

    // ---- classes by rows per lane, tasks longest first ----------------------------------------------
    std::vector<PairRec> recs[N_CLASSES];
    int max_sum = 0;
    for (int64_t p = 0; p < n; ++p) {
        const int32_t l1 = fp->ref_len[p], l2 = fp->alt_len[p];
        int c = 0;
        while (c < N_CLASSES - 1 && 32 * CLASS_KH[c] < l1) ++c;
        PairRec r;
        r.ref_off = fp->ref_off[p] - rlo; r.alt_off = fp->alt_off[p] - alo;
        r.ref_len = l1; r.alt_len = l2; r.out = (int32_t)p; r.pad = 0;
        recs[c].push_back(r);
        max_sum = std::max(max_sum, l1 + l2);
        const int kh = CLASS_KH[c], nw = (kh + 7) / 8;
        w->class_bt_stride[c] = std::max<int64_t>(w->class_bt_stride[c], (int64_t)((l2 + 31 + 7) & ~7) * nw * 32);      // whole tiles of eight steps
        w->class_lr_words[c] = std::max<int32_t>(w->class_lr_words[c], ((l2 + 1) * ((kh + 3) & ~3) + 31) & ~31);
        w->bt_bytes += (int64_t)l2 * ((l1 + kh - 1) / kh) * nw * 4;        // one word per active lane, word slot and step
    }
    w->elems_stride = max_sum + 8;
    std::vector<PairRec> all;
    all.reserve((size_t)n);
    int64_t bt_words = 0, warps_sum = 0;
    std::vector<int32_t> bucket_at(4098);
    for (int c = 0; c < N_CLASSES; ++c) {
        // longest first, by a counting sort over cells / 256 (len <= 1024 each: 4097 buckets): the queue only needs
        // the long tasks early, and a comparison sort of 1e5 32-byte records cost as much as a tenth of the kernels
        w->class_off[c] = (int64_t)all.size();
        w->class_ntasks[c] = (int32_t)recs[c].size();
        std::fill(bucket_at.begin(), bucket_at.end(), 0);
        auto key = [](const PairRec &r) { return 4096 - (int)(((int64_t)r.ref_len * r.alt_len) >> 8); };   // 0 = longest
        for (const PairRec &r : recs[c]) ++bucket_at[(size_t)key(r) + 1];
        for (size_t k = 1; k < bucket_at.size(); ++k) bucket_at[k] += bucket_at[k - 1];
        const size_t base_at = all.size();
        all.resize(base_at + recs[c].size());
        for (const PairRec &r : recs[c]) all[base_at + (size_t)bucket_at[(size_t)key(r)]++] = r;
        const int64_t full = (int64_t)ctx->blocks_per_sm[c] * ctx->sm_count;
        const int64_t want = ((int64_t)recs[c].size() + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
        w->class_grid[c] = (int)std::max<int64_t>(1, std::min(full, want));
        const int64_t warps = (int64_t)w->class_grid[c] * WARPS_PER_BLOCK;
        if (recs[c].empty()) continue;
        w->class_bt_stride[c] += w->class_lr_words[c];
        w->class_bt_off[c] = bt_words;
        w->class_elems_off[c] = warps_sum * w->elems_stride;
        bt_words += warps * w->class_bt_stride[c];
        warps_sum += warps;
    }

    auto fail = [&](int code) { sw_workload_destroy(w); return code; };
    if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->error = "cudaSetDevice failed"; return fail(SW_ERR_CUDA); }
    const size_t np = (size_t)std::max<int64_t>(1, n);
    struct Piece { void **ptr; size_t bytes; const char *name; };
    const Piece pieces[] = {
        {(void **)&w->d_ref, (size_t)(rhi - rlo), "ref pool"}, {(void **)&w->d_alt, (size_t)(ahi - alo), "alt pool"},
        {(void **)&w->d_pairs, np * sizeof(PairRec), "pairs"},
        {(void **)&w->d_score, np * 4, "score"}, {(void **)&w->d_offset, np * 4, "offset"},
        {(void **)&w->d_cigar_len, np * 4, "cigar_len"},
        {(void **)&w->d_counters, 64 * 4, "counters"},
        {(void **)&w->d_cigar, np * (size_t)cigar_stride, "cigar"},
        {(void **)&w->d_bt, (size_t)bt_words * 4, "backtrack words"},
        {(void **)&w->d_elems, (size_t)(warps_sum * w->elems_stride) * sizeof(int2), "cigar elements"},
    };
    const char *canary_env = getenv("SW_CANARY");
    const int canary = canary_env ? atoi(canary_env) : 0;
    const size_t red = canary ? 256 : 0;
    size_t total = 0;
    for (const Piece &p : pieces) total += ((std::max<size_t>(p.bytes, 1) + 255) & ~(size_t)255) + red;
    cudaError_t e = cudaSuccess;
    if (arena_slot) {
        // steady-state calls never pay cudaMalloc / cudaFree (both synchronise the device)
        w->in_arena = true;
        if (arena_slot->arena_cap < total) {
            cudaFree(arena_slot->arena);
            arena_slot->arena = nullptr;
            arena_slot->arena_cap = 0;
            if ((e = cudaMalloc((void **)&arena_slot->arena, total + total / 4)) == cudaSuccess) arena_slot->arena_cap = total + total / 4;
        }
        w->d_base = arena_slot->arena;
    } else {
        e = cudaMalloc((void **)&w->d_base, total);
    }
    if (e != cudaSuccess) {
        ctx->error = std::string("cudaMalloc: ") + cudaGetErrorString(e);
        return fail(e == cudaErrorMemoryAllocation ? SW_ERR_NOMEM : SW_ERR_CUDA);
    }
    size_t off = 0;
    for (const Piece &p : pieces) {
        *p.ptr = w->d_base + off;
        const size_t span = ((std::max<size_t>(p.bytes, 1) + 255) & ~(size_t)255) + red;
        if (red) w->redzones.push_back({p.name, w->d_base + off + p.bytes, span - p.bytes});
        off += span;
    }
    cudaStream_t st = w->slot->stream;
    for (const sw_workload::RedZone &z : w->redzones) cudaMemsetAsync(z.begin, 0xA5, z.len, st);
    if (canary == 2 && !w->redzones.empty()) cudaMemsetAsync(w->redzones.back().begin + 7, 0, 1, st);
    if (n) {
        if ((e = cudaMemcpyAsync(w->d_ref, fp->ref_pool + rlo, (size_t)(rhi - rlo), cudaMemcpyHostToDevice, st)) != cudaSuccess ||
            (e = cudaMemcpyAsync(w->d_alt, fp->alt_pool + alo, (size_t)(ahi - alo), cudaMemcpyHostToDevice, st)) != cudaSuccess ||
            (e = cudaMemcpyAsync(w->d_pairs, all.data(), all.size() * sizeof(PairRec), cudaMemcpyHostToDevice, st)) != cudaSuccess ||
            (e = cudaStreamSynchronize(st)) != cudaSuccess) {
            ctx->error = std::string("upload: ") + cudaGetErrorString(e);
            return fail(SW_ERR_CUDA);
        }
    }
    w->h2d_bytes = (rhi - rlo) + (ahi - alo) + (int64_t)all.size() * (int64_t)sizeof(PairRec);
    *out = w;
    return SW_OK;
}

static int wl_launch(sw_ctx *ctx, sw_workload *w, const sw_params *pm, int strategy)
{
    Slot &sl = *w->slot;
    cudaStream_t st = sl.stream;
    SW_TRY(ctx, cudaSetDevice(ctx->device));
    SW_TRY(ctx, cudaMemsetAsync(w->d_counters, 0, 64 * 4, st));
    SW_TRY(ctx, cudaEventRecord(sl.ev[0], st));
    SW_TRY(ctx, cudaEventRecord(sl.ev_fork, st));
    for (int k = 0; k < Slot::N_SIDE; ++k) SW_TRY(ctx, cudaStreamWaitEvent(sl.side[k], sl.ev_fork, 0));
    w->launches = 0;
    const bool cut = cutoff_can_bind(pm);
    // longest-running class first: its tail is covered by the blocks of the next
    for (int c = N_CLASSES - 1; c >= 0; --c) {
        if (w->class_ntasks[c] == 0) continue;
        Params P;
        memset(&P, 0, sizeof P);
        P.ref_pool = w->d_ref; P.alt_pool = w->d_alt;
        P.pairs = w->d_pairs + w->class_off[c];
        P.ntasks = w->class_ntasks[c];
        P.counter = w->d_counters + c;
        P.match = pm->match; P.mismatch = pm->mismatch; P.open = pm->open; P.extend = pm->extend;
        P.strategy = strategy;
        P.one = 1;
        P.bt = w->d_bt + w->class_bt_off[c]; P.bt_stride = w->class_bt_stride[c]; P.lr_words = w->class_lr_words[c];
        P.elems = w->d_elems + w->class_elems_off[c]; P.elems_stride = w->elems_stride;
        P.score = w->d_score; P.offset = w->d_offset; P.cigar_len = w->d_cigar_len;
        P.cigar = w->d_cigar; P.cigar_stride = w->cigar_stride;
        kernel_for(c, cut)<<<w->class_grid[c], WARPS_PER_BLOCK * 32, WARPS_PER_BLOCK * sizeof(WarpSmem), sl.side[w->launches % Slot::N_SIDE]>>>(P);
        SW_TRY(ctx, cudaGetLastError());
        ++w->launches;
    }
    for (int k = 0; k < Slot::N_SIDE; ++k) {
        SW_TRY(ctx, cudaEventRecord(sl.ev_join[k], sl.side[k]));
        SW_TRY(ctx, cudaStreamWaitEvent(st, sl.ev_join[k], 0));
    }
    SW_TRY(ctx, cudaEventRecord(sl.ev[1], st));
    return SW_OK;
}

static void fill_stats(sw_ctx *ctx, sw_workload *w, sw_stats *stats)
{
    if (!stats) return;
    memset(stats, 0, sizeof *stats);
    stats->n_pairs = w->n_pairs;
    stats->n_cells = w->n_cells;
    stats->h2d_bytes = w->h2d_bytes;
    stats->bt_bytes = w->bt_bytes;
    stats->n_launches = w->launches;
    for (int c = 0; c < N_CLASSES; ++c) stats->n_classes += w->class_ntasks[c] > 0;
    float ms = 0.0f;
    if (w->launches && cudaEventElapsedTime(&ms, w->slot->ev[0], w->slot->ev[1]) == cudaSuccess) stats->ms_kernels = ms;
}

extern "C" int sw_workload_run(sw_ctx *ctx, sw_workload *w, const sw_params *pm, int strategy, sw_stats *stats)
{
    if (!ctx || !w || w->ctx != ctx) return SW_ERR_ARG;
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    int rc = check_params(ctx, pm, strategy);
    if (rc != SW_OK) return rc;
    rc = wl_launch(ctx, w, pm, strategy);
    if (rc != SW_OK) return rc;
    SW_TRY(ctx, cudaStreamSynchronize(w->slot->stream));
    fill_stats(ctx, w, stats);
    return SW_OK;
}

extern "C" int sw_workload_results(sw_ctx *ctx, sw_workload *w, int32_t *score, int32_t *offset, char *cigar,
                                   int32_t *cigar_len)
{
    if (!ctx || !w || w->ctx != ctx) return SW_ERR_ARG;
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    SW_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)w->n_pairs;
    if (n == 0) return SW_OK;
    cudaStream_t st = w->slot->stream;
    if (score) SW_TRY(ctx, cudaMemcpyAsync(score, w->d_score, n * 4, cudaMemcpyDeviceToHost, st));
    if (offset) SW_TRY(ctx, cudaMemcpyAsync(offset, w->d_offset, n * 4, cudaMemcpyDeviceToHost, st));
    if (cigar_len) SW_TRY(ctx, cudaMemcpyAsync(cigar_len, w->d_cigar_len, n * 4, cudaMemcpyDeviceToHost, st));
    if (cigar) SW_TRY(ctx, cudaMemcpyAsync(cigar, w->d_cigar, n * (size_t)w->cigar_stride, cudaMemcpyDeviceToHost, st));
    SW_TRY(ctx, cudaStreamSynchronize(st));
    std::vector<uint8_t> host;
    for (const sw_workload::RedZone &z : w->redzones) {
        host.resize(z.len);
        SW_TRY(ctx, cudaMemcpy(host.data(), z.begin, z.len, cudaMemcpyDeviceToHost));
        for (size_t k = 0; k < z.len; ++k)
            if (host[k] != 0xA5) {
                ctx->error = std::string("SW_CANARY: out-of-bounds write ") + std::to_string(k) +
                             " bytes behind the end of device buffer '" + z.name + "'";
                return SW_ERR_CUDA;
            }
    }
    return SW_OK;
}

extern "C" int sw_align_flat(sw_ctx *ctx, const sw_flat_pairs *fp, const sw_params *pm, int strategy, int32_t *score,
                             int32_t *offset, char *cigar, int32_t cigar_stride, int32_t *cigar_len, sw_stats *stats)
{
    if (!ctx || !fp) return SW_ERR_ARG;
    int rc = check_params(ctx, pm, strategy);
    if (rc != SW_OK) return rc;
    if (fp->n_pairs && (!cigar || !cigar_len)) { ctx->error = "NULL cigar output"; return SW_ERR_ARG; }
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    // Chunks of consecutive pairs, pipelined through the two slots: the (pageable, hence staged) upload
    // and the class planning of chunk j + 1 run under the kernels of chunk j; results of chunk j are
    // copied back when its slot is needed again.  A quarter of the work per chunk, at least 1e9 cells.
    const int64_t n = fp->n_pairs;
    if (n < 0 || n > INT32_MAX) { ctx->error = "bad pair count"; return SW_ERR_ARG; }
    if (n && (!fp->ref_len || !fp->ref_off || !fp->ref_pool || !fp->alt_len || !fp->alt_off || !fp->alt_pool)) {
        ctx->error = "NULL array in sw_flat_pairs";
        return SW_ERR_ARG;
    }
    if (n == 0) { if (stats) memset(stats, 0, sizeof *stats); return SW_OK; }
    int64_t total_cells = 0;
    for (int64_t p = 0; p < n; ++p) {
        if (fp->ref_len[p] <= 0 || fp->alt_len[p] <= 0) { ctx->error = "sequence of length <= 0"; return SW_ERR_EMPTY_SEQ; }
        total_cells += (int64_t)fp->ref_len[p] * fp->alt_len[p];
    }
    // chunk sizes as fractions of the work: a small first chunk starts the device early, a small last one leaves
    // little to copy back after the final kernel; few chunks in between, because every chunk launches every kernel
    // class again with a fraction of its tasks (SW_CHUNK_SCHED overrides; chunks of at least 1e9 cells)
    static const std::vector<double> sched = []() {
        std::vector<double> f;
        if (const char *e = getenv("SW_CHUNK_SCHED")) {
            for (const char *q = e; *q;) { char *end; const double v = strtod(q, &end); if (end == q) break; f.push_back(v); q = *end ? end + 1 : end; }
        }
        if (f.empty()) f = {0.08, 0.32, 0.40, 0.20};
        return f;
    }();
    std::vector<int64_t> cut(1, 0);
    for (int64_t p = 0, acc = 0; p < n; ++p) {
        acc += (int64_t)fp->ref_len[p] * fp->alt_len[p];
        const size_t k = cut.size() - 1;
        const int64_t want = std::max<int64_t>(1000000000ll, (int64_t)(total_cells * (k < sched.size() ? sched[k] : 1.0)));
        if ((acc >= want && k + 1 < sched.size()) || p == n - 1) { cut.push_back(p + 1); acc = 0; }
    }
    if (cut.size() == 1) cut.push_back(0);
    const size_t nchunks = cut.size() - 1;
    sw_workload *wl[sw_ctx::N_SLOTS] = {};
    size_t chunk_of[sw_ctx::N_SLOTS] = {};
    sw_stats agg;
    memset(&agg, 0, sizeof agg);
    auto retire = [&](int slot) -> int {
        sw_workload *w = wl[slot];
        wl[slot] = nullptr;
        const int64_t p0 = cut[chunk_of[slot]];
        int r = SW_OK;
        if (cudaStreamSynchronize(w->slot->stream) != cudaSuccess) { ctx->error = "kernel execution failed"; r = SW_ERR_CUDA; }
        sw_stats s;
        if (r == SW_OK) {
            fill_stats(ctx, w, &s);
            r = sw_workload_results(ctx, w, score ? score + p0 : nullptr, offset ? offset + p0 : nullptr,
                                    cigar + p0 * (int64_t)cigar_stride, cigar_len + p0);
        }
        if (r == SW_OK) {
            agg.n_pairs += s.n_pairs; agg.n_cells += s.n_cells; agg.h2d_bytes += s.h2d_bytes; agg.bt_bytes += s.bt_bytes;
            agg.n_launches += s.n_launches; agg.n_classes = std::max(agg.n_classes, s.n_classes);
            agg.ms_kernels += s.ms_kernels;
        }
        sw_workload_destroy(w);
        return r;
    };
    for (size_t j = 0; j < nchunks && rc == SW_OK; ++j) {
        const int slot = (int)(j % sw_ctx::N_SLOTS);
        if (wl[slot]) rc = retire(slot);
        if (rc != SW_OK) break;
        chunk_of[slot] = j;
        sw_flat_pairs sub = *fp;
        sub.n_pairs = cut[j + 1] - cut[j];
        sub.ref_len = fp->ref_len + cut[j]; sub.ref_off = fp->ref_off + cut[j];
        sub.alt_len = fp->alt_len + cut[j]; sub.alt_off = fp->alt_off + cut[j];
        rc = wl_create(ctx, &sub, cigar_stride, &ctx->slot[slot], &wl[slot]);
        if (rc == SW_OK) rc = wl_launch(ctx, wl[slot], pm, strategy);
    }
    // oldest chunk first
    for (int k = 0; k < sw_ctx::N_SLOTS; ++k) {
        int slot = -1;
        for (int q = 0; q < sw_ctx::N_SLOTS; ++q)
            if (wl[q] && (slot < 0 || chunk_of[q] < chunk_of[slot])) slot = q;
        if (slot < 0) break;
        const int r = retire(slot);
        if (rc == SW_OK) rc = r;
    }
    if (rc != SW_OK) return rc;
    clock_gettime(CLOCK_MONOTONIC, &t1);
    for (int64_t p = 0; p < n; ++p) agg.n_overflow += cigar_len[p] < 0;
    agg.d2h_bytes = n * (12 + (int64_t)cigar_stride);
    agg.ms_total = (float)((t1.tv_sec - t0.tv_sec) * 1e3 + (t1.tv_nsec - t0.tv_nsec) * 1e-6);
    if (stats) *stats = agg;
    return SW_OK;
}

extern "C" int sw_align_one(sw_ctx *ctx, const uint8_t *ref, int32_t ref_len, const uint8_t *alt, int32_t alt_len,
                            const sw_params *pm, int strategy, char *cigar, int32_t cigar_cap, int32_t *score,
                            int32_t *offset)
{
    if (!ctx || !ref || !alt || !cigar) return SW_ERR_ARG;
    const int64_t zero = 0;
    sw_flat_pairs fp;
    fp.n_pairs = 1;
    fp.ref_len = &ref_len; fp.ref_off = &zero; fp.ref_pool = ref;
    fp.alt_len = &alt_len; fp.alt_off = &zero; fp.alt_pool = alt;
    // device slot sized for the longest possible string (one element per traceback step); only
    // the string itself is copied into the caller's buffer
    const int32_t stride = 5 * (ref_len + alt_len) + 16;
    std::vector<char> tmp((size_t)std::max(stride, 16));
    int32_t len = 0;
    int rc = sw_align_flat(ctx, &fp, pm, strategy, score, offset, tmp.data(), stride, &len, nullptr);
    if (rc != SW_OK) return rc;
    if (len < 0 || len + 1 > cigar_cap) { ctx->error = "CIGAR does not fit cigar_cap"; return SW_ERR_ARG; }
    memcpy(cigar, tmp.data(), (size_t)len + 1);
    return SW_OK;
}

// ============================================================================================
// the reference's C entry points (IntelSmithWatermanCSource.cpp:22-57)
// ============================================================================================
namespace {
sw_ctx *g_legacy = nullptr;
std::mutex g_legacy_mu;
}

extern "C" void initNative(void)
{
    std::lock_guard<std::mutex> lk(g_legacy_mu);
    if (g_legacy) return;
    const char *d = getenv("SW_DEVICE");
    if (sw_create(d ? atoi(d) : 0, &g_legacy) != SW_OK) {
        fprintf(stderr, "libsw_cuda: initNative: %s\n", sw_last_error(nullptr));
        abort();
    }
}

extern "C" void align(char *ref, char *alt, char *cigar, int match, int mismatch, int open, int extend, int strategy,
                      int *score, int *offset)
{
    if (!g_legacy) initNative();
    const sw_params pm = {match, mismatch, open, extend};
    const int l1 = (int)strlen(ref), l2 = (int)strlen(alt);
    int32_t sc = 0, off = 0;
    // the reference sprintf()s into `cigar` without a bound: exactly strlen + 1 bytes are written here too
    const int rc = sw_align_one(g_legacy, (const uint8_t *)ref, l1, (const uint8_t *)alt, l2, &pm, strategy, cigar,
                                INT32_MAX, &sc, &off);
    if (rc != SW_OK) {
        fprintf(stderr, "libsw_cuda: align: %s\n", sw_last_error(g_legacy));
        abort();
    }
    if (score) *score = sc;
    if (offset) *offset = off;
}
