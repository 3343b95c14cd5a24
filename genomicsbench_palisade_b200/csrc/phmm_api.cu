// C ABI of libphmm_cuda: context, device-resident workloads, host-buffer entry points.
// See include/phmm_cuda.h for the contract and the reference interfaces each entry replaces.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <string>
#include <unordered_map>
#include <vector>

#include "phmm_host.h"

using namespace phmm;

namespace {
thread_local std::string g_create_error;

// Streams and events one workload in flight needs.  Class kernels of one phase run on the side
// streams so that the tail of one persistent kernel (SMs idling while its last tasks finish) is
// filled by the blocks of the next.
struct StreamSet {
    static const int N_SIDE = 8;       // 4 in use for large workloads, all for small calls (wl_launch)
    cudaStream_t main = nullptr;
    cudaStream_t side[N_SIDE] = {};
    cudaEvent_t ev_fork = nullptr, ev_join[N_SIDE] = {};
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};     // timing: start, FP32 done, FP64 done
    cudaEvent_t ev_uploaded = nullptr;                     // inputs are in HBM (recorded on the upload stream)

    cudaError_t create()
    {
        cudaError_t e;
        if ((e = cudaStreamCreateWithFlags(&main, cudaStreamNonBlocking)) != cudaSuccess) return e;
        for (int k = 0; k < 3; ++k)
            if ((e = cudaEventCreate(&ev[k])) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&ev_uploaded, cudaEventDisableTiming)) != cudaSuccess) return e;
        for (int k = 0; k < N_SIDE; ++k) {
            if ((e = cudaStreamCreateWithFlags(&side[k], cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&ev_join[k], cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        return cudaSuccess;
    }
    void destroy()
    {
        for (int k = 0; k < 3; ++k) if (ev[k]) cudaEventDestroy(ev[k]);
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_uploaded) cudaEventDestroy(ev_uploaded);
        for (int k = 0; k < N_SIDE; ++k) {
            if (ev_join[k]) cudaEventDestroy(ev_join[k]);
            if (side[k]) cudaStreamDestroy(side[k]);
        }
        if (main) cudaStreamDestroy(main);
    }
};

// Grow-only device arena: the host-buffer entry points carve a chunk's buffers out of it instead
// of paying cudaMalloc / cudaFree (which synchronise the device) for every call.
struct Arena {
    uint8_t *base = nullptr;
    size_t cap = 0, off = 0;
    // pinned staging for small chunks (one active region per call): every uploaded piece is laid out in it as
    // in the arena and goes up with one copy instead of a dozen pageable ones (~8 us of driver time each).  Not
    // for large chunks: laying 35 MB out in a staging buffer first and copying then is slower than the driver's
    // own pipelined staging of pageable memory (measured: C2 end to end 0.941 -> 0.935 of resident, the one-GPU
    // CLI 3009 -> 2729 GCUPS with a 128 MB cap)
    uint8_t *stage = nullptr;
    size_t stage_cap = 0;
    static constexpr size_t STAGE_MAX = 4u << 20;
};
}  // namespace

struct phmm_ctx {
    int device = 0;
    int sm_count = 0;
    float spec_band = 0.015625f;   // FP64-first decision band (relative half-width, PHMM_SPEC_BAND: tests widen it)
    int canary = 0;           // PHMM_CANARY=1: red zones behind every device buffer, checked after each run
    // slot 0 also serves the device-resident workload API; the host-buffer entry points pipeline
    // chunks of batches through all slots (plan + upload of one chunk under the kernels of the others)
    static const int N_SLOTS = 4;
    StreamSet ss[N_SLOTS];
    Arena arena[N_SLOTS];
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    // device tables
    float *d_ph2pr_f = nullptr, *d_mm_f = nullptr;
    double *d_ph2pr_d = nullptr, *d_mm_d = nullptr;
    float *d_bench = nullptr;
    // options
    int precision = 0, fp32_fast = 0, fp64_fast = 0, seg_cols = 0;
    int force_kh = 0, force_lanes = 0;
    int no_hp = 0;     // PHMM_NO_HP=1: fast FP32 kernels in the two-reads packing (A/B measurements)
    int64_t chunk_cells = 0;   // host-buffer entry points: cells per pipelined chunk (0 = default)
    int plan_mode = 0;         // PHMM_OPT_PLAN
    std::string error;
    std::mutex mu;     // the legacy symbols may be entered from several OpenMP threads
    // resident blocks per SM of each kernel (asked of the runtime once, not at every launch)
    std::unordered_map<const void *, int> blocks_per_sm;
    // phmm_submit_pairs / phmm_wait: submissions in flight, oldest first
    struct Inflight {
        struct Scatter { double *out; int64_t p0, n; };
        phmm_workload *w = nullptr;
        int slot = 0;
        int64_t ticket = 0;
        std::vector<Scatter> scatter;
    };
    std::deque<Inflight> inflight;
    int64_t next_ticket = 0;
    phmm_stats retired_stats = {};       // of submissions a later submit had to retire itself; handed out by the next phmm_wait
};

struct ClassDev {
    KClass kc;
    int64_t task_off = 0;     // into d_tasks / d_htasks
    int32_t ntasks = 0;
};

struct phmm_workload {
    phmm_ctx *ctx = nullptr;
    Plan plan;                // tasks/stream vectors are released after upload
    std::vector<ClassDev> classes;
    int64_t n_pairs = 0, n_cells = 0, h2d_bytes = 0;
    int64_t resc_total = 0;
    int32_t resc_off[MAX_KH_SCALAR + 1];
    int32_t resc_cap[MAX_KH_SCALAR + 1];
    int precision = 0;
    int planned_hp = 0;               // the plan holds haplotype-pair tasks: needs the fast FP32 kernels
    Arena *arena = nullptr;           // non-NULL: buffers live in a context arena, nothing to free
    // device buffers
    uint8_t *d_pool = nullptr;        // 5 read pools back to back
    int64_t pool_stride = 0;
    uint8_t *d_stream = nullptr;
    uint8_t *d_hapraw = nullptr;      // referenced pieces of the caller's haplotype pool (raw bases)
    HapSrc *d_hapsrc = nullptr;
    uint8_t *d_unused = nullptr;      // placeholder of whichever of the two haplotype-stream pieces is not in use
    int32_t *d_hap_len = nullptr;
    ReadRec *d_reads = nullptr;
    Task *d_tasks = nullptr;
    HTask *d_htasks = nullptr;
    Task *d_resc = nullptr, *d_guard = nullptr;
    // FP64-first reads (Plan::spec_tasks): their tasks as uploaded (copied to the heads of the rescue lists on the
    // device), and the list of pairs inside the decision band (KernelParams::spec_list)
    Task *d_spec_tasks = nullptr, *d_spec = nullptr;
    int32_t spec_ntasks[MAX_KH_SCALAR + 1] = {0}, spec_off[MAX_KH_SCALAR + 1] = {0};
    int64_t spec_pairs = 0, spec_cells = 0;
    int64_t last_spec_direct = 0;
    int32_t *d_ints = nullptr;        // [0..63] queue heads, [64..] ntasks per class, resc/guard counts
    double *d_out = nullptr, *d_raw64 = nullptr;
    float *d_raw32 = nullptr;
    uint8_t *d_rescued = nullptr;
    double *d_scratch = nullptr;      // multi-pass kernels: boundary rows, one slot per resident warp
    int64_t scratch_stride = 0;       // doubles per warp
    int scratch_blocks = 0;           // grid cap of the multi-pass kernels
    int launches = 0;                 // kernels of the last launch
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};   // timing of the last launch: start, FP32 done, FP64 done
    int64_t last_rescued = 0, last_rescued_cells = 0, last_guard = 0;
    // PHMM_CANARY: [begin, end) of the red zone behind each buffer (the rounding slack + 256 bytes of 0xA5)
    struct RedZone { const char *name; uint8_t *begin; size_t len; };
    std::vector<RedZone> redzones;
};

namespace {

// layout of phmm_workload::d_ints
const int INTS_COUNTER = 0;                       // one queue head per launch: [0, MAX_STEP_LAUNCHES) for the step, then
const int COUNTER_SLOTS = 96;                     //   MAX_KH_SCALAR + 1 for the bit-exact FP64 re-run of wl_finish and as many for
const int MAX_STEP_LAUNCHES = COUNTER_SLOTS - 2 * (MAX_KH_SCALAR + 1);   //   its bit-exact FP32 pass over the FP64-first band list
const int SPEC_FIX_SLOT = MAX_STEP_LAUNCHES + MAX_KH_SCALAR + 1;
const int INTS_NTASKS = 96;                       // per FP32 class (64)
const int INTS_RESC = 160;                        // [MAX_KH_SCALAR + 1]
const int INTS_GUARD = 160 + 16;
const int INTS_GUARD64 = 160 + 32;                // fast FP64 kernels: non-finite sums, re-run by the exact FP64 kernel
const int INTS_RESC_PAIRS = 160 + 48;             // pairs in the rescue lists (a task of the fast kernels holds a run of haplotypes)
const int INTS_SPEC = 160 + 49;                   // FP64-first: [9] band-list tasks per class, [+9] rescued as predicted, [+10] finalised in float
const int INTS_LISTS = 64;                        // ints from INTS_RESC on that a step clears and wl_finish reads back
const int INTS_TOTAL = 224;

#define CUDA_TRY(ctx, expr)                                                                   \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            (ctx)->error = std::string(#expr) + ": " + cudaGetErrorString(e__);               \
            return PHMM_ERR_CUDA;                                                             \
        }                                                                                     \
    } while (0)

double wall_ms()
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

}  // namespace

// Builds the haplotype stream on the device: one warp per haplotype turns its raw bases into code bytes at its
// columns (base_code(): ConvertChar + the match rule of precompute_masks_, phmm_types.h); separator and slack
// bytes stay at the 0 the stream was set to.
__global__ void phmm_code_stream_kernel(const uint8_t *__restrict__ raw, const phmm::HapSrc *__restrict__ tab, int n,
                                        uint8_t *__restrict__ stream)
{
    const int warp = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (warp >= n) return;
    const phmm::HapSrc hs = tab[warp];
    const uint8_t *src = raw + hs.src;
    uint8_t *dst = stream + hs.dst;
    for (int k = lane; k < hs.len; k += 32) dst[k] = phmm::base_code(src[k]);
}

namespace phmm {
KernelF32 lookup_f32(const KClass &kc, bool fast)
{
    if (kc.vl != 2) return nullptr;
    if (kc.hp) {
        if (!fast) return nullptr;
        return kc.lanes == 32 ? lookup_f32_hp_l32(kc.kh) : kc.lanes == 16 ? lookup_f32_hp_l16(kc.kh)
             : kc.lanes == 8 ? lookup_f32_hp_l8(kc.kh) : nullptr;
    }
    if (kc.longr) return lookup_f32_long(fast);
    switch (kc.lanes) {
    case 32: return fast ? lookup_f32_l32_f1(kc.kh) : lookup_f32_l32_f0(kc.kh);
    case 16: return fast ? lookup_f32_l16_f1(kc.kh) : lookup_f32_l16_f0(kc.kh);
    default: return nullptr;
    }
}
KernelF64 lookup_f64(int kh, bool fast) { return fast ? lookup_f64_f1(kh) : lookup_f64_f0(kh); }
}  // namespace phmm

// ============================================================================================
// lifetime
// ============================================================================================
extern "C" int phmm_abi_version(void) { return PHMM_ABI_VERSION; }

extern "C" int phmm_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" const char *phmm_last_error(const phmm_ctx *ctx)
{
    return ctx ? ctx->error.c_str() : g_create_error.c_str();
}

extern "C" int phmm_create(int device, phmm_ctx **out)
{
    if (!out) return PHMM_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                         " (libphmm_cuda has no CPU fallback)";
        return PHMM_ERR_CUDA;
    }
    if (device < 0 || device >= n) { g_create_error = "device index out of range"; return PHMM_ERR_ARG; }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) {
        g_create_error = std::string("cudaGetDeviceProperties: ") + cudaGetErrorString(e);
        return PHMM_ERR_CUDA;
    }
    if (prop.major != 10) {
        char buf[160];
        snprintf(buf, sizeof buf, "device %d is sm_%d%d; libphmm_cuda is built for sm_100a (B200) only", device,
                 prop.major, prop.minor);
        g_create_error = buf;
        return PHMM_ERR_CUDA;
    }
    phmm_ctx *ctx = new phmm_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    if (const char *s = getenv("PHMM_FORCE_CLASS")) sscanf(s, "%d,%d", &ctx->force_kh, &ctx->force_lanes);
    if (const char *s = getenv("PHMM_NO_HP")) ctx->no_hp = atoi(s);
    if (const char *s = getenv("PHMM_CANARY")) ctx->canary = atoi(s);
    if (const char *s = getenv("PHMM_SPEC_BAND")) ctx->spec_band = (float)atof(s);
    if (const char *s = getenv("PHMM_CHUNK_CELLS")) ctx->chunk_cells = atoll(s);
    auto bail = [&](const char *what, cudaError_t err) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
        phmm_destroy(ctx);
        return PHMM_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    for (int k = 0; k < phmm_ctx::N_SLOTS; ++k)
        if ((e = ctx->ss[k].create()) != cudaSuccess) return bail("stream / event creation", e);
    if ((e = cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    if ((e = cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    const HostTables &t = host_tables();
    if ((e = cudaMalloc(&ctx->d_ph2pr_f, 128 * sizeof(float))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&ctx->d_ph2pr_d, 128 * sizeof(double))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&ctx->d_mm_f, MM_REACHABLE * sizeof(float))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&ctx->d_mm_d, MM_REACHABLE * sizeof(double))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&ctx->d_bench, 1 << 20)) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMemcpy(ctx->d_ph2pr_f, t.ph2pr_f, 128 * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);
    if ((e = cudaMemcpy(ctx->d_ph2pr_d, t.ph2pr_d, 128 * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);
    if ((e = cudaMemcpy(ctx->d_mm_f, t.mm_f.data(), MM_REACHABLE * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);
    if ((e = cudaMemcpy(ctx->d_mm_d, t.mm_d.data(), MM_REACHABLE * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);
    *out = ctx;
    return PHMM_OK;
}

extern "C" int phmm_destroy(phmm_ctx *ctx)
{
    if (!ctx) return PHMM_OK;
    cudaSetDevice(ctx->device);
    if (!ctx->inflight.empty()) {
        cudaDeviceSynchronize();
        for (phmm_ctx::Inflight &f : ctx->inflight) phmm_workload_destroy(f.w);
        ctx->inflight.clear();
    }
    cudaFree(ctx->d_ph2pr_f); cudaFree(ctx->d_ph2pr_d); cudaFree(ctx->d_mm_f); cudaFree(ctx->d_mm_d);
    cudaFree(ctx->d_bench);
    for (int k = 0; k < phmm_ctx::N_SLOTS; ++k) {
        cudaFree(ctx->arena[k].base);
        if (ctx->arena[k].stage) cudaFreeHost(ctx->arena[k].stage);
        ctx->ss[k].destroy();
    }
    if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
    if (ctx->copy_out) cudaStreamDestroy(ctx->copy_out);
    delete ctx;
    return PHMM_OK;
}

extern "C" int phmm_set_option(phmm_ctx *ctx, int option, int64_t value)
{
    if (!ctx) return PHMM_ERR_ARG;
    switch (option) {
    case PHMM_OPT_PRECISION:   if (value < 0 || value > 2) break; ctx->precision = (int)value; return PHMM_OK;
    case PHMM_OPT_FP32_KERNEL: if (value < 0 || value > 1) break; ctx->fp32_fast = (int)value; return PHMM_OK;
    case PHMM_OPT_FP64_KERNEL: if (value < 0 || value > 1) break; ctx->fp64_fast = (int)value; return PHMM_OK;
    case PHMM_OPT_SEG_COLS:    if (value < 0 || value > (1 << 24)) break; ctx->seg_cols = (int)value; return PHMM_OK;
    case PHMM_OPT_FORCE_CLASS: ctx->force_kh = (int)(value >> 8); ctx->force_lanes = (int)(value & 255); return PHMM_OK;
    case PHMM_OPT_CHUNK_CELLS: if (value < 0) break; ctx->chunk_cells = value; return PHMM_OK;
    case PHMM_OPT_PLAN:        if (value < 0 || value > 2) break; ctx->plan_mode = (int)value; return PHMM_OK;
    default: break;
    }
    ctx->error = "phmm_set_option: unknown option or value out of range";
    return PHMM_ERR_ARG;
}

extern "C" int64_t phmm_get_option(const phmm_ctx *ctx, int option)
{
    if (!ctx) return -1;
    switch (option) {
    case PHMM_OPT_PRECISION: return ctx->precision;
    case PHMM_OPT_FP32_KERNEL: return ctx->fp32_fast;
    case PHMM_OPT_FP64_KERNEL: return ctx->fp64_fast;
    case PHMM_OPT_SEG_COLS: return ctx->seg_cols;
    case PHMM_OPT_FORCE_CLASS: return ((int64_t)ctx->force_kh << 8) | ctx->force_lanes;
    case PHMM_OPT_CHUNK_CELLS: return ctx->chunk_cells;
    case PHMM_OPT_PLAN: return ctx->plan_mode;
    default: return -1;
    }
}

extern "C" int phmm_get_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d, float *consts_f,
                               double *consts_d)
{
    const HostTables &t = host_tables();
    if (ph2pr_f) memcpy(ph2pr_f, t.ph2pr_f, sizeof t.ph2pr_f);
    if (ph2pr_d) memcpy(ph2pr_d, t.ph2pr_d, sizeof t.ph2pr_d);
    if (mm_f) memcpy(mm_f, t.mm_f.data(), MM_REACHABLE * sizeof(float));
    if (mm_d) memcpy(mm_d, t.mm_d.data(), MM_REACHABLE * sizeof(double));
    if (consts_f) { consts_f[0] = t.init_f; consts_f[1] = t.log10_init_f; }
    if (consts_d) { consts_d[0] = t.init_d; consts_d[1] = t.log10_init_d; }
    return MM_REACHABLE;
}
namespace {
// A range of whole batches of a phmm_flat_batches and the reads / haplotypes / pairs it spans.
struct Chunk { int64_t b0, b1, r0, r1, h0, h1, p0, p1; };

// Cuts the batches into consecutive chunks of whole batches.
//   per > 0: about `per` cells each (PHMM_OPT_CHUNK_CELLS);
//   per = 0: a schedule that starts the device early and ends the call early -- the first chunk is a 64th of the
//   work (its plan + upload is all the caller waits for before kernels run), every next one four times the last up
//   to a third of the work (at most max_cells: device memory), and the last ones halve again, so that what is left
//   after the final kernel (copy-back of one chunk's results, the tail of its last tasks) is small.  Fewer, larger
//   chunks in the middle: every chunk has its own phase joins and kernel tails (C2: 6 equal chunks 37.0 ms,
//   this schedule 36.2; profiles/README.md).
int split_batches(phmm_ctx *ctx, const phmm_flat_batches *fb, int64_t per, int64_t min_cells, int64_t max_cells,
                  std::vector<Chunk> *out)
{
    std::vector<Chunk> &chunks = *out;
    chunks.clear();
    std::vector<int64_t> cells((size_t)fb->n_batches, 0);
    int64_t r = 0, h = 0, total = 0;
    for (int64_t b = 0; b < fb->n_batches; ++b) {
        const int nr = fb->batch_num_reads[b], nh = fb->batch_num_haps[b];
        if (nr < 0 || nh < 0 || r + nr > fb->n_reads || h + nh > fb->n_haps) {
            ctx->error = "batch tables do not add up to n_reads / n_haps";
            return PHMM_ERR_ARG;
        }
        int64_t sr = 0, sh = 0;
        for (int k = 0; k < nr; ++k) sr += fb->read_len[r + k];
        for (int k = 0; k < nh; ++k) sh += fb->hap_len[h + k];
        cells[b] = sr * sh;
        total += cells[b];
        r += nr; h += nh;
    }
    static const struct Sched { double first = 64, growth = 4, cap = 3, tail = 4, floor = 4e9; Sched() {
        if (const char *e = getenv("PHMM_CHUNK_SCHED")) sscanf(e, "%lf,%lf,%lf,%lf,%lf", &first, &growth, &cap, &tail, &floor); } } sched;
    const double cap = std::max<double>((double)min_cells, std::min<double>((double)max_cells, total / sched.cap));
    double want = per > 0 ? (double)per : std::max<double>(total / sched.first, (double)min_cells / 4);
    Chunk c{0, 0, 0, 0, 0, 0, 0, 0};
    int64_t acc = 0, p = 0, done = 0;
    r = 0; h = 0;
    for (int64_t b = 0; b < fb->n_batches; ++b) {
        const int nr = fb->batch_num_reads[b], nh = fb->batch_num_haps[b];
        acc += cells[b];
        r += nr; h += nh; p += (int64_t)nr * nh;
        if (acc >= want || b == fb->n_batches - 1) {
            c.b1 = b + 1; c.r1 = r; c.h1 = h; c.p1 = p;
            chunks.push_back(c);
            c = Chunk{b + 1, b + 1, r, r, h, h, p, p};
            done += acc;
            acc = 0;
            if (per <= 0) {
                const double left = (double)(total - done);
                want = std::min(cap, want * sched.growth);
                if (left < 2 * want) want = left > std::max(cap / sched.tail, sched.floor) ? left / 2 : left;      // the call is ending: halve
            }
        }
    }
    if (chunks.empty()) chunks.push_back(Chunk{0, 0, 0, 0, 0, 0, 0, 0});
    return PHMM_OK;
}

phmm_flat_batches sub_batches(const phmm_flat_batches *fb, const Chunk &c)
{
    phmm_flat_batches sub = *fb;
    sub.n_batches = c.b1 - c.b0; sub.n_reads = c.r1 - c.r0; sub.n_haps = c.h1 - c.h0;
    sub.batch_num_reads = fb->batch_num_reads + c.b0; sub.batch_num_haps = fb->batch_num_haps + c.b0;
    sub.read_len = fb->read_len + c.r0; sub.read_off = fb->read_off + c.r0;
    sub.hap_len = fb->hap_len + c.h0; sub.hap_off = fb->hap_off + c.h0;
    return sub;
}
}  // namespace

// ============================================================================================
// device-resident workloads
// ============================================================================================
static PlanOptions plan_options(const phmm_ctx *ctx)
{
    PlanOptions o;
    o.precision = ctx->precision;
    o.seg_cols = ctx->seg_cols;
    o.force_kh = ctx->force_kh;
    o.force_lanes = ctx->force_lanes;
    o.resident_warps = ctx->sm_count * 16;
    o.latency = ctx->plan_mode == 0 ? -1 : ctx->plan_mode - 1;
    o.hp = (ctx->fp32_fast && ctx->precision != 2 && !ctx->no_hp) ? 1 : 0;
    o.spec = (o.hp && ctx->precision == 0 && spec_enabled()) ? -1 : 0;
    return o;
}

extern "C" int phmm_workload_destroy(phmm_workload *w)
{
    if (!w) return PHMM_OK;
    for (int k = 0; k < 3; ++k) if (w->ev[k]) cudaEventDestroy(w->ev[k]);
    if (w->ctx && !w->arena) {
        cudaSetDevice(w->ctx->device);
        cudaFree(w->d_pool);          // one allocation holds every buffer of the workload
    }
    delete w;
    return PHMM_OK;
}

// Plans `fb`, places the workload's buffers (in `arena` when given, which must be idle; else in one
// cudaMalloc) and queues the uploads on `up`.  Does not wait for them: the caller orders later work
// behind an event on `up`, or synchronises.
// `prebuilt`: a plan of `fb` made ahead of time (the planner thread of phmm_compute_flat); it is moved from.
static int wl_create(phmm_ctx *ctx, const phmm_flat_batches *fb, Arena *arena, cudaStream_t up, phmm_workload **out,
                     Plan *prebuilt = nullptr, const PlanOptions *popt = nullptr)
{
    *out = nullptr;
    phmm_workload *w = new phmm_workload();
    w->ctx = ctx;
    w->precision = ctx->precision;
    w->arena = arena;
    std::string err;
    int rc = PHMM_OK;
    if (prebuilt) w->plan = std::move(*prebuilt);
    else rc = build_plan(*fb, popt ? *popt : plan_options(ctx), &w->plan, &err);
    if (rc != PHMM_OK) { ctx->error = err; delete w; return rc; }
    Plan &P = w->plan;
    w->n_pairs = P.n_pairs;
    w->n_cells = P.n_cells;
    if ((int)P.classes.size() + 2 * (MAX_KH_SCALAR + 1) > MAX_STEP_LAUNCHES) { ctx->error = "too many kernel classes"; delete w; return PHMM_ERR_ARG; }

    auto fail = [&](int code) { phmm_workload_destroy(w); return code; };
    if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->error = "cudaSetDevice failed"; return fail(PHMM_ERR_CUDA); }

    for (int k = 0; k < 3; ++k)
        if (cudaEventCreate(&w->ev[k]) != cudaSuccess) { ctx->error = "cudaEventCreate failed"; return fail(PHMM_ERR_CUDA); }

    // rescue / guard lists: one per scalar class, sized for every pair of that class
    int64_t tot = 0;
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) {
        if (tot + P.resc_cap[k] > INT32_MAX) { ctx->error = "workload too large for one device (rescue list)"; return fail(PHMM_ERR_ARG); }
        w->resc_off[k] = (int32_t)tot;
        w->resc_cap[k] = (int32_t)P.resc_cap[k];
        tot += P.resc_cap[k];
    }
    w->resc_total = tot;
    const bool need_resc = true;   // mode 1 needs the fast FP32 kernel's guard list, mode 2 the fast FP64 kernel's

    const int64_t pool_len = P.pool_len;
    w->pool_stride = (pool_len + 255) & ~(int64_t)255;
    int64_t ntasks = 0, nhtasks = 0;
    for (auto &cp : P.classes) {
        ClassDev cd;
        cd.kc = cp.kc;
        cd.task_off = cp.kc.hp ? nhtasks : ntasks;
        cd.ntasks = (int32_t)cp.ntasks();
        (cp.kc.hp ? nhtasks : ntasks) += (int64_t)cp.ntasks();
        if (cp.kc.hp) w->planned_hp = 1;
        w->classes.push_back(cd);
    }
    if (P.long_max_cols > 0 || P.resc_cap[0] > 0) {
        // resc_cap[0]: reads whose rescue / guard class is the multi-pass kernel; long reads may
        // also come back as single-haplotype rescue / guard tasks
        const int64_t cols = std::max<int64_t>(P.long_max_cols, P.max_haplen + SEP_COLS) + 64;
        w->scratch_stride = 6 * cols;
        w->scratch_blocks = ctx->sm_count * 8 / WARPS_PER_BLOCK;     // eight warps per SM
    }

    // FP64-first tasks and the band list of their pairs
    int64_t n_spec_tasks = 0, spec_total = 0;
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) {
        if (spec_total + P.spec_cap[k] > INT32_MAX) { ctx->error = "workload too large for one device (FP64-first list)"; return fail(PHMM_ERR_ARG); }
        w->spec_ntasks[k] = (int32_t)P.spec_tasks[k].size();
        w->spec_off[k] = (int32_t)spec_total;
        n_spec_tasks += w->spec_ntasks[k];
        spec_total += P.spec_cap[k];
    }
    w->spec_pairs = P.spec_pairs;
    w->spec_cells = P.spec_cells;

    // ---- one allocation, carved into 256-byte aligned pieces ------------------------------------
    const size_t np = (size_t)std::max<int64_t>(1, P.n_pairs);
    struct Piece { void **ptr; size_t bytes; const char *name; };
    const bool host_coded = !P.stream.empty();       // small call: the stream is uploaded, not built on the device
    const Piece pieces[] = {
        {(void **)&w->d_pool, (size_t)(5 * w->pool_stride), "read pools"},
        {(void **)&w->d_hapraw, (size_t)P.hap_pool_len, "haplotype bases"},
        {(void **)&w->d_hapsrc, P.hap_src.size() * sizeof(HapSrc), "haplotype table"},
        {(void **)(host_coded ? &w->d_stream : &w->d_unused), host_coded ? (size_t)P.stream_len : 0, "haplotype stream (uploaded)"},
        {(void **)&w->d_hap_len, P.hap_len.size() * sizeof(int32_t), "hap_len"},
        {(void **)&w->d_reads, P.reads.size() * sizeof(ReadRec), "reads"},
        {(void **)&w->d_tasks, (size_t)ntasks * sizeof(Task), "tasks"},
        {(void **)&w->d_htasks, (size_t)nhtasks * sizeof(HTask), "htasks"},
        {(void **)&w->d_spec_tasks, (size_t)n_spec_tasks * sizeof(Task), "FP64-first tasks"},
        {(void **)&w->d_ints, INTS_TOTAL * sizeof(int32_t), "ints"},
        {(void **)&w->d_out, np * sizeof(double), "out"},
        {(void **)(host_coded ? &w->d_unused : &w->d_stream), host_coded ? 0 : (size_t)P.stream_len, "haplotype stream"},
        {(void **)&w->d_raw64, np * sizeof(double), "raw64"},
        {(void **)&w->d_raw32, np * sizeof(float), "raw32"},
        {(void **)&w->d_rescued, np, "rescued"},
        {(void **)&w->d_resc, need_resc ? (size_t)tot * sizeof(Task) : 0, "rescue list"},
        {(void **)&w->d_guard, need_resc ? (size_t)tot * sizeof(Task) : 0, "guard list"},
        {(void **)&w->d_spec, (size_t)spec_total * sizeof(Task), "FP64-first band list"},
        {(void **)&w->d_scratch, (size_t)(w->scratch_stride * w->scratch_blocks * WARPS_PER_BLOCK) * sizeof(double), "scratch"},
    };
    const size_t red = ctx->canary ? 256 : 0;
    size_t total = 0;
    for (const Piece &p : pieces) total += ((std::max<size_t>(p.bytes, 1) + 255) & ~(size_t)255) + red;
    uint8_t *base = nullptr;
    if (arena) {
        if (arena->cap < total) {
            cudaFree(arena->base);
            arena->base = nullptr;
            arena->cap = 0;
            const size_t want = total + total / 4;
            cudaError_t e = cudaMalloc((void **)&arena->base, want);
            if (e != cudaSuccess) {
                ctx->error = std::string("cudaMalloc (arena): ") + cudaGetErrorString(e);
                return fail(e == cudaErrorMemoryAllocation ? PHMM_ERR_NOMEM : PHMM_ERR_CUDA);
            }
            arena->cap = want;
        }
        base = arena->base;
    } else {
        cudaError_t e = cudaMalloc((void **)&base, total);
        if (e != cudaSuccess) {
            ctx->error = std::string("cudaMalloc: ") + cudaGetErrorString(e);
            return fail(e == cudaErrorMemoryAllocation ? PHMM_ERR_NOMEM : PHMM_ERR_CUDA);
        }
    }
    {
        size_t off = 0;
        for (const Piece &p : pieces) {
            *p.ptr = base + off;
            const size_t span = ((std::max<size_t>(p.bytes, 1) + 255) & ~(size_t)255) + red;
            if (red) w->redzones.push_back({p.name, base + off + p.bytes, span - p.bytes});
            off += span;
        }
    }

#define WL_TRY(expr)                                                                          \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess) { ctx->error = std::string(#expr) + ": " + cudaGetErrorString(e__); return fail(PHMM_ERR_CUDA); } \
    } while (0)
    const uint8_t *pools[5] = {fb->rs, fb->q, fb->i, fb->d, fb->c};
    int64_t h2d = 0;
    int32_t ints[INTS_TOTAL];
    memset(ints, 0, sizeof ints);
    for (size_t c = 0; c < P.classes.size(); ++c) ints[INTS_NTASKS + c] = w->classes[c].ntasks;
    // everything that is uploaded sits in front of d_out: small chunks go up in one copy from pinned staging
    const size_t upload_extent = (size_t)((uint8_t *)w->d_out - base);
    uint8_t *stage = nullptr;
    if (arena && upload_extent <= Arena::STAGE_MAX) {
        // the first small chunk of a context gives every slot its staging buffer: a later call whose small first
        // chunk lands in a slot that never staged would pay the page-locked allocation (1-3 ms) inside its pipeline
        if (arena->stage_cap == 0)
            for (Arena &a : ctx->arena)
                if (a.stage_cap == 0 && cudaMallocHost((void **)&a.stage, Arena::STAGE_MAX) == cudaSuccess) a.stage_cap = Arena::STAGE_MAX;
                else if (a.stage_cap == 0) { a.stage = nullptr; cudaGetLastError(); }
        if (arena->stage_cap < upload_extent) {
            if (arena->stage) cudaFreeHost(arena->stage);
            arena->stage = nullptr;
            const size_t want = std::min<size_t>(Arena::STAGE_MAX, std::max<size_t>(4u << 20, upload_extent + upload_extent / 4));
            if (cudaMallocHost((void **)&arena->stage, want) == cudaSuccess) arena->stage_cap = want;
            else { arena->stage = nullptr; arena->stage_cap = 0; cudaGetLastError(); }
        }
        if (arena->stage_cap >= upload_extent) stage = arena->stage;
    }
    // dst: device address; staged: the same offset in the staging buffer (the slot is idle: its previous chunk
    // has been retired, so the last copy out of the buffer is complete)
    auto put = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
        if (bytes == 0) return cudaSuccess;
        if (stage) { memcpy(stage + ((uint8_t *)dst - base), src, bytes); return cudaSuccess; }
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, up);
    };
    for (int k = 0; k < 5; ++k)
        for (const Plan::Segment &sg : P.segments) {
            WL_TRY(put(w->d_pool + k * w->pool_stride + sg.dst, pools[k] + sg.src, (size_t)sg.len));
            h2d += sg.len;
        }
    for (const Plan::Segment &sg : P.hap_segments) {
        WL_TRY(put(w->d_hapraw + sg.dst, fb->hap + sg.src, (size_t)sg.len));
        h2d += sg.len;
    }
    WL_TRY(put(w->d_hapsrc, P.hap_src.data(), P.hap_src.size() * sizeof(HapSrc)));
    if (host_coded) { WL_TRY(put(w->d_stream, P.stream.data(), P.stream.size())); h2d += (int64_t)P.stream.size(); }
    WL_TRY(put(w->d_hap_len, P.hap_len.data(), P.hap_len.size() * sizeof(int32_t)));
    WL_TRY(put(w->d_reads, P.reads.data(), P.reads.size() * sizeof(ReadRec)));
    h2d += (int64_t)P.hap_src.size() * sizeof(HapSrc) + (int64_t)P.hap_len.size() * 4 + (int64_t)P.reads.size() * sizeof(ReadRec);
    for (size_t c = 0; c < P.classes.size(); ++c) {
        WL_TRY(put(w->d_tasks + w->classes[c].task_off, P.classes[c].tasks.data(), P.classes[c].tasks.size() * sizeof(Task)));
        WL_TRY(put(w->d_htasks + w->classes[c].task_off, P.classes[c].htasks.data(), P.classes[c].htasks.size() * sizeof(HTask)));
        h2d += (int64_t)P.classes[c].tasks.size() * sizeof(Task) + (int64_t)P.classes[c].htasks.size() * sizeof(HTask);
    }
    {
        int64_t at = 0;
        for (int k = 0; k <= MAX_KH_SCALAR; ++k) {
            WL_TRY(put(w->d_spec_tasks + at, P.spec_tasks[k].data(), P.spec_tasks[k].size() * sizeof(Task)));
            at += w->spec_ntasks[k];
            h2d += (int64_t)P.spec_tasks[k].size() * sizeof(Task);
        }
    }
    // (copies from pageable host memory return once the source has been staged, so the plan's
    // vectors and `ints` may be released right after)
    WL_TRY(put(w->d_ints, ints, sizeof ints));
    if (stage) WL_TRY(cudaMemcpyAsync(base, stage, upload_extent, cudaMemcpyHostToDevice, up));
    {
        // the FP64-first tasks are the heads of the rescue lists; the kernels append behind them
        int64_t at = 0;
        for (int k = 0; k <= MAX_KH_SCALAR; ++k) {
            if (w->spec_ntasks[k])
                WL_TRY(cudaMemcpyAsync(w->d_resc + w->resc_off[k], w->d_spec_tasks + at, (size_t)w->spec_ntasks[k] * sizeof(Task),
                                       cudaMemcpyDeviceToDevice, up));
            at += w->spec_ntasks[k];
        }
    }
    // the haplotype stream: zero (separator columns, alignment gaps, slack), then the code bytes from the raw bases
    if (!host_coded) WL_TRY(cudaMemsetAsync(w->d_stream, 0, (size_t)P.stream_len, up));
    if (!host_coded && !P.hap_src.empty()) {
        const int nh = (int)P.hap_src.size();
        phmm_code_stream_kernel<<<(nh + 7) / 8, 256, 0, up>>>(w->d_hapraw, w->d_hapsrc, nh, w->d_stream);
        WL_TRY(cudaGetLastError());
    }
    WL_TRY(cudaMemsetAsync(w->d_raw64, 0, np * sizeof(double), up));
    WL_TRY(cudaMemsetAsync(w->d_rescued, 0, np, up));
    // after the uploads (the staged copy covers the upload pieces' zones), before any kernel of this workload
    for (const phmm_workload::RedZone &z : w->redzones) WL_TRY(cudaMemsetAsync(z.begin, 0xA5, z.len, up));
    // PHMM_CANARY=2 (tests): damage one zone on purpose so that the check itself is exercised
    if (ctx->canary == 2 && !w->redzones.empty()) WL_TRY(cudaMemsetAsync(w->redzones.back().begin + 7, 0, 1, up));
#undef WL_TRY
    w->h2d_bytes = h2d;
    // host copies of the bulky plan parts are no longer needed
    std::vector<HapSrc>().swap(P.hap_src);
    std::vector<uint8_t>().swap(P.stream);
    for (auto &cp : P.classes) { std::vector<Task>().swap(cp.tasks); std::vector<HTask>().swap(cp.htasks); }
    for (auto &v : P.spec_tasks) std::vector<Task>().swap(v);
    *out = w;
    return PHMM_OK;
}

extern "C" int phmm_workload_create(phmm_ctx *ctx, const phmm_flat_batches *fb, phmm_workload **out)
{
    if (!ctx || !fb || !out) return PHMM_ERR_ARG;
    int rc = wl_create(ctx, fb, nullptr, ctx->ss[0].main, out);
    if (rc != PHMM_OK) return rc;
    if (cudaStreamSynchronize(ctx->ss[0].main) != cudaSuccess) {
        ctx->error = "upload failed";
        phmm_workload_destroy(*out);
        *out = nullptr;
        return PHMM_ERR_CUDA;
    }
    return PHMM_OK;
}

// Queues every kernel of the workload on the stream set `ss` (after `wait_for`, when given) and
// records the timing events.  Does not wait.
// fallback64 == nullptr: the whole step.  Otherwise only the rare last phase: counts per class of pairs whose
// fast FP64 sum was non-finite (read back by wl_finish), re-run by the bit-exact FP64 kernels.
// specfix: counts per class of FP64-first pairs inside the decision band (KernelParams::spec_list), decided by the
// bit-exact FP32 kernel in mode 3.
static int wl_launch(phmm_ctx *ctx, phmm_workload *w, StreamSet &ss, cudaEvent_t wait_for, const int32_t *fallback64 = nullptr,
                     const int32_t *specfix = nullptr)
{
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ss.main;
    if (wait_for) CUDA_TRY(ctx, cudaStreamWaitEvent(st, wait_for, 0));
    const HostTables &t = host_tables();
    const Plan &P = w->plan;
    const bool fast32 = ctx->fp32_fast != 0, fast64 = ctx->fp64_fast != 0;
    const int mode = w->precision;
    int launches = 0;
    if (w->planned_hp && !fast32) {
        ctx->error = "workload was planned for the fast FP32 kernels; PHMM_OPT_FP32_KERNEL changed since";
        return PHMM_ERR_ARG;
    }

    // queue heads and list counters back to zero
    if (!fallback64 && !specfix) {
        CUDA_TRY(ctx, cudaMemsetAsync(w->d_ints + INTS_COUNTER, 0, COUNTER_SLOTS * sizeof(int32_t), st));
        CUDA_TRY(ctx, cudaMemsetAsync(w->d_ints + INTS_RESC, 0, INTS_LISTS * sizeof(int32_t), st));
        // the rescue lists start with the FP64-first tasks the planner put there
        if (w->spec_pairs > 0)
            CUDA_TRY(ctx, cudaMemcpyAsync(w->d_ints + INTS_RESC, w->spec_ntasks, sizeof w->spec_ntasks, cudaMemcpyHostToDevice, st));
    }

    KernelParams<float> pf;
    memset(&pf, 0, sizeof pf);
    pf.rs = w->d_pool; pf.q = w->d_pool + w->pool_stride; pf.i = w->d_pool + 2 * w->pool_stride;
    pf.d = w->d_pool + 3 * w->pool_stride; pf.c = w->d_pool + 4 * w->pool_stride;
    pf.reads = w->d_reads; pf.stream = w->d_stream; pf.hap_len = w->d_hap_len;
    pf.ph2pr = ctx->d_ph2pr_f; pf.mm = ctx->d_mm_f;
    pf.init_const = t.init_f; pf.log10_init = t.log10_init_f;
    pf.out = w->d_out; pf.raw32 = w->d_raw32; pf.raw64 = w->d_raw64; pf.rescued = w->d_rescued;
    pf.resc = w->d_resc; pf.resc_count = w->d_ints + INTS_RESC; pf.resc_pairs = w->d_ints + INTS_RESC_PAIRS;
    pf.guard = w->d_guard; pf.guard_count = w->d_ints + INTS_GUARD;
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) pf.resc_off[k] = w->resc_off[k];
    pf.mode = mode;
    pf.resc_force = P.resc_force;
    KernelParams<double> pd;
    memset(&pd, 0, sizeof pd);
    pd.rs = pf.rs; pd.q = pf.q; pd.i = pf.i; pd.d = pf.d; pd.c = pf.c;
    pd.reads = pf.reads; pd.stream = pf.stream; pd.hap_len = pf.hap_len;
    pd.ph2pr = ctx->d_ph2pr_d; pd.mm = ctx->d_mm_d;
    pd.init_const = t.init_d; pd.log10_init = t.log10_init_d;
    pd.out = w->d_out; pd.raw32 = w->d_raw32; pd.raw64 = w->d_raw64; pd.rescued = w->d_rescued;
    pd.mode = mode;
    pd.resc_force = P.resc_force;
    pd.spec_list = w->d_spec; pd.spec_count = w->d_ints + INTS_SPEC;
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) pd.spec_off[k] = w->spec_off[k];
    pd.log10_init32 = t.log10_init_f;
    pd.spec_band = ctx->spec_band;
    // fast FP64 kernels hand non-finite sums to the exact ones through the (by then consumed) guard list storage
    pd.guard = fast64 ? w->d_guard : nullptr; pd.guard_count = w->d_ints + INTS_GUARD64;
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) pd.resc_off[k] = w->resc_off[k];
    pf.scratch = reinterpret_cast<float *>(w->d_scratch); pf.scratch_stride = 2 * w->scratch_stride;   // in floats
    pd.scratch = w->d_scratch; pd.scratch_stride = w->scratch_stride;

    // dynamic shared memory (emission table + stream ring + run table per warp) and a grid that
    // fills the device once: the kernels are persistent, warps pull tasks from the queue
    // layout: 0 = phmm_forward_kernel, 1 = haplotype-pair FP32
    auto prepare = [&](const void *kernel, int kh, int64_t ntasks, bool longr, int *grid, int *smem, int layout = 0) -> bool {
        *smem = WARPS_PER_BLOCK * (layout == 1 ? smem_per_warp_hp(kh) : smem_per_warp(kh));     // a function of the kernel
        int nb = 0;
        auto it = ctx->blocks_per_sm.find(kernel);
        if (it != ctx->blocks_per_sm.end()) {
            nb = it->second;
        } else {
            if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, *smem) != cudaSuccess) return false;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, WARPS_PER_BLOCK * 32, *smem) != cudaSuccess || nb < 1) nb = 1;
            ctx->blocks_per_sm[kernel] = nb;
        }
        const int64_t full = (int64_t)nb * ctx->sm_count;
        const int64_t want = (ntasks + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
        *grid = (int)std::max<int64_t>(1, std::min(full, want));
        if (longr) *grid = std::min(*grid, w->scratch_blocks);
        return true;
    };
    int counter_slot = 0, grid = 0, smem = 0;
    // fork: the side streams wait for everything queued on the main stream so far;
    // join: the main stream waits for the side streams that were used
    // Large workloads: four class kernels side by side (the tail of one is filled by the blocks of the next;
    // more of them at once only split the SMs further).  Small calls: every class kernel has a handful of
    // tasks and lasts as long as its longest one, so all of them run at once.
    const int n_side = P.latency ? (int)StreamSet::N_SIDE : 4;
    int n_forked = 0;
    auto fork = [&]() -> cudaError_t {
        n_forked = 0;
        return cudaEventRecord(ss.ev_fork, st);
    };
    auto side_stream = [&](cudaStream_t *out) -> cudaError_t {
        const int k = n_forked % n_side;
        if (n_forked < n_side) {
            cudaError_t e = cudaStreamWaitEvent(ss.side[k], ss.ev_fork, 0);
            if (e != cudaSuccess) return e;
        }
        ++n_forked;
        *out = ss.side[k];
        return cudaSuccess;
    };
    auto join = [&]() -> cudaError_t {
        for (int k = 0; k < std::min(n_forked, n_side); ++k) {
            cudaError_t e = cudaEventRecord(ss.ev_join[k], ss.side[k]);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(st, ss.ev_join[k], 0);
            if (e != cudaSuccess) return e;
        }
        n_forked = 0;
        return cudaSuccess;
    };
    cudaStream_t ks = st;

    if (!fallback64 && !specfix) {
    CUDA_TRY(ctx, cudaEventRecord(w->ev[0], st));
    // ---- main pass over the planned tasks ---------------------------------------------------
    CUDA_TRY(ctx, fork());
    for (size_t c = 0; c < w->classes.size(); ++c) {
        const ClassDev &cd = w->classes[c];
        if (cd.ntasks == 0) continue;
        if (counter_slot >= MAX_STEP_LAUNCHES) { ctx->error = "too many launches"; return PHMM_ERR_ARG; }
        CUDA_TRY(ctx, side_stream(&ks));
        if (mode == 2) {
            KernelF64 k = cd.kc.longr ? lookup_f64_long(fast64) : lookup_f64(cd.kc.kh, fast64);
            if (!k || cd.kc.vl != 1 || cd.kc.lanes != 32) { ctx->error = "no FP64 kernel for class"; return PHMM_ERR_ARG; }
            pd.tasks = w->d_tasks + cd.task_off;
            pd.ntasks_ptr = w->d_ints + INTS_NTASKS + c;
            pd.counter = w->d_ints + INTS_COUNTER + counter_slot++;
            if (!prepare((const void *)k, cd.kc.kh, cd.ntasks, cd.kc.longr, &grid, &smem)) { ctx->error = "cudaFuncSetAttribute failed"; return PHMM_ERR_CUDA; }
            k<<<grid, WARPS_PER_BLOCK * 32, smem, ks>>>(pd);
        } else {
            KernelF32 k = lookup_f32(cd.kc, fast32);
            if (!k) { ctx->error = "no FP32 kernel for class"; return PHMM_ERR_ARG; }
            pf.tasks = cd.kc.hp ? nullptr : w->d_tasks + cd.task_off;
            pf.htasks = cd.kc.hp ? w->d_htasks + cd.task_off : nullptr;
            pf.ntasks_ptr = w->d_ints + INTS_NTASKS + c;
            pf.counter = w->d_ints + INTS_COUNTER + counter_slot++;
            if (!prepare((const void *)k, cd.kc.kh, cd.ntasks, cd.kc.longr, &grid, &smem, cd.kc.hp ? 1 : 0)) { ctx->error = "cudaFuncSetAttribute failed"; return PHMM_ERR_CUDA; }
            k<<<grid, WARPS_PER_BLOCK * 32, smem, ks>>>(pf);
        }
        CUDA_TRY(ctx, cudaGetLastError());
        ++launches;
    }
    CUDA_TRY(ctx, join());
    // ---- fast kernel only: near-threshold and non-finite sums re-decided by the bit-exact kernel ----
    if (mode != 2 && fast32) {
        CUDA_TRY(ctx, fork());
        for (int kh = 0; kh <= MAX_KH_SCALAR; ++kh) {
            if (w->resc_cap[kh] == 0) continue;
            if (counter_slot >= MAX_STEP_LAUNCHES) { ctx->error = "too many launches"; return PHMM_ERR_ARG; }
            CUDA_TRY(ctx, side_stream(&ks));
            KernelF32 k = kh ? lookup_f32_l32_f0(kh) : lookup_f32_long(false);
            pf.tasks = w->d_guard + w->resc_off[kh];
            pf.ntasks_ptr = w->d_ints + INTS_GUARD + kh;
            pf.counter = w->d_ints + INTS_COUNTER + counter_slot++;
            // the exact kernel never takes the guard branch (FAST = false)
            if (!prepare((const void *)k, kh ? kh : LONG_KH, std::min<int64_t>(w->resc_cap[kh], 4096), kh == 0, &grid, &smem)) { ctx->error = "cudaFuncSetAttribute failed"; return PHMM_ERR_CUDA; }
            k<<<grid, WARPS_PER_BLOCK * 32, smem, ks>>>(pf);
            CUDA_TRY(ctx, cudaGetLastError());
            ++launches;
        }
        CUDA_TRY(ctx, join());
    }
    CUDA_TRY(ctx, cudaEventRecord(w->ev[1], st));
    // ---- FP64 rescue (IntelPairHmmCSource.cpp:73-75) --------------------------------------------
    if (mode == 0) {
        CUDA_TRY(ctx, fork());
        for (int kh = 0; kh <= MAX_KH_SCALAR; ++kh) {
            if (w->resc_cap[kh] == 0) continue;
            if (counter_slot >= MAX_STEP_LAUNCHES) { ctx->error = "too many launches"; return PHMM_ERR_ARG; }
            CUDA_TRY(ctx, side_stream(&ks));
            KernelF64 k = kh ? lookup_f64(kh, fast64) : lookup_f64_long(fast64);
            pd.tasks = w->d_resc + w->resc_off[kh];
            pd.ntasks_ptr = w->d_ints + INTS_RESC + kh;
            pd.counter = w->d_ints + INTS_COUNTER + counter_slot++;
            if (!prepare((const void *)k, kh ? kh : LONG_KH, w->resc_cap[kh], kh == 0, &grid, &smem)) { ctx->error = "cudaFuncSetAttribute failed"; return PHMM_ERR_CUDA; }
            k<<<grid, WARPS_PER_BLOCK * 32, smem, ks>>>(pd);
            CUDA_TRY(ctx, cudaGetLastError());
            ++launches;
        }
        CUDA_TRY(ctx, join());
    }
    CUDA_TRY(ctx, cudaEventRecord(w->ev[2], st));
    w->launches = launches;
    return PHMM_OK;
    }
    // ---- fast FP64 kernels only: non-finite sums re-run by the bit-exact FP64 kernel --------------
    // (launched by wl_finish when the counters say so: the common case pays nothing)
    if (specfix) {
        pf.mode = 3;
        CUDA_TRY(ctx, fork());
        for (int kh = 1; kh <= MAX_KH_SCALAR; ++kh) {
            if (specfix[kh] == 0) continue;
            CUDA_TRY(ctx, side_stream(&ks));
            KernelF32 k = lookup_f32_l32_f0(kh);
            pf.tasks = w->d_spec + w->spec_off[kh];
            pf.ntasks_ptr = w->d_ints + INTS_SPEC + kh;
            pf.counter = w->d_ints + INTS_COUNTER + SPEC_FIX_SLOT + kh;
            if (!prepare((const void *)k, kh, std::min<int64_t>(specfix[kh], 4096), false, &grid, &smem)) { ctx->error = "cudaFuncSetAttribute failed"; return PHMM_ERR_CUDA; }
            k<<<grid, WARPS_PER_BLOCK * 32, smem, ks>>>(pf);
            CUDA_TRY(ctx, cudaGetLastError());
            ++launches;
        }
        CUDA_TRY(ctx, join());
        w->launches += launches;
        return PHMM_OK;
    }
    if (mode != 1 && fast64) {
        pd.guard = nullptr;
        counter_slot = MAX_STEP_LAUNCHES;  // queue heads reserved for this re-run: zeroed at the start of the step, never handed out by it
        CUDA_TRY(ctx, fork());
        for (int kh = 0; kh <= MAX_KH_SCALAR; ++kh) {
            if (fallback64[kh] == 0) continue;
            if (counter_slot >= SPEC_FIX_SLOT) { ctx->error = "too many launches"; return PHMM_ERR_ARG; }
            CUDA_TRY(ctx, side_stream(&ks));
            KernelF64 k = kh ? lookup_f64(kh, false) : lookup_f64_long(false);
            pd.tasks = w->d_guard + w->resc_off[kh];
            pd.ntasks_ptr = w->d_ints + INTS_GUARD64 + kh;
            pd.counter = w->d_ints + INTS_COUNTER + counter_slot++;
            if (!prepare((const void *)k, kh ? kh : LONG_KH, std::min<int64_t>(w->resc_cap[kh], 256), kh == 0, &grid, &smem)) { ctx->error = "cudaFuncSetAttribute failed"; return PHMM_ERR_CUDA; }
            k<<<grid, WARPS_PER_BLOCK * 32, smem, ks>>>(pd);
            CUDA_TRY(ctx, cudaGetLastError());
            ++launches;
        }
        CUDA_TRY(ctx, join());
    }
    w->launches += launches;
    return PHMM_OK;
}

// After the stream set has drained: list counters and event times of the last launch.
static int wl_finish(phmm_ctx *ctx, phmm_workload *w, StreamSet &ss, phmm_stats *stats)
{
    int32_t counts[INTS_LISTS];
    CUDA_TRY(ctx, cudaMemcpyAsync(counts, w->d_ints + INTS_RESC, sizeof counts, cudaMemcpyDeviceToHost, ss.main));
    CUDA_TRY(ctx, cudaStreamSynchronize(ss.main));
    int64_t nguard = 0, n64 = 0, nband = 0;
    const int32_t *spec_counts = counts + (INTS_SPEC - INTS_RESC);
    int64_t nres = (int64_t)counts[INTS_RESC_PAIRS - INTS_RESC] + spec_counts[MAX_KH_SCALAR + 1];
    for (int k = 0; k <= MAX_KH_SCALAR; ++k) { nguard += counts[16 + k]; n64 += counts[32 + k]; nband += spec_counts[k]; }
    w->last_spec_direct = spec_counts[MAX_KH_SCALAR + 2];
    if (n64 > 0) {
        // fast FP64 kernels met non-finite sums (exotic gap-continuation penalties): bit-exact re-run
        int rc = wl_launch(ctx, w, ss, nullptr, counts + 32);
        if (rc != PHMM_OK) return rc;
        CUDA_TRY(ctx, cudaStreamSynchronize(ss.main));
    }
    if (nband > 0) {
        // FP64-first pairs whose rescaled sum fell inside the band around MIN_ACCEPTED: the bit-exact FP32 kernel decides
        // (after the re-run above: a non-finite FP64-first sum is in both lists, and the FP32 verdict has the last word)
        int rc = wl_launch(ctx, w, ss, nullptr, nullptr, spec_counts);
        if (rc != PHMM_OK) return rc;
        int32_t decided = 0;
        CUDA_TRY(ctx, cudaMemcpyAsync(&decided, w->d_ints + INTS_RESC_PAIRS, sizeof decided, cudaMemcpyDeviceToHost, ss.main));
        CUDA_TRY(ctx, cudaStreamSynchronize(ss.main));
        nres += decided - counts[INTS_RESC_PAIRS - INTS_RESC];
    }
    w->last_rescued = nres;
    w->last_guard = nguard;
    if (!w->redzones.empty()) {
        // every kernel of the run has finished (ss.main joined the side streams and is synchronised)
        std::vector<uint8_t> host;
        for (const phmm_workload::RedZone &z : w->redzones) {
            host.resize(z.len);
            CUDA_TRY(ctx, cudaMemcpy(host.data(), z.begin, z.len, cudaMemcpyDeviceToHost));
            for (size_t k = 0; k < z.len; ++k)
                if (host[k] != 0xA5) {
                    ctx->error = std::string("PHMM_CANARY: out-of-bounds write ") + std::to_string(k) +
                                 " bytes behind the end of device buffer '" + z.name + "'";
                    return PHMM_ERR_CUDA;
                }
        }
    }
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->n_pairs = w->n_pairs;
        stats->n_cells = w->n_cells;
        stats->n_rescued = nres;
        stats->n_guard = nguard;
        stats->n_spec_pairs = w->spec_pairs;
        stats->n_spec_cells = w->spec_cells;
        stats->n_spec_direct = w->last_spec_direct;
        stats->n_launches = w->launches;
        stats->n_classes = (int32_t)w->classes.size();
        for (auto &cd : w->classes) stats->n_tasks += cd.ntasks;
        stats->h2d_bytes = w->h2d_bytes;
        cudaEventElapsedTime(&stats->ms_fp32, w->ev[0], w->ev[1]);
        cudaEventElapsedTime(&stats->ms_fp64, w->ev[1], w->ev[2]);
        cudaEventElapsedTime(&stats->ms_total, w->ev[0], w->ev[2]);
    }
    return PHMM_OK;
}

extern "C" int phmm_workload_run(phmm_ctx *ctx, phmm_workload *w, phmm_stats *stats)
{
    if (!ctx || !w || w->ctx != ctx) return PHMM_ERR_ARG;
    int rc = wl_launch(ctx, w, ctx->ss[0], nullptr);
    if (rc != PHMM_OK) return rc;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->ss[0].main));
    return wl_finish(ctx, w, ctx->ss[0], stats);
}

extern "C" int phmm_workload_results(phmm_ctx *ctx, phmm_workload *w, double *out, float *raw32, double *raw64,
                                     uint8_t *rescued)
{
    if (!ctx || !w || w->ctx != ctx) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)w->n_pairs;
    if (n == 0) return PHMM_OK;
    cudaStream_t st = ctx->ss[0].main;
    if (out) CUDA_TRY(ctx, cudaMemcpyAsync(out, w->d_out, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (raw32) CUDA_TRY(ctx, cudaMemcpyAsync(raw32, w->d_raw32, n * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (raw64) CUDA_TRY(ctx, cudaMemcpyAsync(raw64, w->d_raw64, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (rescued) CUDA_TRY(ctx, cudaMemcpyAsync(rescued, w->d_rescued, n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return PHMM_OK;
}

// ============================================================================================
// host-buffer entry points
// ============================================================================================
// The batches are cut into chunks of about `chunk_cells` cells.  Chunk c is planned and uploaded
// by the host while the kernels of chunk c - 1 run (its buffers live in the other arena, its
// kernels queue on the other stream set, so they also fill the tail of chunk c - 1), and the
// results of chunk c - 1 are copied back while chunk c runs.  Exposed to the caller are only the
// plan + upload of the first chunk and the copy-back of the last.
extern "C" int phmm_compute_flat(phmm_ctx *ctx, const phmm_flat_batches *fb, double *out, phmm_stats *stats)
{
    if (!ctx || !fb) return PHMM_ERR_ARG;
    const double t_begin = wall_ms();
    static const bool trace = getenv("PHMM_TRACE") != nullptr;
    std::vector<Chunk> chunks;
    {
        int rc0 = split_batches(ctx, fb, ctx->chunk_cells, 4000000000ll, 40000000000ll, &chunks);
        if (rc0 != PHMM_OK) return rc0;
    }
    if (chunks.back().p1 > 0 && !out) { ctx->error = "NULL output"; return PHMM_ERR_ARG; }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));

    phmm_workload *wl[phmm_ctx::N_SLOTS] = {};
    phmm_stats agg;
    memset(&agg, 0, sizeof agg);
    int rc = PHMM_OK;
    double t_first_launch = 0, ms_d2h = 0;
    // copy chunk j's results back (waits for its kernels), collect its statistics, release it
    auto retire = [&](size_t j) -> int {
        const int slot = (int)(j % phmm_ctx::N_SLOTS);
        phmm_workload *w = wl[slot];
        StreamSet &ss = ctx->ss[slot];
        const double t0 = wall_ms();
        // counters first: wl_finish re-runs the (rare) pairs the fast FP64 kernels could not finish
        phmm_stats s;
        int r2 = wl_finish(ctx, w, ss, &s);
        if (r2 != PHMM_OK) return r2;
        if (w->n_pairs > 0) {
            CUDA_TRY(ctx, cudaMemcpyAsync(out + chunks[j].p0, w->d_out, (size_t)w->n_pairs * sizeof(double),
                                          cudaMemcpyDeviceToHost, ss.main));
            CUDA_TRY(ctx, cudaStreamSynchronize(ss.main));
        }
        ms_d2h += wall_ms() - t0;
        if (trace)
            fprintf(stderr, "phmm trace:   chunk %zu done: fp32 %.2f ms fp64 %.2f ms total %.2f ms, %d launches, %lld tasks, t=%.2f\n", j,
                    s.ms_fp32, s.ms_fp64, s.ms_total, s.n_launches, (long long)s.n_tasks, wall_ms() - t_begin);
        agg.n_pairs += s.n_pairs; agg.n_cells += s.n_cells; agg.n_rescued += s.n_rescued;
        agg.n_guard += s.n_guard; agg.n_tasks += s.n_tasks; agg.h2d_bytes += s.h2d_bytes;
        agg.n_spec_pairs += s.n_spec_pairs; agg.n_spec_cells += s.n_spec_cells; agg.n_spec_direct += s.n_spec_direct;
        agg.n_launches += s.n_launches; agg.n_classes = std::max(agg.n_classes, s.n_classes);
        agg.ms_fp32 += s.ms_fp32; agg.ms_fp64 += s.ms_fp64; agg.ms_total += s.ms_total;
        agg.d2h_bytes += w->n_pairs * (int64_t)sizeof(double);
        phmm_workload_destroy(w);
        wl[slot] = nullptr;
        return PHMM_OK;
    };
    // Planner threads run ahead of the loop below: with many small batches (active regions, C1) planning a chunk
    // takes as long as its kernels, the chunk schedule grows fourfold per chunk while the device only just keeps up
    // with one planner, and the loop's own time is better spent on uploads and launches.  Chunks are planned
    // concurrently, at most six ahead of the one being launched (a plan holds the chunk's task lists).
    struct PlanJob { Plan plan; std::string err; int rc = PHMM_OK; bool done = false; };
    std::vector<PlanJob> jobs(chunks.size());
    std::mutex pmu;
    std::condition_variable pcv;
    bool pstop = false;
    size_t next_job = 0, consumed = 0;
    PlanOptions popt = plan_options(ctx);
    popt.latency = plan_latency(*fb, popt);        // one decision for the whole call, not one per chunk
    popt.one_class = plan_one_class(*fb, popt);
    popt.spec = plan_spec(*fb, popt);
    static const int n_planners_env = getenv("PHMM_PLANNERS") ? atoi(getenv("PHMM_PLANNERS")) : 0;
    const int n_planners = chunks.size() > 1
        ? std::max(1, std::min<int>({n_planners_env > 0 ? n_planners_env : 3, (int)chunks.size(),
                                     (int)std::max(1u, std::thread::hardware_concurrency() / 2)}))
        : 0;
    std::vector<std::thread> planners;
    for (int k = 0; k < n_planners; ++k)
        planners.emplace_back([&]() {
            for (;;) {
                size_t j;
                {
                    std::unique_lock<std::mutex> lk(pmu);
                    pcv.wait(lk, [&]() { return pstop || next_job >= chunks.size() || next_job < consumed + 6; });
                    if (pstop || next_job >= chunks.size()) return;
                    j = next_job++;
                }
                const phmm_flat_batches sub = sub_batches(fb, chunks[j]);
                const int prc = build_plan(sub, popt, &jobs[j].plan, &jobs[j].err);
                { std::lock_guard<std::mutex> lk(pmu); jobs[j].rc = prc; jobs[j].done = true; if (prc != PHMM_OK) pstop = true; }
                pcv.notify_all();
                if (prc != PHMM_OK) return;
            }
        });
    struct PlannerGuard {
        std::vector<std::thread> &t; std::mutex &m; std::condition_variable &cv; bool &stop;
        ~PlannerGuard()
        {
            { std::lock_guard<std::mutex> lk(m); stop = true; }
            cv.notify_all();
            for (std::thread &th : t) if (th.joinable()) th.join();
        }
    } planner_guard{planners, pmu, pcv, pstop};
    for (size_t j = 0; j < chunks.size() && rc == PHMM_OK; ++j) {
        const int slot = (int)(j % phmm_ctx::N_SLOTS);
        // the slot's arena must be idle: wait for the chunk that used it and fetch its results
        if (wl[slot]) rc = retire(j - phmm_ctx::N_SLOTS);
        if (rc != PHMM_OK) break;
        Plan *prebuilt = nullptr;
        if (n_planners > 0) {
            std::unique_lock<std::mutex> lk(pmu);
            pcv.wait(lk, [&]() { return jobs[j].done; });
            if (jobs[j].rc != PHMM_OK) { ctx->error = jobs[j].err; rc = jobs[j].rc; break; }
            consumed = j + 1;
            pcv.notify_all();
            prebuilt = &jobs[j].plan;
        }
        const Chunk &c = chunks[j];
        phmm_flat_batches sub = sub_batches(fb, c);
        StreamSet &ss = ctx->ss[slot];
        const double tc0 = wall_ms();
        rc = wl_create(ctx, &sub, &ctx->arena[slot], ctx->copy_in, &wl[slot], prebuilt, &popt);
        if (rc != PHMM_OK) break;
        const double tc1 = wall_ms();
        if (cudaEventRecord(ss.ev_uploaded, ctx->copy_in) != cudaSuccess) { ctx->error = "cudaEventRecord failed"; rc = PHMM_ERR_CUDA; break; }
        rc = wl_launch(ctx, wl[slot], ss, ss.ev_uploaded);
        const double tc2 = wall_ms();
        if (j == 0) t_first_launch = tc2;
        if (trace)
            fprintf(stderr, "phmm trace: chunk %zu/%zu batches %lld cells %.3e  plan+upload %.2f ms  launch %.2f ms  (waited %.2f ms)  t=%.2f\n",
                    j, chunks.size(), (long long)(c.b1 - c.b0), (double)wl[slot]->n_cells, tc1 - tc0, tc2 - tc1, wall_ms() - tc2,
                    wall_ms() - t_begin);
    }
    for (size_t j = chunks.size() > (size_t)phmm_ctx::N_SLOTS ? chunks.size() - phmm_ctx::N_SLOTS : 0;
         j < chunks.size() && rc == PHMM_OK; ++j)
        if (wl[j % phmm_ctx::N_SLOTS]) rc = retire(j);
    if (rc != PHMM_OK) {
        cudaDeviceSynchronize();
        for (int k = 0; k < phmm_ctx::N_SLOTS; ++k) { phmm_workload_destroy(wl[k]); wl[k] = nullptr; }
        return rc;
    }
    if (stats) {
        *stats = agg;
        stats->ms_h2d = (float)(t_first_launch - t_begin);   // exposed: plan + upload of the first chunk
        stats->ms_d2h = (float)ms_d2h;                       // host time in result copies (mostly overlapped)
    }
    return PHMM_OK;
}

extern "C" int phmm_host_register(phmm_ctx *ctx, void *ptr, size_t bytes)
{
    if (!ctx || !ptr || bytes == 0) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return PHMM_OK;
}

extern "C" int phmm_host_unregister(phmm_ctx *ctx, void *ptr)
{
    if (!ctx || !ptr) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaHostUnregister(ptr));
    return PHMM_OK;
}

namespace {
// Growable flat batches built from pointer-based inputs, de-duplicating strings by address.
struct FlatBuilder {
    ParsedBatches b;
    struct ReadKey {
        const void *p[5];
        int len;
        bool operator==(const ReadKey &o) const { return memcmp(p, o.p, sizeof p) == 0 && len == o.len; }
    };
    struct ReadKeyHash {
        size_t operator()(const ReadKey &k) const
        {
            size_t h = (size_t)k.len * 0x9E3779B97F4A7C15ull;
            for (int j = 0; j < 5; ++j) h = (h ^ (size_t)k.p[j]) * 0x100000001B3ull;
            return h;
        }
    };
    struct HapKey {
        const void *p;
        int len;
        bool operator==(const HapKey &o) const { return p == o.p && len == o.len; }
    };
    struct HapKeyHash {
        size_t operator()(const HapKey &k) const { return ((size_t)k.p * 0x9E3779B97F4A7C15ull) ^ (size_t)k.len; }
    };
    std::unordered_map<ReadKey, int64_t, ReadKeyHash> read_at;   // string addresses -> pool offset
    std::unordered_map<HapKey, int64_t, HapKeyHash> hap_at;

    int64_t add_read(const char *rs, const char *q, const char *i, const char *d, const char *c, int len)
    {
        const ReadKey key{{rs, q, i, d, c}, len};
        auto it = read_at.find(key);
        if (it != read_at.end()) return it->second;
        const int64_t off = (int64_t)b.rs.size();
        b.rs.insert(b.rs.end(), (const uint8_t *)rs, (const uint8_t *)rs + len);
        b.q.insert(b.q.end(), (const uint8_t *)q, (const uint8_t *)q + len);
        b.i.insert(b.i.end(), (const uint8_t *)i, (const uint8_t *)i + len);
        b.d.insert(b.d.end(), (const uint8_t *)d, (const uint8_t *)d + len);
        b.c.insert(b.c.end(), (const uint8_t *)c, (const uint8_t *)c + len);
        read_at.emplace(key, off);
        return off;
    }
    int64_t add_hap(const char *hap, int len)
    {
        const HapKey key{hap, len};
        auto it = hap_at.find(key);
        if (it != hap_at.end()) return it->second;
        const int64_t off = (int64_t)b.hap.size();
        b.hap.insert(b.hap.end(), (const uint8_t *)hap, (const uint8_t *)hap + len);
        hap_at.emplace(key, off);
        return off;
    }
};
}  // namespace

extern "C" int phmm_compute_batches(phmm_ctx *ctx, const phmm_batch *batches, int64_t n_batches)
{
    if (!ctx || n_batches < 0 || (n_batches && !batches)) return PHMM_ERR_ARG;
    FlatBuilder fbld;
    ParsedBatches &b = fbld.b;
    int64_t npairs = 0;
    for (int64_t k = 0; k < n_batches; ++k) {
        const phmm_batch &bt = batches[k];
        if (bt.num_reads < 0 || bt.num_haps < 0 || (bt.num_reads && !bt.reads) || (bt.num_haps && !bt.haps)) {
            ctx->error = "malformed batch";
            return PHMM_ERR_ARG;
        }
        b.batch_num_reads.push_back(bt.num_reads);
        b.batch_num_haps.push_back(bt.num_haps);
        for (int r = 0; r < bt.num_reads; ++r) {
            const phmm_read &rd = bt.reads[r];
            if (rd.length <= 0) { ctx->error = "read of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (!rd.bases || !rd.q || !rd.i || !rd.d || !rd.c) { ctx->error = "NULL read field"; return PHMM_ERR_ARG; }
            b.read_len.push_back(rd.length);
            b.read_off.push_back(fbld.add_read(rd.bases, rd.q, rd.i, rd.d, rd.c, rd.length));
        }
        for (int h = 0; h < bt.num_haps; ++h) {
            const phmm_hap &hp = bt.haps[h];
            if (hp.length <= 0) { ctx->error = "haplotype of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (!hp.bases) { ctx->error = "NULL haplotype"; return PHMM_ERR_ARG; }
            b.hap_len.push_back(hp.length);
            b.hap_off.push_back(fbld.add_hap(hp.bases, hp.length));
        }
        if ((int64_t)bt.num_reads * bt.num_haps > 0 && !bt.results) { ctx->error = "NULL results"; return PHMM_ERR_ARG; }
        npairs += (int64_t)bt.num_reads * bt.num_haps;
    }
    b.finish();
    std::vector<double> out((size_t)npairs);
    int rc = phmm_compute_flat(ctx, &b.view, out.data(), nullptr);
    if (rc != PHMM_OK) return rc;
    int64_t p = 0;
    for (int64_t k = 0; k < n_batches; ++k) {
        const int64_t n = (int64_t)batches[k].num_reads * batches[k].num_haps;
        if (n) memcpy(batches[k].results, out.data() + p, (size_t)n * sizeof(double));
        p += n;
    }
    return PHMM_OK;
}

// computelikelihoodsboth's argument list.  The driver fills its slab read-major with pointers
// into per-read / per-haplotype strings (PairHMMUnitTest_orig.cpp:284-297); consecutive pairs
// with the same `rs` pointer become one read's haplotype list, and consecutive reads with the
// same haplotype pointer list fold back into the reads x haplotypes batch they came from.
// Appends the batches of one testcase array to `fbld` (pairs keep their order: batch-major, read-major).
static int append_testcases(phmm_ctx *ctx, FlatBuilder &fbld, const phmm_testcase *tc, int64_t n)
{
    ParsedBatches &b = fbld.b;
    int64_t k = 0;
    int64_t prev_first = -1, prev_nh = 0;      // pair index / size of the previous read's hap list
    while (k < n) {
        const phmm_testcase &t0 = tc[k];
        if (!t0.rs || !t0.q || !t0.i || !t0.d || !t0.c) { ctx->error = "NULL read field in testcase"; return PHMM_ERR_ARG; }
        if (t0.rslen <= 0) { ctx->error = "rslen <= 0"; return PHMM_ERR_EMPTY_SEQ; }
        int64_t k1 = k;
        while (k1 < n && tc[k1].rs == t0.rs && tc[k1].rslen == t0.rslen && tc[k1].q == t0.q && tc[k1].i == t0.i &&
               tc[k1].d == t0.d && tc[k1].c == t0.c)
            ++k1;
        const int64_t nh = k1 - k;
        bool same = (prev_first >= 0 && nh == prev_nh);
        for (int64_t j = 0; same && j < nh; ++j)
            same = tc[k + j].hap == tc[prev_first + j].hap && tc[k + j].haplen == tc[prev_first + j].haplen;
        if (same) {
            b.batch_num_reads.back() += 1;
        } else {
            b.batch_num_reads.push_back(1);
            b.batch_num_haps.push_back((int32_t)nh);
            for (int64_t j = 0; j < nh; ++j) {
                const phmm_testcase &t = tc[k + j];
                if (!t.hap) { ctx->error = "NULL haplotype in testcase"; return PHMM_ERR_ARG; }
                if (t.haplen <= 0) { ctx->error = "haplen <= 0"; return PHMM_ERR_EMPTY_SEQ; }
                b.hap_len.push_back(t.haplen);
                b.hap_off.push_back(fbld.add_hap(t.hap, t.haplen));
            }
        }
        b.read_len.push_back(t0.rslen);
        b.read_off.push_back(fbld.add_read(t0.rs, t0.q, t0.i, t0.d, t0.c, t0.rslen));
        prev_first = k;
        prev_nh = nh;
        k = k1;
    }
    return PHMM_OK;
}

extern "C" int phmm_compute_pairs(phmm_ctx *ctx, const phmm_testcase *tc, double *out, int64_t n)
{
    if (!ctx || n < 0 || (n && (!tc || !out))) return PHMM_ERR_ARG;
    if (n == 0) return PHMM_OK;
    FlatBuilder fbld;
    int rc = append_testcases(ctx, fbld, tc, n);
    if (rc != PHMM_OK) return rc;
    fbld.b.finish();
    return phmm_compute_flat(ctx, &fbld.b.view, out, nullptr);
}

// testcase[] -> flat batches owned by the returned handle (phmm_parsed_batches / phmm_parsed_free): the per-pair part
// of a legacy call, done by the calling thread so that a dispatcher only has to merge and launch.
extern "C" int phmm_flatten_pairs(const phmm_testcase *tc, int64_t n, phmm_parsed **out, char *err, size_t err_len)
{
    if (!out || n < 0 || (n && !tc)) return PHMM_ERR_ARG;
    *out = nullptr;
    phmm_ctx tmp;                       // only its error string is used
    FlatBuilder fbld;
    const int rc = append_testcases(&tmp, fbld, tc, n);
    if (rc != PHMM_OK) { if (err && err_len) snprintf(err, err_len, "%s", tmp.error.c_str()); return rc; }
    phmm_parsed *p = new phmm_parsed();
    p->b = std::move(fbld.b);
    p->b.finish();
    *out = p;
    return PHMM_OK;
}

// ============================================================================================
// asynchronous host-buffer calls: several callers' batches in one planned chunk
// ============================================================================================
// phmm_submit_pairs merges the requests into one set of batches, plans it, queues its uploads and kernels in a
// free pipeline slot and returns; phmm_wait copies every request's results into its own array.  Up to N_SLOTS
// submissions are in flight; a submit that finds no free slot retires the oldest first.
extern "C" int phmm_wait(phmm_ctx *ctx, int64_t ticket, phmm_stats *stats)
{
    if (!ctx) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (stats && stats != &ctx->retired_stats && ctx->retired_stats.n_launches) {
        phmm_stats &a = ctx->retired_stats;
        stats->n_pairs += a.n_pairs; stats->n_cells += a.n_cells; stats->n_rescued += a.n_rescued; stats->n_guard += a.n_guard;
        stats->n_tasks += a.n_tasks; stats->h2d_bytes += a.h2d_bytes; stats->d2h_bytes += a.d2h_bytes;
        stats->n_launches += a.n_launches; stats->n_classes = std::max(stats->n_classes, a.n_classes);
        stats->ms_fp32 += a.ms_fp32; stats->ms_fp64 += a.ms_fp64; stats->ms_total += a.ms_total;
        memset(&a, 0, sizeof a);
    }
    while (!ctx->inflight.empty() && ctx->inflight.front().ticket <= ticket) {
        phmm_ctx::Inflight f = std::move(ctx->inflight.front());
        ctx->inflight.pop_front();
        StreamSet &ss = ctx->ss[f.slot];
        phmm_stats st;
        int rc = wl_finish(ctx, f.w, ss, &st);
        if (rc == PHMM_OK && stats) {
            stats->n_pairs += st.n_pairs; stats->n_cells += st.n_cells; stats->n_rescued += st.n_rescued;
            stats->n_guard += st.n_guard; stats->n_tasks += st.n_tasks; stats->h2d_bytes += st.h2d_bytes;
            stats->d2h_bytes += st.n_pairs * (int64_t)sizeof(double); stats->n_launches += st.n_launches;
            stats->n_classes = std::max(stats->n_classes, st.n_classes);
            stats->ms_fp32 += st.ms_fp32; stats->ms_fp64 += st.ms_fp64; stats->ms_total += st.ms_total;
        }
        if (rc == PHMM_OK) {
            for (const phmm_ctx::Inflight::Scatter &sc : f.scatter)
                if (sc.n > 0 && cudaMemcpyAsync(sc.out, f.w->d_out + sc.p0, (size_t)sc.n * sizeof(double), cudaMemcpyDeviceToHost,
                                                ss.main) != cudaSuccess) { ctx->error = "result copy failed"; rc = PHMM_ERR_CUDA; break; }
            if (rc == PHMM_OK && cudaStreamSynchronize(ss.main) != cudaSuccess) { ctx->error = "result copy failed"; rc = PHMM_ERR_CUDA; }
        }
        phmm_workload_destroy(f.w);
        if (rc != PHMM_OK) return rc;
    }
    return PHMM_OK;
}

// Queues one merged set of batches (owned by the caller of this helper until it returns) in a free slot.
static int submit_merged(phmm_ctx *ctx, const phmm_flat_batches &view, phmm_ctx::Inflight &f, int64_t *ticket,
                         Plan *prebuilt = nullptr)
{
    f.ticket = ++ctx->next_ticket;
    f.slot = (int)(f.ticket % phmm_ctx::N_SLOTS);
    // the slot's arena and streams must be idle
    for (const phmm_ctx::Inflight &g : ctx->inflight)
        if (g.slot == f.slot) { int rc = phmm_wait(ctx, g.ticket, &ctx->retired_stats); if (rc != PHMM_OK) return rc; break; }
    PlanOptions popt = plan_options(ctx);
    if (!prebuilt) { popt.latency = plan_latency(view, popt); popt.one_class = plan_one_class(view, popt); popt.spec = plan_spec(view, popt); }
    StreamSet &ss = ctx->ss[f.slot];
    int rc = wl_create(ctx, &view, &ctx->arena[f.slot], ctx->copy_in, &f.w, prebuilt, &popt);
    if (rc != PHMM_OK) return rc;
    if (cudaEventRecord(ss.ev_uploaded, ctx->copy_in) != cudaSuccess) { ctx->error = "cudaEventRecord failed"; phmm_workload_destroy(f.w); return PHMM_ERR_CUDA; }
    rc = wl_launch(ctx, f.w, ss, ss.ev_uploaded);
    if (rc != PHMM_OK) { cudaDeviceSynchronize(); phmm_workload_destroy(f.w); return rc; }
    *ticket = f.ticket;
    ctx->inflight.push_back(std::move(f));
    return PHMM_OK;
}

extern "C" int phmm_submit_flat(phmm_ctx *ctx, const phmm_flat_request *reqs, int n_reqs, int64_t *ticket)
{
    if (!ctx || n_reqs < 0 || (n_reqs && !reqs) || !ticket) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    phmm_ctx::Inflight f;
    if (n_reqs == 1 && reqs[0].batches) {          // nothing to merge
        const phmm_flat_batches &v = *reqs[0].batches;
        int64_t np = 0;
        for (int64_t b = 0; b < v.n_batches; ++b) np += (int64_t)v.batch_num_reads[b] * v.batch_num_haps[b];
        if (np && !reqs[0].out) { ctx->error = "NULL output"; return PHMM_ERR_ARG; }
        f.scatter.push_back({reqs[0].out, 0, np});
        return submit_merged(ctx, v, f, ticket);
    }
    // concatenate: tables appended, pools copied back to back, offsets rebased
    ParsedBatches m;
    int64_t p = 0;
    size_t tot_r = 0, tot_h = 0, tot_nr = 0, tot_nh = 0, tot_nb = 0;
    for (int k = 0; k < n_reqs; ++k) {
        const phmm_flat_batches *v = reqs[k].batches;
        if (!v || v->n_batches < 0 || v->n_reads < 0 || v->n_haps < 0) { ctx->error = "malformed request"; return PHMM_ERR_ARG; }
        tot_nb += (size_t)v->n_batches; tot_nr += (size_t)v->n_reads; tot_nh += (size_t)v->n_haps;
        for (int64_t r = 0; r < v->n_reads; ++r) tot_r += (size_t)std::max(v->read_len[r], 0);
        for (int64_t h = 0; h < v->n_haps; ++h) tot_h += (size_t)std::max(v->hap_len[h], 0);
    }
    m.batch_num_reads.reserve(tot_nb); m.batch_num_haps.reserve(tot_nb);
    m.read_len.reserve(tot_nr); m.read_off.reserve(tot_nr); m.hap_len.reserve(tot_nh); m.hap_off.reserve(tot_nh);
    m.rs.resize(tot_r); m.q.resize(tot_r); m.i.resize(tot_r); m.d.resize(tot_r); m.c.resize(tot_r); m.hap.resize(tot_h);
    size_t at_r = 0, at_h = 0;
    for (int k = 0; k < n_reqs; ++k) {
        const phmm_flat_batches &v = *reqs[k].batches;
        int64_t np = 0;
        for (int64_t b = 0; b < v.n_batches; ++b) {
            if (v.batch_num_reads[b] < 0 || v.batch_num_haps[b] < 0) { ctx->error = "negative batch size"; return PHMM_ERR_ARG; }
            m.batch_num_reads.push_back(v.batch_num_reads[b]);
            m.batch_num_haps.push_back(v.batch_num_haps[b]);
            np += (int64_t)v.batch_num_reads[b] * v.batch_num_haps[b];
        }
        if (np && !reqs[k].out) { ctx->error = "NULL output"; return PHMM_ERR_ARG; }
        for (int64_t r = 0; r < v.n_reads; ++r) {
            const int len = v.read_len[r];
            if (len <= 0) { ctx->error = "read of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (v.read_off[r] < 0) { ctx->error = "negative read offset"; return PHMM_ERR_ARG; }
            m.read_len.push_back(len);
            m.read_off.push_back((int64_t)at_r);
            memcpy(m.rs.data() + at_r, v.rs + v.read_off[r], (size_t)len);
            memcpy(m.q.data() + at_r, v.q + v.read_off[r], (size_t)len);
            memcpy(m.i.data() + at_r, v.i + v.read_off[r], (size_t)len);
            memcpy(m.d.data() + at_r, v.d + v.read_off[r], (size_t)len);
            memcpy(m.c.data() + at_r, v.c + v.read_off[r], (size_t)len);
            at_r += (size_t)len;
        }
        for (int64_t h = 0; h < v.n_haps; ++h) {
            const int len = v.hap_len[h];
            if (len <= 0) { ctx->error = "haplotype of length <= 0"; return PHMM_ERR_EMPTY_SEQ; }
            if (v.hap_off[h] < 0) { ctx->error = "negative haplotype offset"; return PHMM_ERR_ARG; }
            m.hap_len.push_back(len);
            m.hap_off.push_back((int64_t)at_h);
            memcpy(m.hap.data() + at_h, v.hap + v.hap_off[h], (size_t)len);
            at_h += (size_t)len;
        }
        f.scatter.push_back({reqs[k].out, p, np});
        p += np;
    }
    m.finish();
    return submit_merged(ctx, m.view, f, ticket);
}

// Library-internal (phmm_gatk.cpp): a submission whose plan was built ahead of time on a thread of the caller's, so
// that the thread that owns the context only uploads and launches.  `plan` is moved from.
namespace phmm {
PlanOptions context_plan_options(const phmm_ctx *ctx)
{
    PlanOptions popt = plan_options(ctx);
    return popt;
}
int submit_planned(phmm_ctx *ctx, const phmm_flat_batches *fb, Plan *plan, double *out, int64_t *ticket)
{
    if (!ctx || !fb || !plan || !ticket) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    phmm_ctx::Inflight f;
    if (plan->n_pairs && !out) { ctx->error = "NULL output"; return PHMM_ERR_ARG; }
    f.scatter.push_back({out, 0, plan->n_pairs});
    return submit_merged(ctx, *fb, f, ticket, plan);
}
}  // namespace phmm

extern "C" int phmm_submit_pairs(phmm_ctx *ctx, const phmm_pairs_request *reqs, int n_reqs, int64_t *ticket)
{
    if (!ctx || n_reqs < 0 || (n_reqs && !reqs) || !ticket) return PHMM_ERR_ARG;
    FlatBuilder fbld;
    phmm_ctx::Inflight f;
    int64_t p = 0;
    for (int k = 0; k < n_reqs; ++k) {
        const phmm_pairs_request &r = reqs[k];
        if (r.n < 0 || (r.n && (!r.pairs || !r.out))) { ctx->error = "malformed request"; return PHMM_ERR_ARG; }
        int rc = append_testcases(ctx, fbld, r.pairs, r.n);
        if (rc != PHMM_OK) return rc;
        f.scatter.push_back({r.out, p, r.n});
        p += r.n;
    }
    fbld.b.finish();
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return submit_merged(ctx, fbld.b.view, f, ticket);
}

// ============================================================================================
// measurement helpers
// ============================================================================================
extern "C" int phmm_microbench(phmm_ctx *ctx, int which, double *lane_ops_per_s, float *ms_out)
{
    if (!ctx || !lane_ops_per_s) return PHMM_ERR_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int grid = ctx->sm_count * 8, iters = 4096;
    double ops_per_thread = 0;
    CUDA_TRY(ctx, launch_pipe_bench(which, ctx->d_bench, grid, 64, ctx->ss[0].main, &ops_per_thread));   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        CUDA_TRY(ctx, cudaEventRecord(ctx->ss[0].ev[0], ctx->ss[0].main));
        CUDA_TRY(ctx, launch_pipe_bench(which, ctx->d_bench, grid, iters, ctx->ss[0].main, &ops_per_thread));
        CUDA_TRY(ctx, cudaEventRecord(ctx->ss[0].ev[1], ctx->ss[0].main));
        CUDA_TRY(ctx, cudaStreamSynchronize(ctx->ss[0].main));
        float ms = 0;
        CUDA_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ss[0].ev[0], ctx->ss[0].ev[1]));
        best = std::min(best, ms);
    }
    *lane_ops_per_s = ops_per_thread * 256.0 * grid / (best * 1e-3);
    if (ms_out) *ms_out = best;
    return PHMM_OK;
}

// Plan-only summary (no device needed): used by the CPU test-suite to check the host logic.
extern "C" int phmm_plan_summary(const phmm_flat_batches *fb, int precision, int seg_cols,
                                 int64_t *info /* [8 + 4 * 16 + 2 * 16 + 2] */, uint8_t *pair_cover /* [n_pairs] or NULL */)
{
    if (!fb || !info) return PHMM_ERR_ARG;
    PlanOptions o;
    o.precision = precision & 0xff;
    o.hp = (precision >> 8) & 1;          // plan for the fast FP32 (haplotype-pair) kernels
    o.spec = (precision >> 9) & 1;        // ... with FP64-first reads (info[104] = their pairs, info[105] = their cells)
    if ((precision >> 10) & 1) o.spec = -1;   // ... or decided from the input, as the library does (plan_spec)
    o.seg_cols = seg_cols;
    Plan P;
    std::string err;
    int rc = build_plan(*fb, o, &P, &err);
    if (rc != PHMM_OK) return rc;
    info[0] = P.n_pairs; info[1] = P.n_cells; info[2] = (int64_t)P.classes.size();
    info[3] = P.stream_len; info[4] = P.max_rslen; info[5] = P.max_haplen;
    int64_t ntasks = 0;
    for (size_t c = 0; c < P.classes.size() && c < 16; ++c) {
        info[8 + 4 * c + 0] = P.classes[c].kc.kh;
        info[8 + 4 * c + 1] = P.classes[c].kc.lanes;
        info[8 + 4 * c + 2] = P.classes[c].kc.vl;
        info[8 + 4 * c + 3] = (int64_t)P.classes[c].ntasks();
    }
    for (auto &cp : P.classes) ntasks += (int64_t)cp.ntasks();
    info[6] = ntasks;
    // info[7] = tile cells over all classes; info[72 + 2c ..] = {cells, tile cells} of class c (c < 16): a read of R
    // bases occupies kh x lanes rows of its class's register tile (multi-pass classes: whole blocks of rows), so
    // cells / tile cells is the share of the FP32 work that is not padding
    int64_t tile_total = 0;
    for (size_t c = 0; c < P.classes.size(); ++c) {
        const KClass &kc = P.classes[c].kc;
        int64_t tile = 0;
        auto add = [&](int rid, int64_t cols) {
            if (rid < 0) return;
            const int rows = kc.kh * kc.lanes;
            const int64_t blocks = kc.longr ? (P.reads[rid].len + rows) / rows : 1;
            tile += blocks * rows * cols;
        };
        for (auto &t : P.classes[c].htasks)
            for (int g = 0; g < kc.group(); ++g) add(t.read[g], (int64_t)t.ncols[0] + t.ncols[1]);
        for (auto &t : P.classes[c].tasks)
            for (int g = 0; g < kc.group(); ++g) add(t.read[g], t.ncols);
        tile_total += tile;
        if (c < 16) { info[72 + 2 * c] = P.classes[c].cells; info[72 + 2 * c + 1] = tile; }
    }
    info[7] = tile_total;
    if (o.spec) { info[104] = P.spec_pairs; info[105] = P.spec_cells; }
    else { info[104] = 0; info[105] = 0; }
    if (pair_cover) {
        for (int k = 0; k <= MAX_KH_SCALAR; ++k)
            for (auto &t : P.spec_tasks[k]) {
                int64_t cols = 0;
                for (int j = 0; j < t.nh; ++j) cols += P.hap_len[t.hap0 + j] + SEP_COLS;
                const ReadRec &rr = P.reads[t.read[0]];
                if (cols != t.ncols || t.read[1] != SPEC_TASK || (rr.len + 1 + 31) / 32 != k) return PHMM_ERR_ARG;
                for (int j = 0; j < t.nh; ++j) pair_cover[rr.out_base + t.out0 + j]++;
            }
        // walk every task the way the kernel does and count how often each pair is produced
        for (auto &cp : P.classes)
            for (auto &t : cp.htasks)
                for (int h = 0; h < 2; ++h) {
                    const int nh = t.nh[h];
                    int64_t cols = 0;
                    for (int j = 0; j < nh; ++j) cols += P.hap_len[t.hap0[h] + j] + SEP_COLS;
                    if (cols != t.ncols[h] || nh > MAX_RUN_HAPS || !cp.kc.hp) return PHMM_ERR_ARG;
                    for (int g = 0; g < cp.kc.group(); ++g) {
                        if (t.read[g] < 0) continue;
                        const ReadRec &rr = P.reads[t.read[g]];
                        if (rr.len > cp.kc.capacity()) return PHMM_ERR_ARG;
                        for (int j = 0; j < nh; ++j) pair_cover[rr.out_base + t.out0[h] + j]++;
                    }
                }
        for (auto &cp : P.classes)
            for (auto &t : cp.tasks) {
                const int nh = t.nh;
                int64_t cols = 0;
                for (int j = 0; j < nh; ++j) cols += P.hap_len[t.hap0 + j] + SEP_COLS;
                if (cols != t.ncols || nh > MAX_RUN_HAPS) return PHMM_ERR_ARG;
                for (int g = 0; g < cp.kc.group(); ++g) {
                    if (t.read[g] < 0) continue;
                    const ReadRec &rr = P.reads[t.read[g]];
                    if (rr.len > cp.kc.capacity()) return PHMM_ERR_ARG;
                    for (int j = 0; j < nh; ++j) pair_cover[rr.out_base + t.out0 + j]++;
                }
            }
    }
    return PHMM_OK;
}
