// phmm_forward_region_noalloc: one reads x haplotypes region through the bit-exact pair-HMM kernels with
// NO allocation of any kind -- for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  No counterpart in the reference.
//
// The kernels are the library's own (phmm_forward_kernel<float | double, KH, 32, exact>, phmm_kernels.cuh); what
// differs from phmm_compute_* is where their scratch lives.  Everything those kernels need besides the caller's
// inputs and outputs -- probability tables, read records, the coded haplotype stream, task lists, queue heads, the
// FP64-rescue list, raw sums -- sits in ONE file-scope `static __device__` array of NA_SCRATCH_BYTES (3 MiB), which
// every GPU that loads the library holds until exit.  Planning runs on the device (noalloc_setup_kernel), so the call
// also makes no host allocation and reads nothing back.
//
// In host code the array's NAME is a host-side stand-in, not a device address: a kernel handed `g_noalloc_scratch`
// from the host would fault (or, where the GPU can reach host memory, silently use the wrong memory).  The address
// comes from cudaGetSymbolAddress, once per device, with that device current (dev_setup below).
//
// Every call shares the array and finds the previous call's contents in it: the queue heads, list counts and the
// stream's separator / slack bytes are cleared by the set-up kernel of every call.  Calls that overlap on one GPU
// (two streams, two threads) race on it -- the caller serialises them.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "phmm_host.h"
#include "phmm_kernels.cuh"

namespace phmm {
namespace {

// ---- stated maxima (mirrored in include/phmm_cuda.h) ------------------------------------------------------------
constexpr int NA_MAX_READS = PHMM_NOALLOC_MAX_READS;
constexpr int NA_MAX_HAPS = PHMM_NOALLOC_MAX_HAPS;
constexpr int NA_MAX_PAIRS = PHMM_NOALLOC_MAX_PAIRS;
constexpr int NA_MAX_READ_LEN = PHMM_NOALLOC_MAX_READ_LEN;
constexpr int NA_MAX_HAP_BYTES = PHMM_NOALLOC_MAX_HAP_BYTES;
constexpr int NA_RUN_COLS = 4096;          // a run of haplotypes is closed before it would exceed this many columns
constexpr int NA_RUN_HAPS = MAX_RUN_HAPS;  // ... or this many haplotypes
// two consecutive runs closed by the column rule hold > NA_RUN_COLS columns together:
// runs <= 2 * (65536 + 4 * 128) / 4096 + 1 + 128 / 62 < 40
constexpr int NA_MAX_RUNS = 40;
constexpr int NA_MAX_TASKS = (NA_MAX_READS / 2) * NA_MAX_RUNS;
static_assert(NA_MAX_READ_LEN <= MAX_KH * 32 - 1, "reads must fit the tallest single-pass class");

// ---- layout of the scratch array (bytes, every piece 256-byte aligned) ---------------------------------------
constexpr size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t OFF_PH2PR_F = 0;
constexpr size_t OFF_PH2PR_D = OFF_PH2PR_F + al256(128 * sizeof(float));
constexpr size_t OFF_MM_F = OFF_PH2PR_D + al256(128 * sizeof(double));
constexpr size_t OFF_MM_D = OFF_MM_F + al256(MM_REACHABLE * sizeof(float));
constexpr size_t OFF_INTS = OFF_MM_D + al256(MM_REACHABLE * sizeof(double));
constexpr size_t OFF_READS = OFF_INTS + 256;
constexpr size_t OFF_HAPLEN = OFF_READS + al256(NA_MAX_READS * sizeof(ReadRec));
constexpr size_t OFF_HAPCOL = OFF_HAPLEN + al256((NA_MAX_HAPS + 2) * sizeof(int32_t));
constexpr size_t STREAM_BYTES = al256(NA_MAX_HAP_BYTES + SEP_COLS * NA_MAX_HAPS + 16 + STREAM_SLACK);
constexpr size_t OFF_STREAM = OFF_HAPCOL + al256((NA_MAX_HAPS + 2) * sizeof(int32_t));
constexpr size_t OFF_TASKS = OFF_STREAM + STREAM_BYTES;
constexpr size_t OFF_RESC = OFF_TASKS + al256(NA_MAX_TASKS * sizeof(Task));
constexpr size_t OFF_RAW32 = OFF_RESC + al256(NA_MAX_PAIRS * sizeof(Task));
constexpr size_t OFF_RAW64 = OFF_RAW32 + al256(NA_MAX_PAIRS * sizeof(float));
constexpr size_t OFF_RESCUED = OFF_RAW64 + al256(NA_MAX_PAIRS * sizeof(double));
constexpr size_t OFF_END = OFF_RESCUED + al256(NA_MAX_PAIRS);
constexpr size_t NA_SCRATCH_BYTES = PHMM_NOALLOC_SCRATCH_BYTES;
static_assert(OFF_END <= NA_SCRATCH_BYTES, "scratch layout exceeds the declared size");

// ints[]: what must start at zero is cleared by noalloc_setup_kernel in every call
enum { I_NTASKS = 0, I_HEAD32 = 1, I_NRESC = 2, I_HEAD64 = 3, I_STATUS = 4, I_UNUSED_COUNTS = 16 /* [16..31] */ };

static __device__ __align__(256) uint8_t g_noalloc_scratch[NA_SCRATCH_BYTES];

struct SetupArgs {
    const int32_t *read_off, *hap_off;
    const uint8_t *hap;
    double *out;
    uint8_t *scratch;
    int32_t n_reads, n_haps, max_read_len, hap_bytes;
};

// Plans the region on the device: read records, haplotype columns and lengths, the coded haplotype stream
// (A, C, T, G, N -> 1..5, four zero separator columns after every haplotype, phmm_types.h), runs of haplotypes and
// one task per (pair of reads, run).  One block.  A region whose tables contradict the sizes the host stated gets
// no tasks and NaN in every result.
__global__ void __launch_bounds__(256) noalloc_setup_kernel(const SetupArgs A)
{
    __shared__ int s_run_h0[NA_MAX_RUNS], s_run_nh[NA_MAX_RUNS], s_run_col[NA_MAX_RUNS], s_run_ncols[NA_MAX_RUNS];
    __shared__ int s_nruns, s_bad, s_total;
    uint8_t *S = A.scratch;            // == g_noalloc_scratch (device code could name the array directly as well)
    int32_t *ints = reinterpret_cast<int32_t *>(S + OFF_INTS);
    ReadRec *reads = reinterpret_cast<ReadRec *>(S + OFF_READS);
    int32_t *hap_len = reinterpret_cast<int32_t *>(S + OFF_HAPLEN);
    int32_t *hap_col = reinterpret_cast<int32_t *>(S + OFF_HAPCOL);
    uint8_t *stream = S + OFF_STREAM;
    Task *tasks = reinterpret_cast<Task *>(S + OFF_TASKS);
    const int tid = threadIdx.x, nt = blockDim.x;

    if (tid == 0) {
        int bad = 0, total = 0, nruns = 0;
        const int32_t h_base = A.hap_off[0];
        for (int h = 0; h < A.n_haps; ++h) {
            const int32_t o = A.hap_off[h], len = A.hap_off[h + 1] - o;
            if (o < h_base || len <= 0 || len > A.hap_bytes || o - h_base > A.hap_bytes - len) { bad = 1; break; }
            hap_col[h] = total;
            hap_len[h] = len;
            // greedy runs: close the open run when this haplotype would overfill it
            if (nruns == 0 || s_run_nh[nruns - 1] >= NA_RUN_HAPS ||
                s_run_ncols[nruns - 1] + len + SEP_COLS > NA_RUN_COLS) {
                if (nruns == NA_MAX_RUNS) { bad = 1; break; }
                s_run_h0[nruns] = h; s_run_nh[nruns] = 0; s_run_col[nruns] = total; s_run_ncols[nruns] = 0;
                ++nruns;
            }
            s_run_nh[nruns - 1] += 1;
            s_run_ncols[nruns - 1] += len + SEP_COLS;
            total += len + SEP_COLS;
        }
        hap_len[A.n_haps] = 1;             // the kernels read one entry past a run's last haplotype
        hap_col[A.n_haps] = total;
        s_nruns = nruns; s_bad = bad; s_total = total;
    }
    __syncthreads();
    for (int r = tid; r < A.n_reads; r += nt) {
        const int32_t o = A.read_off[r], len = A.read_off[r + 1] - o;
        if (o < 0 || len < 1 || len > A.max_read_len) atomicOr(&s_bad, 1);
        ReadRec rr;
        rr.off = o; rr.out_base = (int64_t)r * A.n_haps; rr.len = len; rr.batch = 0;
        reads[r] = rr;
    }
    __syncthreads();
    const int bad = s_bad, nruns = s_nruns, total = s_total;
    const int n_pairs = A.n_reads * A.n_haps;
    if (bad) {
        for (int p = tid; p < n_pairs; p += nt) A.out[p] = __longlong_as_double(0x7ff8000000000000ll);
    } else {
        // separators and the slack past the last column are zero; the previous call's stream is still in there
        const int nwords = (total + 16 + STREAM_SLACK + 3) / 4;
        for (int k = tid; k < nwords; k += nt) reinterpret_cast<uint32_t *>(stream)[k] = 0;
        __syncthreads();
        for (int h = 0; h < A.n_haps; ++h) {
            const uint8_t *src = A.hap + A.hap_off[h];
            uint8_t *dst = stream + hap_col[h];
            const int len = hap_len[h];
            for (int k = tid; k < len; k += nt) dst[k] = base_code(src[k]);
        }
        const int nrp = (A.n_reads + 1) / 2;
        for (int t = tid; t < nrp * nruns; t += nt) {
            const int rp = t / nruns, run = t % nruns;
            Task tk;
            tk.stream_off = s_run_col[run];
            tk.ncols = s_run_ncols[run];
            tk.hap0 = s_run_h0[run];
            tk.out0 = s_run_h0[run];
            tk.read[0] = 2 * rp;
            tk.read[1] = (2 * rp + 1 < A.n_reads) ? 2 * rp + 1 : -1;
            tk.read[2] = -1; tk.read[3] = -1;
            tk.nh = s_run_nh[run];
            tasks[t] = tk;
        }
    }
    if (tid < 64) {
        int v = 0;
        if (tid == I_NTASKS) v = bad ? 0 : ((A.n_reads + 1) / 2) * nruns;
        if (tid == I_STATUS) v = bad;
        ints[tid] = v;
    }
}

// IntelPairHmmCSource.cpp:73-75: a pair whose FP32 sum is below MIN_ACCEPTED is run again in FP64.  The FP32 kernel
// ran in float-only mode (it finalised every pair in float); this lists the pairs the FP64 kernel has to overwrite.
__global__ void __launch_bounds__(256) noalloc_collect_kernel(uint8_t *S, int n_reads, int n_haps)
{
    int32_t *ints = reinterpret_cast<int32_t *>(S + OFF_INTS);
    if (ints[I_STATUS]) return;
    const float *raw32 = reinterpret_cast<const float *>(S + OFF_RAW32);
    const int32_t *hap_len = reinterpret_cast<const int32_t *>(S + OFF_HAPLEN);
    const int32_t *hap_col = reinterpret_cast<const int32_t *>(S + OFF_HAPCOL);
    Task *resc = reinterpret_cast<Task *>(S + OFF_RESC);
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_reads * n_haps) return;
    if (raw32[p] < MIN_ACCEPTED_F) {
        const int r = p / n_haps, h = p % n_haps;
        Task tk;
        tk.stream_off = hap_col[h];
        tk.ncols = hap_len[h] + SEP_COLS;
        tk.hap0 = h; tk.out0 = h;
        tk.read[0] = r; tk.read[1] = -1; tk.read[2] = -1; tk.read[3] = -1;
        tk.nh = 1;
        resc[atomicAdd(&ints[I_NRESC], 1)] = tk;
    }
}

// ---- per-device set-up: the scratch array's device address, the tables, the kernels' launch geometry ---------
struct DevState {
    bool ready = false;
    uint8_t *scratch = nullptr;        // cudaGetSymbolAddress(g_noalloc_scratch) on this device
    int sm_count = 0;
    int nb32[MAX_KH + 1] = {0}, nb64[MAX_KH + 1] = {0};   // resident one-warp blocks per SM of each class kernel
};
constexpr int NA_MAX_DEVICES = 64;
DevState g_dev[NA_MAX_DEVICES];
std::mutex g_dev_mutex;

int fail(char *err, size_t cap, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (err && cap) {
        if (e != cudaSuccess) snprintf(err, cap, "phmm_forward_region_noalloc: %s: %s", what, cudaGetErrorString(e));
        else snprintf(err, cap, "phmm_forward_region_noalloc: %s", what);
    }
    return code;
}

// Once per device, with that device current.  Nothing here allocates device memory: the copies go into the static array.
int dev_setup(int device, DevState **out, char *err, size_t cap)
{
    std::lock_guard<std::mutex> lock(g_dev_mutex);
    DevState &D = g_dev[device];
    *out = &D;
    if (D.ready) return PHMM_OK;
    cudaError_t e;
    uint8_t *ptr = nullptr;
    // the array's name is only a handle on the host; this is its address in the current device's memory
    if ((e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch)) != cudaSuccess || !ptr)
        return fail(err, cap, PHMM_ERR_CUDA, "cudaGetSymbolAddress(scratch)", e);
    const HostTables &t = host_tables();
    if ((e = cudaMemcpy(ptr + OFF_PH2PR_F, t.ph2pr_f, 128 * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(ptr + OFF_PH2PR_D, t.ph2pr_d, 128 * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(ptr + OFF_MM_F, t.mm_f.data(), MM_REACHABLE * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(ptr + OFF_MM_D, t.mm_d.data(), MM_REACHABLE * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess)
        return fail(err, cap, PHMM_ERR_CUDA, "table upload into the scratch array", e);
    if ((e = cudaDeviceGetAttribute(&D.sm_count, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess)
        return fail(err, cap, PHMM_ERR_CUDA, "cudaDeviceGetAttribute", e);
    for (int kh = 1; kh <= MAX_KH; ++kh) {
        const int smem = WARPS_PER_BLOCK * smem_per_warp(kh);
        const void *k32 = (const void *)lookup_f32_l32_f0(kh), *k64 = (const void *)lookup_f64_f0(kh);
        if (!k32 || !k64) return fail(err, cap, PHMM_ERR_CUDA, "kernel class missing from the library");
        if ((e = cudaFuncSetAttribute(k32, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess ||
            (e = cudaFuncSetAttribute(k64, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess)
            return fail(err, cap, PHMM_ERR_CUDA, "cudaFuncSetAttribute", e);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&D.nb32[kh], k32, WARPS_PER_BLOCK * 32, smem) != cudaSuccess || D.nb32[kh] < 1) D.nb32[kh] = 1;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&D.nb64[kh], k64, WARPS_PER_BLOCK * 32, smem) != cudaSuccess || D.nb64[kh] < 1) D.nb64[kh] = 1;
    }
    cudaGetLastError();
    D.scratch = ptr;
    D.ready = true;
    return PHMM_OK;
}

}  // namespace
}  // namespace phmm

using namespace phmm;

extern "C" int phmm_forward_region_noalloc(int device, void *cuda_stream,
                                           const uint8_t *d_rs, const uint8_t *d_q, const uint8_t *d_i,
                                           const uint8_t *d_d, const uint8_t *d_c, const int32_t *d_read_off,
                                           int32_t n_reads, int32_t max_read_len,
                                           const uint8_t *d_hap, const int32_t *d_hap_off, int32_t n_haps,
                                           int32_t hap_bytes, double *d_out, uint8_t *d_rescued,
                                           char *err, size_t err_cap)
{
    if (err && err_cap) err[0] = 0;
    if (device < 0 || device >= NA_MAX_DEVICES) return fail(err, err_cap, PHMM_ERR_ARG, "device out of range");
    if (n_reads < 0 || n_haps < 0) return fail(err, err_cap, PHMM_ERR_ARG, "negative count");
    if (n_reads == 0 || n_haps == 0) return PHMM_OK;
    if (!d_rs || !d_q || !d_i || !d_d || !d_c || !d_read_off || !d_hap || !d_hap_off || !d_out)
        return fail(err, err_cap, PHMM_ERR_ARG, "NULL device pointer");
    if (max_read_len < 1 || hap_bytes < 1) return fail(err, err_cap, PHMM_ERR_EMPTY_SEQ, "max_read_len or hap_bytes < 1");
    // the scratch array is static: sizes beyond what it was laid out for are refused, never grown into
    char msg[160];
    if (n_reads > NA_MAX_READS) { snprintf(msg, sizeof msg, "%d reads exceed PHMM_NOALLOC_MAX_READS = %d", n_reads, NA_MAX_READS); return fail(err, err_cap, PHMM_ERR_TOO_LONG, msg); }
    if (n_haps > NA_MAX_HAPS) { snprintf(msg, sizeof msg, "%d haplotypes exceed PHMM_NOALLOC_MAX_HAPS = %d", n_haps, NA_MAX_HAPS); return fail(err, err_cap, PHMM_ERR_TOO_LONG, msg); }
    if ((int64_t)n_reads * n_haps > NA_MAX_PAIRS) { snprintf(msg, sizeof msg, "%lld pairs exceed PHMM_NOALLOC_MAX_PAIRS = %d", (long long)n_reads * n_haps, NA_MAX_PAIRS); return fail(err, err_cap, PHMM_ERR_TOO_LONG, msg); }
    if (max_read_len > NA_MAX_READ_LEN) { snprintf(msg, sizeof msg, "max_read_len %d exceeds PHMM_NOALLOC_MAX_READ_LEN = %d", max_read_len, NA_MAX_READ_LEN); return fail(err, err_cap, PHMM_ERR_TOO_LONG, msg); }
    if (hap_bytes > NA_MAX_HAP_BYTES) { snprintf(msg, sizeof msg, "%d haplotype bytes exceed PHMM_NOALLOC_MAX_HAP_BYTES = %d", hap_bytes, NA_MAX_HAP_BYTES); return fail(err, err_cap, PHMM_ERR_TOO_LONG, msg); }

    int prev = -1;
    cudaError_t e;
    if ((e = cudaGetDevice(&prev)) != cudaSuccess) return fail(err, err_cap, PHMM_ERR_CUDA, "cudaGetDevice", e);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(err, err_cap, PHMM_ERR_CUDA, "cudaSetDevice", e);
    struct Restore { int d; ~Restore() { if (d >= 0) cudaSetDevice(d); } } restore{prev};
    DevState *D = nullptr;
    int rc = dev_setup(device, &D, err, err_cap);
    if (rc != PHMM_OK) return rc;

    cudaStream_t st = (cudaStream_t)cuda_stream;
    uint8_t *S = D->scratch;
    int32_t *ints = reinterpret_cast<int32_t *>(S + OFF_INTS);
    const int n_pairs = n_reads * n_haps;
    const int kh = rescue_class(max_read_len);          // 32 lanes x kh rows hold max_read_len + 1 rows
    const HostTables &t = host_tables();

    SetupArgs A;
    A.read_off = d_read_off; A.hap_off = d_hap_off; A.hap = d_hap; A.out = d_out; A.scratch = S;
    A.n_reads = n_reads; A.n_haps = n_haps; A.max_read_len = max_read_len; A.hap_bytes = hap_bytes;
    noalloc_setup_kernel<<<1, 256, 0, st>>>(A);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, err_cap, PHMM_ERR_CUDA, "set-up kernel launch", e);

    // FP32 pass over every pair, float-only policy (computelikelihoodsfloat): out = log10f(sum) - log10f(2^120)
    KernelParams<float> pf;
    memset(&pf, 0, sizeof pf);
    pf.rs = d_rs; pf.q = d_q; pf.i = d_i; pf.d = d_d; pf.c = d_c;
    pf.reads = reinterpret_cast<const ReadRec *>(S + OFF_READS);
    pf.stream = S + OFF_STREAM;
    pf.hap_len = reinterpret_cast<const int32_t *>(S + OFF_HAPLEN);
    pf.tasks = reinterpret_cast<const Task *>(S + OFF_TASKS);
    pf.ntasks_ptr = ints + I_NTASKS; pf.counter = ints + I_HEAD32;
    pf.ph2pr = reinterpret_cast<const float *>(S + OFF_PH2PR_F);
    pf.mm = reinterpret_cast<const float *>(S + OFF_MM_F);
    pf.init_const = t.init_f; pf.log10_init = t.log10_init_f;
    pf.out = d_out;
    pf.raw32 = reinterpret_cast<float *>(S + OFF_RAW32);
    pf.raw64 = reinterpret_cast<double *>(S + OFF_RAW64);
    pf.rescued = d_rescued ? d_rescued : S + OFF_RESCUED;
    pf.resc = reinterpret_cast<Task *>(S + OFF_RESC);           // not appended to in float-only mode
    pf.resc_count = ints + I_UNUSED_COUNTS; pf.resc_pairs = ints + I_UNUSED_COUNTS + 15;
    pf.mode = 1;
    {
        const int smem = WARPS_PER_BLOCK * smem_per_warp(kh);
        const int64_t want = (int64_t)((n_reads + 1) / 2) * (n_haps < NA_MAX_RUNS ? n_haps : NA_MAX_RUNS);
        const int64_t full = (int64_t)D->nb32[kh] * D->sm_count;
        const int grid = (int)(want < full ? want : full);
        lookup_f32_l32_f0(kh)<<<grid, WARPS_PER_BLOCK * 32, smem, st>>>(pf);
        if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, err_cap, PHMM_ERR_CUDA, "FP32 kernel launch", e);
    }

    // pairs below MIN_ACCEPTED -> FP64 (IntelPairHmmCSource.cpp:73-75)
    noalloc_collect_kernel<<<(n_pairs + 255) / 256, 256, 0, st>>>(S, n_reads, n_haps);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, err_cap, PHMM_ERR_CUDA, "collect kernel launch", e);

    KernelParams<double> pd;
    memset(&pd, 0, sizeof pd);
    pd.rs = d_rs; pd.q = d_q; pd.i = d_i; pd.d = d_d; pd.c = d_c;
    pd.reads = pf.reads; pd.stream = pf.stream; pd.hap_len = pf.hap_len;
    pd.tasks = reinterpret_cast<const Task *>(S + OFF_RESC);
    pd.ntasks_ptr = ints + I_NRESC; pd.counter = ints + I_HEAD64;
    pd.ph2pr = reinterpret_cast<const double *>(S + OFF_PH2PR_D);
    pd.mm = reinterpret_cast<const double *>(S + OFF_MM_D);
    pd.init_const = t.init_d; pd.log10_init = t.log10_init_d;
    pd.out = d_out; pd.raw32 = pf.raw32; pd.raw64 = pf.raw64; pd.rescued = pf.rescued;
    pd.mode = 0;                                                  // marks what it computes as rescued
    {
        const int smem = WARPS_PER_BLOCK * smem_per_warp(kh);
        const int64_t full = (int64_t)D->nb64[kh] * D->sm_count;
        const int grid = (int)(n_pairs < full ? n_pairs : full);
        lookup_f64_f0(kh)<<<grid, WARPS_PER_BLOCK * 32, smem, st>>>(pd);
        if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, err_cap, PHMM_ERR_CUDA, "FP64 kernel launch", e);
    }
    return PHMM_OK;
}
