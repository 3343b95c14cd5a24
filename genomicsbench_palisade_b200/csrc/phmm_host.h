// Internal host-side declarations of libphmm_cuda (planner, tables, kernel lookup).
#ifndef PHMM_HOST_H
#define PHMM_HOST_H

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/phmm_cuda.h"
#include "phmm_types.h"

namespace phmm {

// ---- probability tables (phmm_tables.cpp) ---------------------------------------------------
static const int MM_REACHABLE = (127 * 128) / 2 + 127 + 1;   // 8256: max index with quals <= 127
struct HostTables {
    float ph2pr_f[128];
    double ph2pr_d[128];
    std::vector<float> mm_f;     // [MM_REACHABLE]
    std::vector<double> mm_d;
    float init_f, log10_init_f;
    double init_d, log10_init_d;
};
const HostTables &host_tables();

// ---- planner (phmm_plan.cpp) -------------------------------------------------------------------
struct PlanOptions {
    int precision = 0;        // PHMM_OPT_PRECISION
    int seg_cols = 0;         // 0 = adaptive
    int resident_warps = 2368;   // warps the device keeps in flight (for adaptive run sizing)
    int force_kh = 0, force_lanes = 0;   // debugging / benchmarking: pin the kernel class
    int hp = 0;               // 1 = plan haplotype-pair tasks (fast FP32 kernels) for single-pass reads
    int latency = -1;         // 1 = plan for the latency of a small call (32 lanes per read, one haplotype per run); -1 = decide from the size
    int spec = 0;             // -1 = decide from the input (plan_spec, once per call); 1 = FP64-first: reads predicted to fail the FP32 pass against their whole batch (phmm_cost.cpp) skip it;
                              // their tasks are laid into the rescue lists by the planner (fast FP32 kernels, precision 0, throughput plans)
    int one_class = -1;       // latency plans whose tasks fit the device in one wave: > 0 = every read in the 32-lane class of this
                              // height, all rescue / guard work in that scalar class; 0 = classes by read length; -1 = decide from the input
};

struct ClassPlan {
    KClass kc;
    std::vector<Task> tasks;   // sorted by decreasing cost
    std::vector<HTask> htasks; // haplotype-pair classes (kc.hp) use these instead
    size_t ntasks() const { return kc.hp ? htasks.size() : tasks.size(); }
    int64_t cells = 0;
    int64_t pairs = 0;
};

static const int64_t HOST_CODED_STREAM_MAX = 96 << 10;

struct Plan {
    // Referenced pieces of a caller's byte pool, packed back to back on the device: bytes [src, src + len) go to
    // device pool offset dst.
    struct Segment { int64_t src, len, dst; };
    std::vector<ReadRec> reads;
    int64_t stream_len = 0;              // bytes of the coded haplotype stream (+ STREAM_SLACK), built on the device ...
    std::vector<uint8_t> stream;         // ... unless the call is small: then coded here (stream_len <= HOST_CODED_STREAM_MAX)
    std::vector<HapSrc> hap_src;         // per haplotype: raw bases -> stream columns
    std::vector<Segment> hap_segments;   // referenced pieces of the caller's haplotype pool (declared below)
    int64_t hap_pool_len = 0;
    std::vector<int32_t> hap_len;        // stream order, + 1 slack entry
    std::vector<ClassPlan> classes;
    int64_t n_pairs = 0, n_cells = 0;
    int latency = 0;                     // planned for the latency of a small call (PlanOptions::latency)
    // Referenced pieces of the caller's read pools, packed back to back on the device: bytes
    // [src, src + len) of each of the five pools go to device pool offset dst.
    std::vector<Segment> segments;
    int64_t pool_len = 0;                // sum of segment lengths
    int64_t resc_cap[MAX_KH_SCALAR + 1]; // pairs whose read falls in each scalar class
    // FP64-first tasks per scalar class (Task::read[1] == SPEC_TASK): the head of that class's rescue list
    std::vector<Task> spec_tasks[MAX_KH_SCALAR + 1];
    int64_t spec_cap[MAX_KH_SCALAR + 1]; // pairs of FP64-first reads per class (capacity of the band list)
    int64_t spec_pairs = 0, spec_cells = 0;
    int resc_force = 0;                  // > 0: one-wave call, all rescue / guard work in this scalar class (KernelParams::resc_force)
    int max_rslen = 0, max_haplen = 0;
    int long_max_cols = 0;               // longest run handed to a multi-pass kernel (0 = none)
};

// Chooses the kernel class for a read length; returns false when no single-pass class fits.
bool choose_class(int rslen, const PlanOptions &opt, KClass *kc);
int plan_latency(const phmm_flat_batches &fb, const PlanOptions &opt);
int plan_one_class(const phmm_flat_batches &fb, const PlanOptions &opt);     // opt.latency resolved; like plan_latency, once per call
// Builds the plan; on failure returns a PHMM_ERR_* code and fills `err`.
int build_plan(const phmm_flat_batches &fb, const PlanOptions &opt, Plan *plan, std::string *err);

// ---- library-internal entry of the asynchronous pipeline (phmm_api.cu), for callers that plan on their own threads
PlanOptions context_plan_options(const phmm_ctx *ctx);
int submit_planned(phmm_ctx *ctx, const phmm_flat_batches *fb, Plan *plan, double *out, int64_t *ticket);

// ---- cost model of the scheduler (phmm_cost.cpp) -------------------------------------------------
// Device time of a read against one haplotype column = lanes x (rows per thread + step_rows) row steps of its
// kernel class; an FP64 row step costs f64_row_cost FP32 ones, a row step of the bit-exact FP32 kernels (two reads
// per thread, unfused) exact_row_cost.  Calibrated on B200 (tools/class_sweep.py); PHMM_COST_MODEL overrides.
struct CostModel {
    double step_rows_fast = 1.6;      // per-step overhead of a lane in rows: shuffles, LDS, loop control
    double step_rows_exact = 1.5;     //   (fit to profiles/r1_v4_class_sweep.log: 16 lanes, kh 4 / 8 / 16 = 3114 / 3650 / 4027 GCUPS)
    double step_rows_f64 = 1.0;
    double f64_row_cost = 2.2;        // FP32 cell rate / FP64 cell rate on full tiles (3716 / 1737 GCUPS, C4)
    double exact_row_cost = 0.9;
    double lanes8_cost = 1.08;        // 8-lane classes: 3729 against 4027 GCUPS at kh = 16
    double lanes32_cost = 1.015;
};
CostModel default_cost_model();
// cost[b]: predicted time of batch b (arbitrary unit); read_rescued (may be NULL): 1 where the predictor expects
// most pairs of the read to be re-run in FP64.
int batch_costs(const phmm_flat_batches &fb, int fast, int precision, const CostModel &m, int64_t *cost,
                uint8_t *read_rescued, int spec_plan = -1);
// FP64-first planning on (default) unless PHMM_SPEC64=0
bool spec_enabled();
int plan_spec(const phmm_flat_batches &fb, const PlanOptions &opt);      // opt.latency resolved; once per call

// ---- kernels (phmm_inst_*.cu, phmm_microbench.cu) -------------------------------------------
typedef void (*KernelF32)(const KernelParams<float>);
typedef void (*KernelF64)(const KernelParams<double>);
KernelF32 lookup_f32_l32_f0(int kh);
KernelF32 lookup_f32_l32_f1(int kh);
KernelF32 lookup_f32_l16_f0(int kh);
KernelF32 lookup_f32_l16_f1(int kh);
KernelF32 lookup_f32_hp_l32(int kh);
KernelF32 lookup_f32_hp_l16(int kh);
KernelF32 lookup_f32_hp_l8(int kh);
KernelF64 lookup_f64_f0(int kh);
KernelF64 lookup_f64_f1(int kh);
KernelF32 lookup_f32(const KClass &kc, bool fast);
KernelF64 lookup_f64(int kh, bool fast);
KernelF32 lookup_f32_long(bool fast);
KernelF64 lookup_f64_long(bool fast);
cudaError_t launch_pipe_bench(int which, float *out, int grid, int iters, cudaStream_t st,
                              double *ops_per_thread);

// ---- text parser (phmm_parse.cpp) -----------------------------------------------------------
struct ParsedBatches {
    std::vector<int32_t> batch_num_reads, batch_num_haps, read_len, hap_len;
    std::vector<int64_t> read_off, hap_off;
    std::vector<uint8_t> rs, q, i, d, c, hap;
    phmm_flat_batches view;
    void finish();     // fills `view` from the vectors
};

}  // namespace phmm

struct phmm_parsed {
    phmm::ParsedBatches b;           // owns the data when parsed from text ...
    void *map = nullptr;             // ... or b.view points into this mapping (phmm_cache_open)
    size_t map_len = 0;
    std::vector<double> expected;    // phmm_testdata_*: expected log10 likelihood per pair
    ~phmm_parsed();
};

#endif
