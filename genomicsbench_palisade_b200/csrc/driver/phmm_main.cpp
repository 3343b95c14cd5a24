// `phmm` -- the rewritten GenomicsBench pair-HMM driver on top of libphmm_cuda's C ABI.
//
// Keeps the reference driver's interface (benchmarks/phmm/PairHMMUnitTest_orig.cpp):
//   -f/--testfile <file>   input in the reference's text format (:97-169)
//   -l/--loop <n>          repeat the compute phase n times (:225,:266)
//   -t/--threads <n>       accepted for compatibility; here it caps the number of GPUs used
//   -p/--print             one "%lf" per pair in batch order (the PRINT_OUTPUT build, :316-321)
//   -c/--cache             new: keep a binary cache <file>.phmmsoa beside the input and map it
//                          instead of re-parsing the text when it is at least as new as the input
// and its final line "PairHMM completed. Kernel runtime: %.2f sec" (:327).  As in the original,
// parsing is outside the timed region (:243 vs :259).  What differs: all batches are parsed into
// flat pools (no per-string heap allocation) and the visible B200s pull chunks of consecutive batches, sized by
// predicted device time, from one shared queue, three chunks in flight per GPU (see main).
#include <getopt.h>
#include <malloc.h>
#include <sys/stat.h>
#include <sys/time.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "phmm_cuda.h"

namespace {

double now_sec()
{
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    return tv.tv_sec + tv.tv_usec * 1e-6;
}

void usage(const char *argv0)
{
    fprintf(stderr, "usage: %s -f <testfile> [-l <loops>] [-t <max GPUs>] [-p] [-c]\n", argv0);
}

}  // namespace

int main(int argc, char **argv)
{
    static const struct option long_opts[] = {{"testfile", required_argument, nullptr, 'f'},
                                              {"loop", required_argument, nullptr, 'l'},
                                              {"threads", required_argument, nullptr, 't'},
                                              {"print", no_argument, nullptr, 'p'},
                                              {"cache", no_argument, nullptr, 'c'},
                                              {nullptr, 0, nullptr, 0}};
    const char *path = nullptr;
    int loops = 1, max_gpus = 0, print = 0, use_cache = 0, c;
    while ((c = getopt_long(argc, argv, "f:l:t:pc", long_opts, nullptr)) != -1) {
        switch (c) {
        case 'f': path = optarg; break;
        case 'l': loops = atoi(optarg); break;
        case 't': max_gpus = atoi(optarg); break;
        case 'p': print = 1; break;
        case 'c': use_cache = 1; break;
        default: usage(argv[0]); return 2;
        }
    }
    if (!path) { usage(argv[0]); return 2; }
    if (loops < 1) loops = 1;
    // One host thread per GPU plans chunk after chunk: megabytes of task lists allocated and freed every few
    // milliseconds.  glibc serves such sizes with mmap / munmap, and in ONE process all threads then queue on the
    // address-space lock and fault every page in again (8 GPUs: 11 % slower per GPU than one GPU alone).  Keep
    // the memory in the per-thread heaps instead.
    mallopt(M_MMAP_THRESHOLD, 1 << 30);
    mallopt(M_TRIM_THRESHOLD, 1 << 30);
    mallopt(M_TOP_PAD, 64 << 20);
    setbuf(stdout, nullptr);

    phmm_parsed *parsed = nullptr;
    char perr[256];
    int rc = -1;
    const double tp0 = now_sec();
    const std::string cache_path = std::string(path) + ".phmmsoa";
    bool from_cache = false;
    if (use_cache) {
        struct stat st_in, st_c;
        if (stat(path, &st_in) == 0 && stat(cache_path.c_str(), &st_c) == 0 && st_c.st_mtime >= st_in.st_mtime) {
            rc = phmm_cache_open(cache_path.c_str(), &parsed, perr, sizeof perr);
            if (rc != PHMM_OK) fprintf(stderr, "phmm: ignoring cache: %s\n", perr);
            from_cache = (rc == PHMM_OK);
        }
    }
    if (rc != PHMM_OK) rc = phmm_parse_file(path, &parsed, perr, sizeof perr);
    if (rc == PHMM_OK && use_cache && !from_cache &&
        phmm_cache_write(phmm_parsed_batches(parsed), cache_path.c_str(), perr, sizeof perr) != PHMM_OK)
        fprintf(stderr, "phmm: cache not written: %s\n", perr);
    const double parse_sec = now_sec() - tp0;
    if (rc != PHMM_OK) {
        fprintf(stderr, "phmm: %s: %s\n", path, perr);
        return rc == PHMM_ERR_IO ? 0 : 1;      // the reference exit(0)s on an unopenable file (:196-199)
    }
    const phmm_flat_batches *fb = phmm_parsed_batches(parsed);

    int ngpu = phmm_device_count();
    if (ngpu <= 0) { fprintf(stderr, "phmm: no CUDA device (libphmm_cuda has no CPU fallback)\n"); return 1; }
    if (max_gpus > 0 && max_gpus < ngpu) ngpu = max_gpus;
    if (fb->n_batches < ngpu) ngpu = fb->n_batches > 0 ? (int)fb->n_batches : 1;

    // ---- a shared queue of chunks of consecutive batches -------------------------------------------
    // The reference hands whole batches to OpenMP threads dynamically (schedule(dynamic), :271-276); here the GPUs
    // pull chunks of consecutive batches from one queue.  A chunk is sized by PREDICTED DEVICE TIME
    // (phmm_batch_costs: register-tile geometry of every read's kernel class + predicted FP64-rescue work; the
    // reference's Batch::num_cells, :166, ignores both), about a twelfth of a GPU's share, so that the GPU that
    // draws the dear batches simply draws fewer chunks.  Chunks are views into the parsed pools (no copy), their
    // results land at the batch's own place in one output array, and every GPU keeps three chunks in flight
    // (phmm_submit_flat / phmm_wait), so planning and uploading one chunk hides under the kernels of the others.
    const int64_t nb = fb->n_batches;
    std::vector<int64_t> cost(nb > 0 ? nb : 1), rstart(nb + 1, 0), hstart(nb + 1, 0), ostart(nb + 1, 0);
    for (int64_t b = 0; b < nb; ++b) {
        rstart[b + 1] = rstart[b] + fb->batch_num_reads[b];
        hstart[b + 1] = hstart[b] + fb->batch_num_haps[b];
        ostart[b + 1] = ostart[b] + (int64_t)fb->batch_num_reads[b] * fb->batch_num_haps[b];
    }
    const char *fast_env = getenv("PHMM_FAST");
    const int fast = fast_env ? atoi(fast_env) : 0;
    if (nb > 0 && phmm_batch_costs(fb, fast, 0, cost.data(), nullptr) != PHMM_OK) { fprintf(stderr, "phmm: malformed batches\n"); return 1; }
    double total_cost = 0;
    for (int64_t b = 0; b < nb; ++b) total_cost += (double)cost[b];
    // ~1.2 cost units per cell on typical input: 5e9 .. 5e10 per chunk = 1.5 .. 15 ms of one B200
    // (PHMM_CLI_CHUNK_COST overrides: tests use it to cut small inputs into many chunks)
    const double chunk_cost = getenv("PHMM_CLI_CHUNK_COST") ? atof(getenv("PHMM_CLI_CHUNK_COST"))
                                                            : std::min(5e10, std::max(5e9, total_cost / (12.0 * ngpu)));
    struct Chunk { int64_t b0, b1; };
    std::vector<Chunk> chunks;
    {
        double acc = 0;
        int64_t b0 = 0;
        for (int64_t b = 0; b < nb; ++b) {
            acc += (double)cost[b];
            if (acc >= chunk_cost || b == nb - 1) { chunks.push_back(Chunk{b0, b + 1}); b0 = b + 1; acc = 0; }
        }
    }
    std::vector<double> out((size_t)std::max<int64_t>(ostart[nb], 1));

    // ---- contexts first (table generation + device binding are `initPairHMM`, untimed, :238) ---
    std::vector<phmm_ctx *> ctx(ngpu, nullptr);
    for (int g = 0; g < ngpu; ++g) {
        if (phmm_create(g, &ctx[g]) != PHMM_OK) { fprintf(stderr, "phmm: GPU %d: %s\n", g, phmm_last_error(nullptr)); return 1; }
        phmm_set_option(ctx[g], PHMM_OPT_FP32_KERNEL, fast);
        phmm_set_option(ctx[g], PHMM_OPT_FP64_KERNEL, fast);
    }

    // ---- timed compute phase (:259-313) ---------------------------------------------------------
    struct PerGpu { phmm_stats stats; int rc = PHMM_OK; int64_t chunks = 0; double busy_s = 0; char err[256] = {0}; };
    std::vector<PerGpu> per(ngpu);
    for (PerGpu &p : per) memset(&p.stats, 0, sizeof p.stats);
    const double t0 = now_sec();
    for (int l = 0; l < loops; ++l) {
        std::atomic<size_t> next{0};
        std::vector<std::thread> th;
        for (int g = 0; g < ngpu; ++g)
            th.emplace_back([&, g]() {
                PerGpu &me = per[g];
                const double tb = now_sec();
                std::vector<int64_t> tickets;
                size_t waited = 0;
                while (me.rc == PHMM_OK) {
                    const size_t k = next.fetch_add(1);
                    if (k >= chunks.size()) break;
                    const Chunk &c = chunks[k];
                    phmm_flat_batches v = *fb;
                    v.n_batches = c.b1 - c.b0;
                    v.batch_num_reads = fb->batch_num_reads + c.b0;
                    v.batch_num_haps = fb->batch_num_haps + c.b0;
                    v.n_reads = rstart[c.b1] - rstart[c.b0];
                    v.read_len = fb->read_len + rstart[c.b0];
                    v.read_off = fb->read_off + rstart[c.b0];
                    v.n_haps = hstart[c.b1] - hstart[c.b0];
                    v.hap_len = fb->hap_len + hstart[c.b0];
                    v.hap_off = fb->hap_off + hstart[c.b0];
                    const phmm_flat_request rq{&v, out.data() + ostart[c.b0]};
                    int64_t t = 0;
                    me.rc = phmm_submit_flat(ctx[g], &rq, 1, &t);
                    if (me.rc != PHMM_OK) break;
                    tickets.push_back(t);
                    ++me.chunks;
                    if (tickets.size() - waited >= 3) me.rc = phmm_wait(ctx[g], tickets[waited++], &me.stats);
                }
                if (me.rc == PHMM_OK && !tickets.empty()) me.rc = phmm_wait(ctx[g], tickets.back(), &me.stats);
                if (me.rc != PHMM_OK) snprintf(me.err, sizeof me.err, "%s", phmm_last_error(ctx[g]));
                me.busy_s += now_sec() - tb;
            });
        for (auto &t : th) t.join();
    }
    const double runtime = now_sec() - t0;

    int64_t cells = 0, pairs = 0, rescued = 0;
    double busy_max = 0, busy_sum = 0;
    for (int g = 0; g < ngpu; ++g) {
        if (per[g].rc != PHMM_OK) { fprintf(stderr, "phmm: GPU %d: %s\n", g, per[g].err); return 1; }
        cells += per[g].stats.n_cells;
        pairs += per[g].stats.n_pairs;
        rescued += per[g].stats.n_rescued;
        busy_max = std::max(busy_max, per[g].busy_s);
        busy_sum += per[g].busy_s;
        phmm_destroy(ctx[g]);
    }
    cells /= loops; pairs /= loops; rescued /= loops;

    if (print) {
        // batch order of the input file, r * num_haps + h inside a batch (:316-321)
        for (int64_t j = 0; j < ostart[nb]; ++j) printf("%lf\n", out[(size_t)j]);
    }
    fprintf(stderr, "phmm: %zu chunks of consecutive batches pulled from one queue; per GPU:", chunks.size());
    for (int g = 0; g < ngpu; ++g) fprintf(stderr, " %lld", (long long)per[g].chunks);
    fprintf(stderr, " (all loops); busiest GPU %.3f s, mean %.3f s\n", busy_max, busy_sum / ngpu);
    fprintf(stderr, "phmm: %lld batches, %lld pairs, %lld cells, %lld FP64-rescued, %d GPU(s), %.1f GCUPS incl. planning and copies; input %s in %.3f s\n",
            (long long)nb, (long long)pairs, (long long)cells, (long long)rescued, ngpu,
            runtime > 0 ? cells * (double)loops / runtime / 1e9 : 0.0, from_cache ? "mapped from cache" : "parsed from text", parse_sec);
    printf("PairHMM completed. Kernel runtime: %.2f sec\n", runtime);
    phmm_parsed_free(parsed);
    return 0;
}
