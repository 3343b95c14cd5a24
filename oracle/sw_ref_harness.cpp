/*
 * TEST INFRASTRUCTURE -- extern "C" shim around the UNMODIFIED reference Smith-Waterman objects
 * (tools/GKL/src/main/native/smithwaterman/avx2_impl.cc, avx512_impl.cc -> PairWiseSW.h), built in
 * place by oracle/build_ref.sh into oracle/_ref/libsw_ref.so.  The loop is what a caller of
 * IntelSmithWatermanCSource.cpp:45-57 / IntelSmithWaterman.cc:52-75 does: one runSWOnePairBT per
 * pair through the cpuid-selected function pointer (:22-40), here under OpenMP so the CPU baseline
 * can use every host core.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "avx2_impl.h"
#include "avx512_impl.h"

extern "C" {

/* isa: 0 = as the reference dispatches (AVX-512 when the CPU has F/DQ/BW/VL, else AVX2), 1 = AVX2,
 * 2 = AVX-512.  Returns the ISA used (1 or 2).  score receives the int16 the reference hands back
 * (PairWiseSW.h:441 stores the int32 score through an int16_t*). */
int sw_ref_run(int64_t n, const uint8_t *ref_pool, const int64_t *ref_off, const int32_t *ref_len,
               const uint8_t *alt_pool, const int64_t *alt_off, const int32_t *alt_len, int match, int mismatch,
               int open, int extend, int strategy, int32_t *score, int32_t *offset, char *cigar, int stride,
               int32_t *cigar_len, int nthreads, int isa)
{
    const bool has512 = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512dq") &&
                        __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl");
    if (isa == 0) isa = has512 ? 2 : 1;
    if (isa == 2 && !has512) return -1;
    auto fn = (isa == 2) ? runSWOnePairBT_fp_avx512 : runSWOnePairBT_fp_avx2;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads > 0 ? nthreads : 1)
    for (int64_t p = 0; p < n; ++p) {
        std::vector<char> buf((size_t)4 * (ref_len[p] + alt_len[p]) + 64, 0);
        int16_t count = 0, sc = 0;
        offset[p] = fn(match, mismatch, open, extend, const_cast<uint8_t *>(ref_pool + ref_off[p]),
                       const_cast<uint8_t *>(alt_pool + alt_off[p]), ref_len[p], alt_len[p], (int8_t)strategy,
                       buf.data(), &count, &sc);
        score[p] = sc;
        const int len = (int)strlen(buf.data());
        if (len + 1 > stride) {
            cigar_len[p] = -1 - len;
        } else {
            memcpy(cigar + p * (int64_t)stride, buf.data(), (size_t)len + 1);
            cigar_len[p] = len;
        }
    }
    return isa;
}

} /* extern "C" */
