// FP32 / FP64 pipe microbenchmarks of libphmm_cuda (roofline denominators).
#include <cuda_runtime.h>
#include <stdint.h>

namespace phmm {

// ------------------------------------------------------------------------------------------
// FP32 / FP64 pipe microbenchmarks: the measured roofline denominators.
// Each thread runs `iters` rounds of 8 independent dependency chains.
// ------------------------------------------------------------------------------------------
template <int WHICH>
__global__ void __launch_bounds__(256) pipe_bench_kernel(float *out, int iters, float seed)
{
    float r = 0.f;
    if (WHICH == 0) {            // FFMA
        float a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = seed + k;
        const float b = 1.0000001f, c = 1e-9f;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = __fmaf_rn(a[k], b, c);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) r += a[k];
    } else if (WHICH == 1) {     // FFMA2
        float2 a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = make_float2(seed + k, seed - k);
        const float2 b = make_float2(1.0000001f, 0.9999999f), c = make_float2(1e-9f, 2e-9f);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = __ffma2_rn(a[k], b, c);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) r += a[k].x + a[k].y;
    } else if (WHICH == 2) {     // FMUL + FADD, unfused (the bit-exact scalar mix: 2 mul : 1 add)
        float a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = seed + k;
        const float b = 1.0000001f, c = 0.9999999f;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = __fadd_rn(__fmul_rn(a[k], b), __fmul_rn(a[(k + 1) & 7], c));
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) r += a[k];
    } else if (WHICH == 3) {     // FMUL2 + scalar FADD (the bit-exact packed mix)
        float2 a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = make_float2(seed + k, seed - k);
        const float2 b = make_float2(1.0000001f, 0.9999999f), c = make_float2(0.9999999f, 1.0000001f);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float2 m1 = __fmul2_rn(a[k], b), m2 = __fmul2_rn(a[(k + 1) & 7], c);
                    a[k] = make_float2(__fadd_rn(m1.x, m2.x), __fadd_rn(m1.y, m2.y));
                }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) r += a[k].x + a[k].y;
    } else {                     // DFMA
        double a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = seed + k;
        const double b = 1.0000001, c = 1e-9;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = __fma_rn(a[k], b, c);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) r += (float)a[k];
    }
    if (r == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// lane-operations per thread per iteration for each WHICH (an FMA, a MUL and an ADD each count 1;
// packed instructions count 2 per thread).
cudaError_t launch_pipe_bench(int which, float *out, int grid, int iters, cudaStream_t st, double *ops_per_thread)
{
    switch (which) {
    case 0: pipe_bench_kernel<0><<<grid, 256, 0, st>>>(out, iters, 1.5f); *ops_per_thread = 32.0 * iters; break;
    case 1: pipe_bench_kernel<1><<<grid, 256, 0, st>>>(out, iters, 1.5f); *ops_per_thread = 64.0 * iters; break;
    case 2: pipe_bench_kernel<2><<<grid, 256, 0, st>>>(out, iters, 1.5f); *ops_per_thread = 96.0 * iters; break;
    case 3: pipe_bench_kernel<3><<<grid, 256, 0, st>>>(out, iters, 1.5f); *ops_per_thread = 192.0 * iters; break;
    case 4: pipe_bench_kernel<4><<<grid, 256, 0, st>>>(out, iters, 1.5f); *ops_per_thread = 32.0 * iters; break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace phmm
