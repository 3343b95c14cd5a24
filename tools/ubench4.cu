// Operand-delivery microbenchmarks for Blackwell's packed FP32 (FFMA2 / FMUL2): how many cycles an
// instruction costs as a function of how many DISTINCT register operands it reads and whether an
// operand repeats in the same slot of the next instruction (the operand reuse cache).
// 16 independent float2 accumulators per thread, so no dependency stalls.  Development tool.
#include <cuda_runtime.h>
#include <cstdio>

constexpr int NA = 16;

// MODE 0: acc[i] = fma(x[i], c[i], acc[i])      three distinct operands, constant differs every instruction
// MODE 1: acc[i] = fma(x[i], c[i / 4], acc[i])  the scalar constant repeats in slot B for 4 instructions
// MODE 2: acc[i] = fma(x[i], c[0], acc[i])      same constant always
// MODE 3: acc[i] = fma(acc[i], c[i], acc[i])    two distinct registers (one read twice)
// MODE 4: acc[i] = acc[i] * c[i]                FMUL2, two operands
// MODE 5: acc[i] = fma(x[i], C2[i], acc[i])     constant is a register PAIR (two reads packing), differs every time
// MODE 6: acc[i] = fma(x[i / 4], c[i], acc[i])  the f32x2 multiplicand repeats in slot A for 4 instructions
// MODE 7: acc[i] = fma(x[i], c[i / 4], acc[i]) ; acc2[i] = acc2[i] * c[i / 4]   mixed FFMA2 / FMUL2 sharing slot B
template <int MODE>
__global__ void __launch_bounds__(128) k(float *out, const float *in, int iters)
{
    float2 acc[NA], x[NA], c2[NA];
    float c[NA];
    for (int i = 0; i < NA; ++i) {
        acc[i] = make_float2(in[i] + threadIdx.x, in[i + 1]);
        x[i] = make_float2(in[i + 2] * 1e-3f, in[i + 3] * 1e-3f + threadIdx.x * 1e-9f);
        c[i] = in[i + 4] + threadIdx.x * 1e-7f;
        c2[i] = make_float2(in[i + 5], in[i + 6] + threadIdx.x * 1e-7f);
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
            for (int i = 0; i < NA; ++i) {
                if (MODE == 0) acc[i] = __ffma2_rn(x[i], make_float2(c[i], c[i]), acc[i]);
                if (MODE == 1) acc[i] = __ffma2_rn(x[i], make_float2(c[i / 4], c[i / 4]), acc[i]);
                if (MODE == 2) acc[i] = __ffma2_rn(x[i], make_float2(c[0], c[0]), acc[i]);
                if (MODE == 3) acc[i] = __ffma2_rn(acc[i], make_float2(c[i], c[i]), acc[i]);
                if (MODE == 4) acc[i] = __fmul2_rn(acc[i], make_float2(c[i], c[i]));
                if (MODE == 5) acc[i] = __ffma2_rn(x[i], c2[i], acc[i]);
                if (MODE == 6) acc[i] = __ffma2_rn(x[i / 4], make_float2(c[i], c[i]), acc[i]);
                if (MODE == 7) {
                    if (i & 1) acc[i] = __fmul2_rn(acc[i], make_float2(c[i / 4], c[i / 4]));
                    else acc[i] = __ffma2_rn(x[i], make_float2(c[i / 4], c[i / 4]), acc[i]);
                }
            }
        }
    }
    float s = 0;
    for (int i = 0; i < NA; ++i) s += acc[i].x + acc[i].y;
    if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static void run(const char *name, float *out, const float *in)
{
    for (int bps = 2; bps <= 8; bps *= 2) {       // 8, 16, 32 warps per SM = 2, 4, 8 per scheduler
        const int grid = 148 * bps, iters = 4096;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        k<MODE><<<grid, 128>>>(out, in, 64);
        float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            k<MODE><<<grid, 128>>>(out, in, iters);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double instr = (double)grid * 4 /*warps*/ * iters * 4 * NA;       // warp instructions
        const double cyc = 148.0 * 4 * 1.965e9 * best * 1e-3;                    // scheduler cycles at 1965 MHz
        printf("UB4 %-44s warps/SMSP %d  %7.3f ms  %5.2f cycles per warp instruction (%s)\n", name, bps, best, cyc / instr,
               cudaGetErrorString(cudaGetLastError()));
    }
}

int main()
{
    float *out, *in;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&in, 256);
    float h[64];
    for (int j = 0; j < 64; ++j) h[j] = 0.9f + 1e-3f * j;
    cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
    run<0>("FFMA2 x[i], c[i].F32, acc[i]", out, in);
    run<1>("FFMA2 x[i], c[i/4].F32, acc[i]  (B repeats)", out, in);
    run<2>("FFMA2 x[i], c[0].F32, acc[i]    (B constant)", out, in);
    run<3>("FFMA2 acc[i], c[i].F32, acc[i]  (A == C)", out, in);
    run<4>("FMUL2 acc[i], c[i].F32", out, in);
    run<5>("FFMA2 x[i], c2[i] pair, acc[i]", out, in);
    run<6>("FFMA2 x[i/4], c[i].F32, acc[i]  (A repeats)", out, in);
    run<7>("FFMA2/FMUL2 alternating, B repeats", out, in);
    return 0;
}
