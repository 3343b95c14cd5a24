// Marginal cost of LDS.32 / LDS.128 / scalar FMUL / FMUL2 / SHFL when mixed into a stream of FFMA2
// (pair, scalar, pair operands).  16 FFMA2 per repetition are the base load.  Development tool.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
constexpr int NA = 16;
// MODE 0: base                         1: + 16 LDS.32 (feeding the FFMA2 scalar operand)
// MODE 2: + 4 LDS.128 (same data)      3: + 16 scalar FMUL on register halves
// MODE 4: + 8 FMUL2 (pair, scalar)     5: + 6 SHFL
// MODE 6: + 16 IADD (ALU pipe)         7: + 16 LDS.32 + 16 scalar FMUL + 6 SHFL
template <int MODE>
__global__ void __launch_bounds__(128) k(float *out, const float *in, int iters)
{
    __shared__ __align__(16) float tab[128 * 16 + 64 + 2048];
    for (int i = threadIdx.x; i < 128 * 16 + 64 + 2048; i += 128) tab[i] = 0.9f + 1e-5f * i;
    __syncthreads();
    float2 acc[NA], x[NA], m[8];
    float c[NA];
    int ia[NA];
    for (int i = 0; i < NA; ++i) {
        acc[i] = make_float2(in[i] + threadIdx.x, in[i + 1]);
        x[i] = make_float2(in[i + 2] * 1e-3f, in[i + 3] * 1e-3f + threadIdx.x * 1e-9f);
        c[i] = in[i + 4] + threadIdx.x * 1e-7f;
        ia[i] = threadIdx.x + i;
    }
    for (int i = 0; i < 8; ++i) m[i] = make_float2(1.0f + in[i] * 1e-6f, 1.0f);
    uint32_t base = (uint32_t)__cvta_generic_to_shared(tab) + threadIdx.x * 16;
    float2 sh = make_float2(in[9], in[10]);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
            const uint32_t a = base + (((it + rep) & 3) << 9);     // varies, so the loads stay in the loop
            if (MODE == 1 || MODE == 7) {
#pragma unroll
                for (int i = 0; i < NA; ++i)
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(c[i]) : "r"(a + (i & 3) * 4 + (i >> 2) * 2048));
            }
            if (MODE == 2) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(c[4 * q]), "=f"(c[4 * q + 1]), "=f"(c[4 * q + 2]), "=f"(c[4 * q + 3]) : "r"(a + q * 2048));
            }
#pragma unroll
            for (int i = 0; i < NA; ++i) {
                acc[i] = __ffma2_rn(x[i], make_float2(c[i], c[i]), acc[i]);
                if (MODE == 3 || MODE == 7) { if (i & 1) m[i / 2].y = __fmul_rn(m[i / 2].y, c[i]); else m[i / 2].x = __fmul_rn(m[i / 2].x, c[i]); }
                if (MODE == 4 && (i & 1)) m[i / 2] = __fmul2_rn(m[i / 2], make_float2(c[i], c[i]));
                if ((MODE == 5 || MODE == 7) && i < 6) {
                    if (i & 1) sh.y = __shfl_up_sync(0xffffffffu, sh.y + 0.0f, 1); else sh.x = __shfl_up_sync(0xffffffffu, sh.x, 1);
                }
                if (MODE == 6) ia[i] = ia[i] + (ia[(i + 1) & 15] ^ 0x55);
            }
        }
    }
    float s = sh.x + sh.y;
    for (int i = 0; i < NA; ++i) s += acc[i].x + acc[i].y + c[i] + ia[i];
    for (int i = 0; i < 8; ++i) s += m[i].x + m[i].y;
    if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
static void run(const char *name, float *out, const float *in)
{
    for (int bps = 2; bps <= 4; bps *= 2) {
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
        const double reps = (double)grid * 4 * iters * 4;                  // warp-level repetitions (16 FFMA2 each)
        const double cyc = 148.0 * 4 * 1.965e9 * best * 1e-3;
        printf("UB5 %-40s warps/SMSP %d  %7.3f ms  %6.2f cycles per repetition (%s)\n", name, bps, best, cyc / reps,
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
    run<0>("16 FFMA2", out, in);
    run<1>("16 FFMA2 + 16 LDS.32", out, in);
    run<2>("16 FFMA2 + 4 LDS.128", out, in);
    run<3>("16 FFMA2 + 16 FMUL", out, in);
    run<4>("16 FFMA2 + 8 FMUL2", out, in);
    run<5>("16 FFMA2 + 6 SHFL", out, in);
    run<6>("16 FFMA2 + 16 IADD/LOP", out, in);
    run<7>("16 FFMA2 + 16 LDS.32 + 16 FMUL + 6 SHFL", out, in);
    return 0;
}
