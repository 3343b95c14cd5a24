// FP64 issue model on B200: DFMA latency, DFMA throughput per scheduler at 1 / 2 / 3 warps, and what an LDS, SHFL
// or integer instruction costs when it is mixed into a DFMA stream (does it ride in the second cycle of the
// half-rate FP64 issue or does it take a slot of its own?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench6 tools/ubench6.cu && tools/ubench6
#include <cuda_runtime.h>
#include <cstdio>

template <int MODE, int NMIX>
__global__ void k(double *out, int iters, double a, double b, int one)
{
    __shared__ double sm[1024];
    sm[threadIdx.x & 1023] = a + threadIdx.x;
    __syncthreads();
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    int acc = threadIdx.x, acc2 = one;
    double l = 0;
    unsigned addr = (threadIdx.x & 31) * 8;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (MODE == 0) {                         // one dependent chain: latency
                x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b);
                x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b);
            } else {                                 // eight independent chains
                x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
                x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
            }
#pragma unroll
            for (int m = 0; m < NMIX; ++m) {
                if (MODE == 2) { acc = acc * one + acc2; acc2 ^= acc; }                         // 2 integer ops
                if (MODE == 3) { acc = __shfl_up_sync(0xffffffffu, acc, 1); acc2 = __shfl_up_sync(0xffffffffu, acc2, 1); }
                if (MODE == 4) { l += sm[((addr + 8 * (u * NMIX + m) * one) >> 3) & 1023]; }       // LDS.64 (+ DADD)
                if (MODE == 5) {                                                                // LDS.128
                    double2 v = *reinterpret_cast<double2 *>(&sm[(((addr * 2) + 16 * (u * NMIX + m) * one) >> 3) & 1022]);
                    l += v.x; x7 += v.y;
                }
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + acc + acc2 + l;
}

template <int MODE, int NMIX>
void run(const char *name, double *d, int warps_per_sm)
{
    const int iters = 2000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE, NMIX><<<148, warps_per_sm * 32>>>(d, 10, 1.0000001, 1e-9, 1);
    cudaEventRecord(e0);
    k<MODE, NMIX><<<148, warps_per_sm * 32>>>(d, iters, 1.0000001, 1e-9, 1);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double cyc = ms * 1e-3 * 1.965e9;
    const double dfma_per_warp = iters * 64.0;
    const double wps = warps_per_sm / 4.0;     // warps per scheduler
    printf("%-34s warps/SMSP %.0f  cycles per DFMA per scheduler %.3f  (per warp %.2f)%s\n", name, wps,
           cyc / (dfma_per_warp * wps), cyc / dfma_per_warp, cudaGetLastError() == cudaSuccess ? "" : "  ERROR");
}

int main()
{
    double *d;
    cudaMalloc(&d, 148 * 1024 * sizeof(double));
    for (int w : {4, 8, 12, 16}) {
        if (w == 4) run<0, 0>("dependent chain (latency)", d, w);
        run<1, 0>("8 chains, DFMA only", d, w);
        run<2, 1>("8 chains + 2 int ops per 8 DFMA", d, w);
        run<2, 2>("8 chains + 4 int ops per 8 DFMA", d, w);
        run<2, 4>("8 chains + 8 int ops per 8 DFMA", d, w);
        run<3, 1>("8 chains + 2 SHFL per 8 DFMA", d, w);
        run<3, 2>("8 chains + 4 SHFL per 8 DFMA", d, w);
        run<4, 1>("8 chains + 1 LDS.64 per 8 DFMA", d, w);
        run<4, 2>("8 chains + 2 LDS.64 per 8 DFMA", d, w);
        run<5, 1>("8 chains + 1 LDS.128 per 8 DFMA", d, w);
        printf("\n");
    }
    return 0;
}
