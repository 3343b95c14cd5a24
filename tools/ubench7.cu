// DFMA / DMUL cost against the number of distinct register operands (B200): is the column-at-a-time FP64 kernel
// bound by the register-file read port (3 distinct 64-bit operands = 6 words) rather than by the FP64 pipe?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench7 tools/ubench7.cu && tools/ubench7
#include <cuda_runtime.h>
#include <cstdio>

// MODE 0: x_j = fma(x_j, a, b)        one distinct operand per DFMA (a, b from the reuse cache)
// MODE 1: x_j = fma(x_j, a_j, b)      two distinct
// MODE 2: x_j = fma(x_j, a_j, y_j)    three distinct
// MODE 3: x_j = x_j * a_j             DMUL, two distinct
// MODE 4: pairs: x_j = fma(x_j, a_j, y_j); x_j' = fma(x_j', a_j, y_j')   (second of each pair reuses a_j)
// MODE 5: quads sharing a_j
template <int MODE>
__global__ void k(double *out, int iters, double a, double b)
{
    double x[8], y[8], c[8], z[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { x[j] = threadIdx.x + j; y[j] = b * (j + 1 + threadIdx.x); c[j] = a + 1e-9 * (j + threadIdx.x); z[j] = x[j] + 0.5; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (MODE == 0) x[j] = fma(x[j], a, b);
                if (MODE == 1) x[j] = fma(x[j], c[j], b);
                if (MODE == 2) x[j] = fma(x[j], c[j], y[j]);
                if (MODE == 3) x[j] = x[j] * c[j];
                if (MODE == 4) { if (j & 1) x[j] = fma(x[j], c[j / 2], y[j]); else x[j] = fma(x[j], c[j / 2], y[j]); }
                if (MODE == 5) x[j] = fma(x[j], c[j / 4], y[j]);
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x[j] + z[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double *d, int warps_per_sm)
{
    const int iters = 2000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148, warps_per_sm * 32>>>(d, 10, 1.0000001, 1e-9);
    cudaEventRecord(e0);
    k<MODE><<<148, warps_per_sm * 32>>>(d, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double cyc = ms * 1e-3 * 1.965e9, n = iters * 64.0, wps = warps_per_sm / 4.0;
    printf("%-52s warps/SMSP %.0f  cycles per instruction per scheduler %.3f\n", name, wps, cyc / (n * wps));
}

int main()
{
    double *d;
    cudaMalloc(&d, 148 * 1024 * sizeof(double));
    for (int w : {8, 12}) {
        run<0>("DFMA x, a, b      (a, b repeat: reuse cache)", d, w);
        run<1>("DFMA x, a_j, b    (two distinct operands)", d, w);
        run<2>("DFMA x, a_j, y_j  (three distinct operands)", d, w);
        run<3>("DMUL x, a_j       (two distinct operands)", d, w);
        run<4>("DFMA pairs sharing a_j", d, w);
        run<5>("DFMA quads sharing a_j", d, w);
        printf("\n");
    }
    return 0;
}
