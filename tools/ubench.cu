// Operand-pattern microbenchmarks for the FP32 pipe of sm_100a (B200): how many cycles an
// FFMA / FFMA2 / FMUL2 costs when its sources are distinct registers (no operand reuse), and how
// ALU / LDS instructions co-issue next to them.  Build: nvcc -gencode arch=compute_100a,code=sm_100a
// -O3 -o tools/ubench tools/ubench.cu ; run on the GPU box.  Development tool, not product code.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define NK 8

template <int WHICH>
__global__ void __launch_bounds__(256) k(float *out, int iters, float seed, const float *in)
{
    __shared__ float2 tab[6 * 256];
    for (int i = threadIdx.x; i < 6 * 256; i += 256) tab[i] = make_float2(1.0f + i * 1e-9f, 1.0f - i * 1e-9f);
    __syncthreads();
    float r = 0.f;
    const float tv = 1e-9f * threadIdx.x;
#define IN(j) (in[(j)] + tv * (float)((j) + 1))
    if (WHICH == 0) {   // FFMA, b and c shared by all instructions (reuse cache)
        float a[NK];
        for (int j = 0; j < NK; ++j) a[j] = seed + j;
        const float b = IN(0), c = IN(1);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) a[j] = __fmaf_rn(a[j], b, c);
        for (int j = 0; j < NK; ++j) r += a[j];
    } else if (WHICH == 1) {   // FFMA, three distinct sources per instruction
        float a[NK], b[NK], c[NK];
        for (int j = 0; j < NK; ++j) { a[j] = seed + j; b[j] = IN(j); c[j] = IN(8 + j); }
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) a[j] = __fmaf_rn(b[j], c[j], a[j]);
        for (int j = 0; j < NK; ++j) r += a[j];
    } else if (WHICH == 2) {   // FFMA2, b and c shared
        float2 a[NK];
        for (int j = 0; j < NK; ++j) a[j] = make_float2(seed + j, seed - j);
        const float2 b = make_float2(IN(0), IN(1)), c = make_float2(IN(2), IN(3));
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) a[j] = __ffma2_rn(a[j], b, c);
        for (int j = 0; j < NK; ++j) r += a[j].x + a[j].y;
    } else if (WHICH == 3) {   // FFMA2, three distinct register pairs per instruction
        float2 a[NK], b[NK], c[NK];
        for (int j = 0; j < NK; ++j) { a[j] = make_float2(seed + j, seed - j); b[j] = make_float2(IN(j), IN(j + 1)); c[j] = make_float2(IN(8 + j), IN(9 + j)); }
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) a[j] = __ffma2_rn(b[j], c[j], a[j]);
        for (int j = 0; j < NK; ++j) r += a[j].x + a[j].y;
    } else if (WHICH == 4) {   // FMUL2, two distinct register pairs
        float2 a[NK], b[NK];
        for (int j = 0; j < NK; ++j) { a[j] = make_float2(seed + j, seed - j); b[j] = make_float2(IN(j), IN(j + 1)); }
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) a[j] = __fmul2_rn(a[j], b[j]);
        for (int j = 0; j < NK; ++j) r += a[j].x + a[j].y;
    } else if (WHICH == 5) {   // FFMA2 d = a*b + a'  (two distinct pairs + another accumulator)
        float2 a[NK], b[NK];
        for (int j = 0; j < NK; ++j) { a[j] = make_float2(seed + j, seed - j); b[j] = make_float2(IN(j), IN(j + 1)); }
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) a[j] = __ffma2_rn(a[j], b[j], b[(j + 3) & 7]);
        for (int j = 0; j < NK; ++j) r += a[j].x + a[j].y;
    } else if (WHICH == 6) {   // the fast cell update, 4 rows, as the forward kernel writes it
        float2 M[4], X[4], Y[4], pMM[4], pG[4], pMX[4], pXX[4], pMY[4], e[4];
        for (int j = 0; j < 4; ++j) {
            M[j] = make_float2(seed + j, seed - j); X[j] = make_float2(seed * 2 + j, seed); Y[j] = make_float2(seed * 3, seed + j);
            pMM[j] = make_float2(IN(j), IN(j + 1)); pG[j] = make_float2(IN(j + 2), IN(j + 3)); pMX[j] = make_float2(IN(j + 4), IN(j + 5));
            pXX[j] = make_float2(IN(j + 6), IN(j + 7)); pMY[j] = make_float2(IN(j + 8), IN(j + 9)); e[j] = make_float2(IN(j + 10), IN(j + 11));
        }
        float2 U2m = M[3], U2x = X[3], U2y = Y[3], U1m = M[2], U1x = X[2];
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                float2 nM[4], nX[4], nY[4];
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float2 dm = p ? M[p - 1] : U2m, dx = p ? X[p - 1] : U2x, dy = p ? Y[p - 1] : U2y;
                    nM[p] = __fmul2_rn(__ffma2_rn(dy, pG[p], __ffma2_rn(dx, pG[p], __fmul2_rn(dm, pMM[p]))), e[p]);
                    nY[p] = __ffma2_rn(M[p], pMY[p], __fmul2_rn(Y[p], pXX[p]));
                }
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float2 um = p ? nM[p - 1] : U1m, ux = p ? nX[p - 1] : U1x;
                    nX[p] = __ffma2_rn(ux, pXX[p], __fmul2_rn(um, pMX[p]));
                }
                U2m = U1m; U2x = U1x; U2y = nY[3]; U1m = nM[3]; U1x = nX[3];
#pragma unroll
                for (int p = 0; p < 4; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; }
            }
        }
        for (int j = 0; j < 4; ++j) r += M[j].x + X[j].y + Y[j].x;
    } else if (WHICH == 7) {   // same cell update with (X+Y) pre-added: one 3-source FFMA2 less per cell
        float2 M[4], S[4], X[4], Y[4], pMM[4], pG[4], pMX[4], pXX[4], pMY[4], e[4];
        for (int j = 0; j < 4; ++j) {
            M[j] = make_float2(seed + j, seed - j); X[j] = make_float2(seed * 2 + j, seed); Y[j] = make_float2(seed * 3, seed + j);
            S[j] = __fadd2_rn(X[j], Y[j]);
            pMM[j] = make_float2(IN(j), IN(j + 1)); pG[j] = make_float2(IN(j + 2), IN(j + 3)); pMX[j] = make_float2(IN(j + 4), IN(j + 5));
            pXX[j] = make_float2(IN(j + 6), IN(j + 7)); pMY[j] = make_float2(IN(j + 8), IN(j + 9)); e[j] = make_float2(IN(j + 10), IN(j + 11));
        }
        float2 U2m = M[3], U2s = S[3], U1m = M[2], U1x = X[2];
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                float2 nM[4], nX[4], nY[4], nS[4];
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float2 dm = p ? M[p - 1] : U2m, ds = p ? S[p - 1] : U2s;
                    nM[p] = __fmul2_rn(__ffma2_rn(dm, pMM[p], __fmul2_rn(ds, pG[p])), e[p]);
                    nY[p] = __ffma2_rn(M[p], pMY[p], __fmul2_rn(Y[p], pXX[p]));
                }
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const float2 um = p ? nM[p - 1] : U1m, ux = p ? nX[p - 1] : U1x;
                    nX[p] = __ffma2_rn(ux, pXX[p], __fmul2_rn(um, pMX[p]));
                    nS[p] = __fadd2_rn(nX[p], nY[p]);
                }
                U2m = U1m; U2s = nS[3]; U1m = nM[3]; U1x = nX[3];
#pragma unroll
                for (int p = 0; p < 4; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; S[p] = nS[p]; }
            }
        }
        for (int j = 0; j < 4; ++j) r += M[j].x + X[j].y + Y[j].x;
    } else if (WHICH == 8) {   // FFMA2 (3 distinct) : LOP3/FSEL 1:1
        float2 a[NK], b[NK], c[NK];
        unsigned m[NK];
        for (int j = 0; j < NK; ++j) { a[j] = make_float2(seed + j, seed - j); b[j] = make_float2(IN(j), IN(j + 1)); c[j] = make_float2(IN(8 + j), IN(9 + j)); m[j] = __float_as_uint(IN(j + 3)); }
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < NK; ++j) {
                    a[j] = __ffma2_rn(b[j], c[j], a[j]);
                    m[j] = (m[j] & m[(j + 1) & 7]) ^ m[(j + 5) & 7];
                }
        for (int j = 0; j < NK; ++j) r += a[j].x + a[j].y + __uint_as_float(m[j]);
    } else if (WHICH == 9) {   // FMUL2 by an emission pair fetched with LDS.64 (one LDS per 4 packed FP instructions)
        float2 a[NK], b[NK], c[NK];
        for (int j = 0; j < NK; ++j) { a[j] = make_float2(seed + j, seed - j); b[j] = make_float2(IN(j), IN(j + 1)); c[j] = make_float2(IN(8 + j), IN(9 + j)); }
        unsigned idx = threadIdx.x;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                idx = (idx + 256u) % (6u * 256u);
#pragma unroll
                for (int j = 0; j < NK; j += 4) {
                    const float2 e = tab[(idx + 256u * j) % (6u * 256u)];
                    a[j] = __ffma2_rn(b[j], c[j], a[j]);
                    a[j + 1] = __ffma2_rn(b[j + 1], c[j + 1], a[j + 1]);
                    a[j + 2] = __ffma2_rn(b[j + 2], c[j + 2], a[j + 2]);
                    a[j + 3] = __fmul2_rn(a[j + 3], e);
                }
            }
        for (int j = 0; j < NK; ++j) r += a[j].x + a[j].y;
    }
    if (r == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int W>
static void run(const char *name, double fp_lane_ops_per_iter, float *out, const float *in, int blocks_per_sm)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * blocks_per_sm, iters = 4096;
    k<W><<<grid, 256>>>(out, 64, 1.5f, in);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k<W><<<grid, 256>>>(out, iters, 1.5f, in);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    const double ops = fp_lane_ops_per_iter * iters * 256.0 * grid / (best * 1e-3);
    printf("UB %-34s blocks/SM %d  %8.3f ms  %7.3f T FP-lane-ops/s  (%s)\n", name, blocks_per_sm, best, ops / 1e12, cudaGetErrorString(err));
}

int main()
{
    float *out, *in;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&in, 256);
    float h[64];
    for (int j = 0; j < 64; ++j) h[j] = 1.0f + (j % 2 ? 1e-8f : -1e-8f) * j;
    h[0] = 1.0000001f; h[1] = 1e-9f;
    cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
    for (int bps = 2; bps <= 8; bps *= 2) {
        run<0>("FFMA reuse", 32, out, in, bps);
        run<1>("FFMA 3 distinct", 32, out, in, bps);
        run<2>("FFMA2 reuse", 64, out, in, bps);
        run<3>("FFMA2 3 distinct", 64, out, in, bps);
        run<4>("FMUL2 2 distinct", 64, out, in, bps);
        run<5>("FFMA2 2 distinct + 1", 64, out, in, bps);
        run<6>("cell fast 4 rows (8 ops/cell)", 2 * 4 * 8 * 2, out, in, bps);
        run<7>("cell fast X+Y pre-add (9 ops/cell)", 2 * 4 * 9 * 2, out, in, bps);
        run<8>("FFMA2 3 distinct + LOP3 1:1", 64, out, in, bps);
        run<9>("3 FFMA2 + FMUL2(LDS.64 e)", 64, out, in, bps);
    }
    return 0;
}
