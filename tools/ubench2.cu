// Prototype of the v2 inner loop (two reads per thread in f32x2 halves, emission from a shared-memory
// table, four columns per iteration) on synthetic data: measures the achievable cell rate before the
// full kernel is written.  Development tool, not product code.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

#define FULL 0xffffffffu
constexpr int WARPS = 4;
constexpr int RING = 2048;

template <bool FAST> struct O2 {
    static __device__ __forceinline__ float2 mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
    static __device__ __forceinline__ float2 add(float2 a, float2 b)
    {
        if (FAST) return __fadd2_rn(a, b);
        return make_float2(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y));
    }
    // a * b + c * d
    static __device__ __forceinline__ float2 dot2(float2 a, float2 b, float2 c, float2 d)
    {
        if (FAST) return __ffma2_rn(a, b, __fmul2_rn(c, d));
        return add(__fmul2_rn(a, b), __fmul2_rn(c, d));
    }
};

__device__ __forceinline__ float2 shfl_up2(float2 v)
{
    return make_float2(__shfl_up_sync(FULL, v.x, 1), __shfl_up_sync(FULL, v.y, 1));
}

template <int KH, bool FAST, int VARIANT>
__global__ void __launch_bounds__(WARPS * 32) proto(float *out, const float *in, int ncols)
{
    typedef O2<FAST> O;
    extern __shared__ __align__(16) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // per warp: e-table [KH][6] rows of 32 float2 (256 B), then the haplotype ring
    uint8_t *etab = smem + warp * (KH * 6 * 256 + RING);
    uint8_t *ring = etab + KH * 6 * 256;
    const float tv = 1e-9f * threadIdx.x;
    float2 M[KH], X[KH], Y[KH], pMM[KH], pG[KH], pMX[KH], pXX[KH], pMY[KH];
#pragma unroll
    for (int p = 0; p < KH; ++p) {
        M[p] = make_float2(0.f, 0.f); X[p] = M[p]; Y[p] = make_float2(1e30f, 1e30f);
        pMM[p] = make_float2(in[p] + tv, in[p + 1] - tv); pG[p] = make_float2(in[p + 2] * 0.1f + tv, in[p + 3] * 0.1f);
        pMX[p] = make_float2(1e-3f + tv, 2e-3f); pXX[p] = make_float2(0.1f + tv, 0.1f - tv); pMY[p] = make_float2(1e-3f - tv, 1e-3f);
        for (int code = 0; code < 6; ++code)
            *reinterpret_cast<float2 *>(etab + (p * 6 + code) * 256 + lane * 8) =
                code == 0 ? make_float2(0.f, 0.f) : ((code + p + lane) & 3) == 0 ? make_float2(0.999f, 0.99f) : make_float2(3e-4f, 3e-3f);
    }
    for (int i = lane; i < RING / 4; i += 32) {
        uint32_t w = 0;
        for (int b = 0; b < 4; ++b) {
            const int col = i * 4 + b;
            const uint32_t code = (col % 304) >= 300 ? 0u : 1u + ((col * 2654435761u) >> 30);
            w |= code << (8 * b);
        }
        reinterpret_cast<uint32_t *>(ring)[i] = w;
    }
    __syncwarp();
    float2 U1m = M[0], U1x = M[0], U1y = M[0], U2m, U2x, U2y;
    float2 sumM = M[0], sumX = M[0];
    float acc = 0.f;
    const uint32_t sh = 8 * ((32 - lane) & 3);
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(etab) + lane * 8;
    uint32_t wprev = 0;
    int widx = (RING / 4 - ((lane + 3) >> 2)) & (RING / 4 - 1);
    for (int s0 = 0; s0 < ncols; s0 += 4) {
        const uint32_t wnext = reinterpret_cast<const uint32_t *>(ring)[widx];
        widx = (widx + 1) & (RING / 4 - 1);
        const uint32_t codes = __funnelshift_r(wprev, wnext, sh);
        wprev = wnext;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // table row of this column's code: code << 8 lands in byte 1 of the address offset
            const uint32_t eoff = __byte_perm(codes, 0, 0x4404 | (k << 4));   // (code_k << 8)
            U2m = U1m; U2x = U1x; U2y = U1y;
            U1m = shfl_up2(M[KH - 1]); U1x = shfl_up2(X[KH - 1]); U1y = shfl_up2(Y[KH - 1]);
            float2 e[KH];
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                float2 v;
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(ebase + eoff + p * 6 * 256));
                e[p] = v;
            }
            float2 nM[KH], nX[KH], nY[KH];
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const float2 dm = p ? M[p - 1] : U2m, dx = p ? X[p - 1] : U2x, dy = p ? Y[p - 1] : U2y;
                if (VARIANT == 0) {
                    if (FAST) nM[p] = O::mul(__ffma2_rn(dy, pG[p], __ffma2_rn(dx, pG[p], O::mul(dm, pMM[p]))), e[p]);
                    else nM[p] = O::mul(O::add(O::add(O::mul(dm, pMM[p]), O::mul(dx, pG[p])), O::mul(dy, pG[p])), e[p]);
                } else {
                    nM[p] = O::mul(O::dot2(dm, pMM[p], O::add(dx, dy), pG[p]), e[p]);
                }
                nY[p] = O::dot2(M[p], pMY[p], Y[p], pXX[p]);
            }
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const float2 um = p ? nM[p - 1] : U1m, ux = p ? nX[p - 1] : U1x;
                nX[p] = O::dot2(ux, pXX[p], um, pMX[p]);
            }
#pragma unroll
            for (int p = 0; p < KH; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; }
            sumM = O::add(sumM, M[KH - 1]);
            sumX = O::add(sumX, X[KH - 1]);
        }
        // separator handling once per four columns
        const bool z = (codes >> 24) == 0;
        if (__any_sync(FULL, z)) {
            if (z) {
                if (lane == 31) { acc += sumM.x + sumX.x + sumM.y + sumX.y; sumM = make_float2(0.f, 0.f); sumX = sumM; }
#pragma unroll
                for (int p = 0; p < KH; ++p) Y[p] = make_float2(0.f, 0.f);
            }
        }
    }
    if (acc == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = acc + M[0].x;
}

template <int KH, bool FAST, int VARIANT>
static void run(const char *name, float *out, const float *in, int bps)
{
    const int smem = WARPS * (KH * 6 * 256 + RING);
    cudaFuncSetAttribute(proto<KH, FAST, VARIANT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, proto<KH, FAST, VARIANT>, WARPS * 32, smem);
    if (bps > occ) return;
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, proto<KH, FAST, VARIANT>);
    const int grid = 148 * bps, ncols = 1 << 16;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    proto<KH, FAST, VARIANT><<<grid, WARPS * 32, smem>>>(out, in, 1024);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        proto<KH, FAST, VARIANT><<<grid, WARPS * 32, smem>>>(out, in, ncols);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double cells = (double)grid * WARPS * 32 * 2.0 * KH * ncols;
    printf("P2 %-26s KH %d regs %3d occ %d blocks/SM %d  %8.3f ms  %8.1f Gcells/s (%s)\n", name, KH, fa.numRegs, occ, bps, best,
           cells / (best * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    float *out, *in;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&in, 256);
    float h[64];
    for (int j = 0; j < 64; ++j) h[j] = 0.9f + 1e-3f * j;
    cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
    for (int bps = 2; bps <= 4; ++bps) {
        run<4, true, 0>("fast", out, in, bps);
        run<4, true, 1>("fast x+y", out, in, bps);
        run<4, false, 0>("exact", out, in, bps);
        run<6, true, 0>("fast", out, in, bps);
        run<6, true, 1>("fast x+y", out, in, bps);
        run<6, false, 0>("exact", out, in, bps);
        run<8, true, 0>("fast", out, in, bps);
        run<8, true, 1>("fast x+y", out, in, bps);
        run<8, false, 0>("exact", out, in, bps);
    }
    return 0;
}
