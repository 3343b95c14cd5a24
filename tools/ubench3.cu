// Prototype of the scaled-recurrence inner loop (kernel v3) in two packings, on synthetic data:
//   SC = 0  two READS per thread in the f32x2 halves: every per-row constant is a register pair
//   SC = 1  one read against two HAPLOTYPES per thread: the per-row constants are the same in both
//           halves, so ptxas can use FFMA2's scalar-broadcast operand form (Rn.F32) -- one 32-bit
//           register read instead of a pair -- and the emission factors come from two 32-bit LDS
// Measures the achievable cell rate of each before the full kernel is written.  Development tool.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

#define FULL 0xffffffffu
constexpr int WARPS = 4;
constexpr int RING = 2048;

__device__ __forceinline__ float2 shfl_up2(float2 v)
{
    return make_float2(__shfl_up_sync(FULL, v.x, 1), __shfl_up_sync(FULL, v.y, 1));
}

template <bool SC> struct CT;
template <> struct CT<false> {
    typedef float2 C;
    static __device__ __forceinline__ float2 v(float2 c) { return c; }
    static __device__ __forceinline__ float2 mk(float a, float b) { return make_float2(a, b); }
};
template <> struct CT<true> {
    typedef float C;
    static __device__ __forceinline__ float2 v(float c) { return make_float2(c, c); }
    static __device__ __forceinline__ float mk(float a, float) { return a; }
};

template <int KH, bool SC>
__global__ void __launch_bounds__(WARPS * 32) proto(float *out, const float *in, int ncols)
{
    typedef CT<SC> K;
    typedef typename K::C C;
    constexpr int NTOP = (KH + 3) / 4, GRP = (KH + NTOP - 1) / NTOP;
    constexpr int ESZ = SC ? 128 : 256;          // bytes per (row, code) table line
    extern __shared__ __align__(16) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t *etab = smem + warp * (KH * 6 * ESZ + 2 * RING);
    uint8_t *ring = etab + KH * 6 * ESZ;
    const float tv = 1e-9f * threadIdx.x;
    float2 M[KH], X[KH], Y[KH];
    C pMM[KH], pG[KH], pMX[KH], pXX[KH], pMY[KH], K2[NTOP];
#pragma unroll
    for (int p = 0; p < KH; ++p) {
        M[p] = make_float2(0.f, 0.f); X[p] = M[p]; Y[p] = make_float2(1e30f, 1e30f);
        pMM[p] = K::mk(in[p] + tv, in[p + 1] - tv); pG[p] = K::mk(in[p + 2] * 0.1f + tv, in[p + 3] * 0.1f);
        pMX[p] = K::mk(1e-3f + tv, 2e-3f); pXX[p] = K::mk(0.1f + tv, 0.1f - tv); pMY[p] = K::mk(1e-3f - tv, 1e-3f);
        if (p % GRP == 0) K2[p / GRP] = K::mk(0.1f + 2 * tv, 0.1f);
        for (int code = 0; code < 6; ++code) {
            const float2 ev = code == 0 ? make_float2(0.f, 0.f)
                                        : ((code + p + lane) & 3) == 0 ? make_float2(0.999f, 0.99f) : make_float2(3e-4f, 3e-3f);
            if (SC) *reinterpret_cast<float *>(etab + (p * 6 + code) * ESZ + lane * 4) = ev.x;
            else *reinterpret_cast<float2 *>(etab + (p * 6 + code) * ESZ + lane * 8) = ev;
        }
    }
    for (int i = lane; i < 2 * RING / 4; i += 32) {
        uint32_t w = 0;
        for (int b = 0; b < 4; ++b) {
            const int col = i * 4 + b;
            const uint32_t code = (col % 304) >= 300 ? 0u : 1u + ((col * 2654435761u) >> 30);
            w |= code << (8 * b);
        }
        reinterpret_cast<uint32_t *>(ring)[i] = w;
    }
    __syncwarp();
    float2 U1m = M[0], U1x = M[0], U1z = M[0], U2m, U2z;
    float2 sumM = M[0], sumX = M[0];
    float acc = 0.f;
    const uint32_t sh = 8 * ((32 - lane) & 3);
    const uint32_t ebase = (uint32_t)__cvta_generic_to_shared(etab) + lane * (SC ? 4 : 8);
    uint32_t wprev = 0, wprev2 = 0;
    int widx = (RING / 4 - ((lane + 3) >> 2)) & (RING / 4 - 1);
    for (int s0 = 0; s0 < ncols; s0 += 4) {
        const uint32_t wnext = reinterpret_cast<const uint32_t *>(ring)[widx];
        const uint32_t codes = __funnelshift_r(wprev, wnext, sh);
        wprev = wnext;
        uint32_t codes2 = 0;
        if (SC) {
            const uint32_t wnext2 = reinterpret_cast<const uint32_t *>(ring)[widx + RING / 4];
            codes2 = __funnelshift_r(wprev2, wnext2, sh);
            wprev2 = wnext2;
        }
        widx = (widx + 1) & (RING / 4 - 1);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float2 e[KH];
            if (SC) {
                // code << 7 : two 32-bit loads per row, one per haplotype
                const uint32_t o1 = (__byte_perm(codes, 0, 0x4440 | k) << 7);
                const uint32_t o2 = (__byte_perm(codes2, 0, 0x4440 | k) << 7);
#pragma unroll
                for (int p = 0; p < KH; ++p) {
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(e[p].x) : "r"(ebase + o1 + p * 6 * ESZ));
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(e[p].y) : "r"(ebase + o2 + p * 6 * ESZ));
                }
            } else {
                const uint32_t eoff = __byte_perm(codes, 0, 0x4404 | (k << 4));   // (code_k << 8)
#pragma unroll
                for (int p = 0; p < KH; ++p) {
                    float2 v;
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(ebase + eoff + p * 6 * ESZ));
                    e[p] = v;
                }
            }
            U2m = U1m; U2z = U1z;
            const float2 zb = __ffma2_rn(X[KH - 1], K::v(pG[KH - 1]), Y[KH - 1]);
            U1m = shfl_up2(M[KH - 1]); U1x = shfl_up2(X[KH - 1]); U1z = shfl_up2(zb);
            float2 nM[KH], nX[KH], nY[KH];
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const float2 dm = p ? M[p ? p - 1 : 0] : U2m;
                const float2 dz = p ? __ffma2_rn(X[p ? p - 1 : 0], K::v(pG[p ? p - 1 : 0]), Y[p ? p - 1 : 0]) : U2z;
                nM[p] = __fmul2_rn(__ffma2_rn(dm, K::v(pMM[p]), dz), e[p]);
                nY[p] = __ffma2_rn(M[p], K::v(pMY[p]), __fmul2_rn(Y[p], K::v(pXX[p])));
            }
#pragma unroll
            for (int p = 0; p < KH; ++p) {
                const float2 um = p ? nM[p ? p - 1 : 0] : U1m, ux = p ? nX[p ? p - 1 : 0] : U1x;
                if (p % GRP == 0) nX[p] = __ffma2_rn(um, K::v(pMX[p]), __fmul2_rn(ux, K::v(K2[p / GRP])));
                else nX[p] = __ffma2_rn(um, K::v(pMX[p]), ux);
            }
#pragma unroll
            for (int p = 0; p < KH; ++p) { M[p] = nM[p]; X[p] = nX[p]; Y[p] = nY[p]; }
            sumM = __fadd2_rn(sumM, M[KH - 1]);
            sumX = __fadd2_rn(sumX, X[KH - 1]);
        }
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

template <int KH, bool SC>
static void run(float *out, const float *in, int bps)
{
    const int smem = WARPS * (KH * 6 * (SC ? 128 : 256) + 2 * RING);
    cudaFuncSetAttribute(proto<KH, SC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, proto<KH, SC>, WARPS * 32, smem);
    if (bps > occ) return;
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, proto<KH, SC>);
    const int grid = 148 * bps, ncols = 1 << 16;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    proto<KH, SC><<<grid, WARPS * 32, smem>>>(out, in, 1024);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        proto<KH, SC><<<grid, WARPS * 32, smem>>>(out, in, ncols);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double cells = (double)grid * WARPS * 32 * 2.0 * KH * ncols;
    printf("P3 %s KH %2d regs %3d occ %d blocks/SM %d  %8.3f ms  %8.1f Gcells/s (%s)\n", SC ? "scalar-const" : "pair-const  ", KH,
           fa.numRegs, occ, bps, best, cells / (best * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    float *out, *in;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&in, 256);
    float h[64];
    for (int j = 0; j < 64; ++j) h[j] = 0.9f + 1e-3f * j;
    cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
    for (int bps = 2; bps <= 5; ++bps) {
        run<4, false>(out, in, bps); run<4, true>(out, in, bps);
        run<6, false>(out, in, bps); run<6, true>(out, in, bps);
        run<8, false>(out, in, bps); run<8, true>(out, in, bps);
        run<10, true>(out, in, bps); run<12, true>(out, in, bps);
    }
    return 0;
}
