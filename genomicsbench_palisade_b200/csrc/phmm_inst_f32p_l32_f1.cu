// FP32 packed (f32x2) forward kernels: 32 lanes per read, FMA-contracted.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32p_l32_f1(int kh)
{
    switch (kh) {
    case 1: return f32_kernel<2, 1, 32, true>();
    case 2: return f32_kernel<2, 2, 32, true>();
    case 3: return f32_kernel<2, 3, 32, true>();
    case 4: return f32_kernel<2, 4, 32, true>();
    default: return nullptr;
    }
}
}  // namespace phmm
