// FP32 packed (f32x2) forward kernels: 8 lanes per read, FMA-contracted.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32p_l8_f1(int kh)
{
    switch (kh) {
    case 1: return f32_kernel<2, 1, 8, true>();
    case 2: return f32_kernel<2, 2, 8, true>();
    case 3: return f32_kernel<2, 3, 8, true>();
    case 4: return f32_kernel<2, 4, 8, true>();
    case 5: return f32_kernel<2, 5, 8, true>();
    case 6: return f32_kernel<2, 6, 8, true>();
    default: return nullptr;
    }
}
}  // namespace phmm
