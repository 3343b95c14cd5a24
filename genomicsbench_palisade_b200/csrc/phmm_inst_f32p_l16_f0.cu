// FP32 packed (f32x2) forward kernels: 16 lanes per read, bit-exact.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32p_l16_f0(int kh)
{
    switch (kh) {
    case 1: return f32_kernel<2, 1, 16, false>();
    case 2: return f32_kernel<2, 2, 16, false>();
    case 3: return f32_kernel<2, 3, 16, false>();
    case 4: return f32_kernel<2, 4, 16, false>();
    case 5: return f32_kernel<2, 5, 16, false>();
    case 6: return f32_kernel<2, 6, 16, false>();
    default: return nullptr;
    }
}
}  // namespace phmm
