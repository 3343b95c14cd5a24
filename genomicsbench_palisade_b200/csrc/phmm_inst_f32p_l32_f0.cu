// FP32 packed (f32x2) forward kernels: 32 lanes per read, bit-exact.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32p_l32_f0(int kh)
{
    switch (kh) {
    case 1: return f32_kernel<2, 1, 32, false>();
    case 2: return f32_kernel<2, 2, 32, false>();
    case 3: return f32_kernel<2, 3, 32, false>();
    case 4: return f32_kernel<2, 4, 32, false>();
    default: return nullptr;
    }
}
}  // namespace phmm
