// FP32 scalar forward kernels (guard-band re-decision, non-packed variant).
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32s_f0(int kh)
{
    switch (kh) {
    case 1: return f32_kernel<1, 1, 32, false>();
    case 2: return f32_kernel<1, 2, 32, false>();
    case 3: return f32_kernel<1, 3, 32, false>();
    case 4: return f32_kernel<1, 4, 32, false>();
    case 5: return f32_kernel<1, 5, 32, false>();
    case 6: return f32_kernel<1, 6, 32, false>();
    case 7: return f32_kernel<1, 7, 32, false>();
    case 8: return f32_kernel<1, 8, 32, false>();
    default: return nullptr;
    }
}
}  // namespace phmm
