// FP32 forward kernels (two reads per thread in f32x2 halves): 16 lanes per read pair, bit-exact.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32_l16_f0(int kh)
{
    switch (kh) {
    case 1: return f32_kernel<1, 16, false>();
    case 2: return f32_kernel<2, 16, false>();
    case 3: return f32_kernel<3, 16, false>();
    case 4: return f32_kernel<4, 16, false>();
    case 5: return f32_kernel<5, 16, false>();
    case 6: return f32_kernel<6, 16, false>();
    case 7: return f32_kernel<7, 16, false>();
    case 8: return f32_kernel<8, 16, false>();
    default: return nullptr;
    }
}
}  // namespace phmm
