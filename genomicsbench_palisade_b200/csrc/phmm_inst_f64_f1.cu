// FP64 forward kernels (rescue path), FMA-contracted.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF64 lookup_f64_f1(int kh)
{
    switch (kh) {
    case 1: return f64_kernel<1, true>();
    case 2: return f64_kernel<2, true>();
    case 3: return f64_kernel<3, true>();
    case 4: return f64_kernel<4, true>();
    case 5: return f64_kernel<5, true>();
    case 6: return f64_kernel<6, true>();
    case 7: return f64_kernel<7, true>();
    case 8: return f64_kernel<8, true>();
    default: return nullptr;
    }
}
}  // namespace phmm
