// FP64 forward kernels (rescue path), bit-exact.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF64 lookup_f64_f0(int kh)
{
    switch (kh) {
    case 1: return f64_kernel<1, false>();
    case 2: return f64_kernel<2, false>();
    case 3: return f64_kernel<3, false>();
    case 4: return f64_kernel<4, false>();
    case 5: return f64_kernel<5, false>();
    case 6: return f64_kernel<6, false>();
    case 7: return f64_kernel<7, false>();
    case 8: return f64_kernel<8, false>();
    default: return nullptr;
    }
}
}  // namespace phmm
