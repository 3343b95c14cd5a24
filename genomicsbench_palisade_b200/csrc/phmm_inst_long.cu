// Multi-pass forward kernels for reads longer than any single-pass class.
#include "phmm_kernels.cuh"
namespace phmm {
KernelF32 lookup_f32_long(bool fast)
{
    return fast ? f32_kernel<LONG_KH, 32, true, true>() : f32_kernel<LONG_KH, 32, false, true>();
}
KernelF64 lookup_f64_long(bool fast)
{
    return fast ? f64_kernel<LONG_KH, true, true>() : f64_kernel<LONG_KH, false, true>();
}
}  // namespace phmm
