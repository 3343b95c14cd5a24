// Multi-pass forward kernels for reads longer than any single-pass class.
#include "phmm_kernels.cuh"
namespace phmm {
// which: 0 FP32 packed bit-exact, 1 FP32 packed fast, 2 FP32 scalar bit-exact (guard re-decision)
KernelF32 lookup_f32_long(int which)
{
    switch (which) {
    case 0: return f32_kernel<2, LONG_KH_F32, 32, false, true>();
    case 1: return f32_kernel<2, LONG_KH_F32, 32, true, true>();
    case 2: return f32_kernel<1, LONG_KH_F32S, 32, false, true>();
    case 3: return f32_kernel<1, LONG_KH_F32S, 32, true, true>();
    default: return nullptr;
    }
}
KernelF64 lookup_f64_long(bool fast)
{
    return fast ? f64_kernel<LONG_KH_F64, true, true>() : f64_kernel<LONG_KH_F64, false, true>();
}
}  // namespace phmm
