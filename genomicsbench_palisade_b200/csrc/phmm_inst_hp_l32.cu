// Fast FP32 forward kernels, haplotype-pair packing: 32 lanes per read.
#include "phmm_kernels_hp.cuh"
namespace phmm {
KernelF32 lookup_f32_hp_l32(int kh)
{
    switch (kh) {
    case 1: return f32_hp_kernel<1, 32>();
    case 2: return f32_hp_kernel<2, 32>();
    case 3: return f32_hp_kernel<3, 32>();
    case 4: return f32_hp_kernel<4, 32>();
    case 5: return f32_hp_kernel<5, 32>();
    case 6: return f32_hp_kernel<6, 32>();
    case 7: return f32_hp_kernel<7, 32>();
    case 8: return f32_hp_kernel<8, 32>();
    case 9: return f32_hp_kernel<9, 32>();
    case 10: return f32_hp_kernel<10, 32>();
    case 11: return f32_hp_kernel<11, 32>();
    case 12: return f32_hp_kernel<12, 32>();
    case 13: return f32_hp_kernel<13, 32>();
    case 14: return f32_hp_kernel<14, 32>();
    case 15: return f32_hp_kernel<15, 32>();
    case 16: return f32_hp_kernel<16, 32>();
    default: return nullptr;
    }
}
}  // namespace phmm
