// Fast FP32 forward kernels, haplotype-pair packing: 8 lanes per read.
#include "phmm_kernels_hp.cuh"
namespace phmm {
KernelF32 lookup_f32_hp_l8(int kh)
{
    switch (kh) {
    case 1: return f32_hp_kernel<1, 8>();
    case 2: return f32_hp_kernel<2, 8>();
    case 3: return f32_hp_kernel<3, 8>();
    case 4: return f32_hp_kernel<4, 8>();
    case 5: return f32_hp_kernel<5, 8>();
    case 6: return f32_hp_kernel<6, 8>();
    case 7: return f32_hp_kernel<7, 8>();
    case 8: return f32_hp_kernel<8, 8>();
    case 9: return f32_hp_kernel<9, 8>();
    case 10: return f32_hp_kernel<10, 8>();
    case 11: return f32_hp_kernel<11, 8>();
    case 12: return f32_hp_kernel<12, 8>();
    case 13: return f32_hp_kernel<13, 8>();
    case 14: return f32_hp_kernel<14, 8>();
    case 15: return f32_hp_kernel<15, 8>();
    case 16: return f32_hp_kernel<16, 8>();
    default: return nullptr;
    }
}
}  // namespace phmm
