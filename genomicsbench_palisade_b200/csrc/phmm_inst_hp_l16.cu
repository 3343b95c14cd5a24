// Fast FP32 forward kernels, haplotype-pair packing: 16 lanes per read.
#include "phmm_kernels_hp.cuh"
namespace phmm {
KernelF32 lookup_f32_hp_l16(int kh)
{
    switch (kh) {
    case 1: return f32_hp_kernel<1, 16>();
    case 2: return f32_hp_kernel<2, 16>();
    case 3: return f32_hp_kernel<3, 16>();
    case 4: return f32_hp_kernel<4, 16>();
    case 5: return f32_hp_kernel<5, 16>();
    case 6: return f32_hp_kernel<6, 16>();
    case 7: return f32_hp_kernel<7, 16>();
    case 8: return f32_hp_kernel<8, 16>();
    case 9: return f32_hp_kernel<9, 16>();
    case 10: return f32_hp_kernel<10, 16>();
    case 11: return f32_hp_kernel<11, 16>();
    case 12: return f32_hp_kernel<12, 16>();
    case 13: return f32_hp_kernel<13, 16>();
    case 14: return f32_hp_kernel<14, 16>();
    case 15: return f32_hp_kernel<15, 16>();
    case 16: return f32_hp_kernel<16, 16>();
    default: return nullptr;
    }
}
}  // namespace phmm
