// Step-kernel instantiations, unit 1 of 5 (see phys_kernel.cuh). sm_100a, -fmad=false.
#include "phys_kernel.cuh"

static const mdr::KernelVariant kPart[] = {
    MDR_STEP_VARIANT(128, 6),      // 80 registers,
    MDR_STEP_VARIANT_C64(256, 2),  // 128 registers; + the collisions-64 shape with a compile-time layout
};

const mdr::KernelVariant *mdrStepVariants1(int *n)
{
    *n = (int)(sizeof(kPart) / sizeof(kPart[0]));
    return kPart;
}
