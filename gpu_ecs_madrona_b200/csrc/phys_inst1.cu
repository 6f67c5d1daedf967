// Step-kernel instantiations, unit 1 of 5 (see phys_kernel.cuh). sm_100a, -fmad=false.
#include "phys_kernel.cuh"

static const mdr::KernelVariant kPart[] = {
    MDR_STEP_VARIANT(128, 6),      // 80 registers,
    MDR_STEP_VARIANT(256, 2),      // 128 registers
};

const mdr::KernelVariant *mdrStepVariants1(int *n)
{
    *n = (int)(sizeof(kPart) / sizeof(kPart[0]));
    return kPart;
}
