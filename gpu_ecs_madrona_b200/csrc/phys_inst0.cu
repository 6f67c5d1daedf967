// Step-kernel instantiations, unit 0 of 5 (see phys_kernel.cuh). sm_100a, -fmad=false.
#include "phys_kernel.cuh"

static const mdr::KernelVariant kPart[] = {
    MDR_STEP_VARIANT(64, 8),      // 128 registers,
    MDR_STEP_VARIANT(128, 4),      // 128 registers
};

const mdr::KernelVariant *mdrStepVariants0(int *n)
{
    *n = (int)(sizeof(kPart) / sizeof(kPart[0]));
    return kPart;
}
