// Step-kernel instantiations, unit 2 of 5 (see phys_kernel.cuh). sm_100a, -fmad=false.
#include "phys_kernel.cuh"

static const mdr::KernelVariant kPart[] = {
    MDR_STEP_VARIANT(256, 3),      // 80 registers,
    MDR_STEP_VARIANT(384, 1),      // 168 registers
};

const mdr::KernelVariant *mdrStepVariants2(int *n)
{
    *n = (int)(sizeof(kPart) / sizeof(kPart[0]));
    return kPart;
}
