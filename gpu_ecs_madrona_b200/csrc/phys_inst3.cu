// Step-kernel instantiations, unit 3 of 5 (see phys_kernel.cuh). sm_100a, -fmad=false.
#include "phys_kernel.cuh"

static const mdr::KernelVariant kPart[] = {
    MDR_STEP_VARIANT(384, 2),      // 80 registers,
    MDR_STEP_VARIANT(512, 1),      // 128 registers
};

const mdr::KernelVariant *mdrStepVariants3(int *n)
{
    *n = (int)(sizeof(kPart) / sizeof(kPart[0]));
    return kPart;
}
