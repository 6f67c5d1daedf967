// Step-kernel instantiations, unit 4 of 5 (see phys_kernel.cuh). sm_100a, -fmad=false.
#include "phys_kernel.cuh"

static const mdr::KernelVariant kPart[] = {
    MDR_STEP_VARIANT(640, 1),      // 96 registers,
    MDR_STEP_VARIANT(1024, 1),      // 64 registers
};

const mdr::KernelVariant *mdrStepVariants4(int *n)
{
    *n = (int)(sizeof(kPart) / sizeof(kPart[0]));
    return kPart;
}
