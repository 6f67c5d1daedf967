// Experiment build of the step kernel: one translation unit with just the variants under test
// (profiles/exp/build_exp.sh). Not part of the product library.
#include "../../gpu_ecs_madrona_b200/csrc/phys_kernel.cuh"

#include "exp_list.h"      // written by build_exp.sh from $VARIANTS: #define MDR_EXP_LIST MDR_STEP_VARIANT(T, B), ...
#undef MDR_STEP_VARIANT
// joints variant aliased to the plain one: halves the build
#ifdef MDR_EXP_FIXED
#define MDR_STEP_VARIANT(T, B) { T, B, mdr::physStepKernel<T, B, false>, mdr::physStepKernel<T, B, false>, \
                                 mdr::physStepKernel<T, B, false>, mdr::physStepKernel<T, B, false, 1, mdr::Collisions64Layout>, \
                                 68, 640, 71, 64 }
#else
#define MDR_STEP_VARIANT(T, B) { T, B, mdr::physStepKernel<T, B, false>, mdr::physStepKernel<T, B, false>, \
                                 mdr::physStepKernel<T, B, false>, nullptr, 0, 0, 0, 0 }
#endif
static const mdr::KernelVariant kPart[] = { MDR_EXP_LIST };
const mdr::KernelVariant *mdrStepVariants0(int *n) { *n = (int)(sizeof(kPart) / sizeof(kPart[0])); return kPart; }
const mdr::KernelVariant *mdrStepVariants1(int *n) { *n = 0; return kPart; }
const mdr::KernelVariant *mdrStepVariants2(int *n) { *n = 0; return kPart; }
const mdr::KernelVariant *mdrStepVariants3(int *n) { *n = 0; return kPart; }
const mdr::KernelVariant *mdrStepVariants4(int *n) { *n = 0; return kPart; }
