// step_inst_w1.cu — step-kernel instantiations with 1 warp(s) per environment.
#include "step_kernel.cuh"
#include "step_variants.h"

namespace sparc {
static const StepVariant kTable[] = {
#ifdef SPARC_QUICK   /* experiments: one variant per W, compiles in seconds */
    SPARC_STEP_VARIANT(12, 1),
#else
    SPARC_STEP_VARIANT(4, 1), SPARC_STEP_VARIANT(8, 1), SPARC_STEP_VARIANT(12, 1), SPARC_STEP_VARIANT(14, 1), SPARC_STEP_VARIANT(16, 1), SPARC_STEP_VARIANT(20, 1), SPARC_STEP_VARIANT(24, 1),
#endif
};
const StepVariant *step_variants_w1(int *n)
{
    *n = (int)(sizeof(kTable) / sizeof(kTable[0]));
    return kTable;
}
}  // namespace sparc
