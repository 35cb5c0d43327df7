// step_inst_w4.cu — step-kernel instantiations with 4 warp(s) per environment.
#include "step_kernel.cuh"
#include "step_variants.h"

namespace sparc {
static const StepVariant kTable[] = {
#ifdef SPARC_QUICK   /* experiments: one variant per W, compiles in seconds */
    SPARC_STEP_VARIANT(18, 4),
#else
    SPARC_STEP_VARIANT(18, 4), SPARC_STEP_VARIANT(22, 4), SPARC_STEP_VARIANT(26, 4),
#endif
};
const StepVariant *step_variants_w4(int *n)
{
    *n = (int)(sizeof(kTable) / sizeof(kTable[0]));
    return kTable;
}
}  // namespace sparc
