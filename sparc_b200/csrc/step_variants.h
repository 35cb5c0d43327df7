// step_variants.h — table of step-kernel instantiations, one translation unit per W (warps per
// environment) so that they compile in parallel (sparc_b200/build.py).
#pragma once
#include "../../include/sparc_b200.h"
#include "state_layout.h"
#include <cuda_runtime.h>

namespace sparc {
struct StepArgs;
using StepFn = void (*)(const SparcParams, const SparcStatePtrs, const StepArgs);

struct StepVariant {
    int K;              // cells per lane (even: the stencil is two-wide)
    int W;              // warps per environment (1: warp per env, several envs per block; > 1: block per env)
    StepFn fn[2][2][2]; // [LOG][INJECT][ALIGNED]
};

// mechanics pre-pass of one launch epoch (mech_kernel.cu)
void mech_launch(const SparcParams &P, const SparcStatePtrs &S, const StepArgs &a, cudaStream_t stream);

const StepVariant *step_variants_w1(int *n);
const StepVariant *step_variants_w2(int *n);
const StepVariant *step_variants_w4(int *n);
const StepVariant *step_variants_w8(int *n);
}  // namespace sparc

#define SPARC_STEP_VARIANT(K, W)                                                                                      \
    {K, W,                                                                                                            \
     {{{sparc::step_kernel<K, false, W, false, false>, sparc::step_kernel<K, false, W, false, true>},                  \
       {sparc::step_kernel<K, true, W, false, false>, sparc::step_kernel<K, true, W, false, true>}},                   \
      {{sparc::step_kernel<K, false, W, true, false>, sparc::step_kernel<K, false, W, true, true>},                    \
       {sparc::step_kernel<K, true, W, true, false>, sparc::step_kernel<K, true, W, true, true>}}}}
