// mech_kernel.cu — mechanics pre-pass of one launch epoch (see step_kernel.cuh for its consumer).
#include "step_common.cuh"
#include "step_variants.h"

namespace sparc {

// ---- mechanics pre-pass --------------------------------------------------------------------
// MechanicsModule.update depends on nothing but the latched servo command, which is fixed for a
// launch, so the whole fp64 recurrence of a launch is evaluated BEFORE the step kernel by one
// thread per environment (32 envs per warp instruction instead of one: fp64 instructions cost two
// issue cycles each on B200 and were 27 % of the fused kernel when every warp ran them for its own
// env).  Output: the wire position after every µs (read back 32 µs at a time, one per lane) and
// (velocity, prev_accel) checkpoints for the rare exact rewind when an env terminates.
__global__ void __launch_bounds__(128) mech_kernel(const __grid_constant__ SparcParams P,
                                                   const __grid_constant__ SparcStatePtrs S,
                                                   const __grid_constant__ StepArgs a)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= a.n_envs) return;
    if (S.flags[e] & SPARC_F_DONE) return;
    double wire = S.wire[e], v = S.v[e], pa = S.prev_accel[e];
    double delta = S.target_delta[e];
    const double delta_new = (double)a.act.servo[e];
    const int period = (P.servo_interval > 1) ? P.servo_interval : 1;
    int latch = P.servo_interval - S.t_servo[e];        // launch-µs of the next action latch (wire_edm.py:117)
    if (latch < 0) latch = 0;
    double *traj = S.mech_traj + (size_t)e * SPARC_MECH_EPOCH;
    double *ck = S.mech_ckpt + (size_t)e * SPARC_MECH_NCKPT * 2;
#pragma unroll 1
    for (int it = 0; it < a.n_us; it++) {
        if (it % SPARC_MECH_CKPT == 0) { ck[2 * (it / SPARC_MECH_CKPT)] = v; ck[2 * (it / SPARC_MECH_CKPT) + 1] = pa; }
        if (it == latch) { delta = delta_new; latch += period; }
        mech_step(P, wire, v, pa, delta);
        traj[it] = wire;
    }
    ck[2 * (SPARC_MECH_NCKPT - 1)] = v;                 // end-of-launch pair
    ck[2 * (SPARC_MECH_NCKPT - 1) + 1] = pa;
}

void mech_launch(const SparcParams &P, const SparcStatePtrs &S, const StepArgs &a, cudaStream_t stream)
{
    mech_kernel<<<(a.n_envs + 127) / 128, 128, 0, stream>>>(P, S, a);
}

}  // namespace sparc
