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
// One block of <= SPARC_MECH_CKPT µs without clip handling: the same operation sequence as mech_step
// when no acceleration / jerk / speed clip is active; the clip tests only accumulate a flag (off
// the dependency chain, which is what bounds this latency-bound kernel).  Returns false if any
// clip condition was met: the caller redoes the block with mech_step.
template <int MODE>
__device__ __forceinline__ bool mech_block_fast(const SparcParams &P, const EnvPrm &q, double &wire, double &v, double &pa, const int it0,
                                                const int nb, const int first_latch, const double d_old,
                                                const double d_new, double *traj, double *trajv)
{
    const double dt = P.mech_dt, c1 = q.damping_coeff, c2 = q.stiffness_coeff, nw = -q.omega_n;
    const double amax = q.max_acceleration, jmax = q.max_jerk_dt, vmax = q.max_speed;
    bool bad = false;
#pragma unroll 4
    for (int k = 0; k < nb; k++) {
        const double delta = (it0 + k >= first_latch) ? d_new : d_old;
        double a_nom;
        if (MODE == 0) {
            const double x_err = wire - (wire + delta);
            a_nom = c1 * v + c2 * x_err;
        } else {
            a_nom = nw * (v - delta);
        }
        const double da = a_nom - pa;
        bad = bad || !(fabs(a_nom) <= amax) || !(fabs(da) <= jmax);
        pa = pa + da;
        v += pa * dt;
        bad = bad || !(fabs(v) <= vmax);
        wire += v * dt;
        traj[it0 + k] = wire;
        if (trajv) trajv[it0 + k] = v;      // only while signals are sampled
    }
    return !bad;
}

#ifndef SPARC_MECH_MAX_THREADS
#define SPARC_MECH_MAX_THREADS 256     /* largest pre-pass block (envs per block), see mech_launch */
#endif
__global__ void __launch_bounds__(SPARC_MECH_MAX_THREADS) mech_kernel(const __grid_constant__ SparcParams P,
                                                   const __grid_constant__ SparcStatePtrs S,
                                                   const __grid_constant__ StepArgs a)
{
    // Let the step kernel start NOW (programmatic dependent launch): it is launched once every block
    // of this grid has executed this instruction (or exited), i.e. once all producers are resident,
    // and then follows this kernel block by block through mech_prog — it never waits for the grid.
    asm volatile("griddepcontrol.launch_dependents;");
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= a.n_envs) return;
    if (S.flags[e] & SPARC_F_DONE) return;      // the step kernel does not read the scratch of a frozen env
    double wire = S.wire[e], v = S.v[e], pa = S.prev_accel[e];
    EnvPrm q;                                   // only the mechanics members are read (the rest is never loaded)
    q.damping_coeff = S.prm_damping_coeff[e]; q.stiffness_coeff = S.prm_stiffness_coeff[e]; q.omega_n = S.prm_omega_n[e];
    q.max_acceleration = S.prm_max_acceleration[e]; q.max_jerk_dt = S.prm_max_jerk_dt[e]; q.max_speed = S.prm_max_speed[e];
    const double d_old = S.target_delta[e];
    const double d_new = servo_command(a, S, e);
    // launch-µs of the next action latch (wire_edm.py:117); later latches re-latch the same command
    int first_latch = P.servo_interval - S.t_servo[e];
    if (first_latch < 0) first_latch = 0;
    double *traj = S.mech_traj + (size_t)e * SPARC_MECH_EPOCH;
    double *trajv = a.signals ? S.mech_trajv + (size_t)e * SPARC_MECH_EPOCH : nullptr;
    double *ck = S.mech_ckpt + (size_t)e * SPARC_MECH_NCKPT * 2;
    uint32_t *prog = (uint32_t *)S.mech_prog + e;
    ck[0] = v;                                  // checkpoint c = state before µs c*SPARC_MECH_CKPT
    ck[1] = pa;
#pragma unroll 1
    for (int it0 = 0; it0 < a.n_us; it0 += SPARC_MECH_CKPT) {
        const int nb = (a.n_us - it0 < SPARC_MECH_CKPT) ? a.n_us - it0 : SPARC_MECH_CKPT;
        const double w0 = wire, v0 = v, p0 = pa;
        const bool ok = (P.control_mode == 0)
                            ? mech_block_fast<0>(P, q, wire, v, pa, it0, nb, first_latch, d_old, d_new, traj, trajv)
                            : mech_block_fast<1>(P, q, wire, v, pa, it0, nb, first_latch, d_old, d_new, traj, trajv);
        if (!ok) {                              // a clip was active somewhere in the block: exact path
            wire = w0; v = v0; pa = p0;
#pragma unroll 1
            for (int k = 0; k < nb; k++) {
                mech_step(P, q, wire, v, pa, (it0 + k >= first_latch) ? d_new : d_old);
                traj[it0 + k] = wire;
                if (trajv) trajv[it0 + k] = v;
            }
        }
        const int b = it0 / SPARC_MECH_CKPT + 1;        // blocks finished
        ck[2 * b] = v;                                  // also what a rewind to the block boundary reads
        ck[2 * b + 1] = pa;
        if (it0 + nb >= a.n_us) {                       // end-of-launch pair
            ck[2 * (SPARC_MECH_NCKPT - 1)] = v;
            ck[2 * (SPARC_MECH_NCKPT - 1) + 1] = pa;
        }
        // publish: positions of µs < it0 + nb and checkpoints <= b are final (release at GPU scope)
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(prog), "r"(a.mech_seq + (uint32_t)b) : "memory");
    }
}

// Envs (threads) per pre-pass block.  The threads are independent and latency-bound (~110 cycles per simulated µs),
// so the block shape only decides HOW MANY SMs host pre-pass warps next to the step kernel's.  Measured (same box,
// profiles/r02b_exp_prepass2.txt): 32 blocks of 128 at 4096 envs -4.3 % per control step against 128 blocks of 32,
// 64 blocks of 256 at 16384 envs -2.6 %; 8 blocks of 512 +25 % (sixteen fp64 chains per scheduler).  A 120-register
// step kernel, which leaves room for the pre-pass blocks beside eight resident step blocks, changes nothing: it is
// not the register footprint.  Rule: about 32 blocks, at least one warp, at most SPARC_MECH_MAX_THREADS per block.
static int mech_block_threads(const int n_envs)
{
#ifdef SPARC_MECH_THREADS   /* experiments: fixed block size */
    return SPARC_MECH_THREADS;
#else
    int t = 32;
    while (t < SPARC_MECH_MAX_THREADS && n_envs > 32 * t) t *= 2;
    return t;
#endif
}

void mech_launch(const SparcParams &P, const SparcStatePtrs &S, const StepArgs &a, cudaStream_t stream)
{
    const int T = mech_block_threads(a.n_envs);
    mech_kernel<<<(a.n_envs + T - 1) / T, T, 0, stream>>>(P, S, a);
}

}  // namespace sparc
