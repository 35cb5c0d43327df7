// step_kernel.cuh — the fused step kernel (sm_100a), round-2 organisation.
//
// One warp per environment (W == 1; W warps of one block for wires longer than one warp), the wire
// temperature field in REGISTERS, halos by warp shuffle, the per-µs scalar state machines replaced
// by a lane-parallel LOOK-AHEAD exactly as in round 1 (mechanics from the pre-pass kernel, quiet-run
// detection by ballot, scalar-free stencil runs, exact general µs for everything else).  What is new:
//
//   * SCALAR STATE IN SHARED MEMORY.  Everything that is not the field lives in ONE shared-memory record per
//     warp (WarpRec: the EnvState the general µs works on, the per-env parameters, the launch bookkeeping, the
//     32 wire positions of the current look-ahead block, a window of Philox words).  Across the stencil loops a
//     lane holds the field (K floats), FOUR convection coefficients, two conduction scalars and a loop count —
//     the round-1 kernel held 36 coefficient registers and ~40 scalars and carried 168 bytes of spills at 128
//     registers.  The kernel also compiles to 72 registers (28 warps/SM); measured, more resident warps do not
//     help (the kernel is bound by instruction dispatch, profiles/r02_experiments.txt), so it ships at 128
//     registers / 16 warps per SM with 8 bytes of spills.
//   * REGION LAYOUT (SparcLayout, step_common.cuh).  Zone boundaries, the zero-flux end and the pad
//     slots used to be per-cell coefficient registers (cE, c0, cR per cell).  Now the cells are laid out
//     so that every such boundary falls between two HALVES of a lane: the conduction coefficient is one
//     scalar for the whole wire, the convection pair one value per half, the zero-flux end is a halo
//     override, and pads sit in compile-time slots that mirror their right-hand neighbour.
//   * FOUR packed FP32 instructions per pair of cells and µs: q = fma(cC,T,c0); q = fma(cR,T_next,q);
//     q = fma(cL,T_prev,q); T += q, with cC = cE - (cR + cL) (stencil_apply).  The form with explicit
//     differences (five packed instructions and three scalar subtractions per lane and µs) was the kernel's
//     FMA-pipe bound; the coefficients are rounded so that cC + cR + cL is exactly the convection weight
//     (see "stencil coefficients" below), which keeps the per-µs error of q near 5e-8 K, two to three
//     orders of magnitude under the fp32 rounding of T' = T + q that the reference performs as well.
//   * Philox ignition words are computed for 128 µs at a time (each lane one counter, all four
//     words kept) instead of 32.
//
// Every shortcut is conservative: a µs that is not provably quiet runs the general path, which is
// the reference's logic statement by statement, so results do not depend on where the run boundaries fall.
#pragma once
#include "step_common.cuh"

namespace sparc {

__device__ __forceinline__ float2 sub2(const float2 a, const float2 b)
{
    return __fadd2_rn(a, make_float2(-b.x, -b.y));
}

// ---- wires longer than one warp: W warps (one thread block) per environment -------------------
// Each warp runs the scalar state machine redundantly on its own record (same inputs, same results,
// so control flow is identical in all warps of the block); only three things cross warps: the
// stencil halos (shared memory, double buffered, one barrier per µs), the temperature votes
// (__syncthreads_or) and the zone-mean sum.
template <int W>
struct EnvCtx {
    int lane, wid, gl;      // lane in warp, warp in env, lane in env
    float (*halo)[W][2];    // [parity][warp][first, last]
    int parity;
};

// Where the two halves of this lane sit on the wire (SparcLayout).
struct LaneGeom {
    int cx0, nx, cy0, ny;   // first cell and number of real cells of the x / y half (n == 0: empty)
    int rx, ry;             // region of each half (-1: empty)
};

__device__ __forceinline__ LaneGeom lane_geom(const SparcLayout &L, const int gl, const int H)
{
    LaneGeom g;
    half_geom(L, 2 * gl, H, g.cx0, g.nx, g.rx);
    half_geom(L, 2 * gl + 1, H, g.cy0, g.ny, g.ry);
    return g;
}

// loop-invariant lane predicates of the stencil
struct LaneFlags {
    bool first;             // lane 0 of the env: left neighbour is the spool cell
    bool short_x, short_y;  // the last slot of the half is a pad
    bool end_x, end_y;      // the half ends with cell N-1 (zero flux to the right)
    bool empty_x, empty_y;  // the half holds no cells
};

template <int W>
__device__ __forceinline__ bool vote_any(const bool pred)
{
    if constexpr (W > 1) return __syncthreads_or(pred ? 1 : 0) != 0;
    else return __any_sync(0xffffffffu, pred);
}

// Halo exchange and pad upkeep of one µs.  Out: `left` (cell before the x half), `right` (cell after
// the y half, or the end cell itself at the zero-flux end), `xl` (last real cell of the x half), `yv`
// (cell after the x half).  Pads are set to their right-hand neighbour, so the packed difference
// T[H-1] - T[H-2] of a short half is the real cell's difference to its true neighbour.
template <int H, int W, bool XP>
__device__ __forceinline__ void stencil_edges(EnvCtx<W> &c, float2 (&T)[H], const LaneFlags &f, const float spoolf,
                                              float &left, float &right, float &xl, float &yv)
{
    constexpr unsigned FULL = 0xffffffffu;
    const float t_last = f.short_y ? T[H - 2].y : T[H - 1].y;
    left = __shfl_up_sync(FULL, t_last, 1);
    right = __shfl_down_sync(FULL, T[0].x, 1);
    if constexpr (W > 1) {
        float(*h)[2] = c.halo[c.parity];
        if (c.lane == 0) h[c.wid][0] = T[0].x;
        if (c.lane == 31) h[c.wid][1] = t_last;
        __syncthreads();
        if (c.lane == 0 && c.wid > 0) left = h[c.wid - 1][1];
        if (c.lane == 31 && c.wid < W - 1) right = h[c.wid + 1][0];
        c.parity ^= 1;
    }
    if (f.first) left = spoolf;
    if (f.end_y) right = t_last;
    if (f.short_y) T[H - 1].y = right;
    if constexpr (XP) {                         // general layout: x halves may be short or hold the end cell
        xl = f.short_x ? T[H - 2].x : T[H - 1].x;
        yv = f.end_x ? xl : T[0].y;
        if (f.short_x) T[H - 1].x = yv;
    } else {                                    // aligned layout: pads and the end cell sit in y halves only
        xl = T[H - 1].x;
        yv = T[0].y;
    }
}

// Per-cell coefficients of the update T' = T + q, q = cC*T + c0 + cR*T_next + cL*T_prev, where
//   cC = cE - (cR + cL), cE = -f*h*A (convection, wire.py:108-111), c0 = -cE*T_dielectric,
//   cR = f*k_cond (wire.py:66-75), cL = f*(k_cond + adv) (advection folded in, wire.py:113-119).
// One (cC, c0) pair per half when the layout is aligned, per cell otherwise.  See stencil_coefficients()
// for how the four numbers are rounded so that a uniform field is a fixed point of the conduction part.
template <int H, bool ALIGNED>
struct Coefs {
    float2 cC[ALIGNED ? 1 : H], c0[ALIGNED ? 1 : H];
    __device__ __forceinline__ float2 e(const int j) const { return cC[ALIGNED ? 0 : j]; }
    __device__ __forceinline__ float2 z(const int j) const { return c0[ALIGNED ? 0 : j]; }
};

// maximum over the REAL cells of this lane (pads and empty halves excluded), never below `floor`
template <int H>
__device__ __forceinline__ float lane_max(const float2 (&T)[H], const LaneFlags &f, const float floor_v)
{
    float mx = floor_v, my = floor_v;
#pragma unroll
    for (int j = 0; j < H - 1; j++) { mx = fmaxf(mx, T[j].x); my = fmaxf(my, T[j].y); }
    mx = f.short_x ? mx : fmaxf(mx, T[H - 1].x);
    my = f.short_y ? my : fmaxf(my, T[H - 1].y);
    mx = f.empty_x ? floor_v : mx;
    my = f.empty_y ? floor_v : my;
    return fmaxf(mx, my);
}

// Conduction + advection of one µs on top of acc[j] (= the convection term + sources), then T += q.  The SAME
// function and the same order of accumulation serve quiet and general µs, so a cell's value does not depend
// on which path executed a µs.  Two forms, chosen per env from its conduction weights (DIFF, see "stencil
// coefficients" in step_env):
//   DIFF = false  q = fma(cR, T_next, acc); q = fma(cL, T_prev, q) with acc = fma(cC, T, c0), cC = cE - (cR + cL):
//                 FOUR packed FP32 instructions per pair of cells and µs.  The products are of magnitude
//                 (cR + cL)*T, so q carries a rounding error of ~ulp((cR + cL)*T): used when cR + cL <= 2^-6,
//                 where that is <= 1/64 of the rounding T' = T + q performs anyway (README default: 2^-9).
//   DIFF = true   q = fma(cR, T_next - T, acc); q = fma(-cL, T - T_prev, q) with acc = fma(cE, T, c0): five
//                 packed instructions and three scalar subtractions per lane; exact differences, for fine
//                 segmentations (cR + cL up to 0.5), and the round-1 arithmetic bit for bit.
// The halo-independent terms come first so that the shuffles complete underneath them.
template <int H, bool DIFF>
__device__ __forceinline__ void stencil_apply(float2 (&T)[H], float2 (&acc)[H], const float2 cR2, const float2 cL2,
                                              const float left, const float right, const float xl, const float yv)
{
    if constexpr (!DIFF) {
#pragma unroll
        for (int j = 0; j < H - 1; j++) acc[j] = __ffma2_rn(cR2, T[j + 1], acc[j]);
#pragma unroll
        for (int j = 1; j < H; j++) acc[j] = __ffma2_rn(cL2, T[j - 1], acc[j]);
        acc[H - 1] = __ffma2_rn(cR2, make_float2(yv, right), acc[H - 1]);
        acc[0] = __ffma2_rn(cL2, make_float2(left, xl), acc[0]);
#pragma unroll
        for (int j = 0; j < H; j++) T[j] = __fadd2_rn(T[j], acc[j]);
    } else {                                    // cL2 holds -cL here
        const float d_mid = yv - xl;
        const float2 d0 = sub2(T[1], T[0]);     // dnext of pair 0, dprev of pair 1 (old values)
        float2 dprev = d0;
#pragma unroll
        for (int j = 1; j < H - 1; j++) {
            const float2 dnext = sub2(T[j + 1], T[j]);
            float2 q = __ffma2_rn(cR2, dnext, acc[j]);
            q = __ffma2_rn(cL2, dprev, q);
            dprev = dnext;
            T[j] = __fadd2_rn(T[j], q);
        }
        {
            const float2 dlast = make_float2(d_mid, right - T[H - 1].y);
            float2 q = __ffma2_rn(cR2, dlast, acc[H - 1]);
            q = __ffma2_rn(cL2, dprev, q);
            T[H - 1] = __fadd2_rn(T[H - 1], q);
        }
        {
            const float2 dfirst = make_float2(T[0].x - left, d_mid);
            float2 q = __ffma2_rn(cR2, d0, acc[0]);
            q = __ffma2_rn(cL2, dfirst, q);
            T[0] = __fadd2_rn(T[0], q);
        }
    }
}

// One µs of compute_thermal_update (wire.py:58-123) without heat sources.
template <int H, bool ALIGNED, int W, bool DIFF>
__device__ __forceinline__ void stencil_quiet_us(float2 (&T)[H], const Coefs<H, ALIGNED> &co, const float2 cR2,
                                                 const float2 cL2, const float spoolf, const LaneFlags &f, EnvCtx<W> &ctx)
{
    float left, right, xl, yv;
    stencil_edges<H, W, !ALIGNED>(ctx, T, f, spoolf, left, right, xl, yv);
    float2 acc[H];
#pragma unroll
    for (int j = 0; j < H; j++) acc[j] = __ffma2_rn(co.e(j), T[j], co.z(j));
    stencil_apply<H, DIFF>(T, acc, cR2, cL2, left, right, xl, yv);
}

#ifndef SPARC_COLD_UNROLL
#define SPARC_COLD_UNROLL 2     /* µs per trip of the cold quiet loop (4: one instruction per µs fewer, but the 2.2 kB loop body measured +3 ... +4 % time: no_instruction 0.34 -> 0.62 per issue) */
#endif
// n quiet µs.  HOT additionally runs _check_wire_breaking (wire.py:376-388) every µs and returns the
// number of µs executed (n, or k+1 when the wire broke on µs k; `broke` says which).
template <int H, bool HOT, bool ALIGNED, int W, bool LOG, bool DIFF, class LogTick, class SigCap>
__device__ __forceinline__ int run_quiet_t(float2 (&T)[H], const Coefs<H, ALIGNED> &co, const float cRf, const float cLf,
                                         const float spoolf, const LaneFlags &f, EnvCtx<W> &ctx, const int n,
                                         const float crit_T, const float break_T, int &crit_count, bool &broke,
                                         bool &cold_after, LogTick &&log_tick, SigCap &&sig_cap)
{
    const float2 cL2 = make_float2(cLf, cLf), cR2 = make_float2(cRf, cRf);
    const float gate_T = fminf(crit_T, break_T);
    broke = false;
#ifdef SPARC_ABL_NOSTENCIL
    if (!HOT) { crit_count = 0; return n; }
#endif
    if constexpr (!HOT && !LOG && W == 1) {     // the common case: SPARC_COLD_UNROLL µs per trip (halves the loop overhead)
        int k = 0;
#pragma unroll 1
        for (; k + SPARC_COLD_UNROLL <= n; k += SPARC_COLD_UNROLL) {
#pragma unroll
            for (int u = 0; u < SPARC_COLD_UNROLL; u++) stencil_quiet_us<H, ALIGNED, W, DIFF>(T, co, cR2, cL2, spoolf, f, ctx);
        }
#pragma unroll 1
        for (; k < n; k++) stencil_quiet_us<H, ALIGNED, W, DIFF>(T, co, cR2, cL2, spoolf, f, ctx);
        crit_count = 0;
        return n;
    }
#pragma unroll 1
    for (int k = 0; k < n; k++) {
        stencil_quiet_us<H, ALIGNED, W, DIFF>(T, co, cR2, cL2, spoolf, f, ctx);
        if (LOG) log_tick();
        if (HOT) {
            const float lmax = lane_max<H>(T, f, spoolf);
            bool hot = false, brk = false;
            if (vote_any<W>(lmax > gate_T)) {
                hot = vote_any<W>(lmax > crit_T);
                brk = vote_any<W>(lmax > break_T);
            }
            crit_count = hot ? crit_count + 1 : 0;
            if (brk) { broke = true; return k + 1; }
            if (k == n - 1) cold_after = !vote_any<W>(lmax > gate_T - 1.0f);
        }
        if (LOG) sig_cap(k, HOT ? crit_count : 0);   // sampled signals: this µs's max temperature / critical count
    }
    if (!HOT) crit_count = 0;
    return n;
}

// run_quiet_t in the form this env's stencil uses (`diff`, uniform over the env's warps)
template <int H, bool HOT, bool ALIGNED, int W, bool LOG, class LogTick, class SigCap>
__device__ __forceinline__ int run_quiet(const bool diff, float2 (&T)[H], const Coefs<H, ALIGNED> &co, const float cRf,
                                         const float cLf, const float spoolf, const LaneFlags &f, EnvCtx<W> &ctx, const int n,
                                         const float crit_T, const float break_T, int &crit_count, bool &broke,
                                         bool &cold_after, LogTick &log_tick, SigCap &sig_cap)
{
    if (diff)
        return run_quiet_t<H, HOT, ALIGNED, W, LOG, true>(T, co, cRf, cLf, spoolf, f, ctx, n, crit_T, break_T, crit_count, broke,
                                                          cold_after, log_tick, sig_cap);
    return run_quiet_t<H, HOT, ALIGNED, W, LOG, false>(T, co, cRf, cLf, spoolf, f, ctx, n, crit_T, break_T, crit_count, broke,
                                                       cold_after, log_tick, sig_cap);
}

// resident blocks per SM the register budget is sized for (64-thread blocks: 8 -> 16 warps/SM).  Measured in
// round 2 (profiles/r02_occupancy.txt): the kernel is bound by instruction dispatch, not by latency — 16, 24, 28
// and 32 warps/SM run at the same cycles per env-us — so registers are not traded for occupancy.
#ifndef SPARC_MINB_SMALL
#define SPARC_MINB_SMALL 8
#endif
constexpr int min_blocks_for(int K, bool aligned)
{
    return ((K <= 20 || aligned) ? SPARC_MINB_SMALL : 6) * 64 / SPARC_BLOCK_THREADS;
}

#define SPARC_LN2_UP 0.6931471805599460    /* ln2*(1+2^-50): u3 < fl(ln2/den) implies u3*den <= this */

// Wait until the mechanics pre-pass of THIS launch (tag `seq`) has published at least `need` blocks of
// the env whose progress word is *pp; returns the number published, or 0 after a bounded spin (the
// caller then computes the trajectory itself, mech_self_serve).  Out of line: it runs a handful of times
// per launch and must not cost the step kernel registers.
static __device__ __noinline__ uint32_t mech_wait(const uint32_t *pp, const uint32_t seq, const uint32_t need)
{
    unsigned spins = 0;
    for (;;) {
        uint32_t pv;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(pv) : "l"(pp) : "memory");
        const uint32_t d = pv - seq;
        if (d < 64u && d >= need) return d;
        __nanosleep(100);
        if (++spins > (1u << 21)) return 0;     // ~0.3 s: do not hang and do not kill the context
    }
}

// The hand-off timed out (a producer that never ran: debugger, time-slicing, a failed launch): this
// warp evaluates the recurrence of its own env up to block `need`, from the launch's initial state,
// with the same operation sequence as mech_kernel's exact path, and writes the same scratch.  Both
// writers produce identical values, so a slow-but-alive producer does no harm.
static __device__ __noinline__ void mech_self_serve(const SparcParams &P, const SparcStatePtrs &S, const StepArgs &a,
                                                    EnvState &s, const uint32_t need)
{
    const int e = s.e;
    double *traj = S.mech_traj + (size_t)e * SPARC_MECH_EPOCH;
    double *trajv = a.signals ? S.mech_trajv + (size_t)e * SPARC_MECH_EPOCH : nullptr;
    double *ck = S.mech_ckpt + (size_t)e * SPARC_MECH_NCKPT * 2;
    double wire = s.wire0, v = S.v[e], pa = S.prev_accel[e];
    ck[0] = v; ck[1] = pa;
    const int upto = ((int)need * SPARC_MECH_CKPT < a.n_us) ? (int)need * SPARC_MECH_CKPT : a.n_us;
    for (int k = 0; k < upto; k++) {
        mech_step(P, s.q, wire, v, pa, (k >= s.first_latch_it) ? s.delta_new0 : s.delta_old0);
        traj[k] = wire;
        if (trajv) trajv[k] = v;
        if ((k + 1) % SPARC_MECH_CKPT == 0) { ck[2 * ((k + 1) / SPARC_MECH_CKPT)] = v; ck[2 * ((k + 1) / SPARC_MECH_CKPT) + 1] = pa; }
        if (k + 1 == a.n_us) { ck[2 * (SPARC_MECH_NCKPT - 1)] = v; ck[2 * (SPARC_MECH_NCKPT - 1) + 1] = pa; }
    }
    __threadfence();
    __syncwarp();
    s.mech_self = need;
    if (a.mech_seq) s.flags |= SPARC_F_HANDOFF_SLOW;
}

// One environment, start to end of the launch, on the calling warp (W == 1) or block (W > 1).
template <int K, bool INJECT, int W, bool LOG, bool ALIGNED, int WARPS>
__device__ __forceinline__ void step_env(const SparcParams &P, const SparcStatePtrs &S, const StepArgs &a, const int e,
                                         float (&sm_halo)[2][W][2], double (&sm_red)[W], WarpRec (&sm_rec)[WARPS],
                                         float *sm_stage)
{
    constexpr unsigned FULL = 0xffffffffu;
    static_assert(K % 2 == 0 && K >= 4, "packed stencil: even number of cells per lane, at least two per half");
    constexpr int H = K / 2;
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
#ifdef SPARC_DBG_TIMES
    unsigned long long dbg_t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
#endif
    EnvCtx<W> ctx;
    ctx.lane = lane; ctx.wid = (W == 1) ? 0 : wib; ctx.gl = ctx.wid * 32 + lane;
    ctx.halo = sm_halo; ctx.parity = 0;
    const int gl = ctx.gl;                      // lane within the env
    WarpRec &R = sm_rec[wib];
    EnvState &s = R.s;
    const uint32_t k0 = (uint32_t)a.seed, k1 = (uint32_t)(a.seed >> 32);

    // ---------------- load: scalar state -> the warp's shared-memory record ---------------------
    // every lane executes the same scalar code and writes the same values (no divergence, no sync)
    s.e = e; s.it = 0; s.writer = gl == 0;
    load_env_prm(s.q, S, e);
    s.env_id = a.env_id_base + (uint32_t)e;
    s.episode = (uint32_t)S.episode[e];
    s.debris_density = S.debris_density[e]; s.flow = S.flow[e];
    s.d_last_gap = S.d_last_gap[e]; s.d_last_density = S.d_last_density[e];
    s.w_last_flow = S.w_last_flow[e];
    s.rand_short_rem = S.rand_short_rem[e]; s.deb_short_rem = S.deb_short_rem[e];
    s.current_mode = S.current_mode[e];
    // the per-call window (mean voltage, sparks, short µs of the last step call) carries over the
    // launches of one call (it_base > 0: a later mechanics epoch / sampling chunk of the same call)
    const bool cont = a.it_base > 0;
    s.win_sparks = cont ? S.win_sparks[e] : 0;
    s.win_short0 = cont ? S.win_short_us[e] : 0;
    s.win_steps0 = cont ? S.win_steps[e] : 0;
    s.win_short_us = s.win_short0;
    s.dens_valid = true; s.pend_flush = false; s.fast_ok = false;
    s.tsi_prev = 0; s.plasma_idx = -1;
    s.heat_on = false; s.thermal = false; s.refresh = false; s.freeze = false; s.became_done = false;
    s.jA = 0.f; s.jB = 0.f; s.fP = 0.f; s.hb = 0.f; s.he = 0.f;
    s.gap_lo_f = 0.0; s.gap_hi_f = -1.0; s.lo_c = 0.0; s.hi_c = 0.0; s.flush_amt = 0.0; s.onoff_thr = 0;
    s.ign_latched = S.ign_cached_mode[e] != 0;
    derive_generator(s, P, S.target_voltage[e], S.on_time[e], S.off_time[e]);
    s.wp = S.wp[e]; s.wire = S.wire[e]; s.v = S.v[e]; s.prev_accel = S.prev_accel[e];
    s.fin_wire = s.wire; s.fin_v = s.v; s.fin_pa = s.prev_accel;
    s.target_delta = S.target_delta[e]; s.voltage = S.voltage[e]; s.debris_volume = S.debris_volume[e];
    s.win_vsum = cont ? S.win_vsum[e] : 0.0;
    s.gap_prev = s.wp - s.wire;
    s.time = S.time[e]; s.spark_state = S.spark_state[e]; s.spark_dur = S.spark_dur[e]; s.flags = S.flags[e];
    s.crit_count = S.crit_count[e];
    s.w_zone_counter = S.w_zone_counter[e];
    s.avg_view = P.compute_zone_mean ? S.w_last_zone_mean[e] : S.wire_avg_T[e];
    s.time0 = s.time;
    s.latch_time = s.time - S.t_servo[e];
    s.tse_ref = s.time - S.t_spark_end[e];
    s.ion_end = s.time + S.ion_rem[e];
    if (s.spark_state == 1 && (s.flags & SPARC_F_HAS_Y))
        s.plasma_idx = (P.segment_len != 0.0) ? P.zone_start + (int)py_floordiv(S.spark_y[e], P.segment_len) : P.zone_start;

    const int N = P.n_segments;
    float spoolf = (float)P.spool_T;
    asm volatile("" : "+f"(spoolf));            // keep it in a register (no per-µs F2F rematerialisation)
    const LaneGeom geom = lane_geom(a.lay, gl, H);
    LaneFlags lf;
    lf.first = gl == 0;
    lf.short_x = geom.nx == H - 1; lf.short_y = geom.ny == H - 1;
    lf.empty_x = geom.nx == 0; lf.empty_y = geom.ny == 0;
    lf.end_x = (2 * gl == a.lay.hb[3] - 1); lf.end_y = (2 * gl + 1 == a.lay.hb[3] - 1);
    float *Trow = S.T + (size_t)e * S.n_pad;

    if (s.flags & SPARC_F_DONE) {               // terminated envs are left untouched
        if (s.writer && !cont) { S.win_vsum[e] = 0.0; S.win_sparks[e] = 0; S.win_short_us[e] = 0; S.win_steps[e] = 0; }
        const bool frozen_rows = LOG && a.signals != nullptr && a.sample_every > 0 && a.it_base == 0;
        if (a.obs || frozen_rows) {             // observation / signal rows of a frozen env, from the stored state
            float m = -3.0e38f;
            for (int c = gl; c < N; c += 32 * W) m = fmaxf(m, Trow[c]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL, m, o));
            if constexpr (W > 1) {
                if (lane == 0) sm_red[ctx.wid] = (double)m;
                __syncthreads();
                for (int w = 0; w < W; w++) m = fmaxf(m, (float)sm_red[w]);
            }
            if (s.writer && frozen_rows) {      // every sample of this call sees the frozen state (the reference's
                const size_t ne = (size_t)a.n_envs;                    // logger would read the same EDMState again)
                const double vals[SPARC_N_SIGNALS] = {
                    (double)s.time, s.voltage, S.current[e], s.wire, s.v, s.wp, s.target_delta, s.debris_density, s.flow,
                    (s.flags & SPARC_F_SHORT) ? 1.0 : 0.0, (double)s.spark_state, (double)m, S.wire_avg_T[e], s.debris_volume,
                    (double)s.crit_count, (double)s.flags};
                for (int r = 0; r < a.total_us / a.sample_every; r++)
                    for (int k = 0; k < SPARC_N_SIGNALS; k++) a.signals[((size_t)r * SPARC_N_SIGNALS + k) * ne + e] = vals[k];
            }
            if (s.writer && a.obs) {
                const size_t ne = (size_t)a.n_envs;
                const int steps = cont ? s.win_steps0 : 0;
                a.obs[0 * ne + e] = (float)(s.wp - s.wire);
                a.obs[1 * ne + e] = (float)s.v;
                a.obs[2 * ne + e] = steps > 0 ? (float)(s.win_vsum / (double)steps) : 0.0f;
                a.obs[3 * ne + e] = (float)s.win_sparks;
                a.obs[4 * ne + e] = (float)s.win_short_us;
                a.obs[5 * ne + e] = m;
                a.obs[6 * ne + e] = (float)s.debris_density;
                a.obs[7 * ne + e] = (float)s.flags;
                a.obs[8 * ne + e] = __int_as_float(s.time);
                a.obs[9 * ne + e] = __int_as_float(s.spark_state);
                a.obs[10 * ne + e] = __int_as_float(s.time - s.latch_time);
                a.obs[11 * ne + e] = 0.0f;
            }
        }
        return;
    }

    s.next_latch = s.latch_time + P.servo_interval;
    s.ran_fast = false; s.zone_mean_written = false; s.done = false;
    s.skip_flush = 0;                           // 1: the next quiet run starts with no pending debris flush
    s.tsi_break = -1;
    s.rng_valid = false; s.rng_cbase = 0;
    s.mech_ready = 0; s.mech_self = 0;

    // upper bound of lambda(gap) = ln2/(a g^2 + b g + c) over all gaps (ignition.py:348-364): a
    // draw above it can never ignite, whatever the gap.
    {
        double lam_bound = 2.0;
        if (s.q.ignition_a_coeff > 0.0) {
            double den_min = s.q.ignition_c_coeff - s.q.ignition_b_coeff * s.q.ignition_b_coeff / (4.0 * s.q.ignition_a_coeff);
            if (den_min > 0.0) lam_bound = 0.6931471805599453 / den_min * (1.0 + 1e-9);
        }
        uint32_t w_thresh = 0xffffffffu;        // word >= w_thresh  =>  u01(word) >= lam_bound
        if (lam_bound < 0.999) w_thresh = (uint32_t)(lam_bound * 4294967296.0) + 2u;
        s.lam_bound = lam_bound; s.w_thresh = w_thresh;
    }

    // ---------------- temperature field -> registers (region layout) --------------------------
    float2 T[H];                                // T[j] = (slot j of the x half, slot j of the y half)
#pragma unroll
    for (int j = 0; j < H; j++) {
        T[j].x = (j < geom.nx) ? Trow[geom.cx0 + j] : spoolf;
        T[j].y = (j < geom.ny) ? Trow[geom.cy0 + j] : spoolf;
    }

    // ---------------- stencil coefficients ---------------------------------------------------
    // (forms: stencil_apply.)  Four-instruction form: q = cC*T + c0 + cR*T_next + cL*T_prev, cC = cE - (cR + cL).
    // The three conduction weights must cancel EXACTLY on a uniform field, otherwise every µs adds
    // (cC + cR + cL - cE)*T ~ 1e-10 * T of the same sign to every cell.  So: csum = fl(cR + cL); the smaller of
    // the two weights is replaced by csum - larger (exact in fp32; a relative change of <= 2^-24 * larger/smaller
    // of one conduction weight); cC = fl(cE - csum), and the convection weight the update really applies,
    // cC + csum (exact; differs from cE by <= ulp(csum)/2, ~1e-6 of cE), is the one that c0 = -(cC + csum)*T_diel
    // is formed with, in fp64.  Difference form (fine segmentations, csum > 2^-6): the plain weights.
    const double f = P.temp_update_factor;
    float cRf, cLf, csum;
    bool diff_form;
    {
        const double unwind_v = S.unwind_v[e];
        double adv = 0.0;                       // wire.py:299-309
        if (fabs(unwind_v) > 1e-6) adv = P.density_times_cp * fabs(unwind_v) * P.S;
        if (!(fabs(adv) > 1e-9)) adv = 0.0;     // wire.py:114
        cRf = P.cond_coeff_f32;                  // (float)(f * k_cond_coeff), rounded on the host
        cLf = (float)(f * (P.k_cond_coeff + adv));
        csum = cRf + cLf;
        diff_form = !(csum <= 0.015625f);
#ifdef SPARC_FORCE_DIFF_FORM
        diff_form = true;
#endif
        if (diff_form) cLf = -cLf;              // stencil_apply<DIFF> multiplies T - T_prev
        else if (fabsf(cLf) >= fabsf(cRf)) cRf = csum - cLf;
        else cLf = csum - cRf;
    }
    const float gate_T = fminf(P.critical_temperature_f32, P.breaking_temperature_f32);
    Coefs<H, ALIGNED> co;
    auto conv_pair = [&](float h, float &e_, float &z_) {
        const double hA = (double)h * P.A;      // conv_coeff, wire.py:110
        if (diff_form) {
            e_ = (float)(-(f * hA));
            z_ = (float)(f * hA * P.dielectric_temperature);
        } else {
            e_ = (float)(-(f * hA)) - csum;
            z_ = (float)(-((double)e_ + (double)csum) * P.dielectric_temperature);
        }
    };
    auto refresh_coeffs = [&](float hb, float he) {
        // slots that hold no cell (pads, empty halves) get h = 0: pure conduction, bounded
        if constexpr (ALIGNED) {
            conv_pair(geom.rx >= 0 ? (geom.rx == 1 ? he : hb) : 0.f, co.cC[0].x, co.c0[0].x);
            conv_pair(geom.ry >= 0 ? (geom.ry == 1 ? he : hb) : 0.f, co.cC[0].y, co.c0[0].y);
        } else {
#pragma unroll
            for (int j = 0; j < H; j++) {
                const int ix = geom.cx0 + j, iy = geom.cy0 + j;
                conv_pair(j < geom.nx ? ((ix >= P.actual_zone_start && ix < P.actual_zone_end) ? he : hb) : 0.f, co.cC[j].x, co.c0[j].x);
                conv_pair(j < geom.ny ? ((iy >= P.actual_zone_start && iy < P.actual_zone_end) ? he : hb) : 0.f, co.cC[j].y, co.c0[j].y);
            }
        }
        const float hmax = fmaxf(hb, he), hmin = fminf(hb, he);
        const float ce_max = (float)(f * ((double)hmax * P.A));
        // T' is a convex combination of T and T_diel when no source is on: the maximum cannot grow
        s.convex = hmin >= 0.0f && cRf >= 0.0f && csum >= cRf && (csum + ce_max) <= 0.999f &&
                   fmaxf((float)P.dielectric_temperature, spoolf) < gate_T - 1.0f;
    };
    refresh_coeffs(S.h_base[e], S.h_enh[e]);
    s.cold_ok = false;                          // max T < gate - 1 K after the last executed µs

    const double *drow = INJECT ? a.draws + ((size_t)e * a.total_us + a.it_base) * SPARC_N_DRAW_SLOTS : nullptr;

    // temperature-field logging (SimulationLogger "wire_temperature", utils/logger.py:128-150): row r
    // holds the field after µs (r+1)*log_every - 1 of this call, layout (n_timesteps, n_segments)
    const bool logging = LOG && a.tlog != nullptr && e < a.n_log_envs && a.log_every > 0;
    const int log_rows = logging ? a.total_us / a.log_every : 0;
    int log_ctr = logging ? a.log_every - a.it_base % a.log_every : 0;  // µs until the next sample
    int log_row = logging ? a.it_base / a.log_every : 0;                // next row to write
    // Row staging for the log (LOG kernels only; dynamic shared memory): NBUF buffers of n_stage floats per env,
    // the whole row in cell order.  When the rows are 16-byte aligned in the log (N % 4 == 0) a row leaves as ONE
    // bulk asynchronous copy shared -> global (cp.async.bulk, SASS UBLKCP) issued by one thread, double buffered
    // so that the next µs's stencil does not wait for it; otherwise as coalesced 4-byte stores.
    const int n_stage = (N + 3) & ~3;
    constexpr int NBUF = SPARC_LOG_NBUF(W);    // staging buffers per env (host: same macro)
    float *stage0 = sm_stage + (size_t)((W == 1) ? wib : 0) * NBUF * n_stage;
    const bool bulk_rows = (N % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.tlog) & 15u) == 0);
    int log_buf = 0;
    auto group_sync = [&]() { if constexpr (W > 1) __syncthreads(); else __syncwarp(); };
    // shared-memory byte addresses of this lane's two halves inside staging buffer 0 (buffer 1 is n_stage floats on)
    const uint32_t st_x = (uint32_t)__cvta_generic_to_shared(stage0) + 4u * (uint32_t)geom.cx0;
    const uint32_t st_y = (uint32_t)__cvta_generic_to_shared(stage0) + 4u * (uint32_t)geom.cy0;
    auto write_log_row = [&](int r) {
        const uint32_t boff = 4u * (uint32_t)(log_buf * n_stage);
        if (bulk_rows && W == 1) {
            // the bulk copy issued two rows ago has finished READING this buffer (W > 1: the issuing thread waited
            // right after its last commit, and every thread has passed a halo barrier since)
            if (gl == 0) asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(NBUF - 1) : "memory");
            group_sync();
        }
        // one base register per half, immediate offsets (the generic form recomputed every address and every
        // `j < n` predicate: 3 instructions per stored cell, profiles/r02_experiments.txt §6)
        if (!lf.empty_x) {
            const uint32_t ax = st_x + boff;
#pragma unroll
            for (int j = 0; j < H - 1; j++) asm volatile("st.shared.f32 [%0], %1;" :: "r"(ax + 4u * j), "f"(T[j].x) : "memory");
            if (!lf.short_x) asm volatile("st.shared.f32 [%0], %1;" :: "r"(ax + 4u * (H - 1)), "f"(T[H - 1].x) : "memory");
        }
        if (!lf.empty_y) {
            const uint32_t ay = st_y + boff;
#pragma unroll
            for (int j = 0; j < H - 1; j++) asm volatile("st.shared.f32 [%0], %1;" :: "r"(ay + 4u * j), "f"(T[j].y) : "memory");
            if (!lf.short_y) asm volatile("st.shared.f32 [%0], %1;" :: "r"(ay + 4u * (H - 1)), "f"(T[H - 1].y) : "memory");
        }
        float *stage = stage0 + log_buf * n_stage;
        if (gl == 0) stage[0] = spoolf;
        if (bulk_rows) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> async proxy
            group_sync();
            if (gl == 0) {
                float *row = a.tlog + ((size_t)e * log_rows + r) * N;          // only the issuing thread needs it
                const uint32_t src = (uint32_t)__cvta_generic_to_shared(stage);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n\tcp.async.bulk.commit_group;"
                             :: "l"(row), "r"(src), "r"((uint32_t)(N * 4)) : "memory");
                if (W > 1) asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(NBUF - 1) : "memory");   // frees the NEXT buffer
            }
            log_buf = (log_buf + 1 == NBUF) ? 0 : log_buf + 1;
        } else {
            float *row = a.tlog + ((size_t)e * log_rows + r) * N;
            group_sync();
            for (int c = ((W == 1) ? lane : (int)threadIdx.x); c < N; c += 32 * W) row[c] = stage[c];
            group_sync();
        }
    };
    auto log_tick = [&]() {                     // called after every thermal update of a logged env
        if (!logging) return;
        if (--log_ctr == 0) {
            if (log_row < log_rows) write_log_row(log_row);
            log_row += 1;
            log_ctr = a.log_every;
        }
    };

    // per-signal log (SimulationLogger scalar signals, utils/logger.py:105-160): row r of the CALL holds what the
    // reference's `getattr(state, name)` returns after µs (r+1)*sample_every - 1.  Written by this kernel: quiet runs
    // are cut at the sampling µs, the lazily kept dielectric state is read through dielectric_view().
    __shared__ float sm_cap[(LOG && W > 1) ? W : 1][32];   // W > 1: per-warp maxima of a run's µs, combined after the run
    const bool sig_on = LOG && a.signals != nullptr && a.sample_every > 0;
    const int sig_rows = sig_on ? a.total_us / a.sample_every : 0;
    int done_idx = 0;                           // launch-µs index of the terminating µs (rows from there on: final state)
    float cap_max = spoolf;                     // quiet runs: lane i + k keeps µs k's max temperature (this warp's cells)
    int cap_crit = 0;                           //             and time_in_critical_temp after it
    int cap_base = 0;                           // block index of the run's first µs
    int cap_ctr = 1;                            // µs until the next sampled µs of the run (1 = the next one)
    auto sig_cap = [&](int k, int crit) {
        if (!(LOG && sig_on)) return;
        if (--cap_ctr != 0) return;             // not a sampled µs
        cap_ctr = a.sample_every;
        // temperatures are positive: unsigned order of the bit patterns is float order -> one REDUX instruction
        const float m = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(lane_max<H>(T, lf, spoolf))));
        if (lane == cap_base + k) { cap_max = m; cap_crit = crit; }
    };
    auto env_max_T = [&]() {                    // max over the env's cells; every thread of the env calls it
        float m = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(lane_max<H>(T, lf, spoolf))));   // cell 0 is spoolf
        if constexpr (W > 1) {
            __syncthreads();
            if (lane == 0) sm_red[ctx.wid] = (double)m;
            __syncthreads();
            for (int w = 0; w < W; w++) m = fmaxf(m, (float)sm_red[w]);
        }
        return m;
    };
    auto sig_write = [&](int r, double time_v, double voltage, double current, double wire, double v, double dens, double vol,
                         int flags, int spark_state, float tmax, double avg_T, int crit) {
        if (gl != 0 || r >= sig_rows) return;
        const size_t ne = (size_t)a.n_envs;
        double *row = a.signals + (size_t)r * SPARC_N_SIGNALS * ne + e;
        row[SPARC_SIG_TIME * ne] = time_v;
        row[SPARC_SIG_VOLTAGE * ne] = voltage;
        row[SPARC_SIG_CURRENT * ne] = current;
        row[SPARC_SIG_WIRE_POSITION * ne] = wire;
        row[SPARC_SIG_WIRE_VELOCITY * ne] = v;
        row[SPARC_SIG_WORKPIECE_POSITION * ne] = s.wp;
        row[SPARC_SIG_TARGET_DELTA * ne] = s.target_delta;
        row[SPARC_SIG_DEBRIS_DENSITY * ne] = dens;
        row[SPARC_SIG_FLOW_RATE * ne] = s.flow;
        row[SPARC_SIG_IS_SHORT_CIRCUIT * ne] = (flags & SPARC_F_SHORT) ? 1.0 : 0.0;
        row[SPARC_SIG_SPARK_STATE * ne] = (double)spark_state;
        row[SPARC_SIG_MAX_TEMPERATURE * ne] = (double)tmax;
        row[SPARC_SIG_AVERAGE_TEMPERATURE * ne] = avg_T;
        row[SPARC_SIG_DEBRIS_VOLUME * ne] = vol;
        row[SPARC_SIG_TIME_IN_CRITICAL * ne] = (double)crit;
        row[SPARC_SIG_FLAGS * ne] = (double)flags;
    };
    auto sig_emit = [&](int launch_us, double wire, double v) { // the state after launch-µs `launch_us` (a general µs)
        const int cu = a.it_base + launch_us + 1;
        if (cu % a.sample_every != 0) return;                // uniform over the env
        const float m = env_max_T();
        double dens, vol;
        dielectric_view(s, P, dens, vol);
        const int st = s.spark_state;
        sig_write(cu / a.sample_every - 1, (double)s.time, s.voltage, (st == 1 || st == -1) ? s.ipk : 0.0, wire, v, dens, vol,
                  s.flags, st, m, s.avg_view, s.crit_count);
    };

    // mechanics pre-pass results of this launch (mech_kernel).  The pre-pass is STILL RUNNING while this
    // kernel executes (programmatic dependent launch, sparc_b200.cu): positions and checkpoints are
    // read only after the producer thread of this env has published them through mech_prog
    // (st.release / ld.acquire at GPU scope), and with ld.global.cg — an L1 line fetched before the
    // producer finished it would be stale.  The producer is ~8x faster than this consumer, so after the
    // first blocks one poll covers many blocks (s.mech_ready caches the last value seen; the poll
    // itself is out of line, mech_wait).  A poll that times out makes this warp compute its own
    // trajectory (mech_self_serve) instead of trapping.
    const double *traj = S.mech_traj + (size_t)e * SPARC_MECH_EPOCH;
    auto need_mech = [&](int upto_us) {         // positions of launch-µs < upto_us and checkpoints up to there
        const uint32_t need = (uint32_t)((upto_us + SPARC_MECH_CKPT - 1) / SPARC_MECH_CKPT);
        if (need > s.mech_ready && need > s.mech_self) {
            // mech_seq == 0: a launch too short to be worth a second kernel (the drop-in's 1-µs step()):
            // there is no pre-pass, the warp evaluates the few mechanics steps itself
            const uint32_t got = a.mech_seq ? mech_wait((const uint32_t *)S.mech_prog + e, a.mech_seq, need) : 0u;
            if (got) s.mech_ready = got;
            else mech_self_serve(P, S, a, s, (uint32_t)((a.n_us + SPARC_MECH_CKPT - 1) / SPARC_MECH_CKPT));
        }
    };
    s.wire0 = s.wire; s.delta_old0 = s.target_delta; s.delta_new0 = servo_command(a, S, e);
    s.first_latch_it = (s.next_latch - s.time > 0) ? s.next_latch - s.time : 0;

    // finished inside a block: exact mechanics state after launch-µs `target`, replayed from the
    // pre-pass checkpoint (same operation sequence as mech_kernel)
    auto rewind_mech = [&](int target) {
        const int c = target / SPARC_MECH_CKPT;
        const double *ck = S.mech_ckpt + ((size_t)e * SPARC_MECH_NCKPT + c) * 2;
        double wire = (c == 0) ? s.wire0 : __ldcg(traj + c * SPARC_MECH_CKPT - 1);
        double v = __ldcg(ck), pa = __ldcg(ck + 1);
#pragma unroll 1
        for (int k = c * SPARC_MECH_CKPT; k < target; k++)
            mech_step(P, s.q, wire, v, pa, (k >= s.first_latch_it) ? s.delta_new0 : s.delta_old0);
        s.fin_wire = wire; s.fin_v = v; s.fin_pa = pa;
    };

    int it = 0, prev_nb = 0;
    double nxt_post = 0.0;
#ifdef SPARC_ABL_ALLQUIET
    {   // ablation: the whole launch as ONE cold quiet run (stencil speed inside this kernel)
        bool broke = false, cold_after = true; int cc = 0;
        run_quiet<H, false, ALIGNED, W, false>(diff_form, T, co, cRf, cLf, spoolf, lf, ctx, a.n_us, 0.f, 0.f, cc, broke, cold_after, log_tick, sig_cap);
        s.time += a.n_us; it = a.n_us; need_mech(a.n_us);
    }
#endif
    while (it < a.n_us && !s.done) {
        // ================= block setup ==========================================================
        int nb = a.n_us - it;
        if (nb > 32) nb = 32;
        bool force_general = false;
        {
            int to_latch = s.next_latch - s.time;
            if (to_latch <= 0) {                // the first µs of this block latches the action
                to_latch = (P.servo_interval > 1) ? P.servo_interval : 1;
                force_general = true;
            }
            if (nb > to_latch) nb = to_latch;   // a block never straddles a latch
        }
        // wire positions of this block from the mechanics pre-pass: pos[0] before its first µs,
        // pos[1 + l] after µs l.  They were requested one block AGO (nxt_post: the L2 round trip of
        // ld.global.cg passes under the previous block's stencil runs; it was the largest single stall
        // of the scalar phase, profiles/r02_experiments.txt); only the first block of a launch waits.
        {
            int upto = it + nb + 32;
            if (upto > a.n_us) upto = a.n_us;
            need_mech(upto);                    // this block and the next
        }
        {
            double cur_post, p0;
            if (it == 0) {
                cur_post = __ldcg(traj + ((lane < a.n_us) ? lane : a.n_us - 1));
                p0 = s.wire0;
            } else {
                cur_post = nxt_post;            // requested with index it + lane during the previous block
                p0 = R.pos[prev_nb];            // wire after the last µs of the previous block
            }
            __syncwarp();
            R.pos[lane + 1] = cur_post;
            if (lane == 0) R.pos[0] = p0;
            prev_nb = nb;
        }
        if (sig_on) {                           // velocities after each µs of this block (published with the positions)
            __syncwarp();
            R.vel[lane + 1] = __ldcg(S.mech_trajv + (size_t)e * SPARC_MECH_EPOCH + it + ((lane < nb) ? lane : nb - 1));
        }
        const int blk_time = s.time;
        // ignition rolls (slot 2): word t&3 of Philox block t>>2; each lane computes one counter of a
        // 32-counter window and keeps all four words: one refill serves 125..128 µs
        if (!INJECT) {
            const uint32_t c_lo = (uint32_t)blk_time >> 2, c_hi = (uint32_t)(blk_time + nb - 1) >> 2;
            if (!s.rng_valid || c_lo < s.rng_cbase || c_hi - s.rng_cbase >= 32u) {
                uint32_t w[4];
                philox4x32_10(s.env_id, c_lo + (uint32_t)lane, s.episode, 2u, k0, k1, w);
                *reinterpret_cast<uint4 *>(R.rng[lane]) = make_uint4(w[0], w[1], w[2], w[3]);
                s.rng_cbase = c_lo; s.rng_valid = true;
            }
        }
        __syncwarp();
        {   // next block's positions: requested only now, AFTER this block's were stored.  Issued before that store,
            // the load took the scoreboard the store of cur_post waits on, and the store then sat out the whole L2
            // round trip of a load it does not depend on (7.9 % of the kernel's stall samples on that one STS,
            // profiles/r02b_cfg2_line_profile.txt).
            int idx = it + nb + lane;
            if (idx > a.n_us - 1) idx = a.n_us - 1;
            nxt_post = __ldcg(traj + idx);
        }

        int i = 0;
        while (i < nb) {
            // ================= look-ahead: length of the quiet run starting at µs i ===========
            int n = 0;
            int off_tr = -1;                    // run offset of the rest -> idle transition µs
            const int spark_state = s.spark_state;
            if (!force_general && s.fast_ok && (spark_state == 0 || spark_state == -2)) {
                const double pre_l = R.pos[lane], post_l = R.pos[lane + 1];
                const double wp = s.wp, debris_volume = s.debris_volume;
                const double g = wp - pre_l;
                const int off = lane - i;
                // band conditions: the hysteresis bands of the last flow recomputation
                auto in_band = [&]() {
                    bool b = g >= s.gap_lo_f && g <= s.gap_hi_f;
                    // debris after at most off+1 deferred flushes (sequential rounding <= 32 ulp(1))
                    double deb_lo = debris_volume;
                    if (debris_volume > 0.001) deb_lo = debris_volume - (double)(off + 1) * s.flush_amt - 1e-14;
                    return b && deb_lo >= s.lo_c * g && debris_volume <= s.hi_c * g;
                };
                // everything else that makes a µs quiet
                bool ok = lane < nb && !(post_l > wp + 100.0);  // collision µs (wire_edm.py:173-175)
                if (spark_state == -2) {
                    off_tr = s.onoff_thr - s.spark_dur - 1;
                    if (off_tr < 0) off_tr = 0;
                }
                if (P.compute_zone_mean) ok = ok && (s.w_zone_counter + off + 1 < P.zone_mean_interval);
                if (INJECT) {
                    if (lane < nb) {
                        const double *d = drow + (size_t)(it + lane) * SPARC_N_DRAW_SLOTS;
                        ok = ok && d[0] >= SPARC_U1_SAFE;
                        if (off > off_tr) {
                            const double u3 = d[2];
                            if (!(u3 >= s.lam_bound)) {
                                double den = s.q.ignition_a_coeff * (g * g) + s.q.ignition_b_coeff * g + s.q.ignition_c_coeff;
                                ok = ok && (den > 0.0 && u3 * den > SPARC_LN2_UP);
                            }
                        }
                    }
                } else if (off > off_tr) {
                    const uint32_t tl = (uint32_t)(blk_time + lane);
                    const uint32_t w_l = R.rng[((tl >> 2) - s.rng_cbase) & 31u][tl & 3u];
                    if (w_l < s.w_thresh) {
                        // u3 < fl(ln2/den) can only hold if u3*den <= ln2*(1+2^-50)
                        const double u3 = u01(w_l);
                        double den = s.q.ignition_a_coeff * (g * g) + s.q.ignition_b_coeff * g + s.q.ignition_c_coeff;
                        ok = ok && (den > 0.0 && u3 * den > SPARC_LN2_UP);
                    }
                }
                unsigned stop = __ballot_sync(FULL, !(ok && in_band())) >> i;
                n = stop ? __ffs(stop) - 1 : 32 - i;
                if (n == 0 && ((__ballot_sync(FULL, ok) >> i) & 1u)) {
                    // only a band failed on µs i: refresh the hysteresis state in place (light_refresh)
                    const bool pending = s.ran_fast || s.pend_flush;
                    if (light_refresh(s, P, S, wp - R.pos[i], s.debris_volume, pending)) {
                        s.ran_fast = false;
                        s.skip_flush = 1;       // the refreshed µs has no pending flush before it
                        if (s.fast_ok) {
                            const double dv2 = s.debris_volume;
                            bool b = g >= s.gap_lo_f && g <= s.gap_hi_f;
                            double deb_lo = dv2;
                            if (dv2 > 0.001) deb_lo = dv2 - (double)(off + 1) * s.flush_amt - 1e-14;
                            b = b && deb_lo >= s.lo_c * g && dv2 <= s.hi_c * g;
                            stop = __ballot_sync(FULL, !(ok && b)) >> i;
                            n = stop ? __ffs(stop) - 1 : 32 - i;
                        }
                    }
                }
            }
            force_general = false;

            if (n > 0) {
                // ================= quiet run ===================================================
                bool broke = false, cold_after = s.cold_ok;
                int nx, crit_count = s.crit_count;
                cap_base = i;
                if (LOG && sig_on) cap_ctr = a.sample_every - (a.it_base + it + i) % a.sample_every;
                const bool cold = s.cold_ok && s.convex;
                if (!LOG || !(logging || sig_on)) {
                    if (cold)
                        nx = run_quiet<H, false, ALIGNED, W, false>(diff_form, T, co, cRf, cLf, spoolf, lf, ctx, n, 0.f, 0.f, crit_count,
                                                                    broke, cold_after, log_tick, sig_cap);
                    else
                        nx = run_quiet<H, true, ALIGNED, W, false>(diff_form, T, co, cRf, cLf, spoolf, lf, ctx, n, P.critical_temperature_f32,
                                                                   P.breaking_temperature_f32, crit_count, broke, cold_after, log_tick, sig_cap);
                } else {
                    if (cold)
                        nx = run_quiet<H, false, ALIGNED, W, LOG>(diff_form, T, co, cRf, cLf, spoolf, lf, ctx, n, 0.f, 0.f, crit_count, broke,
                                                                  cold_after, log_tick, sig_cap);
                    else
                        nx = run_quiet<H, true, ALIGNED, W, LOG>(diff_form, T, co, cRf, cLf, spoolf, lf, ctx, n, P.critical_temperature_f32,
                                                                 P.breaking_temperature_f32, crit_count, broke, cold_after, log_tick, sig_cap);
                }
                s.crit_count = crit_count;
                s.cold_ok = cold_after;
                if (LOG && sig_on) {
                    // signal rows of this run, one µs per lane (lane i + l owns run offset l), from the state BEFORE the
                    // bookkeeping below: what the reference holds after each of these quiet step() calls
                    float m_l = cap_max;
                    if constexpr (W > 1) {      // combine the warps' per-µs maxima
                        sm_cap[ctx.wid][lane] = cap_max;
                        __syncthreads();
#pragma unroll
                        for (int w = 0; w < W; w++) m_l = fmaxf(m_l, sm_cap[w][lane]);
                        __syncthreads();
                    }
                    const int l = lane - i, n_rows = broke ? nx - 1 : nx;       // the breaking µs: final state (done path)
                    const int cu = a.it_base + it + lane + 1;
                    if (ctx.wid == 0 && l >= 0 && l < n_rows && cu % a.sample_every == 0 && cu / a.sample_every - 1 < sig_rows) {
                        const bool idle_l = (spark_state == 0) || l >= off_tr;   // rest -> idle on run offset off_tr
                        const int flags_l = (spark_state == -2 && l >= off_tr) ? (s.flags & ~SPARC_F_HAS_Y) : s.flags;
                        const double fa = s.flush_amt;
                        double u = s.debris_volume;                              // volume before the pending flush
                        const int n_before = l + 1 - s.skip_flush;               // flushes applied before µs l's density
                        for (int t = 0; t < n_before; t++) if (fa > 0.0 && u > 0.001) u = u - fa;
                        const double gl_ = s.wp - R.pos[lane];
                        const double gq = (gl_ > 0.001) ? gl_ : 0.001;
                        const double cavity = P.cavity_volume_coeff * (gq * 0.001);
                        double dens = 0.0;
                        if (cavity > 0.0) { const double d = u / cavity; dens = (d < 1.0) ? d : 1.0; }
                        const double vol = (fa > 0.0 && u > 0.001) ? u - fa : u;
                        const size_t ne = (size_t)a.n_envs;
                        double *row = a.signals + (size_t)(cu / a.sample_every - 1) * SPARC_N_SIGNALS * ne + e;
                        row[SPARC_SIG_TIME * ne] = (double)(s.time + l + 1);
                        row[SPARC_SIG_VOLTAGE * ne] = idle_l ? s.tv_eff : 0.0;
                        row[SPARC_SIG_CURRENT * ne] = 0.0;
                        row[SPARC_SIG_WIRE_POSITION * ne] = R.pos[lane + 1];
                        row[SPARC_SIG_WIRE_VELOCITY * ne] = R.vel[lane + 1];
                        row[SPARC_SIG_WORKPIECE_POSITION * ne] = s.wp;
                        row[SPARC_SIG_TARGET_DELTA * ne] = s.target_delta;
                        row[SPARC_SIG_DEBRIS_DENSITY * ne] = dens;
                        row[SPARC_SIG_FLOW_RATE * ne] = s.flow;
                        row[SPARC_SIG_IS_SHORT_CIRCUIT * ne] = 0.0;
                        row[SPARC_SIG_SPARK_STATE * ne] = idle_l ? 0.0 : -2.0;
                        row[SPARC_SIG_MAX_TEMPERATURE * ne] = (double)m_l;
                        row[SPARC_SIG_AVERAGE_TEMPERATURE * ne] = s.avg_view;
                        row[SPARC_SIG_DEBRIS_VOLUME * ne] = vol;
                        row[SPARC_SIG_TIME_IN_CRITICAL * ne] = (double)cap_crit;
                        row[SPARC_SIG_FLAGS * ne] = (double)flags_l;
                    }
                }
                // scalar bookkeeping of the nx quiet µs (ignition.py:302-319, dielectric.py:104-144)
                int n_idle = nx;
                if (spark_state == -2) {
                    if (nx > off_tr) {          // rest -> idle on run offset off_tr
                        s.spark_state = 0; s.spark_dur = 0; s.flags &= ~SPARC_F_HAS_Y;
                        if (s.writer) S.spark_y[e] = -1.0;
                        n_idle = nx - off_tr;
                    } else { s.spark_dur += nx; n_idle = 0; }
                }
                s.voltage = (n_idle > 0) ? s.tv_eff : 0.0;
                s.win_vsum += (double)n_idle * s.tv_eff;
                {
                    double dv = s.debris_volume;
                    if (dv > 0.001 && s.flush_amt > 0.0) {
                        const double fa = s.flush_amt;
#pragma unroll 1
                        for (int k = s.skip_flush; k < nx && dv > 0.001; k++) dv = dv - fa;
                        s.debris_volume = dv;
                    }
                }
                s.skip_flush = 0;
                s.gap_prev = s.wp - R.pos[i + nx - 1];
                s.ran_fast = true;
                if (P.compute_zone_mean) s.w_zone_counter += nx;
                if (broke) {                    // wire_edm.py:129-130 on the last µs of the run
                    s.flags |= SPARC_F_BROKEN | SPARC_F_THERMAL_BREAK;
                    s.tsi_break = 0;
                    s.done = true;
                    s.time += nx - 1;
                    done_idx = it + i + nx - 1;
                    rewind_mech(it + i + nx - 1);
                    break;
                }
                s.time += nx;
                i += nx;
                continue;
            }

            // ================= general µs ======================================================
            s.wire = R.pos[i];
            s.it = it + i;
            if (s.ran_fast) { s.dens_valid = false; s.pend_flush = true; s.ran_fast = false; }
            s.skip_flush = 0;
            general_pre<INJECT>(s, P, S, a);
            s.next_latch = s.latch_time + P.servo_interval;
            if (s.refresh) refresh_coeffs(s.hb, s.he);
            if (s.freeze) {                     // crater-table miss: frozen before the thermal update
                s.done = true;
                done_idx = it + i;
                rewind_mech(it + i);
                break;
            }
            // ---- compute_thermal_update with sources (wire.py:58-123) -------------------------
            {
                float left, right, xl, yv;
                stencil_edges<H, W, !ALIGNED>(ctx, T, lf, spoolf, left, right, xl, yv);
                float2 acc[H];
#pragma unroll
                for (int j = 0; j < H; j++) acc[j] = __ffma2_rn(co.e(j), T[j], co.z(j));
                if (s.heat_on) {
                    const float jA = s.jA, jB = s.jB, fP = s.fP;
                    const int pidx = (s.spark_state == 1 && s.plasma_idx >= 0 && s.plasma_idx < N) ? s.plasma_idx : -(1 << 20);
                    // contact span and plasma cell in slot coordinates of each half
                    const int xlo = P.contact_bottom_idx - geom.cx0, xhi = P.contact_top_idx - geom.cx0, xp = pidx - geom.cx0;
                    const int ylo = P.contact_bottom_idx - geom.cy0, yhi = P.contact_top_idx - geom.cy0, yp = pidx - geom.cy0;
#pragma unroll
                    for (int j = 0; j < H; j++) {
                        if (j < geom.nx) {
                            if (j >= xlo && j <= xhi) acc[j].x += fmaf(jA, T[j].x, jB);
                            if (j == xp) acc[j].x += fP;
                        }
                        if (j < geom.ny) {
                            if (j >= ylo && j <= yhi) acc[j].y += fmaf(jA, T[j].y, jB);
                            if (j == yp) acc[j].y += fP;
                        }
                    }
                }
                if (diff_form) stencil_apply<H, true>(T, acc, make_float2(cRf, cRf), make_float2(cLf, cLf), left, right, xl, yv);
                else stencil_apply<H, false>(T, acc, make_float2(cRf, cRf), make_float2(cLf, cLf), left, right, xl, yv);
            }
            if (LOG && logging) log_tick();
            // ---- _check_wire_breaking (wire.py:376-388) ---------------------------------------
            bool hot = false, brk = false;
            {
                const float lmax = lane_max<H>(T, lf, spoolf);
                s.cold_ok = !vote_any<W>(lmax > gate_T - 1.0f);
                if (!s.cold_ok) {
                    hot = vote_any<W>(lmax > P.critical_temperature_f32);
                    brk = vote_any<W>(lmax > P.breaking_temperature_f32);
                }
            }
            s.crit_count = hot ? s.crit_count + 1 : 0;
            // zone mean (wire.py:338-347, 390-394), only when enabled
            if (P.compute_zone_mean) {
                s.w_zone_counter += 1;
                if (s.w_zone_counter >= P.zone_mean_interval) {
                    int za = P.actual_zone_start, zb = P.actual_zone_end;
                    if (!(zb > za && zb <= N)) { za = 0; zb = N; }
                    double tot = 0.0;
#pragma unroll
                    for (int j = 0; j < H; j++) {
                        const int cx = geom.cx0 + j;
                        if (j < geom.nx && cx >= za && cx < zb) tot += (double)T[j].x;
                    }
#pragma unroll
                    for (int j = 0; j < H; j++) {
                        const int cy = geom.cy0 + j;
                        if (j < geom.ny && cy >= za && cy < zb) tot += (double)T[j].y;
                    }
                    if (gl == 0 && za == 0) tot += (double)spoolf;  // cell 0 when the zone starts at 0
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(FULL, tot, o);
                    if constexpr (W > 1) {      // sum of the warps' partial sums, in warp order
                        if (lane == 0) sm_red[ctx.wid] = tot;
                        __syncthreads();
                        tot = 0.0;
#pragma unroll
                        for (int w = 0; w < W; w++) tot += sm_red[w];
                        __syncthreads();
                    }
                    const double zm = (double)(float)(tot / (double)(zb - za));
                    if (s.writer) { S.w_last_zone_mean[e] = zm; S.wire_avg_T[e] = zm; }
                    s.avg_view = zm;
                    s.zone_mean_written = true;
                    s.w_zone_counter = 0;
                }
            }
            if (brk) {
                // wire_edm.py:129-130: early return, no mechanics, no time bookkeeping
                s.flags |= SPARC_F_BROKEN | SPARC_F_THERMAL_BREAK;
                s.tsi_break = s.tsi_prev;
                s.done = true;
                done_idx = it + i;
                rewind_mech(it + i);
                break;
            }
            // mechanics of this µs is pos[i + 1]; time bookkeeping (wire_edm.py:134-144)
            s.time += 1;
            if (s.spark_state == 1) s.tse_ref = s.time;
            // ---- _check_termination (wire_edm.py:172-179) -------------------------------------
            if (R.pos[i + 1] > s.wp + 100.0) { s.flags |= SPARC_F_BROKEN; s.done = true; }
            else if (s.wp >= s.q.target_position) { s.flags |= SPARC_F_TARGET; s.done = true; }
            if (s.done) { done_idx = it + i; rewind_mech(it + i + 1); break; }
            i += 1;
            if (LOG && sig_on) sig_emit(it + i - 1, R.pos[i], R.vel[i]);
        }
        it += nb;
    }
    if (!s.done) {                              // end-of-launch mechanics state from the pre-pass
        s.fin_wire = __ldcg(traj + a.n_us - 1); // published: the last block of the loop waited for it
        s.fin_v = __ldcg(S.mech_ckpt + ((size_t)e * SPARC_MECH_NCKPT + (SPARC_MECH_NCKPT - 1)) * 2);
        s.fin_pa = __ldcg(S.mech_ckpt + ((size_t)e * SPARC_MECH_NCKPT + (SPARC_MECH_NCKPT - 1)) * 2 + 1);
    }
    if (LOG && logging && s.done) {             // rows sampled at or after termination: the frozen field
        for (int r = log_row; r < log_rows; r++) write_log_row(r);
    }
    if (LOG && logging && bulk_rows) {          // the staging buffers are reused by the next env of this warp / block
        if (gl == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        group_sync();
    }

#ifdef SPARC_DBG_TIMES   /* experiments: when and where each env ran, in the unused tail of its trajectory scratch */
    if (gl == 0 && a.n_us <= 1000) {
        unsigned long long t1; uint32_t smid, wid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        asm volatile("mov.u32 %0, %%warpid;" : "=r"(wid));
        double *dbg = S.mech_traj + (size_t)e * SPARC_MECH_EPOCH + 1016;
        dbg[0] = (double)dbg_t0; dbg[1] = (double)t1; dbg[2] = (double)smid; dbg[3] = (double)wid;
    }
#endif
    // ---------------- store ------------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < H; j++) {
        if (j < geom.nx) Trow[geom.cx0 + j] = T[j].x;
        if (j < geom.ny) Trow[geom.cy0 + j] = T[j].y;
    }
    if (gl == 0) {
        const int flags = s.flags, time = s.time, spark_state = s.spark_state, spark_dur = s.spark_dur;
        const bool thermal_break = (flags & SPARC_F_THERMAL_BREAK) != 0 && s.tsi_break >= 0;   // broke in THIS call
        const int win_steps = (time - s.time0) + (thermal_break ? 1 : 0);
        if (win_steps > 0 || (flags & SPARC_F_MODE_ERROR)) {
            // bring the lazily kept dielectric state to what the reference holds after its last µs
            if (s.ran_fast) { s.dens_valid = false; s.pend_flush = true; }
            materialize(s, P);
        }
        Trow[0] = spoolf;
        S.wp[e] = s.wp; S.wire[e] = s.fin_wire; S.v[e] = s.fin_v; S.prev_accel[e] = s.fin_pa;
        S.voltage[e] = s.voltage;
        S.current[e] = (spark_state == 1 || spark_state == -1) ? s.ipk : 0.0;
        S.target_delta[e] = s.target_delta;
        S.debris_volume[e] = s.debris_volume; S.debris_density[e] = s.debris_density;
        if (win_steps > 0) {
            const double gq = (s.gap_prev > 0.001) ? s.gap_prev : 0.001;
            S.cavity_volume[e] = P.cavity_volume_coeff * (gq * 0.001);     // dielectric.py:91-92
        }
        S.flow[e] = s.flow;
        if (!(spark_state == 1 && spark_dur == 0) && win_steps > 0) S.last_crater_volume[e] = 0.0;
        S.d_last_gap[e] = s.d_last_gap; S.d_last_density[e] = s.d_last_density;
        S.w_last_flow[e] = s.w_last_flow;
        if (P.compute_zone_mean && win_steps > 0 && !s.zone_mean_written) S.wire_avg_T[e] = S.w_last_zone_mean[e];
        S.win_vsum[e] = s.win_vsum;
        S.time[e] = time; S.t_servo[e] = time - s.latch_time; S.t_open_v[e] = time;
        // time_since_spark_ignition (wire_edm.py:139-144): dur+1 while sparking, else 0; the
        // µs that returns early (thermal break / mode error) leaves the previous value
        int tsi = (spark_state == 1) ? spark_dur + 1 : 0;
        if (thermal_break) tsi = s.tsi_break;
        else if (flags & SPARC_F_THERMAL_BREAK) tsi = S.t_spark_ign[e];
        if (flags & SPARC_F_MODE_ERROR) tsi = 0;
        S.t_spark_ign[e] = tsi;
        S.t_spark_end[e] = time - s.tse_ref;
        S.spark_state[e] = spark_state; S.spark_dur[e] = spark_dur;
        S.rand_short_rem[e] = s.rand_short_rem; S.deb_short_rem[e] = s.deb_short_rem;
        {
            int ir = s.ion_end - time - (thermal_break ? 1 : 0);
            S.ion_rem[e] = ir > 0 ? ir : 0;
        }
        S.crit_count[e] = s.crit_count;
        S.current_mode[e] = s.current_mode;
        S.w_zone_counter[e] = s.w_zone_counter; S.flags[e] = flags;
        S.stat_short_us[e] += s.win_short_us - s.win_short0;
        S.stat_steps[e] += win_steps;
        S.win_sparks[e] = s.win_sparks; S.win_short_us[e] = s.win_short_us; S.win_steps[e] = s.win_steps0 + win_steps;
        s.win_steps_out = win_steps;
    }
    if (LOG && sig_on && s.done) {              // samples at or after the terminating µs: the final state, as stored
        const float m = env_max_T();
        if (gl == 0) {
            const int st = s.spark_state;
            const double avg = S.wire_avg_T[e];
            for (int r = (a.it_base + done_idx) / a.sample_every; r < sig_rows; r++)
                sig_write(r, (double)s.time, s.voltage, (st == 1 || st == -1) ? s.ipk : 0.0, s.fin_wire, s.fin_v, s.debris_density,
                          s.debris_volume, s.flags, st, m, avg, s.crit_count);
        }
    }
    // ---------------- observation summary (what sparc_observe computes from the stored state) ----
    if (a.obs) {
        float m = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(lane_max<H>(T, lf, spoolf))));   // cell 0 is spoolf
        if constexpr (W > 1) {
            __syncthreads();
            if (lane == 0) sm_red[ctx.wid] = (double)m;
            __syncthreads();
            for (int w = 0; w < W; w++) m = fmaxf(m, (float)sm_red[w]);
        }
        if (gl == 0) {
            const size_t ne = (size_t)a.n_envs;
            const int steps = s.win_steps0 + s.win_steps_out;
            a.obs[0 * ne + e] = (float)(s.wp - s.fin_wire);
            a.obs[1 * ne + e] = (float)s.fin_v;
            a.obs[2 * ne + e] = steps > 0 ? (float)(s.win_vsum / (double)steps) : 0.0f;
            a.obs[3 * ne + e] = (float)s.win_sparks;
            a.obs[4 * ne + e] = (float)s.win_short_us;
            a.obs[5 * ne + e] = m;
            a.obs[6 * ne + e] = (float)s.debris_density;
            a.obs[7 * ne + e] = (float)s.flags;
            a.obs[8 * ne + e] = __int_as_float(s.time);
            a.obs[9 * ne + e] = __int_as_float(s.spark_state);
            a.obs[10 * ne + e] = __int_as_float(s.time - s.latch_time);
            a.obs[11 * ne + e] = 0.0f;
        }
    }
}

// Each warp (W == 1) or block (W > 1) takes environments e, e + stride, ...; the host launches one per env.
// (A grid capped at the resident blocks was measured and rejected, profiles/r02_experiments.txt: the
// hardware's own refill of freed slots balances equal-length envs just as well, and a capped grid delays the
// blocks that wait for the registers of the pre-pass kernel.)
template <int K, bool INJECT, int W, bool LOG, bool ALIGNED>
#ifndef SPARC_WARPS_MULTI
#define SPARC_WARPS_MULTI 16    /* resident warps per SM the block-per-env variants (W > 1) are sized for */
#endif
#ifdef SPARC_MAXNREG            /* experiments: an explicit register cap instead of a residency target */
__global__ void __maxnreg__(SPARC_MAXNREG)
#else
__global__ void __launch_bounds__(W == 1 ? SPARC_BLOCK_THREADS : 32 * W,
                                  W == 1 ? min_blocks_for(K, ALIGNED) : (SPARC_WARPS_MULTI / W > 0 ? SPARC_WARPS_MULTI / W : 1))
#endif
step_kernel(const __grid_constant__ SparcParams P, const __grid_constant__ SparcStatePtrs S,
            const __grid_constant__ StepArgs a)
{
    // W == 1: one warp per env, SPARC_BLOCK_THREADS/32 envs per block at a time; W > 1: one block of W warps per env
    constexpr int WARPS = (W == 1) ? SPARC_BLOCK_THREADS / 32 : W;
    __shared__ float sm_halo[2][W][2];
    __shared__ double sm_red[W];
    __shared__ WarpRec sm_rec[WARPS];
    extern __shared__ __align__(16) float sm_stage[];   // LOG: row staging, 2 buffers per env (host sizes it)
    if constexpr (W == 1) {
        const int stride = (int)gridDim.x * WARPS;
        for (int e = (int)blockIdx.x * WARPS + (int)(threadIdx.x >> 5); e < a.n_envs; e += stride) {
            step_env<K, INJECT, W, LOG, ALIGNED, WARPS>(P, S, a, e, sm_halo, sm_red, sm_rec, sm_stage);
            __syncwarp();
        }
    } else {
        for (int e = (int)blockIdx.x; e < a.n_envs; e += (int)gridDim.x) {
            step_env<K, INJECT, W, LOG, ALIGNED, WARPS>(P, S, a, e, sm_halo, sm_red, sm_rec, sm_stage);
            __syncthreads();
        }
    }
}

}  // namespace sparc
