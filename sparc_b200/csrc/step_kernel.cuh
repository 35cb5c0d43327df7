// step_kernel.cuh — the fused step kernel, "look-ahead" organisation (sm_100a).
//
// Same mapping as step_kernel.cuh (one warp per environment, the wire temperature field in
// registers, K cells per lane, halos by warp shuffle) and the same general µs (general_pre), but
// the per-µs scalar state machines no longer run on every µs:
//
//   * MECHANICS (mechanics.py:79-114) depends on nothing but the latched servo command, so it is
//     advanced one BLOCK of <= 32 µs ahead of everything else (exact sequential fp64, identical
//     operations); the 32 wire positions go to shared memory and from there one per lane.
//   * LOOK-AHEAD: lane l then evaluates, for µs l of the block, every condition under which the
//     reference's step() is "quiet" (idle/rest, no latch, no short circuit, no ignition, no
//     hysteresis refresh, no collision, no zone-mean sample).  The ignition roll is evaluated
//     here too: lane l computes the Philox word of its own µs.  One ballot gives the length n of
//     the quiet run.
//   * QUIET RUN: n µs of pure stencil — no scalar instruction in the loop.  When the wire is
//     provably cold (max T < critical-1 K and the update is a convex combination, so the maximum
//     cannot grow without a heat source) even the max/break test is skipped.
//   * everything else takes the exact general µs (general_pre + heated stencil + break test +
//     termination), one µs at a time.
//
// Every shortcut is conservative: a µs that is not provably quiet runs the general path, which is
// the reference's logic statement by statement, so results do not depend on where the run boundaries fall.
#pragma once
#include "step_common.cuh"

namespace sparc {

// ---- packed stencil ---------------------------------------------------------------------------
// sm_100 has two-wide fp32 instructions (FFMA2/FADD2, PTX fma.rn.f32x2): one issue slot and one
// pipe pass for two IEEE-rn operations — measured 1.95 FFMA2 vs 2.1 register-operand FFMA per
// clock per SM (tools/microbench2.cu).  Lane cells are split in two halves of H = K/2 cells;
// pair j holds (T[j], T[j+H]) so that neighbour differences of pair j are pair j+1 minus pair j.
// Element-wise the operation sequence is unchanged: q = fma(cE,T,c0); q = fma(cR,dnext,q);
// q = fma(ncL,dprev,q); T = T + q.
__device__ __forceinline__ float2 sub2(const float2 a, const float2 b)
{
    return __fadd2_rn(a, make_float2(-b.x, -b.y));
}

// ---- wires longer than one warp: W warps (one thread block) per environment -------------------
// Each warp runs the scalar state machine redundantly (same inputs, same results, so control flow
// is identical in all warps of the block); only three things cross warps: the stencil halos
// (shared memory, double buffered, one barrier per µs), the temperature votes
// (__syncthreads_or) and the zone-mean sum.
template <int W>
struct EnvCtx {
    int lane, wid, gl;      // lane in warp, warp in env, lane in env
    float (*halo)[W][2];    // [parity][warp][first, last]
    int parity;
};

template <int W>
__device__ __forceinline__ void stencil_halos(EnvCtx<W> &c, const float t_first, const float t_last, float &left,
                                              float &right, const float spoolf)
{
    constexpr unsigned FULL = 0xffffffffu;
    left = __shfl_up_sync(FULL, t_last, 1);
    right = __shfl_down_sync(FULL, t_first, 1);
    if constexpr (W > 1) {
        float(*h)[2] = c.halo[c.parity];
        if (c.lane == 0) h[c.wid][0] = t_first;
        if (c.lane == 31) h[c.wid][1] = t_last;
        __syncthreads();
        if (c.lane == 0 && c.wid > 0) left = h[c.wid - 1][1];
        if (c.lane == 31 && c.wid < W - 1) right = h[c.wid + 1][0];
        c.parity ^= 1;
    }
    if (c.gl == 0) left = spoolf;
}

template <int W>
__device__ __forceinline__ bool vote_any(const bool pred)
{
    if constexpr (W > 1) return __syncthreads_or(pred ? 1 : 0) != 0;
    else return __any_sync(0xffffffffu, pred);
}

// n quiet µs of compute_thermal_update (wire.py:58-123) without heat sources.  HOT additionally
// runs _check_wire_breaking (wire.py:376-388) every µs and returns the number of µs executed
// (n, or j+1 when the wire broke on µs j; `broke` says which).
template <int H, bool HOT, int W, bool LOG, class LogTick>
__device__ __forceinline__ int run_quiet(float2 (&T)[H], const float2 (&cE)[H], const float2 (&c0)[H],
                                         const float2 (&cR)[H], const float ncL, const float spoolf, EnvCtx<W> &ctx,
                                         const int n, const float crit_T, const float break_T, int &crit_count,
                                         bool &broke, bool &cold_after, LogTick &&log_tick)
{
    const float2 ncL2 = make_float2(ncL, ncL);
    broke = false;
#pragma unroll 1
    for (int k = 0; k < n; k++) {
        float left, right;
        stencil_halos<W>(ctx, T[0].x, T[H - 1].y, left, right, spoolf);
        // the three differences that cross a half boundary are scalar; the middle one is shared.
        // Pairs 1..H-2 do not see the halos: they go first so the shuffles complete underneath them;
        // pair H-1 (needs `right`) and pair 0 (needs `left`) come last.  (Measured: +2 % over the
        // natural order; issuing the next µs's shuffles in the middle of the body was slower.)
        const float d_mid = T[0].y - T[H - 1].x;
        const float2 d0 = sub2(T[1], T[0]);     // dnext of pair 0, dprev of pair 1 (old values)
        float lmax = spoolf;
        float2 dprev = d0;
#pragma unroll
        for (int j = 1; j < H - 1; j++) {
            const float2 dnext = sub2(T[j + 1], T[j]);
            float2 q = __ffma2_rn(cE[j], T[j], c0[j]);
            q = __ffma2_rn(cR[j], dnext, q);
            q = __ffma2_rn(ncL2, dprev, q);
            dprev = dnext;
            T[j] = __fadd2_rn(T[j], q);
            if (HOT) lmax = fmaxf(lmax, fmaxf(T[j].x, T[j].y));
        }
        {
            const float2 dlast = make_float2(d_mid, right - T[H - 1].y);
            float2 q = __ffma2_rn(cE[H - 1], T[H - 1], c0[H - 1]);
            q = __ffma2_rn(cR[H - 1], dlast, q);
            q = __ffma2_rn(ncL2, dprev, q);
            T[H - 1] = __fadd2_rn(T[H - 1], q);
            if (HOT) lmax = fmaxf(lmax, fmaxf(T[H - 1].x, T[H - 1].y));
        }
        {
            const float2 dfirst = make_float2(T[0].x - left, d_mid);
            float2 q = __ffma2_rn(cE[0], T[0], c0[0]);
            q = __ffma2_rn(cR[0], d0, q);
            q = __ffma2_rn(ncL2, dfirst, q);
            T[0] = __fadd2_rn(T[0], q);
            if (HOT) lmax = fmaxf(lmax, fmaxf(T[0].x, T[0].y));
        }
        if (LOG) log_tick();
        if (HOT) {
            const bool hot = vote_any<W>(lmax > crit_T);
            crit_count = hot ? crit_count + 1 : 0;
            if (hot && vote_any<W>(lmax > break_T)) { broke = true; return k + 1; }
            if (k == n - 1) cold_after = !vote_any<W>(lmax > crit_T - 1.0f);
        }
    }
    if (!HOT) crit_count = 0;
    return n;
}

// resident blocks per SM the register budget is sized for
#ifndef SPARC_MINB_SMALL
#define SPARC_MINB_SMALL 7
#endif
constexpr int min_blocks_for(int K)
{
    return (K <= 13 ? SPARC_MINB_SMALL : (K <= 20 ? 6 : 4)) * 64 / SPARC_BLOCK_THREADS;
}

#define SPARC_LN2_UP 0.6931471805599460    /* ln2*(1+2^-50): u3 < fl(ln2/den) implies u3*den <= this */

// Wait until the mechanics pre-pass of THIS launch (tag `seq`) has published at least `need` blocks of
// the env whose progress word is *pp; returns the number published.  Out of line: it runs a handful
// of times per launch and must not cost the step kernel registers.
static __device__ __noinline__ uint32_t mech_wait(const uint32_t *pp, const uint32_t seq, const uint32_t need)
{
    unsigned spins = 0;
    for (;;) {
        uint32_t pv;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(pv) : "l"(pp) : "memory");
        const uint32_t d = pv - seq;
        if (d < 64u && d >= need) return d;
        __nanosleep(100);
        if (++spins > (1u << 23)) __trap();     // ~1 s: the producer is gone — fail loudly, never hang
    }
}

template <int K, bool INJECT, int W, bool LOG>
__global__ void __launch_bounds__(W == 1 ? SPARC_BLOCK_THREADS : 32 * W, W == 1 ? min_blocks_for(K) : 1)
step_kernel(const __grid_constant__ SparcParams P, const __grid_constant__ SparcStatePtrs S,
             const __grid_constant__ StepArgs a)
{
    constexpr unsigned FULL = 0xffffffffu;
    // W == 1: one warp per env, SPARC_BLOCK_THREADS/32 envs per block; W > 1: one block of W warps per env
    constexpr int WARPS = (W == 1) ? SPARC_BLOCK_THREADS / 32 : W;
    __shared__ float sm_halo[2][W][2];
    __shared__ double sm_red[W];
    __shared__ uint32_t sm_mech_ready[WARPS];
    __shared__ float sm_log[LOG ? WARPS : 1][LOG ? 32 * (K + 1) : 1];   // field-log staging: coalesced row stores
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int e = (W == 1) ? blockIdx.x * WARPS + wib : blockIdx.x;
    if (e >= a.n_envs) return;                  // whole env group (warp or block)
    EnvCtx<W> ctx;
    ctx.lane = lane; ctx.wid = (W == 1) ? 0 : wib; ctx.gl = ctx.wid * 32 + lane;
    ctx.halo = sm_halo; ctx.parity = 0;
    const int gl = ctx.gl;                      // lane within the env: owns slots gl*K .. gl*K+K-1
    const uint32_t k0 = (uint32_t)a.seed, k1 = (uint32_t)(a.seed >> 32);

    // ---------------- load ---------------------------------------------------------------------
    EnvState s;
    s.e = e; s.it = 0; s.writer = gl == 0;
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
    const int win_short0 = cont ? S.win_short_us[e] : 0, win_steps0 = cont ? S.win_steps[e] : 0;
    s.win_short_us = win_short0;
    s.dens_valid = true; s.pend_flush = false; s.fast_ok = false;
    s.tsi_prev = 0; s.plasma_idx = -1;
    s.heat_on = false; s.thermal = false; s.refresh = false; s.freeze = false; s.became_done = false;
    s.jA = 0.f; s.jB = 0.f; s.fP = 0.f; s.hb = 0.f; s.he = 0.f;
    s.gap_lo_f = 0.0; s.gap_hi_f = -1.0; s.lo_c = 0.0; s.hi_c = 0.0; s.flush_amt = 0.0; s.onoff_thr = 0;
    derive_generator(s, P, S.target_voltage[e], S.on_time[e], S.off_time[e]);

    double wp = S.wp[e], wire = S.wire[e], v = S.v[e], prev_accel = S.prev_accel[e];
    double target_delta = S.target_delta[e], voltage = S.voltage[e], debris_volume = S.debris_volume[e];
    double win_vsum = cont ? S.win_vsum[e] : 0.0, gap_prev = wp - wire;
    int time = S.time[e], spark_state = S.spark_state[e], spark_dur = S.spark_dur[e], flags = S.flags[e];
    int crit_count = S.crit_count[e];
    int w_zone_counter = S.w_zone_counter[e];
    const int time0 = time;
    s.latch_time = time - S.t_servo[e];
    s.tse_ref = time - S.t_spark_end[e];
    s.ion_end = time + S.ion_rem[e];
    if (spark_state == 1 && (flags & SPARC_F_HAS_Y))
        s.plasma_idx = (P.segment_len != 0.0) ? P.zone_start + (int)py_floordiv(S.spark_y[e], P.segment_len) : P.zone_start;

    if (flags & SPARC_F_DONE) {                 // terminated envs are left untouched
        if (s.writer && !cont) { S.win_vsum[e] = 0.0; S.win_sparks[e] = 0; S.win_short_us[e] = 0; S.win_steps[e] = 0; }
        if (a.obs) {                            // observation of a frozen env, from the stored state
            float m = -3.0e38f;
            const float *row = S.T + (size_t)e * S.n_pad;
            for (int c = gl; c < P.n_segments; c += 32 * W) m = fmaxf(m, row[c]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL, m, o));
            if constexpr (W > 1) {
                if (lane == 0) sm_red[ctx.wid] = (double)m;
                __syncthreads();
                for (int w = 0; w < W; w++) m = fmaxf(m, (float)sm_red[w]);
            }
            if (s.writer) {
                const size_t ne = (size_t)a.n_envs;
                const int steps = cont ? win_steps0 : 0;
                a.obs[0 * ne + e] = (float)(wp - wire);
                a.obs[1 * ne + e] = (float)v;
                a.obs[2 * ne + e] = steps > 0 ? (float)(win_vsum / (double)steps) : 0.0f;
                a.obs[3 * ne + e] = (float)s.win_sparks;
                a.obs[4 * ne + e] = (float)s.win_short_us;
                a.obs[5 * ne + e] = m;
                a.obs[6 * ne + e] = (float)s.debris_density;
                a.obs[7 * ne + e] = (float)flags;
            }
        }
        return;
    }

    int next_latch = s.latch_time + P.servo_interval;
    bool fast_ok = false, ran_fast = false, zone_mean_written = false, done = false;
    int skip_flush = 0;                         // 1: the next quiet run starts with no pending debris flush
    int tsi_break = -1;

    // upper bound of lambda(gap) = ln2/(a g^2 + b g + c) over all gaps (ignition.py:348-364): a
    // draw above it can never ignite, whatever the gap.
    double lam_bound = 2.0;
    if (P.ignition_a_coeff > 0.0) {
        double den_min = P.ignition_c_coeff - P.ignition_b_coeff * P.ignition_b_coeff / (4.0 * P.ignition_a_coeff);
        if (den_min > 0.0) lam_bound = 0.6931471805599453 / den_min * (1.0 + 1e-9);
    }
    uint32_t w_thresh = 0xffffffffu;            // word >= w_thresh  =>  u01(word) >= lam_bound
    if (lam_bound < 0.999) w_thresh = (uint32_t)(lam_bound * 4294967296.0) + 2u;

    // ---------------- temperature field -> registers -----------------------------------------
    const int N = P.n_segments;
    const int M = N - 1;                        // dynamic cells 1..N-1; cell 0 == spool_T
    float spoolf = (float)P.spool_T;
    asm volatile("" : "+f"(spoolf));            // keep it in a register (no per-µs F2F rematerialisation)
    static_assert(K % 2 == 0, "packed stencil needs an even number of cells per lane");
    constexpr int H = K / 2;
    float2 T[H];                                // T[j] = (cell j, cell j+H) of this lane's K slots
    float *Trow = S.T + (size_t)e * S.n_pad;
#pragma unroll
    for (int j = 0; j < H; j++) {
        const int sa = gl * K + j, sb = sa + H;
        T[j].x = (sa < M) ? Trow[sa + 1] : spoolf;
        T[j].y = (sb < M) ? Trow[sb + 1] : spoolf;
    }

    // ---------------- stencil coefficients ---------------------------------------------------
    const double f = P.temp_update_factor;
    float cRf, ncL;
    {
        const double unwind_v = S.unwind_v[e];
        double adv = 0.0;                       // wire.py:299-309
        if (fabs(unwind_v) > 1e-6) adv = P.density_times_cp * fabs(unwind_v) * P.S;
        if (!(fabs(adv) > 1e-9)) adv = 0.0;     // wire.py:114
        cRf = (float)(f * P.k_cond_coeff);
        ncL = -(float)(f * (P.k_cond_coeff + adv));
    }
    float2 cE[H], c0[H], cR[H];
    bool convex = false;                        // T' is a convex combination of T, T_diel (no source)
    auto slot_coeffs = [&](int sl, float hb, float he, float &e_, float &z_, float &r_) {
        const int i = sl + 1;
        if (sl < M) {
            float h = (i >= P.actual_zone_start && i < P.actual_zone_end) ? he : hb;
            double hA = (double)h * P.A;        // conv_coeff, wire.py:110
            e_ = (float)(-(f * hA));
            z_ = (float)(f * hA * P.dielectric_temperature);
            r_ = (i < N - 1) ? cRf : 0.0f;      // Neumann end, wire.py:95
        } else {                                // pad slot: pinned near spool_T, feeds nothing
            e_ = -1.0f;
            z_ = spoolf;
            r_ = 0.0f;
        }
    };
    auto refresh_coeffs = [&](float hb, float he) {
#pragma unroll
        for (int j = 0; j < H; j++) {
            slot_coeffs(gl * K + j, hb, he, cE[j].x, c0[j].x, cR[j].x);
            slot_coeffs(gl * K + j + H, hb, he, cE[j].y, c0[j].y, cR[j].y);
        }
        const float hmax = fmaxf(hb, he), hmin = fminf(hb, he);
        const float ce_max = (float)(f * ((double)hmax * P.A));
        convex = hmin >= 0.0f && cRf >= 0.0f && ncL <= 0.0f && (cRf - ncL + ce_max) <= 0.999f &&
                 fmaxf((float)P.dielectric_temperature, spoolf) < P.critical_temperature_f32 - 1.0f;
    };
    refresh_coeffs(S.h_base[e], S.h_enh[e]);
    bool cold_ok = false;                       // max T < critical - 1 K after the last executed µs

    // contact span in slot coordinates of this lane
    const int j_span_lo = P.contact_bottom_idx - 1 - gl * K;
    const int j_span_hi = P.contact_top_idx - 1 - gl * K;
    const double *drow = INJECT ? a.draws + ((size_t)e * a.total_us + a.it_base) * SPARC_N_DRAW_SLOTS : nullptr;

    // temperature-field logging (SimulationLogger "wire_temperature", utils/logger.py:128-150): row r
    // holds the field after µs (r+1)*log_every - 1 of this call, layout (n_timesteps, n_segments)
    const bool logging = LOG && a.tlog != nullptr && e < a.n_log_envs && a.log_every > 0;
    const int log_rows = logging ? a.total_us / a.log_every : 0;
    int log_ctr = logging ? a.log_every - a.it_base % a.log_every : 0;  // µs until the next sample
    int log_row = logging ? a.it_base / a.log_every : 0;                // next row to write
    auto write_log_row = [&](int r) {
        float *row = a.tlog + ((size_t)e * log_rows + r) * N;
        float *stage = sm_log[wib];             // lane-major with an odd pitch: conflict-free both ways
#pragma unroll
        for (int j = 0; j < H; j++) {
            stage[lane * (K + 1) + j] = T[j].x;
            stage[lane * (K + 1) + j + H] = T[j].y;
        }
        __syncwarp();
        const int wbase = ctx.wid * 32 * K;     // first slot of this warp; slot s is cell s + 1
#pragma unroll
        for (int q = 0; q < K; q++) {
            const int sl = q * 32 + lane;
            if (wbase + sl < M) row[wbase + sl + 1] = stage[(sl / K) * (K + 1) + sl % K];
        }
        __syncwarp();
        if (gl == 0) row[0] = spoolf;
    };
    auto log_tick = [&]() {                     // called after every thermal update of a logged env
        if (--log_ctr == 0) {
            if (log_row < log_rows) write_log_row(log_row);
            log_row += 1;
            log_ctr = a.log_every;
        }
    };

    // mechanics pre-pass results of this launch (mech_kernel).  The pre-pass is STILL RUNNING while this
    // kernel executes (programmatic dependent launch, sparc_b200.cu): positions and checkpoints are
    // read only after the producer thread of this env has published them through mech_prog
    // (st.release / ld.acquire at GPU scope), and with ld.global.cg — an L1 line fetched before the
    // producer finished it would be stale.  The producer is ~8x faster than this consumer, so after the
    // first blocks one poll covers many blocks (mech_ready caches the last value seen; the poll
    // itself is out of line, mech_wait).
    const double *traj = S.mech_traj + (size_t)e * SPARC_MECH_EPOCH;
    // pre-pass blocks known to be published, per warp, in shared memory (read once per 32-µs block;
    // a register held across the stencil loops costs more than the LDS)
    volatile uint32_t *mech_ready = &sm_mech_ready[wib];
    *mech_ready = 0;
    auto need_mech = [&](int upto_us) {         // positions of launch-µs < upto_us and checkpoints up to there
        const uint32_t need = (uint32_t)((upto_us + SPARC_MECH_CKPT - 1) / SPARC_MECH_CKPT);
        if (need > *mech_ready) *mech_ready = mech_wait((const uint32_t *)S.mech_prog + e, a.mech_seq, need);
    };
    const double wire0 = wire, delta_old0 = target_delta, delta_new0 = (double)a.act.servo[e];
    const int first_latch_it = (next_latch - time > 0) ? next_latch - time : 0;

    int it = 0;
    while (it < a.n_us && !done) {
        // ================= block setup ==========================================================
        int nb = a.n_us - it;
        if (nb > 32) nb = 32;
        int to_latch = next_latch - time;
        bool force_general = false;
        if (to_latch <= 0) {                    // the first µs of this block latches the action
            to_latch = (P.servo_interval > 1) ? P.servo_interval : 1;
            force_general = true;
        }
        if (nb > to_latch) nb = to_latch;       // a block never straddles a latch
        // wire positions of this block from the mechanics pre-pass: lane l holds µs it + l
        need_mech(it + nb);
        const double post_l = __ldcg(traj + it + ((lane < nb) ? lane : nb - 1));  // wire after µs `lane`
        double pre_l = __shfl_up_sync(FULL, post_l, 1);                  // wire before µs `lane`
        if (lane == 0) pre_l = (it == 0) ? wire0 : __ldcg(traj + it - 1);
        // ignition roll of µs `lane` (slot 2): word (time+lane)&3 of Philox block (time+lane)>>2
        uint32_t w_l = 0;
        if (!INJECT) {
            const uint32_t tl = (uint32_t)(time + lane);
            uint32_t w[4];
            philox4x32_10(s.env_id, tl >> 2, s.episode, 2u, k0, k1, w);
            w_l = pick4(w, tl & 3);
        }
        // finished inside this block: exact mechanics state after `steps` µs of it, replayed from
        // the pre-pass checkpoint (same operation sequence as mech_kernel)
        auto rewind_mech = [&](int steps) {
            const int target = it + steps;
            const int c = target / SPARC_MECH_CKPT;
            const double *ck = S.mech_ckpt + ((size_t)e * SPARC_MECH_NCKPT + c) * 2;
            wire = (c == 0) ? wire0 : __ldcg(traj + c * SPARC_MECH_CKPT - 1);
            v = __ldcg(ck); prev_accel = __ldcg(ck + 1);
#pragma unroll 1
            for (int k = c * SPARC_MECH_CKPT; k < target; k++)
                mech_step(P, wire, v, prev_accel, (k >= first_latch_it) ? delta_new0 : delta_old0);
        };

        int i = 0;
        while (i < nb) {
            // ================= look-ahead: length of the quiet run starting at µs i ===========
            int n = 0;
            int off_tr = -1;                    // run offset of the rest -> idle transition µs
            const double g = wp - pre_l;
            if (!force_general && fast_ok && (spark_state == 0 || spark_state == -2)) {
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
                    off_tr = s.onoff_thr - spark_dur - 1;
                    if (off_tr < 0) off_tr = 0;
                }
                if (P.compute_zone_mean) ok = ok && (w_zone_counter + off + 1 < P.zone_mean_interval);
                if (INJECT) {
                    if (lane < nb) {
                        const double *d = drow + (size_t)(it + lane) * SPARC_N_DRAW_SLOTS;
                        ok = ok && d[0] >= SPARC_U1_SAFE;
                        if (off > off_tr) {
                            const double u3 = d[2];
                            if (!(u3 >= lam_bound)) {
                                double den = P.ignition_a_coeff * (g * g) + P.ignition_b_coeff * g + P.ignition_c_coeff;
                                ok = ok && (den > 0.0 && u3 * den > SPARC_LN2_UP);
                            }
                        }
                    }
                } else if (off > off_tr && w_l < w_thresh) {
                    // u3 < fl(ln2/den) can only hold if u3*den <= ln2*(1+2^-50)
                    const double u3 = u01(w_l);
                    double den = P.ignition_a_coeff * (g * g) + P.ignition_b_coeff * g + P.ignition_c_coeff;
                    ok = ok && (den > 0.0 && u3 * den > SPARC_LN2_UP);
                }
                unsigned stop = __ballot_sync(FULL, !(ok && in_band())) >> i;
                n = stop ? __ffs(stop) - 1 : 32 - i;
                if (n == 0 && ((__ballot_sync(FULL, ok) >> i) & 1u)) {
                    // only a band failed on µs i: refresh the hysteresis state in place (light_refresh)
                    const bool pending = ran_fast || s.pend_flush;
                    if (light_refresh(s, P, S, __shfl_sync(FULL, g, i), debris_volume, pending)) {
                        ran_fast = false;
                        skip_flush = 1;         // the refreshed µs has no pending flush before it
                        fast_ok = s.fast_ok;
                        if (fast_ok) {
                            stop = __ballot_sync(FULL, !(ok && in_band())) >> i;
                            n = stop ? __ffs(stop) - 1 : 32 - i;
                        }
                    }
                }
            }
            force_general = false;

            if (n > 0) {
                // ================= quiet run ===================================================
                bool broke = false;
                int nx;
                const bool cold = cold_ok && convex;
                if (!LOG || !logging) {
                    if (cold)
                        nx = run_quiet<H, false, W, false>(T, cE, c0, cR, ncL, spoolf, ctx, n, 0.f, 0.f, crit_count, broke,
                                                           cold_ok, log_tick);
                    else
                        nx = run_quiet<H, true, W, false>(T, cE, c0, cR, ncL, spoolf, ctx, n, P.critical_temperature_f32,
                                                          P.breaking_temperature_f32, crit_count, broke, cold_ok, log_tick);
                } else {
                    if (cold)
                        nx = run_quiet<H, false, W, LOG>(T, cE, c0, cR, ncL, spoolf, ctx, n, 0.f, 0.f, crit_count, broke,
                                                         cold_ok, log_tick);
                    else
                        nx = run_quiet<H, true, W, LOG>(T, cE, c0, cR, ncL, spoolf, ctx, n, P.critical_temperature_f32,
                                                        P.breaking_temperature_f32, crit_count, broke, cold_ok, log_tick);
                }
                // scalar bookkeeping of the nx quiet µs (ignition.py:302-319, dielectric.py:104-144)
                int n_idle = nx;
                if (spark_state == -2) {
                    if (nx > off_tr) {          // rest -> idle on run offset off_tr
                        spark_state = 0; spark_dur = 0; flags &= ~SPARC_F_HAS_Y;
                        if (s.writer) S.spark_y[e] = -1.0;
                        n_idle = nx - off_tr;
                    } else { spark_dur += nx; n_idle = 0; }
                }
                voltage = (n_idle > 0) ? s.tv_eff : 0.0;
                win_vsum += (double)n_idle * s.tv_eff;
                if (debris_volume > 0.001 && s.flush_amt > 0.0) {
#pragma unroll 1
                    for (int k = skip_flush; k < nx && debris_volume > 0.001; k++) debris_volume = debris_volume - s.flush_amt;
                }
                skip_flush = 0;
                gap_prev = __shfl_sync(FULL, g, i + nx - 1);
                ran_fast = true;
                if (P.compute_zone_mean) w_zone_counter += nx;
                if (broke) {                    // wire_edm.py:129-130 on the last µs of the run
                    flags |= SPARC_F_BROKEN | SPARC_F_THERMAL_BREAK;
                    tsi_break = 0;
                    done = true;
                    time += nx - 1;
                    rewind_mech(i + nx - 1);
                    break;
                }
                time += nx;
                i += nx;
                continue;
            }

            // ================= general µs ======================================================
            s.wp = wp; s.wire = __shfl_sync(FULL, pre_l, i); s.v = 0.0; s.prev_accel = 0.0;
            s.target_delta = target_delta;
            s.voltage = voltage; s.debris_volume = debris_volume; s.win_vsum = win_vsum; s.gap_prev = gap_prev;
            s.time = time; s.spark_state = spark_state; s.spark_dur = spark_dur; s.flags = flags;
            s.it = it + i;
            if (ran_fast) { s.dens_valid = false; s.pend_flush = true; ran_fast = false; }
            skip_flush = 0;
            general_pre<INJECT>(s, P, S, a);
            wp = s.wp; target_delta = s.target_delta; voltage = s.voltage; debris_volume = s.debris_volume;
            win_vsum = s.win_vsum; gap_prev = s.gap_prev;
            spark_state = s.spark_state; spark_dur = s.spark_dur; flags = s.flags;
            next_latch = s.latch_time + P.servo_interval;
            fast_ok = s.fast_ok;
            if (s.refresh) refresh_coeffs(s.hb, s.he);
            if (s.freeze) {                     // crater-table miss: frozen before the thermal update
                done = true;
                rewind_mech(i);
                break;
            }
            // ---- compute_thermal_update with sources (wire.py:58-123) -------------------------
            float lmax = spoolf;
            {
                float left, right;
                stencil_halos<W>(ctx, T[0].x, T[H - 1].y, left, right, spoolf);
                float2 acc[H];
#pragma unroll
                for (int j = 0; j < H; j++) acc[j] = __ffma2_rn(cE[j], T[j], c0[j]);
                if (s.heat_on) {
                    const float jA = s.jA, jB = s.jB, fP = s.fP;
                    const int jp = (s.spark_state == 1 && s.plasma_idx >= 0 && s.plasma_idx < N)
                                       ? s.plasma_idx - 1 - gl * K : -(1 << 20);
#pragma unroll
                    for (int j = 0; j < H; j++) {
                        if (j >= j_span_lo && j <= j_span_hi) acc[j].x += fmaf(jA, T[j].x, jB);
                        if (j == jp) acc[j].x += fP;
                        if (j + H >= j_span_lo && j + H <= j_span_hi) acc[j].y += fmaf(jA, T[j].y, jB);
                        if (j + H == jp) acc[j].y += fP;
                    }
                }
                const float2 ncL2 = make_float2(ncL, ncL);
                const float d_mid = T[0].y - T[H - 1].x;
                float2 dprev = make_float2(T[0].x - left, d_mid);
                const float2 dlast = make_float2(d_mid, right - T[H - 1].y);
#pragma unroll
                for (int j = 0; j < H; j++) {
                    const float2 dnext = (j + 1 < H) ? sub2(T[j + 1], T[j]) : dlast;
                    float2 q = __ffma2_rn(cR[j], dnext, acc[j]);
                    q = __ffma2_rn(ncL2, dprev, q);
                    dprev = dnext;
                    T[j] = __fadd2_rn(T[j], q);
                    lmax = fmaxf(lmax, fmaxf(T[j].x, T[j].y));
                }
            }
            if (LOG && logging) log_tick();
            // ---- _check_wire_breaking (wire.py:376-388) ---------------------------------------
            cold_ok = !vote_any<W>(lmax > P.critical_temperature_f32 - 1.0f);
            bool hot = false, brk = false;
            if (!cold_ok) {
                hot = vote_any<W>(lmax > P.critical_temperature_f32);
                if (hot) brk = vote_any<W>(lmax > P.breaking_temperature_f32);
            }
            crit_count = hot ? crit_count + 1 : 0;
            // zone mean (wire.py:338-347, 390-394), only when enabled
            if (P.compute_zone_mean) {
                w_zone_counter += 1;
                if (w_zone_counter >= P.zone_mean_interval) {
                    int za = P.actual_zone_start, zb = P.actual_zone_end;
                    if (!(zb > za && zb <= N)) { za = 0; zb = N; }
                    float part = 0.0f;
#pragma unroll
                    for (int jj = 0; jj < K; jj++) {     // ascending cell order within the lane
                        int c = gl * K + jj + 1;
                        if (c >= za && c < zb) part += (jj < H) ? T[jj].x : T[jj - H].y;
                    }
                    if (gl == 0 && za == 0) part += spoolf;    // cell 0 when the zone starts at 0
                    double tot = (double)part;
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
                    double zm = (double)(float)(tot / (double)(zb - za));
                    if (s.writer) { S.w_last_zone_mean[e] = zm; S.wire_avg_T[e] = zm; }
                    zone_mean_written = true;
                    w_zone_counter = 0;
                }
            }
            if (brk) {
                // wire_edm.py:129-130: early return, no mechanics, no time bookkeeping
                flags |= SPARC_F_BROKEN | SPARC_F_THERMAL_BREAK;
                tsi_break = s.tsi_prev;
                done = true;
                rewind_mech(i);
                break;
            }
            // mechanics of this µs is post_l of lane i; time bookkeeping (wire_edm.py:134-144)
            time += 1;
            if (spark_state == 1) s.tse_ref = time;
            // ---- _check_termination (wire_edm.py:172-179) -------------------------------------
            const double wire_after = __shfl_sync(FULL, post_l, i);
            if (wire_after > wp + 100.0) { flags |= SPARC_F_BROKEN; done = true; }
            else if (wp >= P.target_position) { flags |= SPARC_F_TARGET; done = true; }
            if (done) { rewind_mech(i + 1); break; }
            i += 1;
        }
        it += nb;
    }
    if (!done) {                                // end-of-launch mechanics state from the pre-pass
        wire = __ldcg(traj + a.n_us - 1);       // published: the last block of the loop waited for it
        v = __ldcg(S.mech_ckpt + ((size_t)e * SPARC_MECH_NCKPT + (SPARC_MECH_NCKPT - 1)) * 2);
        prev_accel = __ldcg(S.mech_ckpt + ((size_t)e * SPARC_MECH_NCKPT + (SPARC_MECH_NCKPT - 1)) * 2 + 1);
    }
    if (LOG && logging && done) {               // rows sampled at or after termination: the frozen field
        for (int r = log_row; r < log_rows; r++) write_log_row(r);
    }

    // ---------------- store ------------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < H; j++) {
        const int sa = gl * K + j, sb = sa + H;
        if (sa < M) Trow[sa + 1] = T[j].x;
        if (sb < M) Trow[sb + 1] = T[j].y;
    }
    if (gl == 0) {
        const bool thermal_break = (flags & SPARC_F_THERMAL_BREAK) != 0 && tsi_break >= 0;   // broke in THIS call
        const int win_steps = (time - time0) + (thermal_break ? 1 : 0);
        if (win_steps > 0 || (flags & SPARC_F_MODE_ERROR)) {
            // bring the lazily kept dielectric state to what the reference holds after its last µs
            s.debris_volume = debris_volume; s.gap_prev = gap_prev;
            if (ran_fast) { s.dens_valid = false; s.pend_flush = true; }
            materialize(s, P);
            debris_volume = s.debris_volume;
        }
        Trow[0] = spoolf;
        S.wp[e] = wp; S.wire[e] = wire; S.v[e] = v; S.prev_accel[e] = prev_accel;
        S.voltage[e] = voltage;
        S.current[e] = (spark_state == 1 || spark_state == -1) ? s.ipk : 0.0;
        S.target_delta[e] = target_delta;
        S.debris_volume[e] = debris_volume; S.debris_density[e] = s.debris_density;
        if (win_steps > 0) {
            double gq = (gap_prev > 0.001) ? gap_prev : 0.001;
            S.cavity_volume[e] = P.cavity_volume_coeff * (gq * 0.001);     // dielectric.py:91-92
        }
        S.flow[e] = s.flow;
        if (!(spark_state == 1 && spark_dur == 0) && win_steps > 0) S.last_crater_volume[e] = 0.0;
        S.d_last_gap[e] = s.d_last_gap; S.d_last_density[e] = s.d_last_density;
        S.w_last_flow[e] = s.w_last_flow;
        if (P.compute_zone_mean && win_steps > 0 && !zone_mean_written) S.wire_avg_T[e] = S.w_last_zone_mean[e];
        S.win_vsum[e] = win_vsum;
        S.time[e] = time; S.t_servo[e] = time - s.latch_time; S.t_open_v[e] = time;
        // time_since_spark_ignition (wire_edm.py:139-144): dur+1 while sparking, else 0; the
        // µs that returns early (thermal break / mode error) leaves the previous value
        int tsi = (spark_state == 1) ? spark_dur + 1 : 0;
        if (thermal_break) tsi = tsi_break;
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
        S.crit_count[e] = crit_count;
        S.current_mode[e] = s.current_mode;
        S.w_zone_counter[e] = w_zone_counter; S.flags[e] = flags;
        S.stat_short_us[e] += s.win_short_us - win_short0;
        S.stat_steps[e] += win_steps;
        S.win_sparks[e] = s.win_sparks; S.win_short_us[e] = s.win_short_us; S.win_steps[e] = win_steps0 + win_steps;
    }
    // ---------------- observation summary (what sparc_observe computes from the stored state) ----
    if (a.obs) {
        float m = spoolf;                       // cell 0; pad slots never exceed the real maximum
#pragma unroll
        for (int j = 0; j < H; j++) m = fmaxf(m, fmaxf(T[j].x, T[j].y));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL, m, o));
        if constexpr (W > 1) {
            __syncthreads();
            if (lane == 0) sm_red[ctx.wid] = (double)m;
            __syncthreads();
            for (int w = 0; w < W; w++) m = fmaxf(m, (float)sm_red[w]);
        }
        if (gl == 0) {
            const size_t ne = (size_t)a.n_envs;
            const int steps = S.win_steps[e];
            a.obs[0 * ne + e] = (float)(wp - wire);
            a.obs[1 * ne + e] = (float)v;
            a.obs[2 * ne + e] = steps > 0 ? (float)(win_vsum / (double)steps) : 0.0f;
            a.obs[3 * ne + e] = (float)s.win_sparks;
            a.obs[4 * ne + e] = (float)s.win_short_us;
            a.obs[5 * ne + e] = m;
            a.obs[6 * ne + e] = (float)s.debris_density;
            a.obs[7 * ne + e] = (float)flags;
        }
    }
}

}  // namespace sparc
