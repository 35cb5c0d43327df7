// step_kernel.cuh — the fused per-µs physics kernel (sm_100a).
//
// One kernel advances every environment through all µs of a step call.  The five reference
// modules (ignition -> material -> dielectric -> wire -> [break early-return] -> mechanics,
// src/wedm/envs/wire_edm.py:116-157) are fused; no state touches HBM between µs.
//
// Mapping: L lanes (32, 16 or 8) cooperate on one environment; 32/L environments share a warp.
// The wire temperature field lives in REGISTERS: lane `sub` owns K consecutive dynamic cells
// (cell 0 is the spool Dirichlet cell and is a constant).  Halos move with two warp shuffles
// per µs, the wire-break test is a warp vote.  The O(1)-per-µs scalar state machines run
// redundantly on every lane of the group in fp64 (bit-compatible with the reference's Python
// floats; compiled with -fmad=false so no contraction changes their rounding).
//
// Stencil arithmetic (reference: Numba kernel, src/wedm/modules/wire.py:58-123): the
// reference stores T and dT/dt in float32 and evaluates each term in float64.  Here the update
// is T' = T + dq with dq accumulated in float32 from exact neighbour differences:
//     dq = cE_j*T_j + c0_j                      convection  (-f*h_j*A*(T_j - T_diel))
//        + cR_j*(T_{j+1}-T_j) - cL*(T_j-T_{j-1})  conduction + upwind advection
//        + [I^2>1e-6, contact span] ja*T_j + jb   Joule      (f*J*(1+alpha*(T_j-T_ref)))
//        + [plasma cell] f*plasma_heat
// so the only rounding at the magnitude of T is the final add, like the reference's store.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sparc_b200.h"
#include "state_layout.h"

namespace sparc {

// ---- Philox4x32-10 (Salmon et al., SC'11); same stream as oracle/wedm_oracle.c ----------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4])
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// uniform in (0,1) from one 32-bit word: (w + 0.5) * 2^-32, exact in fp64
__device__ __forceinline__ double u01(uint32_t w) { return ((double)w + 0.5) * 2.3283064365386963e-10; }

// CPython float floor-division (Objects/floatobject.c float_divmod), used by the reference for
// plasma_idx = zone_start + int(y // segment_len)  (wire.py:286-292)
__device__ __forceinline__ double py_floordiv(double vx, double wx)
{
    double mod = fmod(vx, wx);
    double div = (vx - mod) / wx;
    if (mod != 0.0) {
        if ((wx < 0) != (mod < 0)) { mod += wx; div -= 1.0; }
    }
    double fd;
    if (div != 0.0) {
        fd = floor(div);
        if (div - fd > 0.5) fd += 1.0;
    } else {
        fd = copysign(0.0, vx / wx);
    }
    return fd;
}

template <int L>
__device__ __forceinline__ unsigned group_mask(int grp)
{
    if constexpr (L == 32) return 0xffffffffu;
    else return ((1u << L) - 1u) << (grp * L);
}

// w[idx] without dynamic indexing (keeps the words in registers)
__device__ __forceinline__ uint32_t pick4(const uint32_t w[4], int idx)
{
    uint32_t lo = (idx & 1) ? w[1] : w[0];
    uint32_t hi = (idx & 1) ? w[3] : w[2];
    return (idx & 2) ? hi : lo;
}

}  // namespace sparc


namespace sparc {

// Everything the kernel needs besides the state pointers.
struct StepArgs {
    SparcActions act;        // device pointers
    int n_envs, n_us;
    uint64_t seed;
    uint32_t env_id_base;
    const double *draws;     // [n_envs][n_us][5] when INJECT
};

#define SPARC_BLOCK_THREADS 128

// Sigmoid exponent above which the debris-short probability 1/(1+e^x) < 4.3e-18 can be treated
// as "cannot fire": on-device uniforms are >= 2^-33; injected draws are checked against U1_SAFE.
#define SPARC_EXPO_SAFE 40.0
#define SPARC_U1_SAFE 1.0e-17

// ----------------------------------------------------------------------------------------------
// Per-env scalar state.  The hot loop keeps the HOT block in registers; everything else lives
// here (thread-local memory) and is touched only on the rare "general" µs (sparks, shorts,
// latches, hysteresis refreshes) that run through general_pre().
// ----------------------------------------------------------------------------------------------
struct EnvState {
    // ---- HOT (register-mirrored by the caller) ----
    double wp, wire, v, prev_accel, target_delta, voltage, debris_volume, win_vsum, gap_prev;
    int time, spark_state, spark_dur, flags;
    // ---- cold state ----
    double debris_density, flow, d_last_gap, d_last_density, w_last_flow;
    double tv_eff, on_eff, onoff_eff, ipk;
    int latch_time;          // time at which time_since_servo was 0
    int tse_ref;             // time at which time_since_spark_end was 0
    int tsi_prev;            // time_since_spark_ignition before this µs (for the break early-return)
    int rand_short_rem, deb_short_rem, ion_end, current_mode, win_sparks, win_short_us;
    bool dens_valid;         // debris_density holds last µs's exact value
    bool pend_flush;         // last µs's debris flushing has not been applied to debris_volume yet
    // ---- fast-path constants (valid while fast_ok) ----
    double gap_lo_f, gap_hi_f, lo_c, hi_c, flush_amt;
    int onoff_thr;
    bool fast_ok;
    // ---- outputs of one general µs for the stencil / caller ----
    float jA, jB, fP, hb, he;
    int plasma_idx;
    bool heat_on, thermal, refresh, freeze, became_done;
    // ---- context ----
    int e, it;
    uint32_t env_id, episode;
    bool writer;
    // ---- mechanics state at the start of the current look-ahead block (step_kernel2.cuh) ----
    double blk_wire, blk_v, blk_pa;
};

// generator settings as the ignition module sees them (ignition.py:329-343, 98-113): `x or
// default` (0.0 and None are falsy); peak current of the latched mode.  With fresh-module
// semantics the reference's caches are pure functions of current_mode: None -> the literal
// 60.0 A of ignition.py:83; unknown -> default mode.
__device__ __forceinline__ void derive_generator(EnvState &s, const SparcParams &P, double tv, double on, double off)
{
    s.tv_eff = (tv != 0.0) ? tv : P.default_target_voltage;
    s.on_eff = (on != 0.0) ? on : P.default_on_time;
    s.onoff_eff = s.on_eff + ((off != 0.0) ? off : P.default_off_time);
    int mode = s.current_mode;
    if (mode == 0) s.ipk = 60.0;
    else s.ipk = P.currents[(mode < 1 || mode > SPARC_MAX_MODE) ? P.default_current_mode : mode];
    // smallest integer duration d with (double)d >= onoff_eff (rest -> idle, ignition.py:307-309)
    double c = ceil(s.onoff_eff);
    s.onoff_thr = (c > 2.0e9) ? 2000000000 : ((c < 0.0) ? 0 : (int)c);
}

// Exact density of the previous µs and the deferred flushing (dielectric.py:104-144), needed
// before anything reads debris_density / debris_volume "as the reference has them".
__device__ __forceinline__ void materialize(EnvState &s, const SparcParams &P)
{
    if (!s.dens_valid) {
        double g = (s.gap_prev > 0.001) ? s.gap_prev : 0.001;
        double cavity = P.cavity_volume_coeff * (g * 0.001);
        if (cavity > 0.0) {
            double d = s.debris_volume / cavity;
            s.debris_density = (d < 1.0) ? d : 1.0;
        } else s.debris_density = 0.0;
        s.dens_valid = true;
    }
    if (s.pend_flush) {
        if (s.flow > 0.001 && s.debris_volume > 0.001) {
            double nv = s.debris_volume - P.debris_removal_per_us * s.flow;
            s.debris_volume = (nv > 0.0) ? nv : 0.0;
        }
        s.pend_flush = false;
    }
}

// Decide whether the quiet-µs fast path may run next and precompute its constants.  Every bound
// is shrunk by an absolute 1e-9 so that the reference's own floating-point comparisons
// (|gap-last| > 0.01, |rho-last| > 0.001, sigmoid exponent) provably resolve the same way.
__device__ __forceinline__ void prepare_fast(EnvState &s, const SparcParams &P)
{
    const double M = 1e-9;
    bool ok = (s.rand_short_rem | s.deb_short_rem) == 0 && (s.flags & SPARC_F_DONE) == 0;
    double lo = s.d_last_gap - 0.01 + M, hi = s.d_last_gap + 0.01 - M;
    double need = (P.hard_short_gap > 0.001) ? P.hard_short_gap : 0.001;
    if (P.random_short_max_probability > 0.0 && P.random_short_max_gap > need) need = P.random_short_max_gap;
    need += M;
    if (lo < need) lo = need;
    ok = ok && (lo <= hi) && (hi < 1.0e6);
    // debris-short sigmoid must stay negligible for every gap in [lo,hi] and every density in the band
    double crit = P.base_critical_density + P.gap_coefficient * ((P.gap_coefficient >= 0.0) ? lo : hi);
    crit = (P.max_critical_density < crit) ? P.max_critical_density : crit;
    double dens_hi = s.d_last_density + 0.001;
    ok = ok && (P.sigmoid_steepness > 0.0) && (dens_hi < crit - SPARC_EXPO_SAFE / P.sigmoid_steepness - M);
    ok = ok && (s.debris_density <= dens_hi);
    // density band  |rho - last| <= 0.001  <=>  lo_c*gap <= debris <= hi_c*gap   (rho < 1)
    double k = P.cavity_volume_coeff * 0.001;
    ok = ok && (k > 0.0) && (dens_hi < 1.0);
    s.lo_c = (s.d_last_density - 0.001 + M) * k;
    s.hi_c = (dens_hi - M) * k;
    // deferred flushing: debris - amt must stay positive whenever debris > 0.001
    s.flush_amt = (s.flow > 0.001) ? P.debris_removal_per_us * s.flow : 0.0;
    ok = ok && (s.flush_amt >= 0.0) && (s.flush_amt < 0.001);
    s.gap_lo_f = lo;
    s.gap_hi_f = hi;
    s.fast_ok = ok;
}

// ----------------------------------------------------------------------------------------------
// The general µs: everything before the thermal stencil, exactly in the reference's order
// (wire_edm.py:117-127).  Rare (sparks, shorts, latches, hysteresis refreshes, first µs of a
// call), so it is kept out of line and works on the EnvState in local memory.
// ----------------------------------------------------------------------------------------------
template <bool INJECT>
__device__ __noinline__ void general_pre(EnvState &s, const SparcParams &P, const SparcStatePtrs &S,
                                         const StepArgs &a)
{
    const int e = s.e;
    const uint32_t k0 = (uint32_t)a.seed, k1 = (uint32_t)(a.seed >> 32);
    auto draw = [&](int slot) -> double {
        if (INJECT) return a.draws[((size_t)e * a.n_us + s.it) * SPARC_N_DRAW_SLOTS + slot];
        uint32_t w[4];
        philox4x32_10(s.env_id, (uint32_t)s.time >> 2, s.episode, (uint32_t)slot, k0, k1, w);
        return u01(pick4(w, s.time & 3));
    };
    s.heat_on = false; s.refresh = false; s.freeze = false; s.became_done = false;
    s.jA = 0.0f; s.jB = 0.0f; s.fP = 0.0f;
    s.thermal = true;
    s.tsi_prev = (s.spark_state == 1) ? s.spark_dur + 1 : 0;
    materialize(s, P);

    // ---- step(): action latch (wire_edm.py:117-121, 162-170) ---------------------------------
    if (s.time - s.latch_time >= P.servo_interval) {
        const double a_tv = (double)a.act.target_voltage[e];
        const double a_on = (double)a.act.on_time[e];
        const double a_off = (double)a.act.off_time[e];
        const int a_mode = a.act.current_mode[e];
        s.target_delta = (double)a.act.servo[e];
        s.current_mode = (a_mode == 0) ? -1 : a_mode;       // "I0" is a string, not None
        derive_generator(s, P, a_tv, a_on, a_off);
        if (s.writer) { S.target_voltage[e] = a_tv; S.on_time[e] = a_on; S.off_time[e] = a_off; }
        s.latch_time = s.time;
    }

    // ---- IgnitionModule._update_short_circuit_detection (ignition.py:197-245) ----------------
    const double gap_raw = s.wp - s.wire;
    const double gap = (gap_raw > 0.0) ? gap_raw : 0.0;
    bool is_short;
    if (s.rand_short_rem > 0) { s.rand_short_rem -= 1; is_short = true; }
    else if (s.deb_short_rem > 0) { s.deb_short_rem -= 1; is_short = true; }
    else {
        is_short = false;
        double p_deb;                                       // ignition.py:115-146
        bool tiny = false;                                  // 0 < p_deb < 4.3e-18
        if (gap < P.hard_short_gap) p_deb = 1.0;
        else {
            double crit = P.base_critical_density + P.gap_coefficient * gap;
            crit = (P.max_critical_density < crit) ? P.max_critical_density : crit;
            double expo = -P.sigmoid_steepness * (s.debris_density - crit);
            if (expo > 500.0) p_deb = 0.0;
            else if (expo < -500.0) p_deb = 1.0;
            else if (!INJECT && expo > SPARC_EXPO_SAFE) { p_deb = 0.0; tiny = true; }
            else p_deb = 1.0 / (1.0 + exp(expo));
        }
        (void)tiny;
        double p_rnd;
        if (gap >= P.random_short_max_gap) p_rnd = 0.0;
        else if (gap <= P.random_short_min_gap) p_rnd = P.random_short_max_probability;
        else {
            double gf = 1.0 - (gap - P.random_short_min_gap) / (P.random_short_max_gap - P.random_short_min_gap);
            p_rnd = gf * P.random_short_max_probability;
        }
        if (p_deb > 0.0 && draw(0) < p_deb) {
            s.deb_short_rem = P.debris_short_duration;
            is_short = true;
            if (s.writer) S.stat_short_events[e] += 1;
        } else if (p_rnd > 0.0 && draw(1) < p_rnd) {
            s.rand_short_rem = P.random_short_duration;
            is_short = true;
            if (s.writer) S.stat_short_events[e] += 1;
        }
    }
    if (is_short) s.voltage = 0.0;

    // ---- state machine (ignition.py:186-195, 247-319) ----------------------------------------
    bool fresh_spark = false;
    if (s.spark_state == 0) {
        if (is_short) {
            s.spark_state = -1; s.flags &= ~SPARC_F_HAS_Y; s.spark_dur = 0;
            if (s.writer) { S.stat_pulse_short[e] += 1; S.spark_y[e] = -1.0; }
        } else {
            s.voltage = s.tv_eff;
            const double u3 = draw(2);
            // get_lambda (ignition.py:348-364): ln2 / (a*gap^2 + b*gap + c), raw gap
            double den = P.ignition_a_coeff * (gap_raw * gap_raw) + P.ignition_b_coeff * gap_raw + P.ignition_c_coeff;
            if (u3 < 0.6931471805599453 / den) {
                double y = 0.0 + (P.workpiece_height - 0.0) * draw(3);      // Generator.uniform
                s.spark_state = 1; s.flags |= SPARC_F_HAS_Y; s.spark_dur = 0;
                s.voltage = s.tv_eff * P.spark_voltage_factor;
                s.win_sparks += 1;
                fresh_spark = true;
                s.plasma_idx = (P.segment_len != 0.0) ? P.zone_start + (int)py_floordiv(y, P.segment_len) : P.zone_start;
                if (s.writer) { S.stat_sparks[e] += 1; S.spark_y[e] = y; }
            }
        }
    } else if (s.spark_state == 1) {
        s.spark_dur += 1;
        if ((double)s.spark_dur >= s.on_eff) { s.spark_state = -2; if (!is_short) s.voltage = 0.0; }
        else if (!is_short) s.voltage = s.tv_eff * P.spark_voltage_factor;
    } else if (s.spark_state == -1) {
        s.spark_dur += 1;
        if ((double)s.spark_dur >= s.on_eff) s.spark_state = -2;
    } else {  // -2 rest
        s.spark_dur += 1;
        if ((double)s.spark_dur >= s.onoff_eff) {
            s.spark_state = 0; s.flags &= ~SPARC_F_HAS_Y; s.spark_dur = 0;
            if (s.writer) S.spark_y[e] = -1.0;
            if (!is_short) s.voltage = s.tv_eff;
        } else if (!is_short) s.voltage = 0.0;
    }
    if (is_short) { s.flags |= SPARC_F_SHORT; s.win_short_us += 1; }
    else s.flags &= ~SPARC_F_SHORT;
    s.win_vsum += s.voltage;
    // state.current is reassigned on every path above: Ipk in spark / short-pulse, else 0
    const double current = (s.spark_state == 1 || s.spark_state == -1) ? s.ipk : 0.0;

    // ---- MaterialRemovalModule.update (material.py:79-174) -----------------------------------
    double crater_volume = 0.0;                             // state.last_crater_volume [mm^3]
    if (fresh_spark) {
        const int mode = (s.current_mode == 0) ? 1 : s.current_mode;       // material.py:104-105
        if (mode < 1 || mode > SPARC_MAX_MODE || !P.crater_avail[mode]) {
            s.flags |= SPARC_F_MODE_ERROR;                  // the reference raises ValueError here
            s.became_done = true; s.thermal = false; s.freeze = true;
            s.fast_ok = false;
            return;
        }
        double z;
        if (INJECT) z = draw(4);
        else {
            uint32_t w[4];
            philox4x32_10(s.env_id, (uint32_t)s.time, s.episode, 4u, k0, k1, w);
            z = sqrt(-2.0 * log(u01(w[0]))) * cospi(2.0 * u01(w[1]));
        }
        double vol_um3 = P.crater_mean[mode] + P.crater_std[mode] * z;
        vol_um3 = (vol_um3 > 0.0) ? vol_um3 : 0.0;
        crater_volume = vol_um3 / 1e9;
        if (crater_volume > 0.0) {
            double kh = P.kerf_times_height[mode];
            double dx_um = 0.0;
            if (kh > 0.0) dx_um = (crater_volume / kh) * 1000.0;
            s.wp += dx_um;
        }
        if (s.writer) {
            S.stat_crater_sum[e] += vol_um3;
            S.last_crater_volume[e] = crater_volume;
            if (s.current_mode != 0) S.mat_cached_mode[e] = mode;
            S.mat_info_mode[e] = mode;
        }
    }

    // ---- DielectricModule.update (dielectric.py:82-163) --------------------------------------
    double gap_um = s.wp - s.wire;
    gap_um = (gap_um > 0.001) ? gap_um : 0.001;
    s.gap_prev = gap_um;
    const double cavity_volume = P.cavity_volume_coeff * (gap_um * 0.001);
    if (crater_volume > 0.0) {                              // fresh spark with a crater
        s.debris_volume += crater_volume;
        s.ion_end = s.time + P.ion_channel_duration;        // ion_rem = max(0, ion_end - time)
    }
    if (cavity_volume > 0.0) {
        double d = s.debris_volume / cavity_volume;
        s.debris_density = (d < 1.0) ? d : 1.0;
    } else s.debris_density = 0.0;
    s.dens_valid = true;
    if (fabs(gap_um - s.d_last_gap) > 0.01 || fabs(s.debris_density - s.d_last_density) > 0.001) {
        double r = gap_um / P.reference_gap;
        double gf = r * r * r;                              // (gap/ref)**3
        gf = (gf < 1.0) ? gf : 1.0;
        double x = P.debris_obstruction_coeff * s.debris_density;
        double df;
        if (x < 0.5) df = (1.0 - 0.5 * x) / (1.0 + 0.5 * x);               // fast_exp Pade branch
        else df = exp(-x);
        s.flow = gf * df;
        s.d_last_gap = gap_um;
        s.d_last_density = s.debris_density;
    }
    s.pend_flush = true;                                    // flushing of this µs applied lazily

    // ---- WireModule.update, host part (wire.py:259-309, 349-374) -----------------------------
    if (fabs(s.flow - s.w_last_flow) > 0.01) {
        double ve = P.convection_velocity_factor * S.unwind_v[e];
        ve = (ve > -0.9) ? ve : -0.9;
        double hb = P.base_convection_coefficient * (1.0 + ve);
        double hfloor = 0.1 * P.base_convection_coefficient;
        hb = (hb > hfloor) ? hb : hfloor;
        double he = hb * (1.0 + P.convection_flow_enhancement * s.flow);
        s.hb = (float)hb; s.he = (float)he; s.refresh = true;
        if (s.writer) { S.h_base[e] = s.hb; S.h_enh[e] = s.he; }
        s.w_last_flow = s.flow;
    }
    if (current * current > 1e-6) {                         // wire.py:98-102
        const double f = P.temp_update_factor;
        double jf = f * (P.joule_geom_factor * (current * current) * P.rho_elec);
        s.jA = (float)(jf * P.alpha_rho);
        s.jB = (float)(jf * (1.0 - P.alpha_rho * P.temp_ref));
        s.heat_on = true;
        if (s.spark_state == 1) {                           // plasma spot, wire.py:293-297
            double ph = P.plasma_efficiency * s.voltage * current;
            if (!isfinite(ph)) ph = 0.0;
            s.fP = (float)(f * ph);
        }
    }
    prepare_fast(s, P);
}

// resident blocks per SM the register budget is sized for (65536 regs / (128 thr * blocks))
#ifndef SPARC_MINB_SMALL
#define SPARC_MINB_SMALL 4
#endif
#ifndef SPARC_MINB_MID
#define SPARC_MINB_MID 3
#endif
#ifndef SPARC_MINB_BIG
#define SPARC_MINB_BIG 2
#endif
constexpr int min_blocks_for(int K) { return K <= 13 ? SPARC_MINB_SMALL : (K <= 20 ? SPARC_MINB_MID : SPARC_MINB_BIG); }

template <int K, int L, bool INJECT>
__global__ void __launch_bounds__(SPARC_BLOCK_THREADS, min_blocks_for(K))
step_kernel(const __grid_constant__ SparcParams P, const __grid_constant__ SparcStatePtrs S,
            const __grid_constant__ StepArgs a)
{
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int G = 32 / L;
    const int lane = threadIdx.x & 31;
    const int sub = lane % L;
    const int grp = lane / L;
    const int env = (blockIdx.x * (SPARC_BLOCK_THREADS / 32) + (threadIdx.x >> 5)) * G + grp;
    const bool valid = env < a.n_envs;
    const int e = valid ? env : 0;

    // ---------------- load: cold state into the local-memory struct, hot state into registers --
    EnvState s;
    s.e = e; s.it = 0; s.writer = valid && sub == 0;
    s.env_id = a.env_id_base + (uint32_t)e;
    s.episode = (uint32_t)S.episode[e];
    s.debris_density = S.debris_density[e]; s.flow = S.flow[e];
    s.d_last_gap = S.d_last_gap[e]; s.d_last_density = S.d_last_density[e];
    s.w_last_flow = S.w_last_flow[e];
    s.rand_short_rem = S.rand_short_rem[e]; s.deb_short_rem = S.deb_short_rem[e];
    s.current_mode = S.current_mode[e];
    s.win_sparks = 0; s.win_short_us = 0;
    s.dens_valid = true; s.pend_flush = false; s.fast_ok = false;
    s.tsi_prev = 0; s.plasma_idx = -1;
    s.heat_on = false; s.thermal = false; s.refresh = false; s.freeze = false; s.became_done = false;
    s.jA = 0.f; s.jB = 0.f; s.fP = 0.f; s.hb = 0.f; s.he = 0.f;
    s.gap_lo_f = 0.0; s.gap_hi_f = -1.0; s.lo_c = 0.0; s.hi_c = 0.0; s.flush_amt = 0.0; s.onoff_thr = 0;
    derive_generator(s, P, S.target_voltage[e], S.on_time[e], S.off_time[e]);

    double wp = S.wp[e], wire = S.wire[e], v = S.v[e], prev_accel = S.prev_accel[e];
    double target_delta = S.target_delta[e], voltage = S.voltage[e], debris_volume = S.debris_volume[e];
    double win_vsum = 0.0, gap_prev = wp - wire;
    int time = S.time[e], spark_state = S.spark_state[e], spark_dur = S.spark_dur[e], flags = S.flags[e];
    int crit_count = S.crit_count[e];
    int w_zone_counter = S.w_zone_counter[e];
    const int time0 = time;
    s.latch_time = time - S.t_servo[e];
    s.tse_ref = time - S.t_spark_end[e];
    s.ion_end = time + S.ion_rem[e];
    if (spark_state == 1 && (flags & SPARC_F_HAS_Y))
        s.plasma_idx = (P.segment_len != 0.0) ? P.zone_start + (int)py_floordiv(S.spark_y[e], P.segment_len) : P.zone_start;

    // fast-path constants mirrored in registers (refreshed after every general µs)
    double gap_lo_f = 0.0, gap_hi_f = -1.0, lo_c = 0.0, hi_c = 0.0, flush_amt = 0.0, tv_eff = s.tv_eff;
    int next_latch = s.latch_time + P.servo_interval, onoff_thr = s.onoff_thr;
    bool fast_ok = false, ran_fast = false, first = true, zone_mean_written = false;
    int tsi_break = -1;

    // upper bound of the ignition probability lambda(gap) = ln2/(a g^2 + b g + c) over all gaps
    // (ignition.py:348-364): a draw above it can never ignite, whatever the gap.
    double lam_bound = 2.0;
    if (P.ignition_a_coeff > 0.0) {
        double den_min = P.ignition_c_coeff - P.ignition_b_coeff * P.ignition_b_coeff / (4.0 * P.ignition_a_coeff);
        if (den_min > 0.0) lam_bound = 0.6931471805599453 / den_min * (1.0 + 1e-9);
    }
    uint32_t w_thresh = 0xffffffffu;            // word >= w_thresh  =>  u01(word) >= lam_bound
    if (lam_bound < 0.999) w_thresh = (uint32_t)(lam_bound * 4294967296.0) + 2u;

    const bool done_at_entry = !valid || (flags & SPARC_F_DONE) != 0;
    bool done = done_at_entry;

    // ---------------- temperature field -> registers -----------------------------------------
    const int N = P.n_segments;
    const int M = N - 1;                        // dynamic cells 1..N-1; cell 0 == spool_T
    const float spoolf = (float)P.spool_T;
    float T[K];
    float *Trow = S.T + (size_t)e * S.n_pad;
#pragma unroll
    for (int j = 0; j < K; j++) {
        int sl = sub * K + j;
        T[j] = (sl < M) ? Trow[sl + 1] : spoolf;
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
    float cE[K], c0[K], cR[K];
    auto refresh_coeffs = [&](float hb, float he) {
#pragma unroll
        for (int j = 0; j < K; j++) {
            int sl = sub * K + j;
            int i = sl + 1;
            if (sl < M) {
                float h = (i >= P.actual_zone_start && i < P.actual_zone_end) ? he : hb;
                double hA = (double)h * P.A;    // conv_coeff, wire.py:110
                cE[j] = (float)(-(f * hA));
                c0[j] = (float)(f * hA * P.dielectric_temperature);
                cR[j] = (i < N - 1) ? cRf : 0.0f;     // Neumann end, wire.py:95
            } else {                            // pad slot: pinned near spool_T, feeds nothing
                cE[j] = -1.0f;
                c0[j] = spoolf;
                cR[j] = 0.0f;
            }
        }
    };
    auto freeze = [&]() {                       // env finished: make the stencil a no-op
#pragma unroll
        for (int j = 0; j < K; j++) { cE[j] = 0.0f; c0[j] = 0.0f; cR[j] = 0.0f; }
        ncL = 0.0f;
    };
    refresh_coeffs(S.h_base[e], S.h_enh[e]);
    if (done) freeze();

    // contact span in slot coordinates of this lane
    const int j_span_lo = P.contact_bottom_idx - 1 - sub * K;
    const int j_span_hi = P.contact_top_idx - 1 - sub * K;

    // Philox cache for the ignition roll (slot 2): one call serves 4 consecutive µs
    uint32_t u3w[4] = {0, 0, 0, 0};
    uint32_t u3_block = 0xffffffffu;
    const double *drow = INJECT ? a.draws + (size_t)e * a.n_us * SPARC_N_DRAW_SLOTS : nullptr;

    for (int it = 0; it < a.n_us; ++it) {
        if (__all_sync(FULL, done)) break;
        bool thermal = false;                   // this env runs the thermal update this µs
        bool heat_on = false, quiet = false;
        if (!done) {
            // ================= quiet-µs fast path =============================================
            // Idle / rest µs with no latch, no short circuit, no ignition and no hysteresis
            // refresh: every reference branch is resolved by a few comparisons against bounds
            // precomputed in prepare_fast(); anything else falls through to general_pre().
            const double g = wp - wire;
            const double deb_new = (debris_volume > 0.001) ? debris_volume - flush_amt : debris_volume;
            quiet = fast_ok && !first && time < next_latch && (spark_state == 0 || spark_state == -2) &&
                    g >= gap_lo_f && g <= gap_hi_f && deb_new >= lo_c * g && deb_new <= hi_c * g;
            if (INJECT && quiet) quiet = drow[(size_t)it * SPARC_N_DRAW_SLOTS + 0] >= SPARC_U1_SAFE;
            if (quiet && spark_state == 0) {
                double u3;
                bool maybe;
                if (INJECT) {
                    u3 = drow[(size_t)it * SPARC_N_DRAW_SLOTS + 2];
                    maybe = !(u3 >= lam_bound);
                } else {
                    const uint32_t blk = (uint32_t)time >> 2;
                    if (blk != u3_block) {
                        philox4x32_10(s.env_id, blk, s.episode, 2u, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), u3w);
                        u3_block = blk;
                    }
                    const uint32_t w = pick4(u3w, time & 3);
                    maybe = w < w_thresh;
                    u3 = u01(w);
                }
                if (maybe) {
                    // u3 < fl(ln2/den) can only hold if u3*den <= ln2*(1+2^-50)
                    double den = P.ignition_a_coeff * (g * g) + P.ignition_b_coeff * g + P.ignition_c_coeff;
                    maybe = !(den > 0.0 && u3 * den > 0.6931471805599460);
                }
                quiet = !maybe;
            }
            if (quiet) {
                debris_volume = deb_new;
                gap_prev = g;
                ran_fast = true;
                if (spark_state == 0) voltage = tv_eff;
                else {
                    const int nd = spark_dur + 1;
                    if (nd >= onoff_thr) {      // rest -> idle (ignition.py:307-314)
                        spark_state = 0; spark_dur = 0; flags &= ~SPARC_F_HAS_Y;
                        voltage = tv_eff;
                        if (s.writer) S.spark_y[e] = -1.0;
                    } else { spark_dur = nd; voltage = 0.0; }
                }
                win_vsum += voltage;
                thermal = true;
            } else {
                // ================= general µs ==================================================
                s.wp = wp; s.wire = wire; s.v = v; s.prev_accel = prev_accel; s.target_delta = target_delta;
                s.voltage = voltage; s.debris_volume = debris_volume; s.win_vsum = win_vsum; s.gap_prev = gap_prev;
                s.time = time; s.spark_state = spark_state; s.spark_dur = spark_dur; s.flags = flags;
                s.it = it;
                if (ran_fast) { s.dens_valid = false; s.pend_flush = true; ran_fast = false; }
                general_pre<INJECT>(s, P, S, a);
                wp = s.wp; target_delta = s.target_delta; voltage = s.voltage; debris_volume = s.debris_volume;
                win_vsum = s.win_vsum; gap_prev = s.gap_prev;
                spark_state = s.spark_state; spark_dur = s.spark_dur; flags = s.flags;
                gap_lo_f = s.gap_lo_f; gap_hi_f = s.gap_hi_f; lo_c = s.lo_c; hi_c = s.hi_c;
                flush_amt = s.flush_amt; tv_eff = s.tv_eff; onoff_thr = s.onoff_thr;
                next_latch = s.latch_time + P.servo_interval;
                fast_ok = s.fast_ok;
                first = false;
                thermal = s.thermal;
                heat_on = s.heat_on;
                if (s.refresh) refresh_coeffs(s.hb, s.he);
                if (s.freeze) { done = true; freeze(); }
            }
        }

        // ---- compute_thermal_update (wire.py:58-123): convergent, all lanes ------------------
        float left = __shfl_up_sync(FULL, T[K - 1], 1, L);
        float right = __shfl_down_sync(FULL, T[0], 1, L);
        if (sub == 0) left = spoolf;
        float acc[K];
#pragma unroll
        for (int j = 0; j < K; j++) acc[j] = fmaf(cE[j], T[j], c0[j]);
        if (heat_on) {
            const float jA = s.jA, jB = s.jB, fP = s.fP;
            const int jp = (s.spark_state == 1 && s.plasma_idx >= 0 && s.plasma_idx < N) ? s.plasma_idx - 1 - sub * K
                                                                                         : -(1 << 20);
#pragma unroll
            for (int j = 0; j < K; j++) {
                if (j >= j_span_lo && j <= j_span_hi) acc[j] += fmaf(jA, T[j], jB);
                if (j == jp) acc[j] += fP;
            }
        }
        float lmax = spoolf;
        {
            float dprev = T[0] - left;
#pragma unroll
            for (int j = 0; j < K; j++) {
                float dnext = ((j + 1 < K) ? T[j + 1] : right) - T[j];
                float q = fmaf(cR[j], dnext, acc[j]);
                q = fmaf(ncL, dprev, q);
                dprev = dnext;
                T[j] = T[j] + q;
                lmax = fmaxf(lmax, T[j]);
            }
        }
        // ---- _check_wire_breaking (wire.py:376-388): warp vote --------------------------------
        const unsigned hot_b = __ballot_sync(FULL, lmax > P.critical_temperature_f32);
        unsigned brk_b = 0;
        if (hot_b) brk_b = __ballot_sync(FULL, lmax > P.breaking_temperature_f32);
        // zone mean (wire.py:338-347, 390-394), only when enabled
        if (P.compute_zone_mean) {
            bool need_zm = false;
            if (thermal) {
                w_zone_counter += 1;
                need_zm = w_zone_counter >= P.zone_mean_interval;
            }
            if (__any_sync(FULL, need_zm)) {
                int za = P.actual_zone_start, zb = P.actual_zone_end;
                if (!(zb > za && zb <= N)) { za = 0; zb = N; }
                float part = 0.0f;
#pragma unroll
                for (int j = 0; j < K; j++) {
                    int i = sub * K + j + 1;
                    if (i >= za && i < zb) part += T[j];
                }
                if (sub == 0 && za == 0) part += spoolf;     // cell 0 when the zone starts at 0
                double tot = (double)part;
#pragma unroll
                for (int o = L / 2; o > 0; o >>= 1) tot += __shfl_xor_sync(FULL, tot, o, L);
                if (need_zm) {
                    double zm = (double)(float)(tot / (double)(zb - za));
                    if (s.writer) { S.w_last_zone_mean[e] = zm; S.wire_avg_T[e] = zm; }
                    zone_mean_written = true;
                    w_zone_counter = 0;
                }
            }
        }
        if (thermal) {
            const unsigned gmask = group_mask<L>(grp);
            crit_count = (hot_b & gmask) ? crit_count + 1 : 0;
            if (brk_b & gmask) {
                // wire_edm.py:129-130: early return, no mechanics, no time bookkeeping
                flags |= SPARC_F_BROKEN | SPARC_F_THERMAL_BREAK;
                tsi_break = quiet ? 0 : s.tsi_prev;
                done = true;
                freeze();
            } else {
                // ---- MechanicsModule.update (mechanics.py:79-114) -----------------------------
                double a_nom;
                if (P.control_mode == 0) {
                    double x_err = wire - (wire + target_delta);
                    a_nom = P.damping_coeff * v + P.stiffness_coeff * x_err;
                } else {
                    a_nom = -P.omega_n * (v - target_delta);
                }
                double da = a_nom - prev_accel;
                if (fabs(a_nom) <= P.max_acceleration && fabs(da) <= P.max_jerk_dt) {
                    prev_accel = prev_accel + da;           // no clip active
                } else {
                    if (a_nom > P.max_acceleration) a_nom = P.max_acceleration;
                    else if (a_nom < -P.max_acceleration) a_nom = -P.max_acceleration;
                    da = a_nom - prev_accel;
                    if (da > P.max_jerk_dt) da = P.max_jerk_dt;
                    else if (da < -P.max_jerk_dt) da = -P.max_jerk_dt;
                    prev_accel = prev_accel + da;
                }
                v += prev_accel * P.mech_dt;
                if (!(fabs(v) <= P.max_speed)) {
                    if (v > P.max_speed) v = P.max_speed;
                    else if (v < -P.max_speed) v = -P.max_speed;
                }
                wire += v * P.mech_dt;
                // ---- time bookkeeping (wire_edm.py:134-144): t_servo, t_spark_end are time-based
                time += 1;
                if (!quiet && spark_state == 1) s.tse_ref = time;
                // ---- _check_termination (wire_edm.py:172-179) ---------------------------------
                if (wire > wp + 100.0) { flags |= SPARC_F_BROKEN; done = true; freeze(); }
                else if (!quiet && wp >= P.target_position) { flags |= SPARC_F_TARGET; done = true; freeze(); }
            }
        }
    }

    if (done_at_entry) {
        if (s.writer) { S.win_vsum[e] = 0.0; S.win_sparks[e] = 0; S.win_short_us[e] = 0; S.win_steps[e] = 0; }
        return;
    }
    // ---------------- store ------------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < K; j++) {
        int sl = sub * K + j;
        if (sl < M) Trow[sl + 1] = T[j];
    }
    if (sub == 0) {
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
        S.stat_short_us[e] += s.win_short_us;
        S.stat_steps[e] += win_steps;
        S.win_sparks[e] = s.win_sparks; S.win_short_us[e] = s.win_short_us; S.win_steps[e] = win_steps;
    }
}

}  // namespace sparc
