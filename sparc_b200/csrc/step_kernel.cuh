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

// resident blocks per SM the register budget is sized for (65536 regs / (128 thr * blocks))
constexpr int min_blocks_for(int K) { return K <= 13 ? 4 : (K <= 20 ? 3 : 2); }

template <int K, int L, bool INJECT>
__global__ void __launch_bounds__(SPARC_BLOCK_THREADS, min_blocks_for(K))
step_kernel(const __grid_constant__ SparcParams P, const SparcStatePtrs S, const __grid_constant__ StepArgs a)
{
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int G = 32 / L;
    const int lane = threadIdx.x & 31;
    const int sub = lane % L;
    const int grp = lane / L;
    const int env = (blockIdx.x * (SPARC_BLOCK_THREADS / 32) + (threadIdx.x >> 5)) * G + grp;
    const bool valid = env < a.n_envs;
    const int e = valid ? env : 0;
    const bool writer = valid && sub == 0;      // the lane that owns rare-event global updates

    // ---------------- per-µs scalar state in registers (identical on every lane of the group) --
    double wp = S.wp[e], wire = S.wire[e], v = S.v[e], prev_accel = S.prev_accel[e];
    double target_delta = S.target_delta[e];
    double voltage = S.voltage[e];
    double debris_volume = S.debris_volume[e], debris_density = S.debris_density[e], flow = S.flow[e];
    double d_last_gap = S.d_last_gap[e], d_last_density = S.d_last_density[e];
    double w_last_flow = S.w_last_flow[e];
    double win_vsum = 0.0;
    int time = S.time[e], t_servo = S.t_servo[e];
    int t_spark_ign = S.t_spark_ign[e], t_spark_end = S.t_spark_end[e];
    int spark_state = S.spark_state[e], spark_dur = S.spark_dur[e];
    int rand_short_rem = S.rand_short_rem[e], deb_short_rem = S.deb_short_rem[e];
    int ion_rem = S.ion_rem[e], crit_count = S.crit_count[e];
    int current_mode = S.current_mode[e];
    int w_zone_counter = S.w_zone_counter[e];
    int flags = S.flags[e];
    int win_sparks = 0, win_short_us = 0, win_steps = 0;
    bool zone_mean_written = false;

    // generator settings as the ignition module sees them (ignition.py:329-343, 98-113):
    // `x or default` (0.0 and None are falsy); peak current of the latched mode.  With
    // fresh-module semantics the reference's two caches are pure functions of current_mode:
    // None -> the literal 60.0 A of ignition.py:83; invalid -> default mode.
    double tv_eff, on_eff, onoff_eff, ipk;
    auto derive_generator = [&](double tv, double on, double off, int mode) {
        tv_eff = (tv != 0.0) ? tv : P.default_target_voltage;
        on_eff = (on != 0.0) ? on : P.default_on_time;
        onoff_eff = on_eff + ((off != 0.0) ? off : P.default_off_time);
        if (mode == 0) ipk = 60.0;
        else ipk = P.currents[(mode < 1 || mode > SPARC_MAX_MODE) ? P.default_current_mode : mode];
    };
    derive_generator(S.target_voltage[e], S.on_time[e], S.off_time[e], current_mode);

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
        int s = sub * K + j;
        T[j] = (s < M) ? Trow[s + 1] : spoolf;
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
            int s = sub * K + j;
            int i = s + 1;
            if (s < M) {
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

    // contact span / plasma cell in slot coordinates of this lane
    const int j_span_lo = P.contact_bottom_idx - 1 - sub * K;
    const int j_span_hi = P.contact_top_idx - 1 - sub * K;
    int jp = -(1 << 20);                        // slot of the plasma cell in this lane's frame
    auto set_plasma_slot = [&](double y) {      // wire.py:286-293
        int pidx = (P.segment_len != 0.0) ? P.zone_start + (int)py_floordiv(y, P.segment_len) : P.zone_start;
        jp = (pidx >= 0 && pidx < N) ? pidx - 1 - sub * K : -(1 << 20);
    };
    if (spark_state == 1 && (flags & SPARC_F_HAS_Y)) set_plasma_slot(S.spark_y[e]);

    // Philox cache for the ignition roll (slot 2): one call serves 4 consecutive µs
    uint32_t u3w[4] = {0, 0, 0, 0};
    uint32_t u3_block = 0xffffffffu;
    const uint32_t env_id = a.env_id_base + (uint32_t)e;
    const uint32_t episode = (uint32_t)S.episode[e];
    const double *drow = INJECT ? a.draws + (size_t)e * a.n_us * SPARC_N_DRAW_SLOTS : nullptr;

    for (int it = 0; it < a.n_us; ++it) {
        if (__all_sync(FULL, done)) break;
        bool thermal = false;                   // this env runs the thermal update this µs
        float jA = 0.0f, jB = 0.0f, fP = 0.0f;
        bool heat_on = false;
        if (!done) {
            thermal = true;
            auto draw = [&](int slot) -> double {
                if (INJECT) return drow[(size_t)it * SPARC_N_DRAW_SLOTS + slot];
                uint32_t w[4];
                philox4x32_10(env_id, (uint32_t)time >> 2, episode, (uint32_t)slot,
                              (uint32_t)a.seed, (uint32_t)(a.seed >> 32), w);
                return u01(pick4(w, time & 3));
            };
            // ---- step(): action latch (wire_edm.py:117-121, 162-170) -------------------------
            if (t_servo >= P.servo_interval) {
                const double a_tv = (double)a.act.target_voltage[e];
                const double a_on = (double)a.act.on_time[e];
                const double a_off = (double)a.act.off_time[e];
                const int a_mode = a.act.current_mode[e];
                target_delta = (double)a.act.servo[e];
                current_mode = (a_mode == 0) ? -1 : a_mode;     // "I0" is a string, not None
                derive_generator(a_tv, a_on, a_off, current_mode);
                if (writer) { S.target_voltage[e] = a_tv; S.on_time[e] = a_on; S.off_time[e] = a_off; }
                t_servo = 0;
            }

            // ---- IgnitionModule._update_short_circuit_detection (ignition.py:197-245) --------
            const double gap_raw = wp - wire;
            const double gap = (gap_raw > 0.0) ? gap_raw : 0.0;
            bool is_short;
            if (rand_short_rem > 0) { rand_short_rem -= 1; is_short = true; }
            else if (deb_short_rem > 0) { deb_short_rem -= 1; is_short = true; }
            else {
                is_short = false;
                double p_deb;                               // ignition.py:115-146
                if (gap < P.hard_short_gap) p_deb = 1.0;
                else {
                    double crit = P.base_critical_density + P.gap_coefficient * gap;
                    crit = (P.max_critical_density < crit) ? P.max_critical_density : crit;
                    double expo = -P.sigmoid_steepness * (debris_density - crit);
                    if (expo > 500.0) p_deb = 0.0;
                    else if (expo < -500.0) p_deb = 1.0;
                    // on-device draws are >= 2^-33; 1/(1+e^40) = 4e-18 can never exceed one
                    else if (!INJECT && expo > 40.0) p_deb = 0.0;
                    else p_deb = 1.0 / (1.0 + exp(expo));
                }
                double p_rnd;
                if (gap >= P.random_short_max_gap) p_rnd = 0.0;
                else if (gap <= P.random_short_min_gap) p_rnd = P.random_short_max_probability;
                else {
                    double gf = 1.0 - (gap - P.random_short_min_gap) / (P.random_short_max_gap - P.random_short_min_gap);
                    p_rnd = gf * P.random_short_max_probability;
                }
                if (p_deb > 0.0 && draw(0) < p_deb) {
                    deb_short_rem = P.debris_short_duration;
                    is_short = true;
                    if (writer) S.stat_short_events[e] += 1;
                } else if (p_rnd > 0.0 && draw(1) < p_rnd) {
                    rand_short_rem = P.random_short_duration;
                    is_short = true;
                    if (writer) S.stat_short_events[e] += 1;
                }
            }
            if (is_short) voltage = 0.0;

            // ---- state machine (ignition.py:186-195, 247-319) --------------------------------
            bool fresh_spark = false;
            if (spark_state == 0) {
                if (is_short) {
                    spark_state = -1; flags &= ~SPARC_F_HAS_Y; spark_dur = 0;
                    if (writer) { S.stat_pulse_short[e] += 1; S.spark_y[e] = -1.0; }
                } else {
                    voltage = tv_eff;
                    double u3;
                    if (INJECT) u3 = draw(2);
                    else {
                        uint32_t blk = (uint32_t)time >> 2;
                        if (blk != u3_block) {
                            philox4x32_10(env_id, blk, episode, 2u, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), u3w);
                            u3_block = blk;
                        }
                        u3 = u01(pick4(u3w, time & 3));
                    }
                    // get_lambda (ignition.py:348-364): ln2 / (a*gap^2 + b*gap + c), raw gap
                    double den = P.ignition_a_coeff * (gap_raw * gap_raw) + P.ignition_b_coeff * gap_raw + P.ignition_c_coeff;
                    if (u3 < 0.6931471805599453 / den) {
                        double y = 0.0 + (P.workpiece_height - 0.0) * draw(3);   // Generator.uniform
                        spark_state = 1; flags |= SPARC_F_HAS_Y; spark_dur = 0;
                        voltage = tv_eff * P.spark_voltage_factor;
                        win_sparks += 1;
                        fresh_spark = true;
                        set_plasma_slot(y);
                        if (writer) { S.stat_sparks[e] += 1; S.spark_y[e] = y; }
                    }
                }
            } else if (spark_state == 1) {
                spark_dur += 1;
                if ((double)spark_dur >= on_eff) { spark_state = -2; if (!is_short) voltage = 0.0; }
                else if (!is_short) voltage = tv_eff * P.spark_voltage_factor;
            } else if (spark_state == -1) {
                spark_dur += 1;
                if ((double)spark_dur >= on_eff) spark_state = -2;
            } else {  // -2 rest
                spark_dur += 1;
                if ((double)spark_dur >= onoff_eff) {
                    spark_state = 0; flags &= ~SPARC_F_HAS_Y; spark_dur = 0;
                    if (writer) S.spark_y[e] = -1.0;
                    if (!is_short) voltage = tv_eff;
                } else if (!is_short) voltage = 0.0;
            }
            if (is_short) { flags |= SPARC_F_SHORT; win_short_us += 1; }
            else flags &= ~SPARC_F_SHORT;
            win_vsum += voltage;
            // state.current is reassigned on every path above: Ipk in spark / short-pulse, else 0
            const double current = (spark_state == 1 || spark_state == -1) ? ipk : 0.0;

            // ---- MaterialRemovalModule.update (material.py:79-174) ---------------------------
            double crater_volume = 0.0;                     // state.last_crater_volume [mm^3]
            if (fresh_spark) {
                const int mode = (current_mode == 0) ? 1 : current_mode;   // material.py:104-105
                if (mode < 1 || mode > SPARC_MAX_MODE || !P.crater_avail[mode]) {
                    flags |= SPARC_F_MODE_ERROR;            // the reference raises ValueError here
                    done = true; thermal = false;
                    freeze();
                } else {
                    double z;
                    if (INJECT) z = draw(4);
                    else {
                        uint32_t w[4];
                        philox4x32_10(env_id, (uint32_t)time, episode, 4u, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), w);
                        z = sqrt(-2.0 * log(u01(w[0]))) * cos(6.283185307179586 * u01(w[1]));
                    }
                    double vol_um3 = P.crater_mean[mode] + P.crater_std[mode] * z;
                    vol_um3 = (vol_um3 > 0.0) ? vol_um3 : 0.0;
                    crater_volume = vol_um3 / 1e9;
                    if (crater_volume > 0.0) {
                        double kh = P.kerf_times_height[mode];
                        double dx_um = 0.0;
                        if (kh > 0.0) dx_um = (crater_volume / kh) * 1000.0;
                        wp += dx_um;
                    }
                    if (writer) {
                        S.stat_crater_sum[e] += vol_um3;
                        S.last_crater_volume[e] = crater_volume;
                        S.mat_cached_mode[e] = (current_mode == 0) ? S.mat_cached_mode[e] : mode;
                        S.mat_info_mode[e] = mode;
                    }
                }
            }
            if (!done) {
                // ---- DielectricModule.update (dielectric.py:82-163) --------------------------
                double gap_um = wp - wire;
                gap_um = (gap_um > 0.001) ? gap_um : 0.001;
                const double cavity_volume = P.cavity_volume_coeff * (gap_um * 0.001);
                if (crater_volume > 0.0) {                  // fresh spark with a crater
                    debris_volume += crater_volume;
                    ion_rem = P.ion_channel_duration;
                }
                if (cavity_volume > 0.0) {
                    double d = debris_volume / cavity_volume;
                    debris_density = (d < 1.0) ? d : 1.0;
                } else debris_density = 0.0;
                if (fabs(gap_um - d_last_gap) > 0.01 || fabs(debris_density - d_last_density) > 0.001) {
                    double r = gap_um / P.reference_gap;
                    double gf = r * r * r;                  // (gap/ref)**3
                    gf = (gf < 1.0) ? gf : 1.0;
                    double x = P.debris_obstruction_coeff * debris_density;
                    double df;
                    if (x < 0.5) df = (1.0 - 0.5 * x) / (1.0 + 0.5 * x);   // fast_exp Pade branch
                    else df = exp(-x);
                    flow = gf * df;
                    d_last_gap = gap_um;
                    d_last_density = debris_density;
                }
                if (flow > 0.001 && debris_volume > 0.001) {
                    double nv = debris_volume - P.debris_removal_per_us * flow;
                    debris_volume = (nv > 0.0) ? nv : 0.0;
                }
                if (ion_rem > 0) ion_rem = (ion_rem > 1) ? ion_rem - 1 : 0;

                // ---- WireModule.update, host part (wire.py:259-309, 349-374) -----------------
                if (fabs(flow - w_last_flow) > 0.01) {
                    double ve = P.convection_velocity_factor * S.unwind_v[e];
                    ve = (ve > -0.9) ? ve : -0.9;
                    double hb = P.base_convection_coefficient * (1.0 + ve);
                    double hfloor = 0.1 * P.base_convection_coefficient;
                    hb = (hb > hfloor) ? hb : hfloor;
                    double he = hb * (1.0 + P.convection_flow_enhancement * flow);
                    refresh_coeffs((float)hb, (float)he);
                    if (writer) { S.h_base[e] = (float)hb; S.h_enh[e] = (float)he; }
                    w_last_flow = flow;
                }
                if (current * current > 1e-6) {             // wire.py:98-102
                    double jf = f * (P.joule_geom_factor * (current * current) * P.rho_elec);
                    jA = (float)(jf * P.alpha_rho);
                    jB = (float)(jf * (1.0 - P.alpha_rho * P.temp_ref));
                    heat_on = true;
                    if (spark_state == 1) {                 // plasma spot, wire.py:293-297
                        double ph = P.plasma_efficiency * voltage * current;
                        if (!isfinite(ph)) ph = 0.0;
                        fP = (float)(f * ph);
                    }
                }
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
                    if (writer) { S.w_last_zone_mean[e] = zm; S.wire_avg_T[e] = zm; }
                    zone_mean_written = true;
                    w_zone_counter = 0;
                }
            }
        }
        if (thermal) {
            const unsigned gmask = group_mask<L>(grp);
            if (hot_b & gmask) crit_count += 1; else crit_count = 0;
            win_steps += 1;
            if (brk_b & gmask) {
                // wire_edm.py:129-130: early return, no mechanics, no time bookkeeping
                flags |= SPARC_F_BROKEN | SPARC_F_THERMAL_BREAK;
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
                if (a_nom > P.max_acceleration) a_nom = P.max_acceleration;
                else if (a_nom < -P.max_acceleration) a_nom = -P.max_acceleration;
                double da = a_nom - prev_accel;
                if (da > P.max_jerk_dt) da = P.max_jerk_dt;
                else if (da < -P.max_jerk_dt) da = -P.max_jerk_dt;
                prev_accel = prev_accel + da;
                v += prev_accel * P.mech_dt;
                if (v > P.max_speed) v = P.max_speed;
                else if (v < -P.max_speed) v = -P.max_speed;
                wire += v * P.mech_dt;
                // ---- time bookkeeping (wire_edm.py:134-144) -----------------------------------
                time += 1; t_servo += 1;
                if (spark_state == 1) { t_spark_ign += 1; t_spark_end = 0; }
                else { t_spark_end += 1; t_spark_ign = 0; }
                // ---- _check_termination (wire_edm.py:172-179) ---------------------------------
                if (wire > wp + 100.0) { flags |= SPARC_F_BROKEN; done = true; freeze(); }
                else if (wp >= P.target_position) { flags |= SPARC_F_TARGET; done = true; freeze(); }
            }
        }
    }

    if (done_at_entry) {
        if (writer) { S.win_vsum[e] = 0.0; S.win_sparks[e] = 0; S.win_short_us[e] = 0; S.win_steps[e] = 0; }
        return;
    }
    // ---------------- store ------------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < K; j++) {
        int s = sub * K + j;
        if (s < M) Trow[s + 1] = T[j];
    }
    if (sub == 0) {
        Trow[0] = spoolf;
        S.wp[e] = wp; S.wire[e] = wire; S.v[e] = v; S.prev_accel[e] = prev_accel;
        S.voltage[e] = voltage;
        S.current[e] = (spark_state == 1 || spark_state == -1) ? ipk : 0.0;
        S.target_delta[e] = target_delta;
        S.debris_volume[e] = debris_volume; S.debris_density[e] = debris_density;
        {
            double g = wp - wire;
            g = (g > 0.001) ? g : 0.001;
            // cavity of the last dielectric update; after a completed µs the wire has moved, so
            // recompute from the pre-mechanics gap is not possible here — it is refreshed from
            // the same expression at the next µs; expose the current-gap value.
            S.cavity_volume[e] = P.cavity_volume_coeff * (g * 0.001);
        }
        S.flow[e] = flow;
        if (!(spark_state == 1 && spark_dur == 0) && win_steps > 0) S.last_crater_volume[e] = 0.0;
        S.d_last_gap[e] = d_last_gap; S.d_last_density[e] = d_last_density;
        S.w_last_flow[e] = w_last_flow;
        if (P.compute_zone_mean && win_steps > 0 && !zone_mean_written) S.wire_avg_T[e] = S.w_last_zone_mean[e];
        S.win_vsum[e] = win_vsum;
        S.time[e] = time; S.t_servo[e] = t_servo; S.t_open_v[e] = time;
        S.t_spark_ign[e] = t_spark_ign; S.t_spark_end[e] = t_spark_end;
        S.spark_state[e] = spark_state; S.spark_dur[e] = spark_dur;
        S.rand_short_rem[e] = rand_short_rem; S.deb_short_rem[e] = deb_short_rem;
        S.ion_rem[e] = ion_rem; S.crit_count[e] = crit_count;
        S.current_mode[e] = current_mode;
        S.w_zone_counter[e] = w_zone_counter; S.flags[e] = flags;
        S.stat_short_us[e] += win_short_us;
        S.stat_steps[e] += win_steps;
        S.win_sparks[e] = win_sparks; S.win_short_us[e] = win_short_us; S.win_steps[e] = win_steps;
    }
}

}  // namespace sparc
