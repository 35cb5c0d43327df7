// step_common.cuh — shared pieces of the fused per-µs physics kernel (sm_100a): Philox, the
// per-env scalar state, and the exact general µs (general_pre).  The kernel is in step_kernel.cuh.
//
// One kernel advances every environment through all µs of a step call.  The five reference
// modules (ignition -> material -> dielectric -> wire -> [break early-return] -> mechanics,
// src/wedm/envs/wire_edm.py:116-157) are fused; no state touches HBM between µs.
//
// Mapping: one warp per environment.  The wire temperature field lives in REGISTERS: each lane
// owns K consecutive dynamic cells (cell 0 is the spool Dirichlet cell and is a constant).
// Halos move with two warp shuffles per µs, the wire-break test is a warp vote.  The scalar
// state machines run in fp64 (bit-compatible with the reference's Python floats; compiled with
// -fmad=false so no contraction changes their rounding), redundantly on every lane.
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

// w[idx] without dynamic indexing (keeps the words in registers)
__device__ __forceinline__ uint32_t pick4(const uint32_t w[4], int idx)
{
    uint32_t lo = (idx & 1) ? w[1] : w[0];
    uint32_t hi = (idx & 1) ? w[3] : w[2];
    return (idx & 2) ? hi : lo;
}

}  // namespace sparc


namespace sparc {

// Register layout of the wire inside one environment's 32*W lanes (computed on the host per launch,
// sparc_b200.cu:compute_layout).  Every lane holds two HALVES of H = K/2 slots (the x and the y
// components of its packed registers); half h = 2*lane + {0,1} covers consecutive cells.  The cells
// 1..N-1 are split into up to three REGIONS (below the flushing zone, the zone, above it; WireModule
// gives the zone its own convection coefficient, wire.py:349-374) and every region starts on a half
// boundary, so the two convection coefficients are one value per half instead of one per cell.
// The slack is absorbed by SHORT halves: a short half holds H-1 cells and a pad in its last slot (a
// compile-time register), which the stencil keeps equal to the right-hand neighbour, so that it is
// transparent.  The first p[r] halves of region r are short.  Halves >= hb[3] hold no cells; the last
// real cell of half hb[3]-1 is cell N-1, the zero-flux end (wire.py:95).  aligned == 0: one region,
// per-cell coefficients (geometries whose regions are too small to be aligned).
struct SparcLayout {
    int hb[4];      // first half of region r; hb[3] = number of halves in use
    int p[3];       // short halves of region r
    int cb[3];      // first cell of region r
    int aligned;
};

// Everything the kernel needs besides the state pointers.
struct StepArgs {
    SparcActions act;        // device pointers
    int n_envs, n_us;
    uint64_t seed;
    uint32_t env_id_base;
    const double *draws;     // [n_envs][n_us][5] when INJECT
    float *tlog;             // temperature-field log [n_log_envs][total_us / log_every][n_segments] or NULL
    int log_every, n_log_envs;
    int it_base, total_us;   // this launch covers µs [it_base, it_base + n_us) of a call of total_us µs
    float *obs;              // observation summary [SPARC_N_OBS][n_envs] written by the kernel epilogue, or NULL
    uint32_t mech_seq;       // tag of this launch pair in SparcStatePtrs::mech_prog (multiple of 64)
    double *signals;         // per-signal log [total_us / sample_every][SPARC_N_SIGNALS][n_envs] of this CALL, or NULL
    int sample_every;
    SparcLayout lay;
    SparcController ctrl;    // kind 0: the servo command is act.servo[e]
};

// First cell, number of real cells (H, H-1 or 0) and region of half h under layout L (host + device: the host
// exposes it through sparc_layout_info for the CPU-side layout sweep in tests/test_abi_cpu.py).
__host__ __device__ __forceinline__ void half_geom(const SparcLayout &L, const int h, const int H, int &c_first, int &n, int &region)
{
    if (h >= L.hb[3]) { c_first = 0; n = 0; region = -1; return; }
    const int r = (h >= L.hb[2]) ? 2 : ((h >= L.hb[1]) ? 1 : 0);
    const int k = h - L.hb[r], p = L.p[r];
    int before, sh;
    if (L.aligned) {                            // pads only in y halves: the first p odd halves of the region
        const int cnt = (L.hb[r] & 1) ? ((k + 1) >> 1) : (k >> 1);      // odd halves of the region before h
        before = (cnt < p) ? cnt : p;
        sh = ((h & 1) && cnt < p) ? 1 : 0;
    } else {                                    // the first p halves of the (single) region
        before = (k < p) ? k : p;
        sh = (k < p) ? 1 : 0;
    }
    c_first = L.cb[r] + k * H - before;
    n = H - sh;
    region = r;
}

// The servo command of this launch: the staged action, or the in-kernel controller evaluated from the
// state the launch starts with (SparcController, include/sparc_b200.h).  mech_kernel and step_kernel
// both call it at their start and read the same stored state (the step kernel overwrites wp / wire
// only in its epilogue, after the pre-pass has finished).
__device__ __forceinline__ double servo_command(const StepArgs &a, const SparcStatePtrs &S, const int e)
{
    if (a.ctrl.kind == 1) {
        double d = ((S.wp[e] - S.wire[e]) - a.ctrl.desired_gap) * a.ctrl.gain;
        if (a.ctrl.limit >= 0.0) d = fmin(fmax(d, -a.ctrl.limit), a.ctrl.limit);
        return (double)(float)d;
    }
    return (double)a.act.servo[e];
}

// staging buffers per env of the LOG kernels: block-per-env variants keep two row copies in flight
#define SPARC_LOG_NBUF(W) ((W) > 1 ? 3 : 2)

#ifndef SPARC_BLOCK_THREADS
#define SPARC_BLOCK_THREADS 64      /* W == 1: two envs per block */
#endif

// Sigmoid exponent above which the debris-short probability 1/(1+e^x) < 4.3e-18 can be treated
// as "cannot fire": on-device uniforms are >= 2^-33; injected draws are checked against U1_SAFE.
#define SPARC_EXPO_SAFE 40.0
#define SPARC_U1_SAFE 1.0e-17

// ----------------------------------------------------------------------------------------------
// Per-env scalar state.  The hot loop keeps the HOT block in registers; everything else lives
// here (thread-local memory) and is touched only on the rare "general" µs (sparks, shorts,
// latches, hysteresis refreshes) that run through general_pre().
// ----------------------------------------------------------------------------------------------
// The per-environment parameters of one env (blob fields prm_<name>, state_layout.h), loaded into the warp's
// record at the start of a launch: every use below reads `q.<name>`, never `P.<name>`.
struct EnvPrm {
#define X(name, ctype, dt) ctype name;
    SPARC_ENV_PARAM_FIELDS(X)
#undef X
};

__device__ __forceinline__ void load_env_prm(EnvPrm &q, const SparcStatePtrs &S, const int e)
{
#define X(name, ctype, dt) q.name = S.prm_##name[e];
    SPARC_ENV_PARAM_FIELDS(X)
#undef X
}

struct EnvState {
    EnvPrm q;
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
    // ---- launch bookkeeping (registers in the round-1 kernel; the record lives in shared memory) ----
    double wire0, delta_old0, delta_new0;        // mechanics inputs of this launch (for the exact rewind)
    double fin_wire, fin_v, fin_pa;              // mechanics state to store at the end of the launch
    double lam_bound;                            // upper bound of lambda(gap) over all gaps
    double cfac;                                 // f = temp_update_factor
    double avg_view;                             // EDMState.wire_average_temperature as the reference holds it now
    int crit_count, w_zone_counter, next_latch, skip_flush, tsi_break, time0, first_latch_it;
    int win_short0, win_steps0, win_steps_out;
    uint32_t w_thresh;                           // Philox word >= w_thresh can never ignite
    uint32_t rng_cbase;                          // first Philox counter of the cached window
    uint32_t mech_ready;                         // pre-pass blocks known to be published
    uint32_t mech_self;                          // blocks this warp computed itself after a hand-off timeout
    bool ran_fast, zone_mean_written, cold_ok, convex, rng_valid, done;
    bool ign_latched;        // IgnitionModule._cached_current_mode is not None (ignition.py:82-113): survives a reference-style reset
};

// Per-warp record in shared memory: the scalar state, the wire positions of the current 32-µs block
// (pos[0] before its first µs, pos[1 + l] after µs l) and a window of Philox ignition words
// (rng[c - rng_cbase][w] = word w of counter c: 128 µs per refill instead of 32).
struct __align__(16) WarpRec {
    uint32_t rng[32][4];
    double pos[33];
    double vel[33];          // wire velocity after µs l of the block at vel[1 + l] (only while signals are sampled)
    EnvState s;
};

// generator settings as the ignition module sees them (ignition.py:329-343, 98-113): `x or
// default` (0.0 and None are falsy); peak current of the latched mode.  With fresh-module
// semantics the reference's caches are pure functions of current_mode: None -> the literal
// 60.0 A of ignition.py:83; unknown -> default mode.
__device__ __forceinline__ void derive_generator(EnvState &s, const SparcParams &P, double tv, double on, double off)
{
    s.tv_eff = (tv != 0.0) ? tv : s.q.default_target_voltage;
    s.on_eff = (on != 0.0) ? on : s.q.default_on_time;
    s.onoff_eff = s.on_eff + ((off != 0.0) ? off : s.q.default_off_time);
    int mode = s.current_mode;
    // None: the literal 60.0 A of a fresh module (ignition.py:83); once any mode was cached, None != cache sends it
    // through the default mode instead (ignition.py:101-110) — only visible after a reference-style reset
    if (mode == 0) s.ipk = s.ign_latched ? P.currents[s.q.default_current_mode] : 60.0;
    else s.ipk = P.currents[(mode < 1 || mode > SPARC_MAX_MODE) ? s.q.default_current_mode : mode];
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
            double nv = s.debris_volume - s.q.debris_removal_per_us * s.flow;
            s.debris_volume = (nv > 0.0) ? nv : 0.0;
        }
        s.pend_flush = false;
    }
}

// What the reference holds in (debris_density, debris_volume) after the last executed µs, WITHOUT touching the
// lazily kept state (signal rows sampled in the middle of a launch).  Same arithmetic as materialize().
__device__ __forceinline__ void dielectric_view(const EnvState &s, const SparcParams &P, double &dens, double &vol)
{
    const bool dens_valid = s.dens_valid && !s.ran_fast, pend = s.pend_flush || s.ran_fast;
    dens = s.debris_density;
    if (!dens_valid) {
        const double g = (s.gap_prev > 0.001) ? s.gap_prev : 0.001;
        const double cavity = P.cavity_volume_coeff * (g * 0.001);
        if (cavity > 0.0) { const double d = s.debris_volume / cavity; dens = (d < 1.0) ? d : 1.0; }
        else dens = 0.0;
    }
    vol = s.debris_volume;
    if (pend && s.flow > 0.001 && vol > 0.001) {
        const double nv = vol - s.q.debris_removal_per_us * s.flow;
        vol = (nv > 0.0) ? nv : 0.0;
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
    double need = (s.q.hard_short_gap > 0.001) ? s.q.hard_short_gap : 0.001;
    if (s.q.random_short_max_probability > 0.0 && s.q.random_short_max_gap > need) need = s.q.random_short_max_gap;
    need += M;
    if (lo < need) lo = need;
    ok = ok && (lo <= hi) && (hi < 1.0e6);
    // debris-short sigmoid must stay negligible for every gap in [lo,hi] and every density in the band
    double crit = s.q.base_critical_density + s.q.gap_coefficient * ((s.q.gap_coefficient >= 0.0) ? lo : hi);
    crit = (s.q.max_critical_density < crit) ? s.q.max_critical_density : crit;
    double dens_hi = s.d_last_density + 0.001;
    ok = ok && (s.q.sigmoid_steepness > 0.0) && (dens_hi < crit - SPARC_EXPO_SAFE / s.q.sigmoid_steepness - M);
    ok = ok && (s.debris_density <= dens_hi);
    // density band  |rho - last| <= 0.001  <=>  lo_c*gap <= debris <= hi_c*gap   (rho < 1)
    double k = P.cavity_volume_coeff * 0.001;
    ok = ok && (k > 0.0) && (dens_hi < 1.0);
    s.lo_c = (s.d_last_density - 0.001 + M) * k;
    s.hi_c = (dens_hi - M) * k;
    // deferred flushing: debris - amt must stay positive whenever debris > 0.001
    s.flush_amt = (s.flow > 0.001) ? s.q.debris_removal_per_us * s.flow : 0.0;
    ok = ok && (s.flush_amt >= 0.0) && (s.flush_amt < 0.001);
    s.gap_lo_f = lo;
    s.gap_hi_f = hi;
    s.fast_ok = ok;
}

// DielectricModule.update hysteresis (dielectric.py:118-137): recompute the flow condition only when
// the gap or the debris density moved out of the band around the values of the last recomputation.
__device__ __forceinline__ double flow_condition(const EnvPrm &q, const double gap_um, const double density)
{
    double r = gap_um / q.reference_gap;
    double gf = r * r * r;                                  // (gap/ref)**3
    gf = (gf < 1.0) ? gf : 1.0;
    double x = q.debris_obstruction_coeff * density;
    double df;
    if (x < 0.5) df = (1.0 - 0.5 * x) / (1.0 + 0.5 * x);                   // fast_exp Pade branch
    else df = exp(-x);
    return gf * df;
}

__device__ __forceinline__ bool dielectric_hysteresis(EnvState &s, const SparcParams &P, const double gap_um)
{
    if (fabs(gap_um - s.d_last_gap) > 0.01 || fabs(s.debris_density - s.d_last_density) > 0.001) {
        s.flow = flow_condition(s.q, gap_um, s.debris_density);
        s.d_last_gap = gap_um;
        s.d_last_density = s.debris_density;
        return true;
    }
    return false;
}

// WireModule.update convection refresh (wire.py:278-281, 349-374): new h_base / h_enh when the flow
// condition moved by more than 0.01 since the last refresh; the caller rebuilds its coefficients.
__device__ __forceinline__ void wire_convection_refresh(EnvState &s, const SparcParams &P, const SparcStatePtrs &S)
{
    if (fabs(s.flow - s.w_last_flow) > 0.01) {
        double ve = s.q.convection_velocity_factor * S.unwind_v[s.e];
        ve = (ve > -0.9) ? ve : -0.9;
        double hb = s.q.base_convection_coefficient * (1.0 + ve);
        double hfloor = 0.1 * s.q.base_convection_coefficient;
        hb = (hb > hfloor) ? hb : hfloor;
        double he = hb * (1.0 + s.q.convection_flow_enhancement * s.flow);
        s.hb = (float)hb; s.he = (float)he; s.refresh = true;
        if (s.writer) { S.h_base[s.e] = s.hb; S.h_enh[s.e] = s.he; }
        s.w_last_flow = s.flow;
    }
}

// Hysteresis refresh without the general µs.  The look-ahead stopped at a µs whose only failed
// condition is a band (gap or debris density moved out of the fast-path band) while everything the
// ignition module does on that µs is provably quiet: idle or rest state, no latch, no countdown, no
// ignition, and — checked here against the density bound of the OLD band, which is what
// _update_short_circuit_detection sees — no short circuit at gap g.  Then the reference's µs is a quiet
// µs whose DielectricModule.update recomputes the flow: apply the pending flush of the previous µs,
// evaluate the exact hysteresis condition, refresh, and re-centre the fast-path bands; the µs itself
// then runs as the first µs of the next quiet run (its own flush is applied by that run, hence `no
// pending flush` on return).  Returns false, with nothing modified, when the exact condition does not
// ask for a refresh, when the new flow would also refresh the wire's convection coefficients, or when
// the short-circuit side cannot be proven quiet: the caller takes the general µs.
static __device__ __noinline__ bool light_refresh(EnvState &s, const SparcParams &P, const SparcStatePtrs &S, const double g,
                                           double &debris_volume, const bool pending)
{
    const double M = 1e-9;
    double need = (s.q.hard_short_gap > 0.001) ? s.q.hard_short_gap : 0.001;
    if (s.q.random_short_max_probability > 0.0 && s.q.random_short_max_gap > need) need = s.q.random_short_max_gap;
    if (!(g >= need + M) || !(s.q.sigmoid_steepness > 0.0)) return false;
    double crit = s.q.base_critical_density + s.q.gap_coefficient * g;
    crit = (s.q.max_critical_density < crit) ? s.q.max_critical_density : crit;
    if (!(s.d_last_density + 0.001 < crit - SPARC_EXPO_SAFE / s.q.sigmoid_steepness - M)) return false;
    double deb = debris_volume;
    if (pending && s.flow > 0.001 && deb > 0.001) {         // flush of the previous µs (dielectric.py:139-144)
        double nv = deb - s.q.debris_removal_per_us * s.flow;
        deb = (nv > 0.0) ? nv : 0.0;
    }
    const double gap_um = (g > 0.001) ? g : 0.001;
    const double cavity = P.cavity_volume_coeff * (gap_um * 0.001);
    double dens = 0.0;
    if (cavity > 0.0) { double d = deb / cavity; dens = (d < 1.0) ? d : 1.0; }
    if (!(fabs(gap_um - s.d_last_gap) > 0.01 || fabs(dens - s.d_last_density) > 0.001)) return false;
    const double flow = flow_condition(s.q, gap_um, dens);
    // a flow change that also moves the wire's convection coefficients is left to the general µs
    if (fabs(flow - s.w_last_flow) > 0.01) return false;
    debris_volume = deb;
    s.debris_volume = deb;
    s.debris_density = dens;
    s.dens_valid = true;
    s.pend_flush = false;
    s.flow = flow;
    s.d_last_gap = gap_um;
    s.d_last_density = dens;
    prepare_fast(s, P);
    return true;
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
        if (INJECT) return a.draws[((size_t)e * a.total_us + a.it_base + s.it) * SPARC_N_DRAW_SLOTS + slot];
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
        s.target_delta = s.delta_new0;                      // act.servo[e] or the in-kernel controller's command
        s.current_mode = (a_mode == 0) ? -1 : a_mode;       // "I0" is a string, not None
        s.ign_latched = true;                               // the next _get_current_from_mode() caches a mode
        if (s.writer) S.ign_cached_mode[e] = (a_mode < 1 || a_mode > SPARC_MAX_MODE) ? s.q.default_current_mode : a_mode;
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
        if (gap < s.q.hard_short_gap) p_deb = 1.0;
        else {
            double crit = s.q.base_critical_density + s.q.gap_coefficient * gap;
            crit = (s.q.max_critical_density < crit) ? s.q.max_critical_density : crit;
            double expo = -s.q.sigmoid_steepness * (s.debris_density - crit);
            if (expo > 500.0) p_deb = 0.0;
            else if (expo < -500.0) p_deb = 1.0;
            else if (!INJECT && expo > SPARC_EXPO_SAFE) { p_deb = 0.0; tiny = true; }
            else p_deb = 1.0 / (1.0 + exp(expo));
        }
        (void)tiny;
        double p_rnd;
        if (gap >= s.q.random_short_max_gap) p_rnd = 0.0;
        else if (gap <= s.q.random_short_min_gap) p_rnd = s.q.random_short_max_probability;
        else {
            double gf = 1.0 - (gap - s.q.random_short_min_gap) / (s.q.random_short_max_gap - s.q.random_short_min_gap);
            p_rnd = gf * s.q.random_short_max_probability;
        }
        if (p_deb > 0.0 && draw(0) < p_deb) {
            s.deb_short_rem = s.q.debris_short_duration;
            is_short = true;
            if (s.writer) S.stat_short_events[e] += 1;
        } else if (p_rnd > 0.0 && draw(1) < p_rnd) {
            s.rand_short_rem = s.q.random_short_duration;
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
            double den = s.q.ignition_a_coeff * (gap_raw * gap_raw) + s.q.ignition_b_coeff * gap_raw + s.q.ignition_c_coeff;
            if (u3 < 0.6931471805599453 / den) {
                double y = 0.0 + (P.workpiece_height - 0.0) * draw(3);      // Generator.uniform
                s.spark_state = 1; s.flags |= SPARC_F_HAS_Y; s.spark_dur = 0;
                s.voltage = s.tv_eff * s.q.spark_voltage_factor;
                s.win_sparks += 1;
                fresh_spark = true;
                s.plasma_idx = (P.segment_len != 0.0) ? P.zone_start + (int)py_floordiv(y, P.segment_len) : P.zone_start;
                if (s.writer) { S.stat_sparks[e] += 1; S.spark_y[e] = y; }
            }
        }
    } else if (s.spark_state == 1) {
        s.spark_dur += 1;
        if ((double)s.spark_dur >= s.on_eff) { s.spark_state = -2; if (!is_short) s.voltage = 0.0; }
        else if (!is_short) s.voltage = s.tv_eff * s.q.spark_voltage_factor;
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
        s.ion_end = s.time + s.q.ion_channel_duration;        // ion_rem = max(0, ion_end - time)
    }
    if (cavity_volume > 0.0) {
        double d = s.debris_volume / cavity_volume;
        s.debris_density = (d < 1.0) ? d : 1.0;
    } else s.debris_density = 0.0;
    s.dens_valid = true;
    dielectric_hysteresis(s, P, gap_um);
    s.pend_flush = true;                                    // flushing of this µs applied lazily

    // ---- WireModule.update, host part (wire.py:259-309, 349-374) -----------------------------
    wire_convection_refresh(s, P, S);
    if (current * current > 1e-6) {                         // wire.py:98-102
        const double f = P.temp_update_factor;
        double jf = f * (P.joule_geom_factor * (current * current) * P.rho_elec);
        s.jA = (float)(jf * P.alpha_rho);
        s.jB = (float)(jf * (1.0 - P.alpha_rho * P.temp_ref));
        s.heat_on = true;
        if (s.spark_state == 1) {                           // plasma spot, wire.py:293-297
            double ph = s.q.plasma_efficiency * s.voltage * current;
            if (!isfinite(ph)) ph = 0.0;
            s.fP = (float)(f * ph);
        }
    }
    prepare_fast(s, P);
}

// MechanicsModule.update (mechanics.py:79-114), one µs.
__device__ __forceinline__ void mech_step(const SparcParams &P, const EnvPrm &q, double &wire, double &v,
                                          double &prev_accel, const double target_delta)
{
    double a_nom;
    if (P.control_mode == 0) {
        double x_err = wire - (wire + target_delta);
        a_nom = q.damping_coeff * v + q.stiffness_coeff * x_err;
    } else {
        a_nom = -q.omega_n * (v - target_delta);
    }
    double da = a_nom - prev_accel;
    if (fabs(a_nom) <= q.max_acceleration && fabs(da) <= q.max_jerk_dt) {
        prev_accel = prev_accel + da;           // no clip active
    } else {
        if (a_nom > q.max_acceleration) a_nom = q.max_acceleration;
        else if (a_nom < -q.max_acceleration) a_nom = -q.max_acceleration;
        da = a_nom - prev_accel;
        if (da > q.max_jerk_dt) da = q.max_jerk_dt;
        else if (da < -q.max_jerk_dt) da = -q.max_jerk_dt;
        prev_accel = prev_accel + da;
    }
    v += prev_accel * P.mech_dt;
    if (!(fabs(v) <= q.max_speed)) {
        if (v > q.max_speed) v = q.max_speed;
        else if (v < -q.max_speed) v = -q.max_speed;
    }
    wire += v * P.mech_dt;
}


}  // namespace sparc
