/* wedm_oracle.c — CPU restatement of the reference hot path (TEST INFRASTRUCTURE ONLY).
 *
 * Follows, function by function, geduardo/SPARC @ /root/reference:
 *   src/wedm/envs/wire_edm.py:116-179        step / _apply_action / _check_termination
 *   src/wedm/modules/ignition.py:98-146,175-364
 *   src/wedm/modules/material.py:79-174
 *   src/wedm/modules/dielectric.py:34-41,82-163
 *   src/wedm/modules/wire.py:58-123 (Numba kernel), 129-257 (geometry), 259-394
 *   src/wedm/modules/mechanics.py:69-114
 * Arithmetic is restated in the reference's own evaluation order and precision: Python
 * floats are C doubles; the Numba kernel's fp32 arrays with fp64 scalars are restated with
 * the casts Numba's type inference produces (see thermal_update()).  Compile WITHOUT
 * -ffast-math / -ffp-contract=fast so no contraction happens here.
 *
 * Parity: pinned against the running reference by oracle/validate_oracle.py and against the
 * committed fixtures in tests/golden/ (tests/test_oracle_golden.py).
 */
#include "wedm_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <pthread.h>

/* CPython float.__floordiv__ (Objects/floatobject.c: float_divmod / _float_div_mod), the
 * algorithm behind `30.0 // 0.2 == 149.0` in wire.py:151-154,289. */
static double py_floordiv(double vx, double wx)
{
    double mod = fmod(vx, wx);
    double div = (vx - mod) / wx;
    if (mod != 0.0) {
        if ((wx < 0) != (mod < 0)) {
            mod += wx;
            div -= 1.0;
        }
    }
    double floordiv;
    if (div != 0.0) {
        floordiv = floor(div);
        if (div - floordiv > 0.5) floordiv += 1.0;
    } else {
        floordiv = copysign(0.0, vx / wx);
    }
    return floordiv;
}

struct OrEnv {
    OrParams p;
    /* ---- derived geometry / constants (wire.py:143-252) ---- */
    int n, zone_start, zone_end, actual_zone_start, actual_zone_end, zone_size;
    int contact_bottom_idx, contact_top_idx;
    double r_wire, delta_y, S, A, k_cond_coeff, denominator, joule_geom_factor;
    double temp_update_factor, critical_temperature, breaking_temperature;
    /* dielectric.py:57-66 */
    double cavity_volume_coeff, debris_removal_per_us;
    /* mechanics.py:48-57 */
    double mech_dt, damping_coeff, stiffness_coeff, max_jerk_dt;
    /* ---- EDMState (core/state.py:25-91) ---- */
    int64_t time, t_servo, t_open_v, t_spark_ign, t_spark_end;
    double voltage, current;                      /* None -> 0.0 (wire.py:272,294) */
    double target_voltage, on_time, off_time;     /* None -> 0.0 (falsy, ignition.py:329-343) */
    int current_mode;                             /* 0 = None, else the int of "I<k>" */
    double wp, wire, v, unwind_v;
    float *T, *q, *h;
    int crit_count;
    double wire_avg_T;
    int spark_state, spark_dur;
    double spark_y;
    int has_y;
    double dielectric_T, debris_volume, debris_density, cavity_volume, flow_rate, last_crater_volume;
    int ion_rem;
    int is_short, broken, target_reached;
    double target_delta, target_position;
    /* ---- hidden module state ---- */
    int rand_short_rem, deb_short_rem;            /* ignition.py:75-76 */
    int ign_cached_mode;                          /* ignition.py:82-83; 0 = None */
    double ign_cached_value;
    int mat_cached_mode, mat_info_mode;           /* material.py:46-47 */
    double prev_accel;                            /* mechanics.py:60 */
    double d_debris_volume, d_cavity, d_density, d_flow;           /* dielectric.py:69-72 */
    double d_last_gap, d_last_density, d_last_flow;                /* dielectric.py:78-80 */
    double w_last_flow, w_last_zone_mean;                          /* wire.py:223-225 */
    int w_zone_counter;
    int mode_error;
    /* ---- statistics (not in the reference; counted here for the parity of the GPU's) ---- */
    double n_sparks, n_short_us, n_short_events, crater_sum, n_pulse_short;
};

static void derive(OrEnv *e)
{
    const OrParams *p = &e->p;
    /* wire.py:144-156 */
    double total_L = p->buffer_len_bottom + p->workpiece_height + p->buffer_len_top;
    int n = (int)(total_L / p->segment_len);
    e->n = n > 1 ? n : 1;
    e->zone_start = (int)py_floordiv(p->buffer_len_bottom, p->segment_len);
    e->zone_end = e->zone_start + (int)py_floordiv(p->workpiece_height, p->segment_len);
    if (e->zone_end > e->n) e->zone_end = e->n;
    if (e->zone_start > e->zone_end) e->zone_start = e->zone_end;
    e->r_wire = p->wire_diameter / 2.0;
    /* wire.py:171-196 */
    e->delta_y = p->segment_len * 1e-3;
    e->S = M_PI * pow(e->r_wire * 1e-3, 2.0);
    e->A = 2 * M_PI * (e->r_wire * 1e-3) * e->delta_y;
    e->k_cond_coeff = p->thermal_conductivity * e->S / e->delta_y;
    e->denominator = p->density * p->specific_heat * e->S * e->delta_y;
    e->joule_geom_factor = e->S != 0 ? e->delta_y / e->S : 0.0;
    e->temp_update_factor = 1e-6 / e->denominator;
    /* wire.py:208-220 */
    e->actual_zone_start = e->zone_start < e->n - 1 ? e->zone_start : e->n - 1;
    e->actual_zone_end = e->zone_end < e->n ? e->zone_end : e->n;
    e->zone_size = e->actual_zone_end > e->actual_zone_start ? e->actual_zone_end - e->actual_zone_start : 1;
    e->critical_temperature = p->melting_point * p->critical_temp_threshold;
    e->breaking_temperature = p->breaking_temperature;
    /* wire.py:229-252 */
    double cb_mm = p->buffer_len_bottom - p->contact_offset_bottom;
    double ct_mm = p->buffer_len_bottom + p->workpiece_height + p->contact_offset_top;
    int cb = (int)(cb_mm / p->segment_len);
    if (cb < 0) cb = 0;
    int ct = (int)(ct_mm / p->segment_len);
    if (ct > e->n - 1) ct = e->n - 1;
    int t = cb < e->zone_start - 1 ? cb : e->zone_start - 1;
    cb = t > 0 ? t : 0;
    t = ct > e->zone_end ? ct : e->zone_end;
    ct = t < e->n - 1 ? t : e->n - 1;
    e->contact_bottom_idx = cb;
    e->contact_top_idx = ct;
    /* dielectric.py:57-66 */
    e->cavity_volume_coeff = M_PI * (p->wire_diameter / 2.0) * p->workpiece_height;
    e->debris_removal_per_us = p->debris_removal_efficiency * p->base_flow_rate * 1e-6;
    /* mechanics.py:48-57 */
    e->mech_dt = p->dt * 1e-6;
    e->damping_coeff = -2.0 * p->zeta * p->omega_n;
    e->stiffness_coeff = -pow(p->omega_n, 2.0);
    e->max_jerk_dt = p->max_jerk * e->mech_dt;
}

OrEnv *or_create(const OrParams *p)
{
    OrEnv *e = (OrEnv *)calloc(1, sizeof(OrEnv));
    if (!e) return NULL;
    e->p = *p;
    derive(e);
    e->T = (float *)malloc(sizeof(float) * e->n);
    e->q = (float *)malloc(sizeof(float) * e->n);
    e->h = (float *)malloc(sizeof(float) * e->n);
    or_reset(e);
    return e;
}

void or_destroy(OrEnv *e)
{
    if (!e) return;
    free(e->T);
    free(e->q);
    free(e->h);
    free(e);
}

/* Fresh-env semantics: EDMState() defaults + reset() (wire_edm.py:106-114) + the module
 * constructors' internal state.  (The reference's reset() leaves module state untouched —
 * SURVEY §0.6; parity tests always use a fresh env, which is what this reproduces.) */
void or_reset(OrEnv *e)
{
    float *T = e->T, *q = e->q, *h = e->h;
    OrParams p = e->p;
    int n = e->n;
    memset(e, 0, sizeof(*e));
    e->p = p;
    e->T = T; e->q = q; e->h = h;
    derive(e);
    for (int i = 0; i < n; i++) { T[i] = (float)p.spool_T; q[i] = 0.f; h[i] = 0.f; }
    e->unwind_v = 0.2;                      /* state.py:53 */
    e->wp = p.initial_gap;                  /* wire_edm.py:111 */
    e->target_position = p.target_cutting_distance;
    e->ign_cached_mode = 0;
    e->ign_cached_value = 60.0;             /* ignition.py:83 */
    e->mat_cached_mode = 0;
    e->mat_info_mode = 1;                   /* material.py:47: crater_data["I1"] */
    e->d_last_gap = -1.0;                   /* dielectric.py:78-80 */
    e->d_last_density = -1.0;
    e->w_last_zone_mean = p.spool_T;        /* wire.py:223 */
    e->spark_y = -1.0;
}

/* The REFERENCE's reset() (wire_edm.py:106-114): a new EDMState() with workpiece_position = initial_gap and
 * target_position = target_cutting_distance — and nothing else.  Everything the module objects keep on
 * themselves survives: ignition's short-circuit countdowns and current cache (ignition.py:75-83),
 * mechanics' prev_accel (mechanics.py:60), the dielectric's debris volume / flow condition / ion channel /
 * hysteresis memory (dielectric.py:69-80), the wire's convection array, flow memory and zone-mean cache
 * (wire.py:204-225), material's crater cache (material.py:46-47).  The new state's wire_temperature is empty
 * and is re-filled with spool_T by the next WireModule.update (wire.py:264-269). */
void or_reset_reference(OrEnv *e)
{
    const OrParams *p = &e->p;
    e->time = e->t_servo = e->t_open_v = e->t_spark_ign = e->t_spark_end = 0;
    e->voltage = e->current = 0.0;
    e->target_voltage = e->on_time = e->off_time = 0.0;
    e->current_mode = 0;
    e->wp = p->initial_gap; e->wire = 0.0; e->v = 0.0; e->unwind_v = 0.2;
    for (int i = 0; i < e->n; i++) e->T[i] = (float)p->spool_T;
    e->crit_count = 0;
    e->wire_avg_T = 0.0;
    e->spark_state = 0; e->spark_dur = 0; e->spark_y = -1.0; e->has_y = 0;
    /* debris_volume is the DielectricModule's own accumulator here (dielectric.py:69): it survives */
    e->dielectric_T = 0.0; e->debris_density = 0.0; e->cavity_volume = 0.0;
    e->flow_rate = 0.0; e->last_crater_volume = 0.0;
    e->is_short = 0; e->broken = 0; e->target_reached = 0;
    e->target_delta = 0.0; e->target_position = p->target_cutting_distance;
    e->mode_error = 0;
    e->n_sparks = e->n_short_us = e->n_short_events = e->crater_sum = e->n_pulse_short = 0.0;
}

int or_n_segments(const OrEnv *e) { return e->n; }
void or_geometry(const OrEnv *e, int32_t out[5])
{
    out[0] = e->n; out[1] = e->zone_start; out[2] = e->zone_end;
    out[3] = e->contact_bottom_idx; out[4] = e->contact_top_idx;
}
float *or_temperature(OrEnv *e) { return e->T; }
void or_set_unwind_velocity(OrEnv *e, double v) { e->unwind_v = v; }
void or_set_initial(OrEnv *e, double wp, double wire, double v) { e->wp = wp; e->wire = wire; e->v = v; }

/* ------------------------------------------------------------------------------------ */
/* ignition.py                                                                          */
/* ------------------------------------------------------------------------------------ */

/* ignition.py:98-113 (_get_current_from_mode with its cache) */
static double ign_peak_current(OrEnv *e)
{
    int mode = e->current_mode;
    if (mode != e->ign_cached_mode) {
        if (mode == 0) mode = e->p.default_current_mode;
        if (mode < 1 || mode > OR_MAX_MODE) mode = e->p.default_current_mode;
        e->ign_cached_value = e->p.currents[mode];
        e->ign_cached_mode = mode;
    }
    return e->ign_cached_value;
}
/* ignition.py:329-343: `x or default` — 0.0 and None are both falsy */
static double ign_target_voltage(const OrEnv *e) { return e->target_voltage != 0.0 ? e->target_voltage : e->p.default_target_voltage; }
static double ign_on_time(const OrEnv *e) { return e->on_time != 0.0 ? e->on_time : e->p.default_on_time; }
static double ign_off_time(const OrEnv *e) { return e->off_time != 0.0 ? e->off_time : e->p.default_off_time; }

/* ignition.py:115-146 */
static double ign_debris_short_probability(const OrEnv *e, double gap, double debris_density)
{
    const OrParams *p = &e->p;
    if (gap < p->hard_short_gap) return 1.0;
    double critical_density = p->base_critical_density + p->gap_coefficient * gap;
    critical_density = critical_density < p->max_critical_density ? critical_density : p->max_critical_density;
    double delta_density = debris_density - critical_density;
    double exponent = -p->sigmoid_steepness * delta_density;
    if (exponent > 500) return 0.0;
    else if (exponent < -500) return 1.0;
    return 1.0 / (1.0 + exp(exponent));
}

/* ignition.py:197-245 */
static void ign_update_short_circuit_detection(OrEnv *e, const double *draws)
{
    const OrParams *p = &e->p;
    double gap = e->wp - e->wire;
    gap = gap > 0.0 ? gap : 0.0;                       /* max(0.0, ...) */
    double debris_density = e->debris_density;
    if (e->rand_short_rem > 0) { e->rand_short_rem -= 1; e->is_short = 1; return; }
    if (e->deb_short_rem > 0) { e->deb_short_rem -= 1; e->is_short = 1; return; }
    double debris_short_prob = ign_debris_short_probability(e, gap, debris_density);
    double random_short_prob;
    if (gap >= p->random_short_max_gap) random_short_prob = 0.0;
    else if (gap <= p->random_short_min_gap) random_short_prob = p->random_short_max_probability;
    else {
        double gap_factor = 1.0 - (gap - p->random_short_min_gap) / (p->random_short_max_gap - p->random_short_min_gap);
        random_short_prob = gap_factor * p->random_short_max_probability;
    }
    if (draws[0] < debris_short_prob) {
        e->deb_short_rem = p->debris_short_duration;
        e->is_short = 1;
        e->n_short_events += 1;
        return;
    }
    if (draws[1] < random_short_prob) {
        e->rand_short_rem = p->random_short_duration;
        e->is_short = 1;
        e->n_short_events += 1;
        return;
    }
    e->is_short = 0;
}

/* ignition.py:348-364 (the lambda_cache memo is value-transparent and dropped) */
static double ign_lambda(const OrEnv *e)
{
    double gap = e->wp - e->wire;
    double denominator = e->p.ignition_a_coeff * pow(gap, 2.0) + e->p.ignition_b_coeff * gap + e->p.ignition_c_coeff;
    return 0.6931471805599453 /* np.log(2) */ / denominator;
}

/* ignition.py:175-195, 247-319 */
static void ign_update(OrEnv *e, const double *draws)
{
    ign_update_short_circuit_detection(e, draws);
    if (e->is_short) e->voltage = 0;
    int s = e->spark_state;
    if (s == 0) {                                              /* _handle_idle_state */
        e->current = 0;
        if (e->is_short) {
            e->spark_state = -1; e->has_y = 0; e->spark_y = -1.0; e->spark_dur = 0;
            e->current = ign_peak_current(e);
            e->n_pulse_short += 1;
        } else {
            e->voltage = ign_target_voltage(e);
            if (draws[2] < ign_lambda(e)) {                    /* _should_ignite */
                double loc = 0 + (e->p.workpiece_height - 0) * draws[3];   /* Generator.uniform */
                e->spark_state = 1; e->has_y = 1; e->spark_y = loc; e->spark_dur = 0;
                e->voltage = ign_target_voltage(e) * e->p.spark_voltage_factor;
                e->current = ign_peak_current(e);
                e->n_sparks += 1;
            }
        }
    } else if (s == 1) {                                       /* _handle_spark_state */
        int duration = e->spark_dur + 1;
        e->spark_dur = duration;
        if (duration >= ign_on_time(e)) {
            e->spark_state = -2;
            e->current = 0;
            if (!e->is_short) e->voltage = 0;
        } else {
            e->current = ign_peak_current(e);
            if (!e->is_short) e->voltage = ign_target_voltage(e) * e->p.spark_voltage_factor;
        }
    } else if (s == -1) {                                      /* _handle_short_state */
        int duration = e->spark_dur + 1;
        e->spark_dur = duration;
        if (duration >= ign_on_time(e)) { e->spark_state = -2; e->current = 0; }
        else e->current = ign_peak_current(e);
    } else if (s == -2) {                                      /* _handle_rest_state */
        int duration = e->spark_dur + 1;
        e->spark_dur = duration;
        double total_cycle_time = ign_on_time(e) + ign_off_time(e);
        if (duration >= total_cycle_time) {
            e->spark_state = 0; e->has_y = 0; e->spark_y = -1.0; e->spark_dur = 0;
            e->current = 0;
            if (!e->is_short) e->voltage = ign_target_voltage(e);
        } else {
            e->current = 0;
            if (!e->is_short) e->voltage = 0;
        }
    }
    if (e->is_short) e->n_short_us += 1;
}

/* ------------------------------------------------------------------------------------ */
/* material.py:79-174                                                                   */
/* ------------------------------------------------------------------------------------ */
static int mat_update(OrEnv *e, const double *draws)
{
    const OrParams *p = &e->p;
    if (e->spark_state == 1 && e->spark_dur == 0) {
        int mode = e->current_mode;
        if (mode != e->mat_cached_mode) {                       /* material.py:103-117 */
            if (mode == 0) mode = 1;
            if (mode < 1 || mode > OR_MAX_MODE || !p->crater_avail[mode]) return -1;   /* ValueError */
            e->mat_info_mode = mode;
            e->mat_cached_mode = mode;
        }
        double mean_volume = p->crater_mean[e->mat_info_mode];
        double std_volume = p->crater_std[e->mat_info_mode];
        double sampled_um3 = mean_volume + std_volume * draws[4];   /* Generator.normal */
        sampled_um3 = sampled_um3 > 0 ? sampled_um3 : 0;            /* max(0, x) */
        e->crater_sum += sampled_um3;
        double crater_volume = sampled_um3 / 1e9;
        e->last_crater_volume = crater_volume;
        if (crater_volume > 0) {                                    /* material.py:140-174 */
            double crater_depth_mm = p->crater_depth[e->mat_info_mode] / 1000.0;
            double kerf_width_mm = p->base_overcut + p->wire_diameter + crater_depth_mm;
            double delta_x_um;
            if (kerf_width_mm > 0 && p->workpiece_height > 0) {
                double delta_x_mm = crater_volume / (kerf_width_mm * p->workpiece_height);
                delta_x_um = delta_x_mm * 1000.0;
            } else delta_x_um = 0.0;
            e->wp += delta_x_um;
        }
    } else {
        e->last_crater_volume = 0.0;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------ */
/* dielectric.py:34-41, 82-163                                                          */
/* ------------------------------------------------------------------------------------ */
static double fast_exp_neg(double x)      /* dielectric.py:34-41: approximates exp(-x) */
{
    if (x < 0.5) return (1 - 0.5 * x) / (1 + 0.5 * x);
    return exp(-x);
}

static void diel_update(OrEnv *e)
{
    const OrParams *p = &e->p;
    e->dielectric_T = p->dielectric_temperature;
    double gap_um = e->wp - e->wire;
    gap_um = gap_um > 0.001 ? gap_um : 0.001;                   /* max(0.001, ...) */
    double gap_mm = gap_um * 0.001;
    e->cavity_volume = e->cavity_volume_coeff * gap_mm;
    if (e->spark_state == 1 && e->spark_dur == 0) {
        double crater_volume = e->last_crater_volume;
        if (crater_volume > 0) {
            e->debris_volume += crater_volume;
            e->ion_rem = p->ion_channel_duration;               /* (y, duration) tuple */
        }
    }
    if (e->cavity_volume > 0) {
        double d = e->debris_volume / e->cavity_volume;
        e->debris_density = d < 1.0 ? d : 1.0;                  /* min(1.0, d) */
    } else e->debris_density = 0.0;
    if (fabs(gap_um - e->d_last_gap) > 0.01 || fabs(e->debris_density - e->d_last_density) > 0.001) {
        double gf = pow(gap_um / p->reference_gap, 3.0);
        double gap_factor = gf < 1.0 ? gf : 1.0;                /* min(1.0, gf) */
        double debris_factor;
        double x = p->debris_obstruction_coeff * e->debris_density;
        if (x < 2.0) debris_factor = fast_exp_neg(x);
        else debris_factor = exp(-x);
        e->flow_rate = gap_factor * debris_factor;
        e->d_last_gap = gap_um;
        e->d_last_density = e->debris_density;
        e->d_last_flow = e->flow_rate;
    } else {
        e->flow_rate = e->d_last_flow;
    }
    if (e->flow_rate > 0.001 && e->debris_volume > 0.001) {
        double debris_removed = e->debris_removal_per_us * e->flow_rate;
        double nv = e->debris_volume - debris_removed;
        e->debris_volume = nv > 0.0 ? nv : 0.0;                 /* max(0.0, nv) */
    }
    if (e->ion_rem > 0) {                                       /* dielectric.py:147-150 */
        e->ion_rem = e->ion_rem > 1 ? e->ion_rem - 1 : 0;
    }
}

/* ------------------------------------------------------------------------------------ */
/* wire.py                                                                              */
/* ------------------------------------------------------------------------------------ */

/* wire.py:58-123 compute_thermal_update.  Types as Numba infers them: T, dT_dt (q), h are
 * float32 arrays, every scalar argument is float64 (ints for the indices).
 *   `2 * T[i]`            int64*float32 -> float64, so the whole Laplacian is float64
 *   `T[n-2] - T[n-1]`     float32 - float32 -> float32 (then * float64)
 *   `T[i-1] - T[i]` (adv) float32
 *   every `dT_dt[i] op= x` computes in float64 and rounds to float32 on the store
 * Numba compiles this with fastmath=True, so LLVM may contract/reassociate the float64
 * intermediates; the restatement uses plain IEEE evaluation (differences are far below the
 * float32 rounding of the stores; measured in oracle/validate_oracle.py). */
static void thermal_update(float *T, float *q, int n, double spool_T, double k_cond_coeff,
                           double I_squared, double joule_geom_factor, double rho_elec,
                           double alpha_rho, double temp_ref, int plasma_idx, double plasma_heat,
                           const float *h_eff_zone, double dielectric_temp, double A,
                           double adv_coeff, double temp_update_factor, int cb, int ct)
{
    T[0] = (float)spool_T;
    for (int i = 0; i < n; i++) q[i] = 0.0f;
    if (n > 1) {
        for (int i = 1; i < n - 1; i++)
            q[i] = (float)(k_cond_coeff * (((double)T[i - 1] - 2.0 * (double)T[i]) + (double)T[i + 1]));
        q[n - 1] = (float)(k_cond_coeff * (double)(float)(T[n - 2] - T[n - 1]));
    }
    if (I_squared > 1e-6) {
        double joule_factor = joule_geom_factor * I_squared * rho_elec;
        for (int i = cb; i < ct + 1; i++) {
            double rho_T = 1.0 + alpha_rho * ((double)T[i] - temp_ref);
            q[i] = (float)((double)q[i] + joule_factor * rho_T);
        }
    }
    if (plasma_idx >= 0 && plasma_idx < n) q[plasma_idx] = (float)((double)q[plasma_idx] + plasma_heat);
    for (int i = 0; i < n; i++) {
        double conv_coeff = (double)h_eff_zone[i] * A;
        q[i] = (float)((double)q[i] - conv_coeff * ((double)T[i] - dielectric_temp));
    }
    if (fabs(adv_coeff) > 1e-9) {
        for (int i = 1; i < n; i++) q[i] = (float)((double)q[i] + adv_coeff * (double)(float)(T[i - 1] - T[i]));
    }
    for (int i = 0; i < n; i++) T[i] = (float)((double)T[i] + (double)q[i] * temp_update_factor);
    T[0] = (float)spool_T;
}

/* wire.py:349-374 */
static void wire_update_convection(OrEnv *e, double wire_unwind_vel, double flow_condition)
{
    const OrParams *p = &e->p;
    double ve = p->convection_velocity_factor * wire_unwind_vel;
    double velocity_enhancement = ve > -0.9 ? ve : -0.9;                       /* max(-0.9, ve) */
    double h_eff_base = p->base_convection_coefficient * (1.0 + velocity_enhancement);
    double floor_h = 0.1 * p->base_convection_coefficient;
    h_eff_base = h_eff_base > floor_h ? h_eff_base : floor_h;                  /* max(h, floor) */
    double h_eff_enhanced = h_eff_base * (1.0 + p->convection_flow_enhancement * flow_condition);
    for (int i = 0; i < e->n; i++) e->h[i] = (float)h_eff_base;
    if (e->actual_zone_start < e->actual_zone_end)
        for (int i = e->actual_zone_start; i < e->actual_zone_end; i++) e->h[i] = (float)h_eff_enhanced;
}

/* wire.py:259-347, 376-394 */
static void wire_update(OrEnv *e)
{
    const OrParams *p = &e->p;
    if (e->broken) return;
    double I = e->current;                  /* `state.current or 0.0` */
    double I_squared = I * I;
    double dielectric_temp = e->dielectric_T;
    double wire_unwind_vel = e->unwind_v;
    double flow_condition = e->flow_rate;
    if (fabs(flow_condition - e->w_last_flow) > 0.01) {
        wire_update_convection(e, wire_unwind_vel, flow_condition);
        e->w_last_flow = flow_condition;
    }
    int plasma_idx = -1;
    double plasma_heat = 0.0;
    if (e->spark_state == 1 && e->has_y) {
        plasma_idx = p->segment_len != 0 ? e->zone_start + (int)py_floordiv(e->spark_y, p->segment_len)
                                         : e->zone_start;
        if (0 <= plasma_idx && plasma_idx < e->n) {
            plasma_heat = p->plasma_efficiency * e->voltage * I;
            if (!isfinite(plasma_heat)) plasma_heat = 0.0;
        }
    }
    double adv_coeff;
    if (fabs(wire_unwind_vel) > 1e-6) {
        double v_wire = fabs(wire_unwind_vel);
        adv_coeff = p->density * p->specific_heat * v_wire * e->S;
    } else adv_coeff = 0.0;

    thermal_update(e->T, e->q, e->n, p->spool_T, e->k_cond_coeff, I_squared, e->joule_geom_factor,
                   p->electrical_resistivity, p->temperature_coefficient, 293.15 /* wire.py:193 */,
                   plasma_idx, plasma_heat, e->h, dielectric_temp, e->A, adv_coeff,
                   e->temp_update_factor, e->contact_bottom_idx, e->contact_top_idx);

    /* _check_wire_breaking (wire.py:376-388).  np.max(T) is a numpy float32 scalar; under
     * NumPy >= 2 (NEP 50) comparing it with a Python float rounds the Python float to
     * float32 first. */
    float max_temp = e->T[0];
    for (int i = 1; i < e->n; i++) if (e->T[i] > max_temp) max_temp = e->T[i];
    if (max_temp > (float)e->critical_temperature) e->crit_count += 1;
    else e->crit_count = 0;
    if (max_temp > (float)e->breaking_temperature) e->broken = 1;

    if (p->compute_zone_mean) {                                  /* wire.py:338-347 */
        e->w_zone_counter += 1;
        if (e->w_zone_counter >= p->zone_mean_interval) {
            int a = e->actual_zone_start, b = e->actual_zone_end;
            if (!(e->zone_size > 0 && b <= e->n)) { a = 0; b = e->n; }
            double s = 0.0;                                      /* np.mean(float32[...]) */
            for (int i = a; i < b; i++) s += (double)e->T[i];
            e->w_last_zone_mean = (double)(float)(s / (double)(b - a));
            e->w_zone_counter = 0;
        }
        e->wire_avg_T = e->w_last_zone_mean;
    }
}

/* ------------------------------------------------------------------------------------ */
/* mechanics.py:69-114                                                                  */
/* ------------------------------------------------------------------------------------ */
static void mech_update(OrEnv *e)
{
    const OrParams *p = &e->p;
    double x = e->wire, v = e->v, a_nom;
    if (p->control_mode == 0) {
        double x_error = x - (x + e->target_delta);
        a_nom = e->damping_coeff * v + e->stiffness_coeff * x_error;
    } else {
        double v_error = v - e->target_delta;
        a_nom = -p->omega_n * v_error;
    }
    if (a_nom > p->max_acceleration) a_nom = p->max_acceleration;
    else if (a_nom < -p->max_acceleration) a_nom = -p->max_acceleration;
    double da = a_nom - e->prev_accel;
    if (da > e->max_jerk_dt) da = e->max_jerk_dt;
    else if (da < -e->max_jerk_dt) da = -e->max_jerk_dt;
    double a = e->prev_accel + da;
    e->prev_accel = a;
    v += a * e->mech_dt;
    if (v > p->max_speed) v = p->max_speed;
    else if (v < -p->max_speed) v = -p->max_speed;
    x += v * e->mech_dt;
    e->v = v;
    e->wire = x;
}

/* ------------------------------------------------------------------------------------ */
/* wire_edm.py:116-179                                                                  */
/* ------------------------------------------------------------------------------------ */
int or_step(OrEnv *e, const OrAction *a, const double draws[OR_N_SLOTS])
{
    const OrParams *p = &e->p;
    int flags = 0;
    int is_ctrl = e->t_servo >= p->servo_interval;
    if (is_ctrl) {                                   /* _apply_action, wire_edm.py:162-170 */
        e->target_delta = a->servo;
        e->target_voltage = a->target_voltage;
        e->current_mode = a->current_mode == 0 ? -1 : a->current_mode;  /* "I0" is a string, not None */
        e->on_time = a->on_time;
        e->off_time = a->off_time;
        e->t_servo = 0;
        flags |= 4;
    }
    ign_update(e, draws);
    if (mat_update(e, draws) < 0) { e->mode_error = 1; return flags | 8 | 1; }
    diel_update(e);
    wire_update(e);
    if (e->broken) return flags | 2 | 1;             /* wire_edm.py:129-130 */
    mech_update(e);
    e->time += p->dt;
    e->t_servo += p->dt;
    e->t_open_v += p->dt;
    if (e->spark_state == 1) { e->t_spark_ign += p->dt; e->t_spark_end = 0; }
    else { e->t_spark_end += p->dt; e->t_spark_ign = 0; }
    /* _check_termination, wire_edm.py:172-179 */
    if (e->wire > e->wp + 100) { e->broken = 1; flags |= 1; }
    else if (e->wp >= e->target_position) { e->target_reached = 1; flags |= 1; }
    return flags;
}

void or_record(const OrEnv *e, int terminated, double *f, int32_t *i)
{
    float mx = e->T[0];
    for (int k = 1; k < e->n; k++) if (e->T[k] > mx) mx = e->T[k];
    if (f) {
        f[0] = e->wp; f[1] = e->wire; f[2] = e->v; f[3] = e->prev_accel; f[4] = e->voltage;
        f[5] = e->current; f[6] = e->debris_volume; f[7] = e->debris_density; f[8] = e->flow_rate;
        f[9] = e->last_crater_volume; f[10] = e->has_y ? e->spark_y : -1.0; f[11] = (double)mx;
    }
    if (i) {
        i[0] = (int32_t)e->time; i[1] = (int32_t)e->t_servo; i[2] = (int32_t)e->t_spark_ign;
        i[3] = (int32_t)e->t_spark_end; i[4] = e->spark_state; i[5] = e->spark_dur; i[6] = e->is_short;
        i[7] = e->rand_short_rem; i[8] = e->deb_short_rem; i[9] = e->ion_rem; i[10] = e->crit_count;
        i[11] = e->broken; i[12] = e->target_reached; i[13] = terminated;
    }
}

void or_stats(const OrEnv *e, double out[8])
{
    out[0] = e->n_sparks; out[1] = e->n_short_us; out[2] = e->n_short_events; out[3] = e->crater_sum;
    out[4] = e->n_pulse_short; out[5] = (double)e->time; out[6] = e->broken; out[7] = e->target_reached;
}

/* ------------------------------------------------------------------------------------ */
/* Counter-based draws (Philox4x32-10, Salmon et al. SC'11) — the stream the CUDA kernel  */
/* uses for its on-device RNG, restated here so device trajectories can be checked.       */
/*   slots 0..3 (u1,u2,u3,u4): ctr = {env_id, t>>2, episode, slot}, word [t&3],            */
/*                             u = (word + 0.5) * 2^-32                                    */
/*   slot 4 (z5):              ctr = {env_id, t, episode, 4}, Box-Muller on words 0,1      */
/* ------------------------------------------------------------------------------------ */
void or_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void or_philox_draws(uint64_t seed, uint32_t env_id, uint32_t episode, uint32_t t, double out[OR_N_SLOTS])
{
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t w[4];
    for (uint32_t s = 0; s < 4; s++) {
        uint32_t ctr[4] = {env_id, t >> 2, episode, s};
        or_philox4x32_10(ctr, key, w);
        out[s] = ((double)w[t & 3] + 0.5) * 2.3283064365386963e-10;
    }
    uint32_t ctr[4] = {env_id, t, episode, 4};
    or_philox4x32_10(ctr, key, w);
    double ua = ((double)w[0] + 0.5) * 2.3283064365386963e-10;
    double ub = ((double)w[1] + 0.5) * 2.3283064365386963e-10;
    out[4] = sqrt(-2.0 * log(ua)) * cos(6.283185307179586 * ub);
}

int64_t or_run(OrEnv *e, int64_t n_us, const OrAction *actions, int64_t n_actions,
               const double *draws, uint64_t seed, uint32_t env_id,
               double *rec_f64, int32_t *rec_i32, int64_t rec_every, int64_t *n_rec)
{
    int64_t idx = 0, nr = 0, k;
    double d[OR_N_SLOTS];
    for (k = 0; k < n_us; k++) {
        const double *dk;
        if (draws) dk = draws + k * OR_N_SLOTS;
        else { or_philox_draws(seed, env_id, 0, (uint32_t)e->time, d); dk = d; }
        const OrAction *a = &actions[idx < n_actions ? idx : n_actions - 1];
        int flags = or_step(e, a, dk);
        int term = flags & 1;
        if (rec_every > 0 && (((k + 1) % rec_every) == 0 || term)) {
            or_record(e, term, rec_f64 ? rec_f64 + nr * OR_N_F64 : NULL, rec_i32 ? rec_i32 + nr * OR_N_I32 : NULL);
            nr++;
        }
        if (term) { k++; break; }
        if (flags & 4) idx++;
    }
    if (n_rec) *n_rec = nr;
    return k;
}

typedef struct {
    const OrParams *p; const OrAction *a; int n_envs, n_threads, tid; int64_t n_us; uint64_t seed;
    int64_t total; double acc[8];
} BenchJob;

static void *bench_worker(void *arg)
{
    BenchJob *j = (BenchJob *)arg;
    for (int i = j->tid; i < j->n_envs; i += j->n_threads) {
        OrEnv *e = or_create(j->p);
        j->total += or_run(e, j->n_us, j->a, 1, NULL, j->seed, (uint32_t)i, NULL, NULL, 0, NULL);
        double s[8];
        or_stats(e, s);
        for (int k = 0; k < 8; k++) j->acc[k] += s[k];
        or_destroy(e);
    }
    return NULL;
}

int64_t or_bench(const OrParams *p, const OrAction *a, int n_envs, int64_t n_us, int n_threads,
                 uint64_t seed, double *stats_out8)
{
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 1024) n_threads = 1024;
    BenchJob *jobs = (BenchJob *)calloc((size_t)n_threads, sizeof(BenchJob));
    pthread_t *th = (pthread_t *)calloc((size_t)n_threads, sizeof(pthread_t));
    for (int t = 0; t < n_threads; t++) {
        jobs[t].p = p; jobs[t].a = a; jobs[t].n_envs = n_envs; jobs[t].n_threads = n_threads;
        jobs[t].tid = t; jobs[t].n_us = n_us; jobs[t].seed = seed;
        pthread_create(&th[t], NULL, bench_worker, &jobs[t]);
    }
    int64_t total = 0;
    double acc[8] = {0};
    for (int t = 0; t < n_threads; t++) {
        pthread_join(th[t], NULL);
        total += jobs[t].total;
        for (int k = 0; k < 8; k++) acc[k] += jobs[t].acc[k];
    }
    free(jobs); free(th);
    if (stats_out8) for (int k = 0; k < 8; k++) stats_out8[k] = acc[k];
    return total;
}
