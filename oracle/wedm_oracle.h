/* wedm_oracle.h — CPU restatement of the reference hot path (TEST INFRASTRUCTURE ONLY).
 *
 * This is the parity oracle for sparc_b200: a plain-C restatement of
 * geduardo/SPARC's `WireEDMEnv.step` (reference src/wedm/envs/wire_edm.py:116-179) and the
 * five module `update()`s it calls.  It is NOT part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Parity status: PINNED against the running reference.  `oracle/validate_oracle.py`
 * (build container only) runs the unmodified reference with injected draws and compares
 * every per-µs field; the same trajectories are committed under tests/golden/ and
 * re-checked by tests/test_oracle_golden.py on every run.  The reference's own tests
 * hold no golden vectors for this path.
 */
#ifndef WEDM_ORACLE_H
#define WEDM_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OR_MAX_MODE 19
#define OR_N_SLOTS 5
#define OR_N_F64 12
#define OR_N_I32 14

/* Raw (un-derived) parameters: the reference's dataclass fields, verbatim. */
typedef struct {
    /* EnvironmentConfig (core/env_config.py:18-30) */
    double workpiece_height, wire_diameter;
    int32_t dt, servo_interval;
    double initial_gap, target_cutting_distance;
    int32_t control_mode;          /* 0 = "position", 1 = "velocity" (mechanics.py:38-41) */
    int32_t pad0;
    /* WireMaterial (data/wire_materials.json) */
    double density, specific_heat, thermal_conductivity, electrical_resistivity,
        temperature_coefficient, melting_point, breaking_temperature;
    /* IgnitionModuleParameters (modules/ignition.py:17-57) */
    double base_critical_density, gap_coefficient, max_critical_density, hard_short_gap,
        sigmoid_steepness;
    int32_t debris_short_duration, random_short_duration;
    double random_short_min_gap, random_short_max_gap, random_short_max_probability;
    double ignition_a_coeff, ignition_b_coeff, ignition_c_coeff;
    double default_target_voltage, default_on_time, default_off_time;
    int32_t default_current_mode;  /* "I5" -> 5 */
    int32_t pad1;
    double spark_voltage_factor;
    /* WireModuleParameters (modules/wire.py:16-54) */
    double buffer_len_bottom, buffer_len_top, segment_len, spool_T;
    double contact_offset_bottom, contact_offset_top;
    double base_convection_coefficient, plasma_efficiency, convection_velocity_factor,
        convection_flow_enhancement;
    int32_t compute_zone_mean, zone_mean_interval;
    double critical_temp_threshold;
    /* MaterialModuleParameters (modules/material.py:17-22) */
    double base_overcut;
    /* DielectricModuleParameters (modules/dielectric.py:15-31) */
    double base_flow_rate, debris_removal_efficiency, debris_obstruction_coeff, reference_gap,
        dielectric_temperature;
    int32_t ion_channel_duration;
    int32_t pad2;
    /* MechanicsModuleParameters (modules/mechanics.py:12-23) */
    double omega_n, zeta, max_acceleration, max_jerk, max_speed;
    /* tables, index 1..19 (modules/currents.json, modules/area_corrected.json) */
    double currents[OR_MAX_MODE + 1];
    double crater_mean[OR_MAX_MODE + 1];   /* ellipsoid_volume_half */
    double crater_std[OR_MAX_MODE + 1];    /* ellipsoid_volume_std */
    double crater_depth[OR_MAX_MODE + 1];  /* depth [um] */
    int32_t crater_avail[OR_MAX_MODE + 1];
} OrParams;

typedef struct {
    double servo, target_voltage, on_time, off_time;
    int32_t current_mode;
    int32_t pad;
} OrAction;

typedef struct OrEnv OrEnv;

/* lifecycle */
OrEnv *or_create(const OrParams *p);
void or_destroy(OrEnv *e);
void or_reset(OrEnv *e);                     /* fresh env semantics (all module state cleared) */
void or_reset_reference(OrEnv *e);           /* the reference's own reset(): new EDMState only, module state survives */
int or_n_segments(const OrEnv *e);
void or_geometry(const OrEnv *e, int32_t out[5]); /* n, zone_start, zone_end, contact_bottom, contact_top */
float *or_temperature(OrEnv *e);             /* live T field, n_segments floats */
void or_set_unwind_velocity(OrEnv *e, double v);
void or_set_initial(OrEnv *e, double wp, double wire, double v);

/* One reference step() call with the 5 draws of this µs.  Returns flags:
 * bit0 terminated, bit1 wire_broken(early return), bit2 control_step, bit3 mode error. */
int or_step(OrEnv *e, const OrAction *a, const double draws[OR_N_SLOTS]);

/* Run up to n_us steps.  The action index starts at 0 and advances after every step whose
 * control_step flag is set — the reference driver loop (README.md:101-106,
 * experiments/run_simulation.py:245-297); the last action is reused when the list runs out.
 * draws: [n_us][5] tape, or NULL to use the counter-based Philox stream (seed, env_id).
 * rec_f64/rec_i32: optional per-record output rows (OR_N_F64 / OR_N_I32 wide), written every
 * `rec_every` µs (and on the terminating µs).  Returns number of steps executed. */
int64_t or_run(OrEnv *e, int64_t n_us, const OrAction *actions, int64_t n_actions,
               const double *draws, uint64_t seed, uint32_t env_id,
               double *rec_f64, int32_t *rec_i32, int64_t rec_every, int64_t *n_rec);

/* snapshot of the same record row for the current state */
void or_record(const OrEnv *e, int terminated, double *f64, int32_t *i32);

/* episode statistics: sparks, short pulse µs, short events, crater sum(um3) */
void or_stats(const OrEnv *e, double out[8]);

/* Counter-based draws shared with the CUDA kernel (same function, same bits). */
void or_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void or_philox_draws(uint64_t seed, uint32_t env_id, uint32_t episode, uint32_t t,
                     double out[OR_N_SLOTS]);

/* Multi-threaded throughput run (cpu_baseline): n_envs fresh envs, each advanced n_us µs
 * with Philox draws and one constant action; returns total env-µs executed. */
int64_t or_bench(const OrParams *p, const OrAction *a, int n_envs, int64_t n_us, int n_threads,
                 uint64_t seed, double *stats_out8);

#ifdef __cplusplus
}
#endif
#endif
