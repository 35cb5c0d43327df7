// state_layout.h — SoA layout of the per-env state inside the caller-owned blob.
//
// One X-macro drives the field table (sparc_field_info), the pointer struct the kernels
// receive, and the reset kernel.  Fields mirror the reference's EDMState
// (src/wedm/core/state.py:25-91) plus the module-internal state that the reference keeps on
// the module objects (ignition.py:75-83, material.py:46-47, dielectric.py:69-80,
// wire.py:204-225, mechanics.py:60) — SURVEY.md §8 a5/a6.
#pragma once
#include <stddef.h>
#include <stdint.h>

// X(name, ctype, dtype_char, reset_value_expression)   — `P` is the SparcParams in scope.
#define SPARC_STATE_FIELDS(X)                                                                     \
    /* positions / mechanics (fp64: per-µs increments are ~1e-5 µm on 1e2 µm) */     \
    X(wp, double, 'd', env_initial_gap)             /* EDMState.workpiece_position [µm] */          \
    X(wire, double, 'd', 0.0)                      /* EDMState.wire_position [µm] */               \
    X(v, double, 'd', 0.0)                         /* EDMState.wire_velocity [µm/s] */             \
    X(prev_accel, double, 'd', 0.0)                /* MechanicsModule.prev_accel */                \
    /* electrical */                                                                              \
    X(voltage, double, 'd', 0.0)                   /* EDMState.voltage (None -> 0) */              \
    X(current, double, 'd', 0.0)                   /* EDMState.current (None -> 0) */              \
    /* latched action (0.0 == None == "use default", ignition.py:329-343) */                      \
    X(target_voltage, double, 'd', 0.0)                                                           \
    X(on_time, double, 'd', 0.0)                                                                  \
    X(off_time, double, 'd', 0.0)                                                                 \
    X(target_delta, double, 'd', 0.0)                                                             \
    X(unwind_v, double, 'd', 0.2)                  /* EDMState.wire_unwinding_velocity */          \
    /* dielectric */                                                                              \
    X(debris_volume, double, 'd', 0.0)                                                            \
    X(debris_density, double, 'd', 0.0)                                                           \
    X(cavity_volume, double, 'd', 0.0)                                                            \
    X(flow, double, 'd', 0.0)                      /* EDMState.flow_rate == module flow_condition */\
    X(last_crater_volume, double, 'd', 0.0)                                                       \
    X(spark_y, double, 'd', -1.0)                  /* spark_status[1] [mm] */                      \
    X(d_last_gap, double, 'd', -1.0)               /* DielectricModule._last_gap_um */             \
    X(d_last_density, double, 'd', -1.0)           /* DielectricModule._last_debris_density */     \
    X(w_last_flow, double, 'd', 0.0)               /* WireModule._last_flow_condition */           \
    X(wire_avg_T, double, 'd', 0.0)                /* EDMState.wire_average_temperature */         \
    X(w_last_zone_mean, double, 'd', P.spool_T)    /* WireModule._last_zone_mean */                \
    X(ign_cached_value, double, 'd', 60.0)         /* IgnitionModule._cached_current_value */      \
    X(stat_crater_sum, double, 'd', 0.0)           /* sum of crater volumes [µm³] */               \
    X(win_vsum, double, 'd', 0.0)                  /* sum of voltage over the last step call */    \
    /* convection coefficients: the two distinct values of WireModule.h_eff_zone (float32) */     \
    X(h_base, float, 'f', 0.0f)                                                                   \
    X(h_enh, float, 'f', 0.0f)                                                                    \
    /* integer state */                                                                           \
    X(time, int32_t, 'i', 0)                                                                      \
    X(t_servo, int32_t, 'i', 0)                                                                   \
    X(t_open_v, int32_t, 'i', 0)                                                                  \
    X(t_spark_ign, int32_t, 'i', 0)                                                               \
    X(t_spark_end, int32_t, 'i', 0)                                                               \
    X(spark_state, int32_t, 'i', 0)                /* spark_status[0]: 0 idle, 1 spark, -1 short, -2 rest */ \
    X(spark_dur, int32_t, 'i', 0)                  /* spark_status[2] */                           \
    X(rand_short_rem, int32_t, 'i', 0)                                                            \
    X(deb_short_rem, int32_t, 'i', 0)                                                             \
    X(ion_rem, int32_t, 'i', 0)                                                                   \
    X(crit_count, int32_t, 'i', 0)                 /* EDMState.time_in_critical_temp */            \
    X(current_mode, int32_t, 'i', 0)               /* 0 None, k = "I<k>", -1 = "I0" */             \
    X(ign_cached_mode, int32_t, 'i', 0)                                                           \
    X(mat_cached_mode, int32_t, 'i', 0)                                                           \
    X(mat_info_mode, int32_t, 'i', 1)              /* crater_data["I1"], material.py:47 */         \
    X(w_zone_counter, int32_t, 'i', 0)                                                            \
    X(flags, int32_t, 'i', 0)                                                                     \
    X(stat_sparks, int32_t, 'i', 0)                                                               \
    X(stat_short_us, int32_t, 'i', 0)                                                             \
    X(stat_short_events, int32_t, 'i', 0)                                                         \
    X(stat_pulse_short, int32_t, 'i', 0)                                                          \
    X(stat_steps, int32_t, 'i', 0)                 /* step() calls executed this episode */        \
    X(win_sparks, int32_t, 'i', 0)                                                                \
    X(win_short_us, int32_t, 'i', 0)                                                              \
    X(win_steps, int32_t, 'i', 0)

// Per-environment PARAMETERS (a different EnvironmentConfig / *ModuleParameters for every env in
// ONE launch).  Every scalar of SparcParams that does not change the wire geometry lives in the blob as one
// array per parameter, filled from the engine's SparcParams by sparc_init and left alone by sparc_reset; the
// kernels read the arrays, never the struct, so overwriting prm_<name>[e] changes env e only.
// X(name, ctype, dtype_char)  — the blob field is called prm_<name>, the default is P.<name>.
#define SPARC_ENV_PARAM_FIELDS(X) \
    X(initial_gap, double, 'd') \
    X(target_position, double, 'd') \
    X(base_critical_density, double, 'd') \
    X(gap_coefficient, double, 'd') \
    X(max_critical_density, double, 'd') \
    X(hard_short_gap, double, 'd') \
    X(sigmoid_steepness, double, 'd') \
    X(random_short_min_gap, double, 'd') \
    X(random_short_max_gap, double, 'd') \
    X(random_short_max_probability, double, 'd') \
    X(ignition_a_coeff, double, 'd') \
    X(ignition_b_coeff, double, 'd') \
    X(ignition_c_coeff, double, 'd') \
    X(default_target_voltage, double, 'd') \
    X(default_on_time, double, 'd') \
    X(default_off_time, double, 'd') \
    X(spark_voltage_factor, double, 'd') \
    X(reference_gap, double, 'd') \
    X(debris_obstruction_coeff, double, 'd') \
    X(debris_removal_per_us, double, 'd') \
    X(base_convection_coefficient, double, 'd') \
    X(convection_velocity_factor, double, 'd') \
    X(convection_flow_enhancement, double, 'd') \
    X(plasma_efficiency, double, 'd') \
    X(damping_coeff, double, 'd') \
    X(stiffness_coeff, double, 'd') \
    X(omega_n, double, 'd') \
    X(max_acceleration, double, 'd') \
    X(max_jerk_dt, double, 'd') \
    X(max_speed, double, 'd') \
    X(debris_short_duration, int32_t, 'i') \
    X(random_short_duration, int32_t, 'i') \
    X(default_current_mode, int32_t, 'i') \
    X(ion_channel_duration, int32_t, 'i')

// Mechanics look-ahead scratch (written by mech_kernel, read by the step kernel WHILE the pre-pass is
// still running — the two kernels overlap, see mech_kernel.cu; carries nothing between calls): wire position after each µs of a launch epoch, and
// (velocity, prev_accel) checkpoints every SPARC_MECH_CKPT µs plus the end-of-epoch pair.
#define SPARC_MECH_EPOCH 1024
#define SPARC_MECH_CKPT 32
#define SPARC_MECH_NCKPT (SPARC_MECH_EPOCH / SPARC_MECH_CKPT + 1)

// `episode` is the only field that survives a reset (it is incremented instead).
struct SparcStatePtrs {
#define X(name, ctype, dt, init) ctype *name;
    SPARC_STATE_FIELDS(X)
#undef X
#define X(name, ctype, dt) ctype *prm_##name;
    SPARC_ENV_PARAM_FIELDS(X)
#undef X
    int32_t *episode;
    float *T;       // [n_envs][n_pad]
    double *mech_traj;   // [n_envs][SPARC_MECH_EPOCH]
    double *mech_trajv;  // [n_envs][SPARC_MECH_EPOCH] wire velocity after each µs — written only while signals are sampled
    double *mech_ckpt;   // [n_envs][SPARC_MECH_NCKPT][2]
    int32_t *mech_prog;  // [n_envs] launch tag + number of SPARC_MECH_CKPT-µs blocks the pre-pass has published
    int32_t n_pad;
};

static inline size_t sparc_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
