/* sparc_b200.h — C-ABI of the B200-native batched WireEDMEnv step.
 *
 * Drop-in boundary for ONE path of geduardo/SPARC: the per-µs physics loop of
 * `WireEDMEnv.step` (reference src/wedm/envs/wire_edm.py:116-157), i.e. the five
 * `EDMModule.update(state)` plug-ins (src/wedm/core/module.py:25-29) it calls in fixed order:
 *   IgnitionModule.update        src/wedm/modules/ignition.py:175-195
 *   MaterialRemovalModule.update src/wedm/modules/material.py:79-96
 *   DielectricModule.update      src/wedm/modules/dielectric.py:82-163
 *   WireModule.update            src/wedm/modules/wire.py:259-347 (+ Numba kernel :58-123)
 *   MechanicsModule.update       src/wedm/modules/mechanics.py:79-114
 * plus `_apply_action` (:162-170), the time bookkeeping (:134-144) and `_check_termination`
 * (:172-179).  The reference has no FFI; the binding a maintainer adds is the ctypes stub in
 * INTEGRATION.md.  All five modules are fused into one sm_100a kernel that advances every
 * environment through all µs of a control step.
 *
 * Conventions
 *  - plain C: pointers, sizes, POD structs.  No torch / CUDA types in signatures; a CUDA
 *    stream is passed as `void*` (cudaStream_t), NULL = default stream.
 *  - every function returns 0 on success, a negative SPARC_E* code otherwise;
 *    `sparc_last_error()` gives a thread-local message.  Nothing throws across the boundary.
 *  - the library owns no device memory: the caller allocates the state blob
 *    (`sparc_state_bytes`) and passes it to every call.  The library never falls back to the
 *    CPU; without a CUDA device every compute entry point returns SPARC_E_CUDA.
 */
#ifndef SPARC_B200_H
#define SPARC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPARC_ABI_VERSION 5
#define SPARC_MAX_MODE 19          /* current modes I1..I19, modules/currents.json */
#define SPARC_N_DRAW_SLOTS 5       /* u1,u2,u3,u4,z5 per µs — see sparc_step_injected */
#define SPARC_N_STATS 8
#define SPARC_N_OBS 12

/* rows of the per-signal log (sparc_step_sampled); names are the reference's EDMState attributes */
enum {
    SPARC_SIG_TIME = 0,              /* time */
    SPARC_SIG_VOLTAGE = 1,           /* voltage */
    SPARC_SIG_CURRENT = 2,           /* current */
    SPARC_SIG_WIRE_POSITION = 3,     /* wire_position */
    SPARC_SIG_WIRE_VELOCITY = 4,     /* wire_velocity */
    SPARC_SIG_WORKPIECE_POSITION = 5,/* workpiece_position */
    SPARC_SIG_TARGET_DELTA = 6,      /* target_delta */
    SPARC_SIG_DEBRIS_DENSITY = 7,    /* debris_density (= legacy debris_concentration, dielectric.py:159-163) */
    SPARC_SIG_FLOW_RATE = 8,         /* flow_rate (= legacy dielectric_flow_rate) */
    SPARC_SIG_IS_SHORT_CIRCUIT = 9,  /* is_short_circuit */
    SPARC_SIG_SPARK_STATE = 10,      /* spark_status[0] */
    SPARC_SIG_MAX_TEMPERATURE = 11,  /* max(wire_temperature) */
    SPARC_SIG_AVERAGE_TEMPERATURE = 12, /* wire_average_temperature */
    SPARC_SIG_DEBRIS_VOLUME = 13,    /* debris_volume */
    SPARC_SIG_TIME_IN_CRITICAL = 14, /* time_in_critical_temp */
    SPARC_SIG_FLAGS = 15,            /* SPARC_F_* bits */
    SPARC_N_SIGNALS = 16
};

enum {
    SPARC_OK = 0,
    SPARC_E_INVALID = -1,   /* bad argument (NULL pointer, negative size, unsupported geometry) */
    SPARC_E_CUDA = -2,      /* CUDA runtime error (message in sparc_last_error) */
    SPARC_E_UNSUPPORTED = -3
};

/* Per-env flag bits (state field "flags"). */
enum {
    SPARC_F_SHORT = 1,          /* EDMState.is_short_circuit */
    SPARC_F_BROKEN = 2,         /* EDMState.is_wire_broken (thermal break or collision, wire_edm.py:173-175) */
    SPARC_F_TARGET = 4,         /* EDMState.is_target_distance_reached */
    SPARC_F_MODE_ERROR = 8,     /* crater table miss: the reference raises ValueError, material.py:108-113 */
    SPARC_F_HAS_Y = 16,         /* spark_status[1] is not None */
    SPARC_F_THERMAL_BREAK = 32, /* break came from WireModule._check_wire_breaking (early return, wire_edm.py:129-130) */
    SPARC_F_HANDOFF_SLOW = 64   /* diagnostic, not an error: the mechanics pre-pass of some launch did not publish in time
                                   and the step kernel computed this env's trajectory itself (same values) */
};
#define SPARC_F_DONE (SPARC_F_BROKEN | SPARC_F_TARGET | SPARC_F_MODE_ERROR)

/* Derived constants of one engine.  Computed on the HOST in Python doubles with the
 * reference's own expressions (geometry: wire.py:143-252; dielectric.py:57-66;
 * mechanics.py:48-57; material.py:152-169) so that index quirks such as
 * `30.0 // 0.2 == 149` are reproduced exactly; the kernel receives numbers, not formulas. */
typedef struct SparcParams {
    /* geometry (wire.py:149-156, 208-252) */
    int32_t n_segments, zone_start, zone_end, actual_zone_start, actual_zone_end;
    int32_t contact_bottom_idx, contact_top_idx;
    /* env (env_config.py:18-30) */
    int32_t servo_interval;
    int32_t control_mode;             /* 0 position, 1 velocity */
    /* ignition (ignition.py:17-57) */
    int32_t debris_short_duration, random_short_duration, default_current_mode;
    /* dielectric / wire ints */
    int32_t ion_channel_duration, compute_zone_mean, zone_mean_interval;
    int32_t reserved0;
    double initial_gap, target_position, workpiece_height, segment_len;
    double base_critical_density, gap_coefficient, max_critical_density, hard_short_gap,
        sigmoid_steepness;
    double random_short_min_gap, random_short_max_gap, random_short_max_probability;
    double ignition_a_coeff, ignition_b_coeff, ignition_c_coeff;
    double default_target_voltage, default_on_time, default_off_time, spark_voltage_factor;
    /* material: per-mode tables (currents.json, area_corrected.json), index 1..19 */
    double currents[SPARC_MAX_MODE + 1];
    double crater_mean[SPARC_MAX_MODE + 1], crater_std[SPARC_MAX_MODE + 1];
    double kerf_times_height[SPARC_MAX_MODE + 1];   /* (base_overcut + d_wire + depth/1000) * h */
    int32_t crater_avail[SPARC_MAX_MODE + 1];
    /* dielectric (dielectric.py:57-66) */
    double cavity_volume_coeff, reference_gap, debris_obstruction_coeff, debris_removal_per_us,
        dielectric_temperature;
    /* wire (wire.py:171-220) */
    double spool_T, k_cond_coeff, joule_geom_factor, rho_elec, alpha_rho, temp_ref, A, S,
        temp_update_factor, density_times_cp, plasma_efficiency, base_convection_coefficient,
        convection_velocity_factor, convection_flow_enhancement;
    float critical_temperature_f32, breaking_temperature_f32;
    /* (float)(temp_update_factor * k_cond_coeff): the conduction coefficient of the stencil, a launch constant
     * the kernel reads straight from the parameter bank (uniform-register operand) */
    float cond_coeff_f32, reserved1;
    /* mechanics (mechanics.py:48-57) */
    double mech_dt, damping_coeff, stiffness_coeff, omega_n, max_acceleration, max_jerk_dt,
        max_speed;
} SparcParams;

/* Batched action, SoA device (or host, for *_host entry points) arrays of length n_envs:
 * the reference's action dict (wire_edm.py:84-98) one array per leaf. */
typedef struct SparcActions {
    const float *servo;            /* action["servo"][0] */
    const float *target_voltage;   /* generator_control.target_voltage[0] */
    const int32_t *current_mode;   /* generator_control.current_mode[0], 1..19 */
    const float *on_time;          /* generator_control.ON_time[0] */
    const float *off_time;         /* generator_control.OFF_time[0] */
} SparcActions;

/* Optional controller evaluated INSIDE the kernels, at the start of a step call, instead of reading
 * `actions->servo` (the generator settings still come from `actions`).  kind 1 is the proportional gap
 * controller of the reference's driver (experiments/run_simulation.py:24-39, README.md:82-99):
 *     servo = (float) clip(gain * ((workpiece_position - wire_position) - desired_gap), -limit, +limit)
 * evaluated in float64 from the state the call starts with, rounded to the float32 the action space
 * carries — the arithmetic of sparc_b200/controllers.py:GapController, so closed-loop trajectories are
 * identical to evaluating it with torch ops between calls, with zero extra launches.  limit < 0: no clip.
 * Only for calls of at most one launch epoch (1024 µs). */
typedef struct SparcController {
    int32_t kind;                  /* 0 none (use actions->servo), 1 proportional gap controller */
    int32_t reserved;
    double desired_gap, gain, limit;
} SparcController;

/* ---- state blob ------------------------------------------------------------------- */
/* The state of n_envs environments is ONE caller-owned device allocation, laid out SoA:
 * one contiguous array per field, temperature field as [n_envs][n_pad] float32.
 * Field table: name, element type ('d' f64, 'f' f32, 'i' i32), byte offset, element count. */
int sparc_abi_version(void);
size_t sparc_params_bytes(void);             /* sizeof(SparcParams) of the built library (binding self-check) */
const char *sparc_last_error(void);

int sparc_num_fields(void);
/* name/dtype of field `idx`; offset/count for a given (n_envs, n_segments). */
int sparc_field_info(int idx, int n_envs, int n_segments, const char **name, char *dtype,
                     size_t *byte_offset, size_t *count);
size_t sparc_state_bytes(int n_envs, int n_segments);
int sparc_padded_segments(int n_segments);   /* row pitch (floats) of the temperature field */

/* ---- compute entry points (all asynchronous on `stream`) --------------------------- */

/* First-time initialisation of a freshly allocated blob: every env gets the state of a newly
 * constructed WireEDMEnv after reset() and episode counter 0. */
int sparc_init(void *state, const SparcParams *p, int n_envs, void *stream);

/* EDMState() + WireEDMEnv.reset() (wire_edm.py:106-114) + fresh module state for the envs whose
 * mask byte is non-zero (mask == NULL: all).  `episode` increments for reset envs. */
int sparc_reset(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream);

/* The reference's OWN reset semantics (strict_reference_reset): `WireEDMEnv.reset()` re-creates only
 * EDMState (wire_edm.py:106-114); everything its module objects keep on themselves survives into the next episode —
 * short-circuit countdowns and the peak-current cache (ignition.py:75-83), prev_accel (mechanics.py:60), debris volume,
 * flow condition, ion channel and hysteresis memory (dielectric.py:69-80), convection coefficients and the zone-mean
 * cache (wire.py:204-225).  sparc_reset clears all of it (a freshly constructed env); this entry point clears only what
 * the reference clears.  Pinned against the running reference: tests/golden/reference_reset.npz. */
int sparc_reset_reference(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream);

/* sparc_reset with the episode counter of the reset envs set back to 0: the state a freshly
 * initialised blob has.  This is `reset(seed=s)` (wire_edm.py:106-107 re-seeds np_random through
 * gym.Env.reset): the same seed must replay the same episode, so the Philox counter word that
 * separates successive episodes of one env starts over. */
int sparc_reseed(void *state, const SparcParams *p, int n_envs, const uint8_t *mask_dev, void *stream);

/* n_us reference step() calls for every env, on-device Philox4x32-10 draws keyed by
 * (seed, env_id_base + env, episode, µs).  Envs already terminated are left untouched.
 * `actions` is latched inside the loop exactly when the reference latches it
 * (time_since_servo >= servo_interval, wire_edm.py:117-121). */
int sparc_step(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
               uint64_t seed, uint32_t env_id_base, void *stream);

/* Same, with the random draws injected: draws_dev[env][t][5] doubles, t = 0..n_us-1 for THIS
 * call; slots: u1 debris-short roll (ignition.py:233), u2 random-short roll (:239), u3 ignition
 * roll (:327), u4 spark location uniform (:261), z5 crater standard normal (material.py:127). */
int sparc_step_injected(void *state, const SparcParams *p, const SparcActions *actions, int n_envs,
                        int n_us, const double *draws_dev, void *stream);

/* Same as sparc_step (draws_dev == NULL) or sparc_step_injected (draws_dev != NULL), and the wire
 * temperature field of envs 0..n_log_envs-1 is sampled every `log_every` µs of this call into
 * tlog_dev[n_log_envs][n_us / log_every][n_segments] float32 — the reference's
 * SimulationLogger("wire_temperature", interval) array layout (n_timesteps, n_segments)
 * (utils/logger.py:128-150, experiments/animate_wire_heating.py:44-47).  Row r is the field after µs
 * (r+1)*log_every - 1; rows at or after an env's termination hold its final field; envs already
 * terminated at entry leave their rows untouched. */
int sparc_step_logged(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                      uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *tlog_dev, int log_every,
                      int n_log_envs, void *stream);

/* Step with per-signal logging: the device-side SimulationLogger (utils/logger.py:105-160,
 * "every_step" with sample_every = 1, "interval" with sample_every = value).  After every
 * `sample_every` µs of this call one row of the scalar signals above is written:
 * signals_dev[n_us / sample_every][SPARC_N_SIGNALS][n_envs] doubles, row r = the state after µs
 * (r+1)*sample_every - 1, i.e. what the reference's logger reads from EDMState after that step().
 * The call is cut into launches at the sampling µs (results are independent of the cut).  draws_dev
 * and tlog_dev / log_every / n_log_envs as in sparc_step_logged, both optional (NULL). */
int sparc_step_sampled(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                       uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *tlog_dev, int log_every,
                       int n_log_envs, double *signals_dev, int sample_every, void *stream);

/* sparc_step / sparc_step_injected (draws_dev NULL or not) with the observation summary below written
 * by the step kernel's epilogue into obs_dev[SPARC_N_OBS][n_envs] — the same values sparc_observe
 * would read back from the state, without the second launch. */
int sparc_step_observe(void *state, const SparcParams *p, const SparcActions *actions, int n_envs, int n_us,
                       uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *obs_dev, void *stream);

/* The general form of the step entry points: draws_dev (injected draws), obs_dev (observation written by
 * the kernel epilogue) and ctrl (in-kernel controller) are each optional (NULL). */
int sparc_step_ex(void *state, const SparcParams *p, const SparcActions *actions, const SparcController *ctrl,
                  int n_envs, int n_us, uint64_t seed, uint32_t env_id_base, const double *draws_dev, float *obs_dev,
                  void *stream);

/* Observation summary of the last step call, obs_dev[SPARC_N_OBS][n_envs] float32:
 * 0 gap [µm], 1 wire_velocity [µm/s], 2 mean voltage over the call [V], 3 sparks ignited in the
 * call, 4 short-circuit µs in the call, 5 max wire temperature [K], 6 debris_density, 7 flags;
 * rows 8..10 carry int32 BIT PATTERNS (read them through an int32 view): 8 time [µs], 9 spark_status[0],
 * 10 time_since_servo — with flags, everything the reference's `info` dict and the control-step test
 * of `WireEDMEnv.step` need (wire_edm.py:117,150-156) without a second read-back; 11 reserved (0). */
int sparc_observe(const void *state, int n_envs, int n_segments, float *obs_dev, void *stream);

/* Sum over envs of [sparks, short µs, short events, crater µm³, short pulses, env-µs simulated,
 * broken, target reached] -> out_dev[8] doubles (for the NCCL all-reduce of episode stats). */
int sparc_reduce_stats(const void *state, int n_envs, int n_segments, double *out_dev8, void *stream);

/* Host-buffer convenience (the reference-facing call): H2D of the five action arrays, n_us
 * steps, D2H of the observation summary, then a stream sync.  Pointers are HOST pointers
 * (pinned for full speed); `scratch_dev` is caller-owned device scratch of
 * sparc_host_scratch_bytes(n_envs) bytes.  With at most 32 envs the host arrays MUST be pinned
 * (cudaHostAlloc / cudaHostRegister / torch pin_memory): the kernel then reads and writes them in place. */
size_t sparc_host_scratch_bytes(int n_envs);
/* If the five host action arrays lie in ONE host buffer, array k at byte offset
 * k * sparc_host_action_pitch(n_envs) (order: servo, target_voltage, current_mode, on_time,
 * off_time), sparc_step_host moves them with a single copy instead of five. */
size_t sparc_host_action_pitch(int n_envs);
int sparc_step_host(void *state, const SparcParams *p, const SparcActions *actions_host, int n_envs,
                    int n_us, uint64_t seed, uint32_t env_id_base, void *scratch_dev,
                    float *obs_host, void *stream);

/* Allocation-free variant of sparc_step_host.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and results as sparc_step_host without `scratch_dev`: the action
 * staging arrays and the observation summary live in a file-scope `static __device__` array inside the library
 * (1 114 112 bytes, held by every GPU that loads the library until exit), so the call — and everything it calls —
 * uses no cudaMalloc / cudaFree / allocating routine.  Limit: n_envs <= SPARC_STATIC_MAX_ENVS, more returns
 * SPARC_E_UNSUPPORTED with a message naming the limit.  The array is shared by EVERY call on a device and keeps
 * the last call's contents (nothing in it has to start at zero: each call overwrites what it reads); two calls
 * that overlap on the GPU (different streams or threads) race on it — serialise them, or use sparc_step_host.
 * Unlike sparc_step_host it always stages through the device array (no in-place path for <= 32 envs), so the
 * host arrays may be pageable. */
#define SPARC_STATIC_MAX_ENVS 16384
int sparc_static_max_envs(void);
int sparc_step_host_static(void *state, const SparcParams *p, const SparcActions *actions_host, int n_envs,
                           int n_us, uint64_t seed, uint32_t env_id_base, float *obs_host, void *stream);

/* Measured FP32 FMA issue peak of the current device [TFLOP/s] (dense FFMA micro-kernel, a few
 * ms; synchronous).  Roofline denominator for the compute-bound step kernel; scratch_dev needs
 * >= 4 MiB of device memory. */
int sparc_measure_fp32_peak(void *scratch_dev, size_t scratch_bytes, double *tflops_out);

/* Where every wire cell lives in the registers of an environment's lanes (for tests and diagnostics; pure host code).
 * Each lane holds two "halves" of K/2 slots; half h = 2*lane + {0,1}.  hdr = {K cells per lane, W warps per env,
 * aligned (1: per-half convection coefficients, 0: per-cell), halves in use}; for every half h < 64*W:
 * first cell, number of real cells (K/2, K/2-1 with one pad, or 0 when empty) and region (0 below the flushing
 * zone, 1 zone, 2 above, -1 empty; WireModule gives the zone its own convection coefficient, wire.py:349-374). */
int sparc_layout_info(const SparcParams *p, int32_t hdr[4], int32_t *half_first, int32_t *half_cells,
                      int32_t *half_region, int max_halves);

/* Kernel launch geometry chosen for (n_envs, n_segments): lanes per env, slots per lane,
 * threads per block, blocks.  For reporting / tests. */
int sparc_launch_info(int n_envs, int n_segments, int32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif /* SPARC_B200_H */
