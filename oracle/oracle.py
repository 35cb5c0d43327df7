"""ctypes wrapper over the C parity oracle (`oracle/wedm_oracle.c`).  TEST INFRASTRUCTURE ONLY.

Only tests/, `__graft_entry__.smoke()` and bench.py's cpu_baseline / `--impl reference`
legs may import this module; the product package `sparc_b200` never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libwedm_oracle.so")

MAX_MODE = 19
N_SLOTS = 5
F64_FIELDS = ("wp", "wire", "v", "prev_accel", "voltage", "current", "debris_volume",
              "debris_density", "flow", "last_crater_volume", "spark_y", "tmax")
I32_FIELDS = ("time", "t_servo", "t_spark_ign", "t_spark_end", "spark_state", "spark_dur",
              "is_short", "rand_short_rem", "deb_short_rem", "ion_rem", "crit_count",
              "broken", "target_reached", "terminated")

# Reference data tables, restated (modules/currents.json:1-21, modules/area_corrected.json:1-73,
# data/wire_materials.json:1-11).
CURRENTS = {1: 30, 2: 35, 3: 40, 4: 50, 5: 60, 6: 68, 7: 80, 8: 95, 9: 110, 10: 130, 11: 155,
            12: 180, 13: 215, 14: 255, 15: 305, 16: 360, 17: 425, 18: 500, 19: 600}
# mode: (ellipsoid_volume_half, ellipsoid_volume_std, depth)
CRATERS = {1: (2163.6174, 447.5072, 2.9967), 3: (2376.5835, 523.6196, 3.0443),
           5: (4866.9691, 899.0243, 3.5302), 7: (5556.8153, 1167.296, 3.6478),
           9: (6219.2736, 1284.4152, 3.7556), 11: (6029.8571, 1005.804, 3.7252),
           13: (8913.8402, 2949.127, 4.1537), 15: (26468.9924, 6472.3303, 5.9966),
           17: (59549.9184, 8997.5034, 6.2647)}
BRASS = dict(density=8400, specific_heat=377, thermal_conductivity=120, electrical_resistivity=6.4e-8,
             temperature_coefficient=0.0039, melting_point=1173, breaking_temperature=1500)

# dataclass defaults of the reference (env_config.py:18-30; ignition.py:17-57; wire.py:16-54;
# material.py:17-22; dielectric.py:15-31; mechanics.py:12-23)
DEFAULTS = dict(
    workpiece_height=20.0, wire_diameter=0.2, dt=1, servo_interval=1000, initial_gap=50.0,
    target_cutting_distance=500.0, control_mode=0,
    **BRASS,
    base_critical_density=0.3, gap_coefficient=0.02, max_critical_density=0.95, hard_short_gap=2.0,
    sigmoid_steepness=500.0, debris_short_duration=50, random_short_duration=100,
    random_short_min_gap=2.0, random_short_max_gap=50.0, random_short_max_probability=0.0,
    ignition_a_coeff=0.48, ignition_b_coeff=-3.69, ignition_c_coeff=14.05,
    default_target_voltage=80.0, default_on_time=3.0, default_off_time=80.0, default_current_mode=5,
    spark_voltage_factor=0.3,
    buffer_len_bottom=30.0, buffer_len_top=30.0, segment_len=0.2, spool_T=293.15,
    contact_offset_bottom=10.0, contact_offset_top=10.0, base_convection_coefficient=14000,
    plasma_efficiency=0.1, convection_velocity_factor=0.5, convection_flow_enhancement=1.0,
    compute_zone_mean=0, zone_mean_interval=100, critical_temp_threshold=0.9,
    base_overcut=0.12,
    base_flow_rate=100.0, debris_removal_efficiency=0.01, debris_obstruction_coeff=1.0,
    reference_gap=25.0, dielectric_temperature=293.15, ion_channel_duration=6,
    omega_n=235.0, zeta=0.38, max_acceleration=3.0e5, max_jerk=1.0e8, max_speed=3.0e4,
)

_D, _I = C.c_double, C.c_int32


class OrParams(C.Structure):
    _fields_ = [
        ("workpiece_height", _D), ("wire_diameter", _D), ("dt", _I), ("servo_interval", _I),
        ("initial_gap", _D), ("target_cutting_distance", _D), ("control_mode", _I), ("pad0", _I),
        ("density", _D), ("specific_heat", _D), ("thermal_conductivity", _D),
        ("electrical_resistivity", _D), ("temperature_coefficient", _D), ("melting_point", _D),
        ("breaking_temperature", _D),
        ("base_critical_density", _D), ("gap_coefficient", _D), ("max_critical_density", _D),
        ("hard_short_gap", _D), ("sigmoid_steepness", _D),
        ("debris_short_duration", _I), ("random_short_duration", _I),
        ("random_short_min_gap", _D), ("random_short_max_gap", _D), ("random_short_max_probability", _D),
        ("ignition_a_coeff", _D), ("ignition_b_coeff", _D), ("ignition_c_coeff", _D),
        ("default_target_voltage", _D), ("default_on_time", _D), ("default_off_time", _D),
        ("default_current_mode", _I), ("pad1", _I), ("spark_voltage_factor", _D),
        ("buffer_len_bottom", _D), ("buffer_len_top", _D), ("segment_len", _D), ("spool_T", _D),
        ("contact_offset_bottom", _D), ("contact_offset_top", _D),
        ("base_convection_coefficient", _D), ("plasma_efficiency", _D),
        ("convection_velocity_factor", _D), ("convection_flow_enhancement", _D),
        ("compute_zone_mean", _I), ("zone_mean_interval", _I), ("critical_temp_threshold", _D),
        ("base_overcut", _D),
        ("base_flow_rate", _D), ("debris_removal_efficiency", _D), ("debris_obstruction_coeff", _D),
        ("reference_gap", _D), ("dielectric_temperature", _D), ("ion_channel_duration", _I), ("pad2", _I),
        ("omega_n", _D), ("zeta", _D), ("max_acceleration", _D), ("max_jerk", _D), ("max_speed", _D),
        ("currents", _D * (MAX_MODE + 1)), ("crater_mean", _D * (MAX_MODE + 1)),
        ("crater_std", _D * (MAX_MODE + 1)), ("crater_depth", _D * (MAX_MODE + 1)),
        ("crater_avail", _I * (MAX_MODE + 1)),
    ]


class OrAction(C.Structure):
    _fields_ = [("servo", _D), ("target_voltage", _D), ("on_time", _D), ("off_time", _D),
                ("current_mode", _I), ("pad", _I)]


def make_params(**overrides) -> OrParams:
    """Raw reference parameters (dataclass field names); `mechanics_control_mode` accepted too."""
    vals = dict(DEFAULTS)
    if "mechanics_control_mode" in overrides:
        m = overrides.pop("mechanics_control_mode")
        if m not in ("position", "velocity"):
            raise ValueError(f"mechanics_control_mode must be 'position' or 'velocity', got {m}")
        overrides["control_mode"] = 0 if m == "position" else 1
    if "default_current_mode" in overrides and isinstance(overrides["default_current_mode"], str):
        overrides["default_current_mode"] = int(overrides["default_current_mode"][1:])
    unknown = set(overrides) - set(vals)
    if unknown:
        raise KeyError(f"unknown oracle parameters: {sorted(unknown)}")
    vals.update(overrides)
    p = OrParams()
    for k, v in vals.items():
        setattr(p, k, int(v) if isinstance(getattr(p, k), int) else float(v))
    for m, cur in CURRENTS.items():
        p.currents[m] = float(cur)
    for m, (mean, std, depth) in CRATERS.items():
        p.crater_mean[m], p.crater_std[m], p.crater_depth[m], p.crater_avail[m] = mean, std, depth, 1
    return p


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (oracle/Makefile).  Building the checker is not using it."""
    src = os.path.join(_HERE, "wedm_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, "wedm_oracle.h"))):
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.or_create.restype = C.c_void_p
        L.or_create.argtypes = [C.POINTER(OrParams)]
        L.or_destroy.argtypes = [C.c_void_p]
        L.or_reset.argtypes = [C.c_void_p]
        L.or_reset_reference.argtypes = [C.c_void_p]
        L.or_n_segments.argtypes = [C.c_void_p]
        L.or_geometry.argtypes = [C.c_void_p, C.POINTER(_I)]
        L.or_temperature.restype = C.POINTER(C.c_float)
        L.or_temperature.argtypes = [C.c_void_p]
        L.or_set_unwind_velocity.argtypes = [C.c_void_p, _D]
        L.or_set_initial.argtypes = [C.c_void_p, _D, _D, _D]
        L.or_step.argtypes = [C.c_void_p, C.POINTER(OrAction), C.POINTER(_D)]
        L.or_run.restype = C.c_int64
        L.or_run.argtypes = [C.c_void_p, C.c_int64, C.POINTER(OrAction), C.c_int64, C.c_void_p,
                             C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int64,
                             C.POINTER(C.c_int64)]
        L.or_record.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.or_stats.argtypes = [C.c_void_p, C.POINTER(_D)]
        L.or_philox4x32_10.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.or_philox_draws.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(_D)]
        L.or_bench.restype = C.c_int64
        L.or_bench.argtypes = [C.POINTER(OrParams), C.POINTER(OrAction), C.c_int, C.c_int64, C.c_int,
                               C.c_uint64, C.POINTER(_D)]
        _lib = L
    return _lib


def _actions_array(actions) -> "C.Array":
    """actions: [n,5] rows (servo, target_voltage, current_mode, ON_time, OFF_time)."""
    a = np.atleast_2d(np.asarray(actions, dtype=np.float64))
    arr = (OrAction * len(a))()
    for k, row in enumerate(a):
        arr[k].servo, arr[k].target_voltage = row[0], row[1]
        arr[k].current_mode = int(row[2])
        arr[k].on_time, arr[k].off_time = row[3], row[4]
    return arr


class OracleEnv:
    """One CPU oracle environment (fresh-env semantics)."""

    def __init__(self, params: Optional[OrParams] = None, **overrides):
        self.params = params if params is not None else make_params(**overrides)
        self._h = lib().or_create(C.byref(self.params))
        if not self._h:
            raise MemoryError("or_create failed")
        self.n_segments = lib().or_n_segments(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().or_destroy(self._h)
            self._h = None

    def reset(self, reference_semantics: bool = False):
        """Fresh-env reset, or (reference_semantics=True) the reference's own reset(): only EDMState is
        re-created, module-internal state survives (wire_edm.py:106-114)."""
        (lib().or_reset_reference if reference_semantics else lib().or_reset)(self._h)

    def geometry(self):
        out = (_I * 5)()
        lib().or_geometry(self._h, out)
        return dict(zip(("n", "zone_start", "zone_end", "contact_bottom", "contact_top"), list(out)))

    @property
    def temperature(self) -> np.ndarray:
        return np.ctypeslib.as_array(lib().or_temperature(self._h), shape=(self.n_segments,))

    def set_unwind_velocity(self, v: float):
        lib().or_set_unwind_velocity(self._h, v)

    def set_initial(self, wp: float, wire: float = 0.0, v: float = 0.0):
        lib().or_set_initial(self._h, wp, wire, v)

    def record(self, terminated: int = 0):
        f = np.zeros(len(F64_FIELDS))
        i = np.zeros(len(I32_FIELDS), dtype=np.int32)
        lib().or_record(self._h, terminated, f.ctypes.data, i.ctypes.data)
        return f, i

    def stats(self) -> np.ndarray:
        out = (_D * 8)()
        lib().or_stats(self._h, out)
        return np.array(out)

    def step(self, action_row: Sequence[float], draws5: Sequence[float]) -> int:
        a = _actions_array([action_row])
        d = (_D * N_SLOTS)(*[float(x) for x in draws5])
        return lib().or_step(self._h, a, d)

    def run(self, n_us: int, actions, draws: Optional[np.ndarray] = None, seed: int = 0,
            env_id: int = 0, rec_every: int = 1, t_every: int = 0):
        """Advance up to n_us µs; returns dict like oracle.ref_runner.run_reference."""
        acts = _actions_array(actions)
        if draws is not None:
            draws = np.ascontiguousarray(draws, dtype=np.float64)
            assert draws.shape[0] >= n_us and draws.shape[1] == N_SLOTS
        chunks = [n_us] if not t_every else [t_every] * (n_us // t_every) + ([n_us % t_every] if n_us % t_every else [])
        f_all, i_all, t_rows, t_times = [], [], [], []
        done = 0
        # the action cursor lives in or_run; when chunking we re-derive it from control steps
        act_idx = 0
        for n in chunks:
            nrec_max = (n // rec_every + 2) if rec_every else 0
            f = np.zeros((max(nrec_max, 1), len(F64_FIELDS)))
            i = np.zeros((max(nrec_max, 1), len(I32_FIELDS)), dtype=np.int32)
            nrec = C.c_int64(0)
            sub = (OrAction * (len(acts) - act_idx))(*acts[act_idx:]) if act_idx < len(acts) else (OrAction * 1)(acts[len(acts) - 1])
            dptr = draws[done:].ctypes.data if draws is not None else None
            t0 = self.record()[1][0]
            k = lib().or_run(self._h, n, sub, len(sub), dptr, seed, env_id,
                             f.ctypes.data if rec_every else None, i.ctypes.data if rec_every else None,
                             rec_every, C.byref(nrec))
            done += k
            f_all.append(f[: nrec.value])
            i_all.append(i[: nrec.value])
            frec, irec = self.record()
            terminated = bool(irec[11] or irec[12]) or k < n
            # control steps in this chunk: entry times that are multiples of servo_interval (>0)
            si = self.params.servo_interval
            lo, hi = int(t0), int(t0) + int(k) - 1
            if k > 0:
                act_idx += hi // si - (max(lo, 1) - 1) // si
            if t_every:
                t_rows.append(self.temperature.copy())
                t_times.append(done)
            if terminated:
                break
        return {
            "f64": np.concatenate(f_all) if f_all else np.zeros((0, len(F64_FIELDS))),
            "i32": np.concatenate(i_all) if i_all else np.zeros((0, len(I32_FIELDS)), dtype=np.int32),
            "T": np.asarray(t_rows, dtype=np.float32),
            "T_times": np.asarray(t_times, dtype=np.int64),
            "n_done": done,
        }


def philox_draws(seed: int, env_id: int, episode: int, t: int) -> np.ndarray:
    out = (_D * N_SLOTS)()
    lib().or_philox_draws(seed, env_id, episode, t, out)
    return np.array(out)


def philox4x32_10(ctr, key) -> np.ndarray:
    c = (C.c_uint32 * 4)(*[int(x) for x in ctr])
    k = (C.c_uint32 * 2)(*[int(x) for x in key])
    o = (C.c_uint32 * 4)()
    lib().or_philox4x32_10(c, k, o)
    return np.array(o, dtype=np.uint32)


def bench(params: OrParams, action_row, n_envs: int, n_us: int, n_threads: int, seed: int = 0):
    """Multi-threaded throughput of the oracle; returns (env_us_done, stats[8])."""
    a = _actions_array([action_row])
    st = (_D * 8)()
    tot = lib().or_bench(C.byref(params), a, n_envs, n_us, n_threads, seed, st)
    return int(tot), np.array(st)
