"""ctypes binding of include/sparc_b200.h.  Fails loudly when the CUDA library is missing."""
from __future__ import annotations

import ctypes as C
import os

from .params import SparcParams

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPARC_LIB_PATH") or os.path.join(HERE, "_lib", "libsparc_b200.so")

ABI_VERSION = 5        # include/sparc_b200.h SPARC_ABI_VERSION this binding was written against
N_OBS = 12         # rows 0..7 float32; rows 8..10 int32 bit patterns (time, spark_state, time_since_servo)
N_STATS = 8
N_DRAW_SLOTS = 5
# rows of the per-signal log (include/sparc_b200.h SPARC_SIG_*), by the reference's EDMState names
SIGNALS = ("time", "voltage", "current", "wire_position", "wire_velocity", "workpiece_position", "target_delta",
           "debris_density", "flow_rate", "is_short_circuit", "spark_state", "max_wire_temperature",
           "wire_average_temperature", "debris_volume", "time_in_critical_temp", "flags")
N_SIGNALS = len(SIGNALS)

F_SHORT, F_BROKEN, F_TARGET, F_MODE_ERROR, F_HAS_Y, F_THERMAL_BREAK = 1, 2, 4, 8, 16, 32
F_DONE = F_BROKEN | F_TARGET | F_MODE_ERROR


class SparcActions(C.Structure):
    _fields_ = [("servo", C.c_void_p), ("target_voltage", C.c_void_p), ("current_mode", C.c_void_p),
                ("on_time", C.c_void_p), ("off_time", C.c_void_p)]


class SparcController(C.Structure):
    _fields_ = [("kind", C.c_int32), ("reserved", C.c_int32), ("desired_gap", C.c_double), ("gain", C.c_double),
                ("limit", C.c_double)]


class SparcError(RuntimeError):
    pass


_lib = None

# every symbol include/sparc_b200.h declares
EXPORTS = ("sparc_abi_version", "sparc_last_error", "sparc_num_fields", "sparc_field_info",
           "sparc_state_bytes", "sparc_padded_segments", "sparc_init", "sparc_reset", "sparc_reseed", "sparc_reset_reference",
           "sparc_params_bytes", "sparc_step",
           "sparc_step_injected", "sparc_step_ex", "sparc_step_logged", "sparc_step_sampled", "sparc_step_observe", "sparc_observe", "sparc_reduce_stats", "sparc_host_scratch_bytes",
           "sparc_step_host", "sparc_step_host_static", "sparc_static_max_envs", "sparc_host_action_pitch", "sparc_launch_info", "sparc_layout_info", "sparc_measure_fp32_peak")


def lib():
    """Load libsparc_b200.so (built in-tree by `python -m sparc_b200.build`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SparcError(f"{LIB_PATH} is missing: run `python -m sparc_b200.build` "
                         "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, u32, u64, sz = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_size_t
    PP, PA = C.POINTER(SparcParams), C.POINTER(SparcActions)
    L.sparc_abi_version.restype = i32
    L.sparc_last_error.restype = C.c_char_p
    L.sparc_num_fields.restype = i32
    L.sparc_field_info.argtypes = [i32, i32, i32, C.POINTER(C.c_char_p), C.POINTER(C.c_char),
                                   C.POINTER(sz), C.POINTER(sz)]
    L.sparc_state_bytes.restype = sz
    L.sparc_state_bytes.argtypes = [i32, i32]
    L.sparc_padded_segments.argtypes = [i32]
    L.sparc_init.argtypes = [vp, PP, i32, vp]
    L.sparc_reset.argtypes = [vp, PP, i32, vp, vp]
    L.sparc_reseed.argtypes = [vp, PP, i32, vp, vp]
    L.sparc_reset_reference.argtypes = [vp, PP, i32, vp, vp]
    L.sparc_params_bytes.restype = sz
    L.sparc_step.argtypes = [vp, PP, PA, i32, i32, u64, u32, vp]
    L.sparc_step_injected.argtypes = [vp, PP, PA, i32, i32, vp, vp]
    L.sparc_step_ex.argtypes = [vp, PP, PA, C.POINTER(SparcController), i32, i32, u64, u32, vp, vp, vp]
    L.sparc_step_logged.argtypes = [vp, PP, PA, i32, i32, u64, u32, vp, vp, i32, i32, vp]
    L.sparc_step_sampled.argtypes = [vp, PP, PA, i32, i32, u64, u32, vp, vp, i32, i32, vp, i32, vp]
    L.sparc_step_observe.argtypes = [vp, PP, PA, i32, i32, u64, u32, vp, vp, vp]
    L.sparc_observe.argtypes = [vp, i32, i32, vp, vp]
    L.sparc_reduce_stats.argtypes = [vp, i32, i32, vp, vp]
    L.sparc_host_scratch_bytes.restype = sz
    L.sparc_host_scratch_bytes.argtypes = [i32]
    L.sparc_host_action_pitch.restype = sz
    L.sparc_host_action_pitch.argtypes = [i32]
    L.sparc_step_host.argtypes = [vp, PP, PA, i32, i32, u64, u32, vp, vp, vp]
    L.sparc_step_host_static.argtypes = [vp, PP, PA, i32, i32, u64, u32, vp, vp]
    L.sparc_static_max_envs.restype = i32
    L.sparc_measure_fp32_peak.argtypes = [vp, sz, C.POINTER(C.c_double)]
    L.sparc_layout_info.argtypes = [PP, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), i32]
    L.sparc_launch_info.argtypes = [i32, i32, C.POINTER(C.c_int32)]
    # a stale or foreign library (SPARC_LIB_PATH) must not be fed a SparcParams of another layout
    if L.sparc_abi_version() != ABI_VERSION or L.sparc_params_bytes() != C.sizeof(SparcParams):
        raise SparcError(f"{LIB_PATH}: ABI version {L.sparc_abi_version()} / sizeof(SparcParams) "
                         f"{L.sparc_params_bytes()} do not match this binding ({ABI_VERSION} / "
                         f"{C.sizeof(SparcParams)}): rebuild with `python -m sparc_b200.build --force`")
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise SparcError(f"sparc_b200 error {rc}: {lib().sparc_last_error().decode()}")


def field_table(n_envs: int, n_segments: int):
    """[(name, dtype_char, byte_offset, count)] for the state blob."""
    L = lib()
    out = []
    for idx in range(L.sparc_num_fields()):
        name, dt, off, cnt = C.c_char_p(), C.c_char(), C.c_size_t(), C.c_size_t()
        check(L.sparc_field_info(idx, n_envs, n_segments, C.byref(name), C.byref(dt), C.byref(off), C.byref(cnt)))
        out.append((name.value.decode(), dt.value.decode(), off.value, cnt.value))
    return out
