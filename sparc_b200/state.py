"""`EDMState`: the reference's state container (core/state.py:25-91) and its device-backed view.

`EDMState` keeps the reference dataclass (same field names and defaults) for code that builds or
inspects states on the host.  `DeviceStateView` exposes the SAME attribute names for one
environment of a batched engine: reads are lazy device->host copies of the SoA fields, writes go
back to the device — so `env.state.workpiece_position = 80.0` keeps working.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np

crater_dtype = np.dtype([("radius", "f4"), ("y_position", "f4"), ("time_formed", "i4"), ("depth", "f4")])


@dataclass
class EDMState:
    time: int = 0
    time_since_servo: int = 0
    time_since_open_voltage: int = 0
    time_since_spark_ignition: int = 0
    time_since_spark_end: int = 0
    voltage: Optional[float] = None
    current: Optional[float] = None
    target_voltage: Optional[float] = None
    current_mode: Optional[str] = None
    OFF_time: Optional[float] = None
    ON_time: Optional[float] = None
    workpiece_position: float = 0.0
    wire_position: float = 0.0
    wire_velocity: float = 0.0
    wire_unwinding_velocity: float = 0.2
    wire_temperature: np.ndarray = field(default_factory=lambda: np.array([], dtype=np.float32))
    time_in_critical_temp: int = 0
    wire_average_temperature: Optional[float] = None
    spark_status: List[Optional[float]] = field(default_factory=lambda: [0, None, 0])
    dielectric_conductivity: float = 0.0
    dielectric_temperature: float = 0.0
    debris_concentration: float = 0.0
    dielectric_flow_rate: float = 0.0
    ionized_channel: Optional[Tuple[float, int]] = None
    debris_volume: float = 0.0
    debris_density: float = 0.0
    cavity_volume: float = 0.0
    flow_rate: float = 0.0
    last_crater_volume: float = 0.0
    is_short_circuit: bool = False
    is_wire_broken: bool = False
    is_wire_colliding: bool = False
    is_target_distance_reached: bool = False
    target_delta: float = 0.0
    target_position: float = 500.0


# reference attribute -> engine field (plain scalar mappings)
_SCALAR_MAP = {
    "time": "time", "time_since_servo": "t_servo", "time_since_open_voltage": "t_open_v",
    "time_since_spark_ignition": "t_spark_ign", "time_since_spark_end": "t_spark_end",
    "voltage": "voltage", "current": "current", "workpiece_position": "wp", "wire_position": "wire",
    "wire_velocity": "v", "wire_unwinding_velocity": "unwind_v", "time_in_critical_temp": "crit_count",
    "debris_volume": "debris_volume", "debris_density": "debris_density", "cavity_volume": "cavity_volume",
    "flow_rate": "flow", "last_crater_volume": "last_crater_volume", "target_delta": "target_delta",
    "debris_concentration": "debris_density",
}
_FALSY_NONE = {"target_voltage": "target_voltage", "ON_time": "on_time", "OFF_time": "off_time"}
_FLAG_MAP = {"is_short_circuit": 1, "is_wire_broken": 2, "is_target_distance_reached": 4}


class DeviceStateView:
    """Attribute-compatible view of environment `index` of an `Engine`."""

    def __init__(self, engine, index: int = 0, base_flow_rate: float = 100.0, dielectric_temperature: float = 293.15):
        object.__setattr__(self, "_eng", engine)
        object.__setattr__(self, "_i", int(index))
        object.__setattr__(self, "_base_flow_rate", base_flow_rate)
        object.__setattr__(self, "_diel_T", dielectric_temperature)

    def _get(self, fname):
        return self._eng[fname][self._i].item()

    def __getattr__(self, name):
        if name in _SCALAR_MAP:
            return self._get(_SCALAR_MAP[name])
        if name in _FALSY_NONE:
            v = self._get(_FALSY_NONE[name])
            return None if (v == 0.0 and self._get("current_mode") == 0) else v
        if name in _FLAG_MAP:
            return bool(self._get("flags") & _FLAG_MAP[name])
        if name == "current_mode":
            m = self._get("current_mode")
            return None if m == 0 else f"I{0 if m < 0 else m}"
        if name == "wire_temperature":
            return self._eng.T[self._i].detach().cpu().numpy().copy()
        if name == "wire_average_temperature":
            return self._get("wire_avg_T") if self._eng.params.compute_zone_mean and self._get("time") > 0 else None
        if name == "spark_status":
            y = self._get("spark_y")
            return [self._get("spark_state"), (y if self._get("flags") & 16 else None), self._get("spark_dur")]
        if name == "ionized_channel":
            r = self._get("ion_rem")
            return (self._get("spark_y"), r) if r > 0 else None
        if name == "dielectric_temperature":
            return self._diel_T if self._get("time") > 0 else 0.0
        if name == "dielectric_flow_rate":
            return self._get("flow") * self._base_flow_rate / 1e9
        if name == "target_position":
            return float(self._eng.params.target_position)
        if name in ("dielectric_conductivity",):
            return 0.0
        if name == "is_wire_colliding":
            return False
        raise AttributeError(name)

    def __setattr__(self, name, value):
        eng, i = self._eng, self._i
        if name in _SCALAR_MAP:
            eng[_SCALAR_MAP[name]][i] = 0 if value is None else value
        elif name in _FALSY_NONE:
            eng[_FALSY_NONE[name]][i] = 0.0 if value is None else float(value)
        elif name in _FLAG_MAP:
            f = int(eng["flags"][i].item())
            eng["flags"][i] = (f | _FLAG_MAP[name]) if value else (f & ~_FLAG_MAP[name])
        elif name == "current_mode":
            eng["current_mode"][i] = 0 if value is None else (int(str(value)[1:]) or -1)
        elif name == "wire_temperature":
            import torch

            arr = np.asarray(value, dtype=np.float32)
            if arr.shape != (eng.n_segments,):
                raise ValueError(f"wire_temperature must have shape ({eng.n_segments},)")
            eng.T[i].copy_(torch.from_numpy(arr))
        elif name == "spark_status":
            st, y, dur = value
            eng["spark_state"][i] = int(st)
            eng["spark_dur"][i] = int(dur)
            eng["spark_y"][i] = -1.0 if y is None else float(y)
            f = int(eng["flags"][i].item())
            eng["flags"][i] = (f | 16) if y is not None else (f & ~16)
        elif name == "target_position":
            raise AttributeError("target_position is fixed by EnvironmentConfig.target_cutting_distance")
        else:
            raise AttributeError(f"cannot set {name!r} on a device-backed EDMState")

    def to_host(self) -> EDMState:
        """Materialise a plain reference-style EDMState (one D2H per field)."""
        out = EDMState()
        for f in out.__dataclass_fields__:
            try:
                setattr(out, f, getattr(self, f))
            except AttributeError:
                pass
        return out
