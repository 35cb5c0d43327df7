"""Host-side configuration: the reference's parameter carriers and the constant derivation.

The dataclasses below keep the reference's names, fields and defaults so user code that
builds an `EnvironmentConfig` or `*ModuleParameters` for geduardo/SPARC works unchanged
(reference: src/wedm/core/env_config.py:10-90, modules/ignition.py:17-57, modules/wire.py:16-54,
modules/material.py:17-22, modules/dielectric.py:15-31, modules/mechanics.py:12-23).

`derive_params()` turns them into the flat `SparcParams` struct of include/sparc_b200.h.  All
geometry is computed here, in Python floats, with the reference's own operators (`//` vs `/`)
because `30.0 // 0.2 == 149.0` while `30.0 / 0.2 == 150.0` (wire.py:149-154) — the kernel gets
integers and coefficients, never formulas.
"""
from __future__ import annotations

import ctypes as C
import json
import math
from dataclasses import asdict, dataclass, fields
from pathlib import Path
from typing import Any, Dict, Optional

MAX_MODE = 19
_DATA = Path(__file__).parent / "data"


# ----------------------------------------------------------------------------------------------
# parameter carriers (API-compatible with the reference)
# ----------------------------------------------------------------------------------------------
@dataclass
class EnvironmentConfig:
    """Fixed set-up of one cut (reference core/env_config.py:10-35)."""

    workpiece_height: float = 20.0          # mm
    wire_diameter: float = 0.2              # mm
    wire_material: str = "brass"
    dt: int = 1                             # µs
    servo_interval: int = 1000              # µs
    initial_gap: float = 50.0               # µm
    target_cutting_distance: float = 500.0  # µm
    max_wire_temperature: float = 1500.0    # K   (never read by the loop)
    min_gap_for_operation: float = 2.0      # µm  (never read)
    max_cutting_force: float = 100.0        # N   (never read)

    @classmethod
    def from_dict(cls, config_dict: Dict[str, Any]) -> "EnvironmentConfig":
        known = {f.name for f in fields(cls)}
        return cls(**{k: v for k, v in config_dict.items() if k in known})

    @classmethod
    def from_json(cls, json_path) -> "EnvironmentConfig":
        return cls.from_dict(json.loads(Path(json_path).read_text()))

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    def to_json(self, json_path) -> None:
        Path(json_path).write_text(json.dumps(self.to_dict(), indent=2))

    def validate(self) -> None:
        """Same checks and messages as the reference (env_config.py:73-90)."""
        for name in ("workpiece_height", "wire_diameter", "initial_gap", "target_cutting_distance",
                     "dt", "servo_interval"):
            if getattr(self, name) <= 0:
                raise ValueError(f"{name} must be positive")
        if self.max_wire_temperature <= 293.15:
            raise ValueError("max_wire_temperature must be greater than room temperature")


@dataclass
class IgnitionModuleParameters:
    base_critical_density: float = 0.3
    gap_coefficient: float = 0.02
    max_critical_density: float = 0.95
    hard_short_gap: float = 2.0
    sigmoid_steepness: float = 500.0
    debris_short_duration: int = 50
    random_short_duration: int = 100
    random_short_min_gap: float = 2.0
    random_short_max_gap: float = 50.0
    random_short_max_probability: float = 0.000
    ignition_a_coeff: float = 0.48
    ignition_b_coeff: float = -3.69
    ignition_c_coeff: float = 14.05
    default_target_voltage: float = 80.0
    default_on_time: float = 3.0
    default_off_time: float = 80.0
    default_current_mode: str = "I5"
    spark_voltage_factor: float = 0.3


@dataclass
class WireModuleParameters:
    buffer_len_bottom: float = 30.0
    buffer_len_top: float = 30.0
    segment_len: float = 0.2
    spool_T: float = 293.15
    contact_offset_bottom: float = 10.0
    contact_offset_top: float = 10.0
    base_convection_coefficient: float = 14000
    plasma_efficiency: float = 0.1
    convection_velocity_factor: float = 0.5
    convection_flow_enhancement: float = 1.0
    compute_zone_mean: bool = False
    zone_mean_interval: int = 100
    critical_temp_threshold: float = 0.9
    wire_breaking_temp_factor: float = 1.1   # unused by the reference loop as well


@dataclass
class MaterialModuleParameters:
    base_overcut: float = 0.12
    # Extension (not in the reference dataclass): which crater record a current mode samples from.
    #   "reference"    the record with the mode's own key; modes without one (even modes, I19) raise
    #                  ValueError at their first spark (material.py:100-113) — the reference's behaviour;
    #   "fill_missing" as above where a record exists, the documented current scaling elsewhere;
    #   "documented"   every mode uses the scaling of docs/theory/material_theory.md:77-89:
    #                  record = available[floor(r*(N-1))], r = (I_machine - I_min)/(I_max - I_min).
    crater_mode_mapping: str = "reference"


@dataclass
class DielectricModuleParameters:
    base_flow_rate: float = 100.0
    debris_removal_efficiency: float = 0.01
    debris_obstruction_coeff: float = 1.0
    reference_gap: float = 25.0
    dielectric_temperature: float = 293.15
    ion_channel_duration: int = 6


@dataclass
class MechanicsModuleParameters:
    omega_n: float = 235.0
    zeta: float = 0.38
    max_acceleration: float = 3.0e5
    max_jerk: float = 1.0e8
    max_speed: float = 3.0e4


# ----------------------------------------------------------------------------------------------
# material database (reference core/material_db.py:10-107)
# ----------------------------------------------------------------------------------------------
@dataclass
class WireMaterial:
    name: str
    density: float
    specific_heat: float
    thermal_conductivity: float
    electrical_resistivity: float
    temperature_coefficient: float
    melting_point: float
    breaking_temperature: float


class MaterialDatabase:
    def __init__(self, data_dir: Optional[Path] = None):
        self.data_dir = Path(data_dir) if data_dir is not None else _DATA
        self._wire_materials: Dict[str, WireMaterial] = {}
        path = self.data_dir / "wire_materials.json"
        raw = json.loads(path.read_text()) if path.exists() else json.loads((_DATA / "wire_materials.json").read_text())
        for name, props in raw.items():
            self._wire_materials[name] = WireMaterial(name=name, **props)

    def get_wire_material(self, name: str) -> WireMaterial:
        if name not in self._wire_materials:
            raise ValueError(f"Unknown wire material: {name}. Available: {list(self._wire_materials.keys())}")
        return self._wire_materials[name]

    def save_materials(self) -> None:
        self.data_dir.mkdir(parents=True, exist_ok=True)
        out = {n: {k: v for k, v in asdict(m).items() if k != "name"} for n, m in self._wire_materials.items()}
        (self.data_dir / "wire_materials.json").write_text(json.dumps(out, indent=2))


_material_db: Optional[MaterialDatabase] = None


def get_material_db() -> MaterialDatabase:
    global _material_db
    if _material_db is None:
        _material_db = MaterialDatabase()
    return _material_db


def load_currents() -> Dict[str, dict]:
    path = _DATA / "currents.json"
    if not path.exists():
        raise FileNotFoundError(f"Could not find currents data file at {path}")
    return json.loads(path.read_text())


def load_craters() -> Dict[str, dict]:
    path = _DATA / "craters.json"
    if not path.exists():
        raise FileNotFoundError(f"Could not find crater data file at {path}")
    return json.loads(path.read_text())


# ----------------------------------------------------------------------------------------------
# the flat struct of include/sparc_b200.h
# ----------------------------------------------------------------------------------------------
_D, _I, _F = C.c_double, C.c_int32, C.c_float
_TAB = _D * (MAX_MODE + 1)


class SparcParams(C.Structure):
    _fields_ = [
        ("n_segments", _I), ("zone_start", _I), ("zone_end", _I), ("actual_zone_start", _I),
        ("actual_zone_end", _I), ("contact_bottom_idx", _I), ("contact_top_idx", _I),
        ("servo_interval", _I), ("control_mode", _I),
        ("debris_short_duration", _I), ("random_short_duration", _I), ("default_current_mode", _I),
        ("ion_channel_duration", _I), ("compute_zone_mean", _I), ("zone_mean_interval", _I),
        ("reserved0", _I),
        ("initial_gap", _D), ("target_position", _D), ("workpiece_height", _D), ("segment_len", _D),
        ("base_critical_density", _D), ("gap_coefficient", _D), ("max_critical_density", _D),
        ("hard_short_gap", _D), ("sigmoid_steepness", _D),
        ("random_short_min_gap", _D), ("random_short_max_gap", _D), ("random_short_max_probability", _D),
        ("ignition_a_coeff", _D), ("ignition_b_coeff", _D), ("ignition_c_coeff", _D),
        ("default_target_voltage", _D), ("default_on_time", _D), ("default_off_time", _D),
        ("spark_voltage_factor", _D),
        ("currents", _TAB), ("crater_mean", _TAB), ("crater_std", _TAB), ("kerf_times_height", _TAB),
        ("crater_avail", _I * (MAX_MODE + 1)),
        ("cavity_volume_coeff", _D), ("reference_gap", _D), ("debris_obstruction_coeff", _D),
        ("debris_removal_per_us", _D), ("dielectric_temperature", _D),
        ("spool_T", _D), ("k_cond_coeff", _D), ("joule_geom_factor", _D), ("rho_elec", _D),
        ("alpha_rho", _D), ("temp_ref", _D), ("A", _D), ("S", _D), ("temp_update_factor", _D),
        ("density_times_cp", _D), ("plasma_efficiency", _D), ("base_convection_coefficient", _D),
        ("convection_velocity_factor", _D), ("convection_flow_enhancement", _D),
        ("critical_temperature_f32", _F), ("breaking_temperature_f32", _F),
        ("cond_coeff_f32", _F), ("reserved1", _F),
        ("mech_dt", _D), ("damping_coeff", _D), ("stiffness_coeff", _D), ("omega_n", _D),
        ("max_acceleration", _D), ("max_jerk_dt", _D), ("max_speed", _D),
    ]


def _mode_index(mode: str) -> int:
    """"I7" -> 7; anything unparsable -> 0 (treated as unknown)."""
    try:
        return int(str(mode)[1:]) if str(mode).startswith("I") else 0
    except ValueError:
        return 0


def wire_geometry(config: EnvironmentConfig, wp: WireModuleParameters) -> Dict[str, int]:
    """Segment indices exactly as WireModule.__init__ computes them (wire.py:143-156, 208-252)."""
    total_L = wp.buffer_len_bottom + config.workpiece_height + wp.buffer_len_top
    n = max(1, int(total_L / wp.segment_len))
    zone_start = int(wp.buffer_len_bottom // wp.segment_len)
    zone_end = min(zone_start + int(config.workpiece_height // wp.segment_len), n)
    zone_start = min(zone_start, zone_end)
    cb = max(0, int((wp.buffer_len_bottom - wp.contact_offset_bottom) / wp.segment_len))
    ct = min(n - 1, int((wp.buffer_len_bottom + config.workpiece_height + wp.contact_offset_top) / wp.segment_len))
    cb = max(0, min(cb, zone_start - 1))
    ct = min(n - 1, max(ct, zone_end))
    return dict(n_segments=n, zone_start=zone_start, zone_end=zone_end,
                actual_zone_start=min(zone_start, n - 1), actual_zone_end=min(zone_end, n),
                contact_bottom_idx=cb, contact_top_idx=ct)


def derive_params(config: EnvironmentConfig, mechanics_control_mode: str = "position",
                  ignition: Optional[IgnitionModuleParameters] = None,
                  wire: Optional[WireModuleParameters] = None,
                  material: Optional[MaterialModuleParameters] = None,
                  dielectric: Optional[DielectricModuleParameters] = None,
                  mechanics: Optional[MechanicsModuleParameters] = None) -> SparcParams:
    if mechanics_control_mode not in ("position", "velocity"):
        raise ValueError(f"mechanics_control_mode must be 'position' or 'velocity', got {mechanics_control_mode}")
    config.validate()
    if config.dt != 1:
        # the reference is only self-consistent for dt == 1 (wire.py:192 hard-codes 1e-6 s,
        # the µs counters assume 1); SURVEY a19
        raise ValueError("only dt == 1 µs is supported")
    ig = ignition or IgnitionModuleParameters()
    wr = wire or WireModuleParameters()
    mt = material or MaterialModuleParameters()
    di = dielectric or DielectricModuleParameters()
    me = mechanics or MechanicsModuleParameters()
    mat = get_material_db().get_wire_material(config.wire_material)

    p = SparcParams()
    for k, v in wire_geometry(config, wr).items():
        setattr(p, k, v)
    p.servo_interval = int(config.servo_interval)
    p.control_mode = 0 if mechanics_control_mode == "position" else 1
    p.debris_short_duration = int(ig.debris_short_duration)
    p.random_short_duration = int(ig.random_short_duration)
    dm = _mode_index(ig.default_current_mode)
    p.default_current_mode = dm if 1 <= dm <= MAX_MODE else 5
    p.ion_channel_duration = int(di.ion_channel_duration)
    p.compute_zone_mean = int(bool(wr.compute_zone_mean))
    p.zone_mean_interval = int(wr.zone_mean_interval)
    p.initial_gap = float(config.initial_gap)
    p.target_position = float(config.target_cutting_distance)
    p.workpiece_height = float(config.workpiece_height)
    p.segment_len = float(wr.segment_len)
    for name in ("base_critical_density", "gap_coefficient", "max_critical_density", "hard_short_gap",
                 "sigmoid_steepness", "random_short_min_gap", "random_short_max_gap",
                 "random_short_max_probability", "ignition_a_coeff", "ignition_b_coeff",
                 "ignition_c_coeff", "default_target_voltage", "default_on_time", "default_off_time",
                 "spark_voltage_factor"):
        setattr(p, name, float(getattr(ig, name)))

    # per-mode tables (ignition.py:98-113; material.py:100-123, 152-169)
    currents = load_currents()
    craters = load_craters()
    mapping = getattr(mt, "crater_mode_mapping", "reference")
    if mapping not in ("reference", "fill_missing", "documented"):
        raise ValueError(f"crater_mode_mapping must be 'reference', 'fill_missing' or 'documented', got {mapping}")
    avail = sorted(craters, key=lambda name: int(name[1:]))            # I1, I3, ..., I17
    amps = [float(v["Current"]) for v in currents.values()]
    i_min, i_max = min(amps), max(amps)
    for k in range(1, MAX_MODE + 1):
        key = f"I{k}"
        if key in currents:
            p.currents[k] = float(currents[key]["Current"])
        src = key if key in craters else None
        if key in currents and (mapping == "documented" or (mapping == "fill_missing" and src is None)):
            r = (float(currents[key]["Current"]) - i_min) / (i_max - i_min)
            src = avail[min(int(math.floor(r * (len(avail) - 1))), len(avail) - 1)]
        if src is not None:
            c = craters[src]
            p.crater_mean[k] = float(c["ellipsoid_volume_half"])
            p.crater_std[k] = float(c["ellipsoid_volume_std"])
            kerf_width_mm = mt.base_overcut + config.wire_diameter + c["depth"] / 1000.0
            kh = kerf_width_mm * config.workpiece_height
            p.kerf_times_height[k] = kh if (kerf_width_mm > 0 and config.workpiece_height > 0) else 0.0
            p.crater_avail[k] = 1

    # dielectric.py:57-66
    p.cavity_volume_coeff = math.pi * (config.wire_diameter / 2.0) * config.workpiece_height
    p.reference_gap = float(di.reference_gap)
    p.debris_obstruction_coeff = float(di.debris_obstruction_coeff)
    p.debris_removal_per_us = di.debris_removal_efficiency * di.base_flow_rate * 1e-6
    p.dielectric_temperature = float(di.dielectric_temperature)

    # wire.py:158-220
    r_wire = config.wire_diameter / 2.0
    delta_y = wr.segment_len * 1e-3
    S = math.pi * (r_wire * 1e-3) ** 2
    A = 2 * math.pi * (r_wire * 1e-3) * delta_y
    denominator = mat.density * mat.specific_heat * S * delta_y
    if denominator == 0:
        raise ValueError("Denominator for dT/dt is zero. Check wire/segment properties.")
    p.spool_T = float(wr.spool_T)
    p.k_cond_coeff = mat.thermal_conductivity * S / delta_y
    p.joule_geom_factor = delta_y / S if S != 0 else 0.0
    p.rho_elec = float(mat.electrical_resistivity)
    p.alpha_rho = float(mat.temperature_coefficient)
    p.temp_ref = 293.15
    p.A = A
    p.S = S
    p.temp_update_factor = 1e-6 / denominator
    p.density_times_cp = mat.density * mat.specific_heat
    p.plasma_efficiency = float(wr.plasma_efficiency)
    p.base_convection_coefficient = float(wr.base_convection_coefficient)
    p.convection_velocity_factor = float(wr.convection_velocity_factor)
    p.convection_flow_enhancement = float(wr.convection_flow_enhancement)
    # np.max(T) is float32; NumPy >= 2 compares it against the Python float rounded to float32
    p.critical_temperature_f32 = mat.melting_point * wr.critical_temp_threshold
    p.breaking_temperature_f32 = mat.breaking_temperature
    p.cond_coeff_f32 = p.temp_update_factor * p.k_cond_coeff        # rounded to float32 by the ctypes field

    # mechanics.py:48-57
    p.mech_dt = config.dt * 1e-6
    p.damping_coeff = -2.0 * me.zeta * me.omega_n
    p.stiffness_coeff = -(me.omega_n ** 2)
    p.omega_n = float(me.omega_n)
    p.max_acceleration = float(me.max_acceleration)
    p.max_jerk_dt = me.max_jerk * p.mech_dt
    p.max_speed = float(me.max_speed)
    return p
