"""The reference's five physics modules as PARAMETER CARRIERS.

In geduardo/SPARC each module mutates the shared `EDMState` in `update(state)`
(core/module.py:25-29) and the env calls them once per µs.  Here the five updates are fused into
one CUDA kernel (sparc_b200/csrc/step_kernel.cuh); these classes keep the reference's names and
constructor signatures, carry the parameters, expose the derived geometry the reference exposes
(`env.wire.n_segments`, ...) and read module-internal state back from the device on request.
Calling `update()` raises: the per-µs loop no longer runs in Python.
"""
from __future__ import annotations

from typing import ClassVar, Dict

from ..params import (DielectricModuleParameters, IgnitionModuleParameters, MaterialModuleParameters,
                      MechanicsModuleParameters, WireModuleParameters, get_material_db, load_craters,
                      load_currents, wire_geometry)


class EDMModule:
    registry: ClassVar[Dict[str, type]] = {}

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        EDMModule.registry[cls.__name__] = cls

    def __init__(self, env):
        self.env = env

    def update(self, state) -> None:
        raise NotImplementedError(
            f"{type(self).__name__}.update is fused into the sparc_b200 step kernel; "
            "advance the simulation with env.step(...)")

    def _field(self, name):
        return self.env._engine[name][self.env._index].item()


class IgnitionModule(EDMModule):
    def __init__(self, env, parameters: IgnitionModuleParameters = None):
        super().__init__(env)
        self.params = parameters or IgnitionModuleParameters()
        self.currents_data = load_currents()

    @property
    def random_short_remaining(self) -> int:
        return self._field("rand_short_rem")

    @property
    def debris_short_remaining(self) -> int:
        return self._field("deb_short_rem")

    def get_critical_density_for_gap(self, gap: float) -> float:
        p = self.params
        if gap < p.hard_short_gap:
            return 0.0
        return min(p.base_critical_density + p.gap_coefficient * gap, p.max_critical_density)

    def get_lambda_for_gap(self, gap: float) -> float:
        """ln2 / (a g^2 + b g + c) — the per-µs ignition probability (ignition.py:348-364)."""
        p = self.params
        return 0.6931471805599453 / (p.ignition_a_coeff * gap ** 2 + p.ignition_b_coeff * gap + p.ignition_c_coeff)

    def get_short_circuit_status(self) -> dict:
        r, d = self.random_short_remaining, self.debris_short_remaining
        return {"has_random_short": r > 0, "random_short_remaining_us": r, "has_debris_short": d > 0,
                "debris_short_remaining_us": d, "total_short_remaining_us": max(r, d)}


class WireModule(EDMModule):
    def __init__(self, env, parameters: WireModuleParameters = None):
        super().__init__(env)
        self.params = parameters or WireModuleParameters()
        self.wire_material = get_material_db().get_wire_material(env.config.wire_material)
        g = wire_geometry(env.config, self.params)
        self.n_segments = g["n_segments"]
        self.zone_start, self.zone_end = g["zone_start"], g["zone_end"]
        self.actual_zone_start, self.actual_zone_end = g["actual_zone_start"], g["actual_zone_end"]
        self.contact_bottom_idx, self.contact_top_idx = g["contact_bottom_idx"], g["contact_top_idx"]
        self.total_L = self.params.buffer_len_bottom + env.config.workpiece_height + self.params.buffer_len_top
        self.r_wire = env.config.wire_diameter / 2.0
        self.critical_temperature = self.wire_material.melting_point * self.params.critical_temp_threshold
        self.breaking_temperature = self.wire_material.breaking_temperature

    def compute_zone_mean_temperature(self, temperature_field) -> float:
        import numpy as np

        a, b = self.actual_zone_start, self.actual_zone_end
        t = np.asarray(temperature_field)
        return float(np.mean(t[a:b])) if b > a and b <= len(t) else float(np.mean(t))


class MaterialRemovalModule(EDMModule):
    def __init__(self, env, parameters: MaterialModuleParameters = None):
        super().__init__(env)
        self.params = parameters or MaterialModuleParameters()
        self.crater_data = load_craters()
        self.currents_data = load_currents()

    def get_crater_statistics(self) -> dict:
        """Count and mean of the crater volumes sampled so far (the reference keeps the full
        list, material.py:51; the kernel keeps count and sum)."""
        n = self._field("stat_sparks")
        s = self._field("stat_crater_sum")
        return {"total_craters": n, "mean_volume_um3": (s / n if n else 0), "sum_volume_um3": s}


class DielectricModule(EDMModule):
    def __init__(self, env, parameters: DielectricModuleParameters = None):
        super().__init__(env)
        self.params = parameters or DielectricModuleParameters()

    def get_debris_statistics(self) -> dict:
        d = self._field("debris_density")
        return {"debris_volume_mm3": self._field("debris_volume"), "debris_density": d,
                "cavity_volume_mm3": self._field("cavity_volume"), "flow_condition": self._field("flow"),
                "debris_fill_percentage": d * 100.0}


class MechanicsModule(EDMModule):
    def __init__(self, env, control_mode: str = "position", parameters: MechanicsModuleParameters = None):
        super().__init__(env)
        if control_mode not in ["position", "velocity"]:
            raise ValueError(f"control_mode must be 'position' or 'velocity', got {control_mode}")
        self.control_mode = control_mode
        self.params = parameters or MechanicsModuleParameters()

    @property
    def prev_accel(self) -> float:
        return self._field("prev_accel")


__all__ = ["EDMModule", "IgnitionModule", "IgnitionModuleParameters", "WireModule", "WireModuleParameters",
           "MaterialRemovalModule", "MaterialModuleParameters", "DielectricModule", "DielectricModuleParameters",
           "MechanicsModule", "MechanicsModuleParameters"]
