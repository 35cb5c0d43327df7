"""`WireEDMEnv`: drop-in for the reference Gymnasium environment (envs/wire_edm.py:17-200).

Same constructor, `reset`, `step`, action space, `info` keys, public attributes and error
behaviour.  One `step(action)` still means ONE microsecond, exactly like the reference — it is a
1-env, 1-µs launch of the batched engine, meant for compatibility and tests.  The fast path is
`step_control(action)` (one launch for a whole control step) or `BatchedWireEDMEnv`.

Randomness: the reference draws from `env.np_random` (NumPy PCG64); here draws come from
on-device Philox4x32-10 streams keyed by the reset seed.  Trajectories are therefore
statistically, not bitwise, equivalent to the reference's unless draws are injected
(`BatchedWireEDMEnv.step(..., draws=...)`).
"""
from __future__ import annotations

import numpy as np

from .. import _native as N
from ..engine import Engine
from ..modules import (DielectricModule, IgnitionModule, MaterialRemovalModule, MechanicsModule, WireModule)
from ..params import (DielectricModuleParameters, EnvironmentConfig, IgnitionModuleParameters,
                      MaterialModuleParameters, MechanicsModuleParameters, WireModuleParameters, derive_params)
from ..spaces import Box, Dict, EnvBase
from ..state import DeviceStateView


class WireEDMEnv(EnvBase):
    """Main-cut Wire-EDM environment (1 µs base step, 1 ms control step)."""

    metadata = {"render_modes": ["human"], "render_fps": 300}

    def __init__(self, *, render_mode: str | None = None, mechanics_control_mode: str = "position",
                 config: EnvironmentConfig = None, ignition_params: IgnitionModuleParameters = None,
                 wire_params: WireModuleParameters = None, material_params: MaterialModuleParameters = None,
                 dielectric_params: DielectricModuleParameters = None,
                 mechanics_params: MechanicsModuleParameters = None, device="cuda:0"):
        super().__init__()
        self.render_mode = render_mode
        if mechanics_control_mode not in ["position", "velocity"]:
            raise ValueError(f"mechanics_control_mode must be 'position' or 'velocity', got {mechanics_control_mode}")
        self.mechanics_control_mode = mechanics_control_mode
        self.config = config or EnvironmentConfig()
        self.config.validate()
        self.dt = self.config.dt
        self.servo_interval = self.config.servo_interval
        self.np_random = np.random.default_rng()

        # module objects: parameter carriers + read-back of module-internal state
        self._index = 0
        self.ignition = IgnitionModule(self, ignition_params)
        self.wire = WireModule(self, wire_params)
        self.material = MaterialRemovalModule(self, material_params)
        self.dielectric = DielectricModule(self, dielectric_params)
        self.mechanics = MechanicsModule(self, control_mode=mechanics_control_mode, parameters=mechanics_params)
        self.modules = {"ignition": self.ignition, "material": self.material, "dielectric": self.dielectric,
                        "wire": self.wire, "mechanics": self.mechanics}

        self._params = derive_params(self.config, mechanics_control_mode, self.ignition.params, self.wire.params,
                                     self.material.params, self.dielectric.params, self.mechanics.params)
        self._engine = Engine(1, self._params, device=device, seed=int(self.np_random.integers(0, 2**63 - 1)))
        self.state = DeviceStateView(self._engine, 0, self.dielectric.params.base_flow_rate,
                                     self.dielectric.params.dielectric_temperature)

        self.action_space = Dict({
            "servo": Box(low=-1.0, high=1.0, shape=(1,), dtype=np.float32),
            "generator_control": Dict({
                "target_voltage": Box(0.0, 200.0, (1,), np.float32),
                "current_mode": Box(1, 19, (1,), dtype=np.int32),
                "ON_time": Box(0.0, 5.0, (1,), np.float32),
                "OFF_time": Box(0.0, 100.0, (1,), np.float32),
            }),
        })
        self.observation_space = Dict({})

    # ------------------------------------------------------------------ Gym API
    def reset(self, *, seed: int | None = None, options=None):
        super().reset(seed=seed)
        if seed is not None:
            self.np_random = np.random.default_rng(seed)
            self._engine.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self._engine.reset(None, reseed=seed is not None)
        return self._get_obs(), {}

    def _stage(self, action):
        gc = action["generator_control"]
        self._engine.set_actions(np.asarray(action["servo"]).reshape(-1)[:1], np.asarray(gc["target_voltage"]).reshape(-1)[:1],
                                 np.asarray(gc["current_mode"]).reshape(-1)[:1], np.asarray(gc["ON_time"]).reshape(-1)[:1],
                                 np.asarray(gc["OFF_time"]).reshape(-1)[:1])

    def _readback(self):
        e = self._engine
        import torch

        v = torch.stack([e["flags"][0], e["time"][0], e["spark_state"][0], e["t_servo"][0]]).cpu().tolist()
        return int(v[0]), int(v[1]), int(v[2]), int(v[3])

    def _raise_mode_error(self):
        m = int(self._engine["current_mode"][0].item())
        avail = list(self.material.crater_data.keys())
        raise ValueError(f"Current mode I{max(m, 0)} is not available in crater data. Available modes: {avail}")

    def step(self, action):
        """ONE microsecond (wire_edm.py:116-157)."""
        flags0, _, _, t_servo0 = self._readback()
        is_ctrl_step = t_servo0 >= self.servo_interval
        broken_before = bool(flags0 & N.F_BROKEN)       # wire.update early-outs, step returns at :129-130
        self._stage(action)
        self._engine.step(1)
        flags, time, spark_state, _ = self._readback()
        if flags & N.F_MODE_ERROR:
            self._raise_mode_error()
        if flags & N.F_THERMAL_BREAK or broken_before:
            return None, 0.0, True, False, {"wire_broken": True}
        terminated = bool(flags & N.F_DONE)
        obs = self._get_obs() if is_ctrl_step else None
        reward = self._calc_reward() if is_ctrl_step else 0.0
        info = {"wire_broken": bool(flags & N.F_BROKEN), "target_reached": bool(flags & N.F_TARGET),
                "spark_state": spark_state, "time": time, "control_step": is_ctrl_step}
        return obs, reward, terminated, False, info

    def step_control(self, action, n_us: int | None = None):
        """`n_us` (default servo_interval) microseconds in ONE kernel launch — the loop
        `for _ in range(1000): env.step(action)` of the reference's drivers (README.md:101-106)."""
        n_us = self.servo_interval if n_us is None else int(n_us)
        self._stage(action)
        self._engine.step(n_us)
        flags, time, spark_state, _ = self._readback()
        if flags & N.F_MODE_ERROR:
            self._raise_mode_error()
        info = {"wire_broken": bool(flags & N.F_BROKEN), "target_reached": bool(flags & N.F_TARGET),
                "spark_state": spark_state, "time": time, "control_step": True}
        return self._get_obs(), self._calc_reward(), bool(flags & N.F_DONE), False, info

    # ------------------------------------------------------------------ helpers
    def _get_obs(self):
        return {}          # placeholder in the reference too (wire_edm.py:181-183)

    def _calc_reward(self):
        return 0.0         # wire_edm.py:185-187

    @property
    def workpiece_height(self) -> float:
        return self.config.workpiece_height

    @property
    def wire_diameter(self) -> float:
        return self.config.wire_diameter
