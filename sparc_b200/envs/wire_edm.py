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
                 mechanics_params: MechanicsModuleParameters = None, device="cuda:0",
                 strict_reference_reset: bool = False):
        super().__init__()
        self.render_mode = render_mode
        # True: reset() keeps module-internal state like the reference does; default: a fresh env
        self.strict_reference_reset = bool(strict_reference_reset)
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
        self._bind()

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
        self._engine.reset(None, reseed=seed is not None, reference_semantics=self.strict_reference_reset)
        if self.strict_reference_reset and seed is not None:
            self._engine["episode"].zero_()
        self._bind()
        return self._get_obs(), {}

    def _bind(self):
        """(Re)build the per-call fast path: numpy views of the pinned action / observation buffers and
        the prebuilt C-ABI calls.  The reference's step() costs ~25 µs of host time per simulated µs
        (wire_edm.py:116-157); a 1-µs launch can only compete if the Python side adds almost nothing:
        five scalar writes into pinned memory, ONE `sparc_step_host` (one H2D copy, one kernel — a
        launch this short skips the mechanics pre-pass kernel —, one D2H copy of the observation row,
        one stream sync), four scalar reads."""
        e = self._engine
        ha = e.host_actions
        self._h_servo, self._h_tv = ha["servo"].numpy(), ha["target_voltage"].numpy()
        self._h_mode, self._h_on, self._h_off = ha["current_mode"].numpy(), ha["on_time"].numpy(), ha["off_time"].numpy()
        self._obs_f = e.obs_host.numpy()
        self._obs_i = self._obs_f.view(np.int32)
        self._step1 = e.host_stepper(1)
        self._stepN = e.host_stepper(self.servo_interval)
        self._flags, self._t_servo = 0, 0           # fresh EDMState: no flags, time_since_servo = 0

    @staticmethod
    def _first(x):
        try:
            return x[0]
        except (TypeError, IndexError):
            return x

    def _stage(self, action):
        gc = action["generator_control"]
        f = self._first
        self._h_servo[0] = f(action["servo"])
        self._h_tv[0] = f(gc["target_voltage"])
        self._h_mode[0] = f(gc["current_mode"])
        self._h_on[0] = f(gc["ON_time"])
        self._h_off[0] = f(gc["OFF_time"])

    def _raise_mode_error(self):
        m = int(self._engine["current_mode"][0].item())
        avail = list(self.material.crater_data.keys())
        raise ValueError(f"Current mode I{max(m, 0)} is not available in crater data. Available modes: {avail}")

    def step(self, action):
        """ONE microsecond (wire_edm.py:116-157)."""
        is_ctrl_step = self._t_servo >= self.servo_interval
        broken_before = self._flags & N.F_BROKEN        # wire.update early-outs, step returns at :129-130
        self._stage(action)
        self._step1()
        flags = int(self._obs_f[7, 0])
        self._flags, self._t_servo = flags, int(self._obs_i[10, 0])
        if flags & N.F_MODE_ERROR:
            self._raise_mode_error()
        if flags & N.F_THERMAL_BREAK or broken_before:
            return None, 0.0, True, False, {"wire_broken": True}
        terminated = bool(flags & N.F_DONE)
        obs = self._get_obs() if is_ctrl_step else None
        reward = self._calc_reward() if is_ctrl_step else 0.0
        info = {"wire_broken": bool(flags & N.F_BROKEN), "target_reached": bool(flags & N.F_TARGET),
                "spark_state": int(self._obs_i[9, 0]), "time": int(self._obs_i[8, 0]), "control_step": is_ctrl_step}
        return obs, reward, terminated, False, info

    def step_control(self, action, n_us: int | None = None):
        """`n_us` (default servo_interval) microseconds in ONE call — the loop
        `for _ in range(1000): env.step(action)` of the reference's drivers (README.md:101-106)."""
        self._stage(action)
        if n_us is None or int(n_us) == self.servo_interval:
            self._stepN()
        else:
            self._engine.step_host(int(n_us))
        flags = int(self._obs_f[7, 0])
        self._flags, self._t_servo = flags, int(self._obs_i[10, 0])
        if flags & N.F_MODE_ERROR:
            self._raise_mode_error()
        info = {"wire_broken": bool(flags & N.F_BROKEN), "target_reached": bool(flags & N.F_TARGET),
                "spark_state": int(self._obs_i[9, 0]), "time": int(self._obs_i[8, 0]), "control_step": True}
        return self._get_obs(), self._calc_reward(), bool(flags & N.F_DONE), False, info

    # ------------------------------------------------------------------ helpers
    @property
    def base_flow_rate(self) -> float:
        return float(self.dielectric.params.base_flow_rate)

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
