"""`BatchedWireEDMEnv`: the vectorised N-environment surface (new; the reference has none).

One `step(actions)` = one control step: the fused kernel advances every environment through
`servo_interval` µs (1000 reference `WireEDMEnv.step()` calls per env) in one launch.  The
action is latched inside the kernel exactly when the reference latches it
(time_since_servo >= servo_interval, wire_edm.py:117-121), so with the reference's
"first millisecond" quirk the very first control step still runs on the generator defaults
unless `latch_first_action=True` (default) arms the latch at reset.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from .. import _native as N
from ..controllers import GapController
from ..engine import Engine
from ..params import (DielectricModuleParameters, EnvironmentConfig, IgnitionModuleParameters,
                      MaterialModuleParameters, MechanicsModuleParameters, WireModuleParameters,
                      derive_params, load_craters)

OBS_KEYS = ("gap", "wire_velocity", "mean_voltage", "sparks", "short_us", "max_temperature", "debris_density")


class BatchedWireEDMEnv:
    """N independent Wire-EDM environments on one GPU.

    actions: dict with the reference's structure (wire_edm.py:84-98); every leaf is an array /
    tensor of length n_envs (or a scalar, broadcast):
        {"servo": ..., "generator_control": {"target_voltage": ..., "current_mode": ...,
                                              "ON_time": ..., "OFF_time": ...}}
    """

    def __init__(self, n_envs: int, *, device="cuda:0", mechanics_control_mode: str = "position",
                 config: EnvironmentConfig = None, ignition_params: IgnitionModuleParameters = None,
                 wire_params: WireModuleParameters = None, material_params: MaterialModuleParameters = None,
                 dielectric_params: DielectricModuleParameters = None,
                 mechanics_params: MechanicsModuleParameters = None, seed: int = 0, env_id_base: int = 0,
                 latch_first_action: bool = True, auto_reset: bool = False, strict_reference_reset: bool = False):
        if mechanics_control_mode not in ["position", "velocity"]:
            raise ValueError(f"mechanics_control_mode must be 'position' or 'velocity', got {mechanics_control_mode}")
        self.config = config or EnvironmentConfig()
        self.config.validate()
        self.mechanics_control_mode = mechanics_control_mode
        self.n_envs = int(n_envs)
        self.servo_interval = int(self.config.servo_interval)
        self.dt = self.config.dt
        self.latch_first_action = bool(latch_first_action)
        self.auto_reset = bool(auto_reset)
        # False (default): reset() gives the state of a freshly CONSTRUCTED env.  True: the reference's own reset(), which
        # re-creates only EDMState and lets module-internal state survive into the next episode
        self.strict_reference_reset = bool(strict_reference_reset)
        self._module_params = dict(ignition=ignition_params, wire=wire_params, material=material_params,
                                   dielectric=dielectric_params, mechanics=mechanics_params)
        self.params = derive_params(self.config, mechanics_control_mode, ignition_params, wire_params,
                                    material_params, dielectric_params, mechanics_params)
        self.base_flow_rate = float((dielectric_params or DielectricModuleParameters()).base_flow_rate)
        self.engine = Engine(self.n_envs, self.params, device=device, seed=seed, env_id_base=env_id_base)
        self._valid_modes = sorted(int(k[1:]) for k in load_craters())
        self._arm_latch(None)

    # ------------------------------------------------------------------------------------------
    @property
    def device(self):
        return self.engine.device

    @property
    def n_segments(self) -> int:
        return self.engine.n_segments

    def _arm_latch(self, mask: Optional[torch.Tensor]):
        if not self.latch_first_action:
            return
        ts = self.engine["t_servo"]
        if mask is None:
            ts.fill_(self.servo_interval)
        else:
            ts[mask.to(self.device).bool()] = self.servo_interval

    def reset(self, *, seed: Optional[int] = None, mask: Optional[torch.Tensor] = None, options=None):
        """Fresh-env state for all envs (or the masked ones).  A seed re-keys the Philox streams."""
        if seed is not None:
            self.engine.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self._reset_engine(mask, seed is not None)
        self._arm_latch(mask)
        return self._obs(), {}

    # ------------------------------------------------------------------------------------------
    def per_env_parameter_names(self):
        """Names of the parameters that exist once per env in the state blob (fields `prm_<name>`)."""
        return [k[4:] for k in self.engine.fields if k.startswith("prm_")]

    def set_env_params(self, env_ids, *, config: EnvironmentConfig = None,
                       ignition_params: IgnitionModuleParameters = None, wire_params: WireModuleParameters = None,
                       dielectric_params: DielectricModuleParameters = None,
                       mechanics_params: MechanicsModuleParameters = None) -> None:
        """Give the envs `env_ids` their own configuration (domain randomisation: a different
        EnvironmentConfig / *ModuleParameters for every env of ONE batch, ONE launch; reference
        core/env_config.py:10-35, modules/ignition.py:17-57, dielectric.py:15-31, mechanics.py:12-23).

        Every scalar that does not change the wire geometry is a per-env array in the state blob
        (`per_env_parameter_names()`); arguments left out keep this env's construction-time value.  A
        configuration that would change the geometry or a shared table (workpiece height, wire diameter,
        segment length, buffers, servo interval, material, crater / kerf tables, control mode) raises
        ValueError — batches of different geometry are `GroupedWireEDMEnv`'s job.  `initial_gap` takes
        effect at the env's next reset."""
        import ctypes as C

        mp = self._module_params
        p = derive_params(config or self.config, self.mechanics_control_mode, ignition_params or mp["ignition"],
                          wire_params or mp["wire"], mp["material"], dielectric_params or mp["dielectric"],
                          mechanics_params or mp["mechanics"])
        per_env = set(self.per_env_parameter_names())
        for name, ctype in p._fields_:
            if name in per_env or name.startswith("reserved"):
                continue
            a, b = getattr(p, name), getattr(self.params, name)
            if isinstance(a, C.Array):
                a, b = list(a), list(b)
            if a != b:
                raise ValueError(f"set_env_params: '{name}' differs from the batch ({a} vs {b}); only parameters that "
                                 "leave the wire geometry and the shared tables unchanged can vary inside one batch")
        idx = torch.as_tensor(env_ids, device=self.device, dtype=torch.long).reshape(-1)
        for name in per_env:
            self.engine["prm_" + name][idx] = getattr(p, name)

    def _reset_engine(self, mask, reseed: bool):
        self.engine.reset(mask, reseed=reseed, reference_semantics=self.strict_reference_reset)
        if self.strict_reference_reset and reseed:          # a seeded reset replays the same Philox episode
            ep = self.engine["episode"]
            if mask is None:
                ep.zero_()
            else:
                ep[mask.to(self.device).bool()] = 0

    def _stage(self, actions: Dict):
        gc = actions["generator_control"]
        self.engine.set_actions(actions["servo"], gc["target_voltage"], gc["current_mode"], gc["ON_time"], gc["OFF_time"])

    def _obs(self) -> Dict[str, torch.Tensor]:
        o = self.engine.observe()
        return {k: o[i] for i, k in enumerate(OBS_KEYS)}

    def step(self, actions: Dict, n_us: Optional[int] = None, draws: Optional[torch.Tensor] = None,
             log_temperature_every: Optional[int] = None, n_log_envs: Optional[int] = None,
             log_signals_every: Optional[int] = None):
        """Advance every env by `n_us` µs (default: one control step).  Returns the Gymnasium
        5-tuple with batched tensors: obs dict of [n_envs] float32, reward [n_envs] (0.0 — the
        reference's reward is a placeholder, wire_edm.py:185-187), terminated, truncated, info.

        `log_signals_every=k` runs the device-side signal logger: `info["signal_log"]` maps each
        name of `_native.SIGNALS` to a [n_us // k, n_envs] float64 tensor (k=1: the reference
        logger's "every_step")."""
        n_us = self.servo_interval if n_us is None else int(n_us)
        if isinstance(actions, GapController) and (log_signals_every or log_temperature_every or n_us > 1024):
            actions = actions(self)             # the logging entry points take the command from the staged array
        if isinstance(actions, GapController):
            # closed loop inside the kernels: servo = clip(gain * (gap - desired), +-limit) from the state this
            # call starts with (same arithmetic as `actions(env)`), no torch op between control steps
            gc = actions.generator_control(self.n_envs, self.device)
            self.engine.set_actions(0.0, gc["target_voltage"], gc["current_mode"], gc["ON_time"], gc["OFF_time"])
            self.engine.set_controller(**actions.kernel_spec())
        else:
            self.engine.set_controller(0)
            self._stage(actions)
        tlog, sig, fused = None, None, None
        if log_signals_every:
            sig, tlog = self.engine.step_sampled(n_us, int(log_signals_every), draws,
                                                 int(log_temperature_every or 0), n_log_envs)
        elif log_temperature_every:
            tlog = self.engine.step_logged(n_us, int(log_temperature_every), n_log_envs, draws)
        else:
            fused = self.engine.step(n_us, draws, observe=True)     # observation from the kernel epilogue
        obs = self._obs() if fused is None else {k: fused[i] for i, k in enumerate(OBS_KEYS)}
        flags = self.engine["flags"]
        terminated = (flags & N.F_DONE) != 0
        info = {
            "wire_broken": (flags & N.F_BROKEN) != 0,
            "target_reached": (flags & N.F_TARGET) != 0,
            "mode_error": (flags & N.F_MODE_ERROR) != 0,
            "spark_state": self.engine["spark_state"],
            "time": self.engine["time"],
            "control_step": True,
        }
        if tlog is not None:
            info["wire_temperature_log"] = tlog     # [n_log_envs, n_us // every, n_segments]
        if sig is not None:
            info["signal_log"] = {name: sig[:, i] for i, name in enumerate(N.SIGNALS)}
        reward = torch.zeros(self.n_envs, dtype=torch.float32, device=self.device)
        truncated = torch.zeros(self.n_envs, dtype=torch.bool, device=self.device)
        if self.auto_reset:
            info["final_terminated"] = terminated.clone()
            self._reset_engine(terminated, False)
            self._arm_latch(terminated)
        return obs, reward, terminated, truncated, info

    def mean_voltage(self) -> torch.Tensor:
        """float64 [n_envs]: mean gap voltage over the µs of the last step call (the 1 ms voltage
        history the reference's PI controller averages, run_simulation.py:68-72); the instantaneous
        voltage when nothing has been stepped yet."""
        steps = self.engine["win_steps"]
        return torch.where(steps > 0, self.engine["win_vsum"] / steps.clamp(min=1).to(torch.float64),
                           self.engine["voltage"])

    def step_host(self, n_us: Optional[int] = None) -> torch.Tensor:
        """Host-buffer control step through `sparc_step_host`: fill `self.engine.host_actions`
        (pinned) first; returns the pinned [8, n_envs] observation array."""
        return self.engine.step_host(self.servo_interval if n_us is None else int(n_us))

    def check_mode_errors(self) -> None:
        """Raise the reference's ValueError (material.py:108-113) if any env sparked with a current
        mode that has no crater data."""
        bad = ((self.engine["flags"] & N.F_MODE_ERROR) != 0).nonzero().flatten()
        if bad.numel():
            m = int(self.engine["current_mode"][bad[0]].item())
            avail = [f"I{k}" for k in self._valid_modes]
            raise ValueError(f"Current mode I{max(m, 0)} is not available in crater data. Available modes: {avail}")

    def episode_stats(self, all_reduce: bool = True) -> Dict[str, float]:
        """Episode statistics summed over envs; all-reduced over ranks (NCCL) when
        torch.distributed is initialised — the only collective on this path."""
        st = self.engine.reduce_stats().clone()
        if all_reduce and torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.all_reduce(st)
        v = st.cpu().tolist()
        keys = ("sparks", "short_us", "short_events", "crater_um3", "short_pulses", "env_us", "broken", "target_reached")
        return dict(zip(keys, v))

    # convenience views of the live SoA state (zero-copy device tensors)
    @property
    def wire_temperature(self) -> torch.Tensor:
        return self.engine.T

    def state_tensor(self, name: str) -> torch.Tensor:
        return self.engine[name]

    @staticmethod
    def constant_action(servo=0.0, target_voltage=80.0, current_mode=5, ON_time=3.0, OFF_time=80.0) -> Dict:
        return {"servo": np.float32(servo),
                "generator_control": {"target_voltage": np.float32(target_voltage), "current_mode": np.int32(current_mode),
                                      "ON_time": np.float32(ON_time), "OFF_time": np.float32(OFF_time)}}
