"""`WireEDMVectorEnv`: Gymnasium `VectorEnv`-shaped adapter over `BatchedWireEDMEnv`.

The reference has no vector surface, and leaves observation and reward as placeholders
(`wire_edm.py:181-187`: `_get_obs` returns `{}`, `_calculate_reward` returns 0.0).  This adapter
gives RL code the step either side of the hot path:

* `num_envs`, `single_action_space` / `action_space`, `single_observation_space` /
  `observation_space`, `reset(seed=, options=)`, `step(actions)`, `close()` with Gymnasium 1.x
  semantics, one `step` = one control step (`servo_interval` µs of the fused kernel);
* a real observation vector `[gap, wire_velocity, mean_voltage, sparks, short_us,
  max_temperature, debris_density]` (float32[7], produced by `sparc_observe` on the device);
* rewards: `"zero"` (the reference's placeholder, default) or `"progress"` (µm of workpiece
  advance in the control step minus `break_penalty` when the wire breaks);
* auto-reset in Gymnasium's `"next_step"` (default), `"same_step"` or `"disabled"` modes and
  `max_episode_steps` truncation, all as masked device operations;
* `to_numpy=True` returns numpy arrays (one D2H per step); `False` keeps torch device tensors.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from .. import _native as N
from ..spaces import Box, Dict as DictSpace
from .batched import OBS_KEYS, BatchedWireEDMEnv

AUTORESET_MODES = ("next_step", "same_step", "disabled")


def _single_action_space() -> DictSpace:
    # the reference's action space (wire_edm.py:84-98)
    return DictSpace({
        "servo": Box(low=-1.0, high=1.0, shape=(1,), dtype=np.float32),
        "generator_control": DictSpace({
            "target_voltage": Box(low=0.0, high=200.0, shape=(1,), dtype=np.float32),
            "current_mode": Box(low=1, high=19, shape=(1,), dtype=np.int32),
            "ON_time": Box(low=0.0, high=5.0, shape=(1,), dtype=np.float32),
            "OFF_time": Box(low=0.0, high=100.0, shape=(1,), dtype=np.float32),
        }),
    })


def _batched_action_space(n: int) -> DictSpace:
    return DictSpace({
        "servo": Box(low=-1.0, high=1.0, shape=(n,), dtype=np.float32),
        "generator_control": DictSpace({
            "target_voltage": Box(low=0.0, high=200.0, shape=(n,), dtype=np.float32),
            "current_mode": Box(low=1, high=19, shape=(n,), dtype=np.int32),
            "ON_time": Box(low=0.0, high=5.0, shape=(n,), dtype=np.float32),
            "OFF_time": Box(low=0.0, high=100.0, shape=(n,), dtype=np.float32),
        }),
    })


class WireEDMVectorEnv:
    metadata = {"render_modes": [], "autoreset_mode": "next_step"}
    render_mode = None
    spec = None

    def __init__(self, num_envs: int, *, device="cuda:0", reward: str = "zero", break_penalty: float = 100.0,
                 autoreset_mode: str = "next_step", max_episode_steps: Optional[int] = None,
                 to_numpy: bool = True, **env_kwargs):
        if autoreset_mode not in AUTORESET_MODES:
            raise ValueError(f"autoreset_mode must be one of {AUTORESET_MODES}, got {autoreset_mode}")
        if reward not in ("zero", "progress"):
            raise ValueError(f"reward must be 'zero' or 'progress', got {reward}")
        env_kwargs.pop("auto_reset", None)
        self.env = BatchedWireEDMEnv(num_envs, device=device, auto_reset=False, **env_kwargs)
        self.num_envs = int(num_envs)
        self.reward_kind = reward
        self.break_penalty = float(break_penalty)
        self.autoreset_mode = autoreset_mode
        self.metadata = dict(self.metadata, autoreset_mode=autoreset_mode)
        self.max_episode_steps = max_episode_steps
        self.to_numpy = bool(to_numpy)
        self.closed = False
        n_obs = len(OBS_KEYS)
        self.single_action_space = _single_action_space()
        self.action_space = _batched_action_space(self.num_envs)
        self.single_observation_space = Box(low=-np.inf, high=np.inf, shape=(n_obs,), dtype=np.float32)
        self.observation_space = Box(low=-np.inf, high=np.inf, shape=(self.num_envs, n_obs), dtype=np.float32)
        dev = self.env.device
        self._elapsed = torch.zeros(self.num_envs, dtype=torch.int32, device=dev)
        self._pending = torch.zeros(self.num_envs, dtype=torch.bool, device=dev)   # next_step: reset due
        self._wp_prev = self.env.state_tensor("wp").clone()

    # ------------------------------------------------------------------------------------------
    @property
    def device(self):
        return self.env.device

    @property
    def unwrapped(self):
        return self.env

    def _obs_matrix(self) -> torch.Tensor:
        return self.env.engine.observe()[: len(OBS_KEYS)].t().contiguous()      # [num_envs, 7]

    def _out(self, t: torch.Tensor):
        return t.cpu().numpy() if self.to_numpy else t

    def reset(self, *, seed: Optional[int] = None, options: Optional[Dict] = None):
        """Reset every env, or only `options["reset_mask"]` (bool[num_envs]) as in Gymnasium 1.x."""
        mask = None
        if options and options.get("reset_mask") is not None:
            mask = torch.as_tensor(options["reset_mask"], device=self.device).bool()
        self.env.reset(seed=seed, mask=mask)
        if mask is None:
            self._elapsed.zero_()
            self._pending.zero_()
        else:
            self._elapsed[mask] = 0
            self._pending[mask] = False
        self._wp_prev = self.env.state_tensor("wp").clone()
        return self._out(self._obs_matrix()), {}

    def _stage(self, actions: Dict):
        gc = actions["generator_control"]

        def leaf(x):
            if isinstance(x, torch.Tensor):
                return x.reshape(-1)
            a = np.asarray(x)
            return a.reshape(-1) if a.size > 1 else a.reshape(())
        return {"servo": leaf(actions["servo"]),
                "generator_control": {k: leaf(gc[k]) for k in ("target_voltage", "current_mode", "ON_time", "OFF_time")}}

    def step(self, actions: Dict):
        env = self.env
        pending = self._pending
        env.step(self._stage(actions))
        if self.autoreset_mode == "next_step" and bool(pending.any()):
            # Gymnasium NEXT_STEP: the call after a terminal one ignores the action of those envs,
            # resets them and returns the reset observation with reward 0 (a terminated env is
            # frozen inside the kernel; whatever a truncated one did in this launch is discarded)
            env.reset(mask=pending)
            self._elapsed[pending] = -1
            self._wp_prev = torch.where(pending, env.state_tensor("wp"), self._wp_prev)
        flags = env.state_tensor("flags")
        wp = env.state_tensor("wp")
        terminated = (flags & N.F_DONE) != 0
        broken = (flags & N.F_BROKEN) != 0
        if self.reward_kind == "progress":
            reward = (wp - self._wp_prev).to(torch.float32) - self.break_penalty * broken.to(torch.float32)
        else:
            reward = torch.zeros(self.num_envs, dtype=torch.float32, device=self.device)
        self._wp_prev = wp.clone()
        self._elapsed += 1
        truncated = torch.zeros_like(terminated)
        if self.max_episode_steps is not None:
            truncated = (self._elapsed >= int(self.max_episode_steps)) & ~terminated
        done = terminated | truncated
        obs = self._obs_matrix()
        info: Dict = {"wire_broken": broken, "target_reached": (flags & N.F_TARGET) != 0,
                      "mode_error": (flags & N.F_MODE_ERROR) != 0, "time": env.state_tensor("time").clone()}
        if self.autoreset_mode == "same_step" and bool(done.any()):
            info["final_obs"] = obs.clone()
            info["_final_obs"] = done.clone()
            env.reset(mask=done)
            self._elapsed[done] = 0
            self._wp_prev = env.state_tensor("wp").clone()
            obs = self._obs_matrix()
        elif self.autoreset_mode == "next_step":
            self._pending = done.clone()
        if self.to_numpy:
            info = {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in info.items()}
        return self._out(obs), self._out(reward), self._out(terminated), self._out(truncated), info

    def close(self, **kwargs):
        self.closed = True

    def episode_stats(self, all_reduce: bool = True):
        return self.env.episode_stats(all_reduce)
