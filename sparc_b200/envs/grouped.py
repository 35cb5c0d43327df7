"""`GroupedWireEDMEnv`: heterogeneous batches as groups of homogeneous ones.

The fused kernel takes its parameters as one grid constant, so envs that share an
`EnvironmentConfig` / module parameters / control mode form a group with its own
`BatchedWireEDMEnv` (own SoA blob, own kernel variant for its segment count).  One `step(actions)`
slices the `[n_total]` action leaves by group, launches every group on its own CUDA stream so small
groups overlap on the GPU, and joins the streams before it concatenates the results.  Global env ids
run through the groups in order, so each env's Philox stream is the one it would have in a single
homogeneous batch with the same `seed` / `env_id_base`.

This covers domain randomisation over a finite set of configurations (heights, wire diameters,
materials, control modes).  A different configuration for every env degenerates into one-env groups;
per-env parameter arrays inside the kernel are the (not yet built) answer to that.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .batched import OBS_KEYS, BatchedWireEDMEnv

_LEAVES = ("target_voltage", "current_mode", "ON_time", "OFF_time")


class GroupedWireEDMEnv:
    """groups: sequence of dicts, each `{"n_envs": int, **BatchedWireEDMEnv keyword arguments}`
    (config, mechanics_control_mode, *_params, latch_first_action, auto_reset)."""

    def __init__(self, groups: Sequence[Dict], *, device="cuda:0", seed: int = 0, env_id_base: int = 0,
                 concurrent: bool = True):
        if not groups:
            raise ValueError("GroupedWireEDMEnv needs at least one group")
        self.envs: List[BatchedWireEDMEnv] = []
        self.offsets: List[int] = [0]
        for g in groups:
            kw = dict(g)
            n = int(kw.pop("n_envs"))
            self.envs.append(BatchedWireEDMEnv(n, device=device, seed=seed, env_id_base=env_id_base + self.offsets[-1],
                                               **kw))
            self.offsets.append(self.offsets[-1] + n)
        self.n_envs = self.offsets[-1]
        self.device = self.envs[0].device
        self.concurrent = bool(concurrent) and len(self.envs) > 1
        self.streams = [torch.cuda.Stream(self.device) for _ in self.envs] if self.concurrent else []
        self.servo_interval = self.envs[0].servo_interval
        if any(e.servo_interval != self.servo_interval for e in self.envs):
            raise ValueError("all groups must share servo_interval (one step = one control step for every env)")

    # ------------------------------------------------------------------------------------------
    def group_slices(self) -> List[slice]:
        return [slice(a, b) for a, b in zip(self.offsets[:-1], self.offsets[1:])]

    def group_of(self, env_index: int) -> int:
        return int(np.searchsorted(self.offsets, env_index, side="right") - 1)

    def _leaf(self, x, sl: slice):
        if isinstance(x, torch.Tensor):
            return x.reshape(-1)[sl] if x.numel() > 1 else x
        a = np.asarray(x)
        return a.reshape(-1)[sl] if a.size > 1 else a

    def _slice_actions(self, actions: Dict, sl: slice) -> Dict:
        gc = actions["generator_control"]
        return {"servo": self._leaf(actions["servo"], sl),
                "generator_control": {k: self._leaf(gc[k], sl) for k in _LEAVES}}

    def _fan_out(self, fn):
        """Run fn(group_index, env) for every group, each on its own stream, and join."""
        if not self.concurrent:
            return [fn(i, e) for i, e in enumerate(self.envs)]
        cur = torch.cuda.current_stream(self.device)
        outs = []
        for i, (e, st) in enumerate(zip(self.envs, self.streams)):
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                outs.append(fn(i, e))
        for st in self.streams:
            cur.wait_stream(st)
        return outs

    def reset(self, *, seed: Optional[int] = None, mask: Optional[torch.Tensor] = None, options=None):
        sls = self.group_slices()
        if mask is not None:
            mask = torch.as_tensor(mask, device=self.device).bool()
        outs = self._fan_out(lambda i, e: e.reset(seed=seed, mask=None if mask is None else mask[sls[i]]))
        return {k: torch.cat([o[0][k] for o in outs]) for k in OBS_KEYS}, {}

    def step(self, actions: Dict, n_us: Optional[int] = None):
        """One control step of every group; returns the batched 5-tuple of `BatchedWireEDMEnv.step`
        over all `n_envs` envs in group order."""
        sls = self.group_slices()
        outs = self._fan_out(lambda i, e: e.step(self._slice_actions(actions, sls[i]), n_us))
        obs = {k: torch.cat([o[0][k] for o in outs]) for k in OBS_KEYS}
        reward = torch.cat([o[1] for o in outs])
        terminated = torch.cat([o[2] for o in outs])
        truncated = torch.cat([o[3] for o in outs])
        info = {k: torch.cat([o[4][k] for o in outs])
                for k in ("wire_broken", "target_reached", "mode_error", "spark_state", "time")}
        info["control_step"] = True
        if "final_terminated" in outs[0][4]:
            info["final_terminated"] = torch.cat([o[4]["final_terminated"] for o in outs])
        return obs, reward, terminated, truncated, info

    def state_tensor(self, name: str) -> torch.Tensor:
        """Concatenated copy of one scalar state field over all groups (group order)."""
        return torch.cat([e.state_tensor(name) for e in self.envs])

    def check_mode_errors(self) -> None:
        for e in self.envs:
            e.check_mode_errors()

    def episode_stats(self, all_reduce: bool = True) -> Dict[str, float]:
        tot: Dict[str, float] = {}
        for e in self.envs:
            for k, v in e.episode_stats(all_reduce).items():
                tot[k] = tot.get(k, 0.0) + v
        return tot
