"""Batched on-device controllers: what produces the action of each control step.

The reference ships two single-env controllers as closures in its driver script
(`experiments/run_simulation.py:24-109`): a proportional gap controller and a PI controller on the
1 ms mean gap voltage.  Here they are evaluated for all N environments at once with torch ops on the
device the state lives on, so a closed-loop control step is `actions = ctrl(env); env.step(actions)`
with no host round trip.  Arithmetic follows the reference literally: the error is formed in
float64 (Python floats there), scaled, clipped, and only then rounded to the float32 the action
space carries (`np.array([delta], dtype=np.float32)`), so with injected draws the closed-loop
trajectories are the reference's (tests/test_gpu_vector_env.py).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

# generator settings the reference controllers return every step (run_simulation.py:40-52)
REFERENCE_GENERATOR = dict(target_voltage=80.0, current_mode=7, ON_time=2.0, OFF_time=33.0)


def _generator_control(n: int, device, generator: Optional[Dict]) -> Dict[str, torch.Tensor]:
    g = dict(REFERENCE_GENERATOR)
    g.update(generator or {})
    return {
        "target_voltage": torch.full((n,), float(g["target_voltage"]), dtype=torch.float32, device=device),
        "current_mode": torch.full((n,), int(g["current_mode"]), dtype=torch.int32, device=device),
        "ON_time": torch.full((n,), float(g["ON_time"]), dtype=torch.float32, device=device),
        "OFF_time": torch.full((n,), float(g["OFF_time"]), dtype=torch.float32, device=device),
    }


class GapController:
    """`create_gap_controller` (run_simulation.py:24-54) for a batch.

    position mode: servo = 0.1 * (gap - desired) [µm increment]; velocity mode:
    servo = clip(50 * (gap - desired), -1000, 1000) [µm/s].  `gain`/`limit` override the
    reference's constants (BASELINE cfg 3 uses desired_gap=25, limit=1 in position mode)."""

    def __init__(self, desired_gap: float = 5.0, control_mode: str = "position", *, gain: Optional[float] = None,
                 limit: Optional[float] = None, generator: Optional[Dict] = None):
        if control_mode not in ("position", "velocity"):
            raise ValueError(f"control_mode must be 'position' or 'velocity', got {control_mode}")
        self.desired_gap = float(desired_gap)
        self.control_mode = control_mode
        self.gain = float(gain) if gain is not None else (0.1 if control_mode == "position" else 50.0)
        self.limit = float(limit) if limit is not None else (None if control_mode == "position" else 1000.0)
        self.generator = generator
        self._gc = None

    def servo(self, gap: torch.Tensor) -> torch.Tensor:
        """gap: float64 [n] (workpiece_position - wire_position) -> float32 [n] servo command."""
        delta = (gap.to(torch.float64) - self.desired_gap) * self.gain
        if self.limit is not None:
            delta = torch.clamp(delta, -self.limit, self.limit)
        return delta.to(torch.float32)

    def reset(self, mask: Optional[torch.Tensor] = None) -> None:   # stateless
        return None

    def kernel_spec(self) -> Dict:
        """Arguments of `Engine.set_controller`: the same law evaluated inside the kernels at the start of a
        control step (include/sparc_b200.h SparcController) — pass the controller itself to
        `BatchedWireEDMEnv.step(ctrl)` and no torch op runs between control steps."""
        return dict(kind=1, desired_gap=self.desired_gap, gain=self.gain, limit=-1.0 if self.limit is None else self.limit)

    def generator_control(self, n: int, device) -> Dict[str, torch.Tensor]:
        if self._gc is None or self._gc["target_voltage"].numel() != n or self._gc["target_voltage"].device != torch.device(device):
            self._gc = _generator_control(n, device, self.generator)
        return self._gc

    def __call__(self, env) -> Dict:
        gap = env.state_tensor("wp") - env.state_tensor("wire")
        if self._gc is None or self._gc["target_voltage"].numel() != gap.numel() or \
                self._gc["target_voltage"].device != gap.device:
            self._gc = _generator_control(gap.numel(), gap.device, self.generator)
        return {"servo": self.servo(gap), "generator_control": self._gc}


class VoltageController:
    """`create_voltage_controller` (run_simulation.py:57-109) for a batch: PI on the mean gap
    voltage of the last control step (the kernel's per-launch voltage average replaces the
    reference's 1000-sample Python list), integral clipped to ±100, output negated, then clipped to
    ±5 µm (position) or scaled by 100 and clipped to ±1000 µm/s (velocity)."""

    KP = 0.05
    KI = 0.1

    def __init__(self, target_voltage: float = 30.0, control_mode: str = "position", *,
                 generator: Optional[Dict] = None):
        if control_mode not in ("position", "velocity"):
            raise ValueError(f"control_mode must be 'position' or 'velocity', got {control_mode}")
        self.target_voltage = float(target_voltage)
        self.control_mode = control_mode
        self.generator = generator
        self.integral: Optional[torch.Tensor] = None
        self._gc = None

    def reset(self, mask: Optional[torch.Tensor] = None) -> None:
        """Zero the integral state of the masked envs (all when mask is None)."""
        if self.integral is None:
            return
        if mask is None:
            self.integral.zero_()
        else:
            self.integral[mask.to(self.integral.device).bool()] = 0.0

    def servo(self, mean_voltage: torch.Tensor) -> torch.Tensor:
        v = mean_voltage.to(torch.float64)
        if self.integral is None or self.integral.shape != v.shape or self.integral.device != v.device:
            self.integral = torch.zeros_like(v)
        error = self.target_voltage - v
        self.integral = torch.clamp(self.integral + error, -100.0, 100.0)
        out = -(self.KP * error + self.KI * self.integral * 0.001)
        if self.control_mode == "position":
            out = torch.clamp(out, -5.0, 5.0)
        else:
            out = torch.clamp(out * 100.0, -1000.0, 1000.0)
        return out.to(torch.float32)

    def __call__(self, env) -> Dict:
        mv = env.mean_voltage()
        if self._gc is None or self._gc["target_voltage"].numel() != mv.numel() or \
                self._gc["target_voltage"].device != mv.device:
            self._gc = _generator_control(mv.numel(), mv.device, self.generator)
        return {"servo": self.servo(mv), "generator_control": self._gc}
