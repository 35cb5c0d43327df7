"""`SimulationLogger` for the batched environment (reference: src/wedm/utils/logger.py:54-237).

Same configuration dict, same signal names (EDMState attribute names), same backends
("memory": dict of lists; "numpy": one `.npz` with one array per signal).  The reference logger is
driven by `collect(state, info)` after every 1-µs `step()`; here `collect(info)` is called after
every `BatchedWireEDMEnv.step()` (one control step), and each sample is a device tensor over the
batch: scalars -> [n_envs], `wire_temperature` -> [n_envs, n_segments].  Sub-control-step
sampling of the temperature field comes from the kernel itself (`sparc_step_logged`): pass
`log_temperature_every=k` to `BatchedWireEDMEnv.step` and the [n_envs, n_us/k, n_segments] block in
`info["wire_temperature_log"]` is appended sample by sample.  Sub-control-step sampling of the scalar
signals ("every_step" / "interval" below one control step) comes from `sparc_step_sampled`: pass
`log_signals_every=k` to `BatchedWireEDMEnv.step` (`step_kwargs()` gives both arguments for this
logger's configuration) and the rows of `info["signal_log"]` are appended one per sampled µs.

`.npz` layout: signal -> array [n_samples, n_envs(, n_segments)]; with one env the env axis is
dropped, which is exactly the reference's `(n_timesteps, n_segments)` (animate_wire_heating.py:44-47).
"""
from __future__ import annotations

import pathlib
from collections import defaultdict
from typing import Any, Dict, List, Optional, TypedDict

import numpy as np

from ..state import _FLAG_MAP, _SCALAR_MAP


class LogFrequencyConfig(TypedDict, total=False):
    type: str            # "every_step" | "control_step" | "interval"
    value: int           # interval in µs (reference: number of step() calls)


class BackendConfig(TypedDict, total=False):
    type: str            # "memory" | "numpy"
    filepath: str
    compress: bool


class LoggerConfig(TypedDict, total=False):
    signals_to_log: List[str]
    log_frequency: LogFrequencyConfig
    backend: BackendConfig


class BatchedSimulationLogger:
    def __init__(self, config: LoggerConfig, env_reference=None):
        self.config = config
        self.env = env_reference
        self._validate_config()
        self.log_data: Dict[str, List[Any]] = defaultdict(list)
        self.step_counter = 0                      # µs, like the reference's call counter

    # reference: utils/logger.py:73-104 (same messages)
    def _validate_config(self):
        if not self.config.get("signals_to_log"):
            raise ValueError("LoggerConfig: 'signals_to_log' must be provided and non-empty.")
        if not self.config.get("log_frequency"):
            raise ValueError("LoggerConfig: 'log_frequency' must be provided.")
        if not self.config.get("backend"):
            raise ValueError("LoggerConfig: 'backend' must be provided.")
        backend_type = self.config["backend"]["type"]
        if backend_type not in ["memory", "numpy"]:
            raise NotImplementedError(f"Backend type '{backend_type}' is not yet implemented.")
        if backend_type == "numpy":
            if "filepath" not in self.config["backend"]:
                raise ValueError("LoggerConfig: 'filepath' must be provided for 'numpy' backend.")
            if not isinstance(self.config["backend"]["filepath"], str):
                raise ValueError("LoggerConfig: 'filepath' for 'numpy' backend must be a string.")
            self.config["backend"].setdefault("compress", False)

    def _engine(self):
        eng = getattr(self.env, "engine", None)
        return eng if eng is not None else self.env._engine      # the drop-in WireEDMEnv keeps it private

    def _flow_scale(self) -> float:
        """state.dielectric_flow_rate = flow_condition * base_flow_rate / 1e9 (dielectric.py:158)."""
        return float(getattr(self.env, "base_flow_rate", 100.0)) / 1e9

    def _signal(self, name: str):
        """Device tensor of one EDMState attribute over the batch (a copy: the state is live)."""
        eng = self._engine()
        if name == "dielectric_flow_rate":
            return eng["flow"] * self._flow_scale()
        if name == "wire_temperature":
            return eng.T.clone()
        if name in _SCALAR_MAP:
            return eng[_SCALAR_MAP[name]].clone()
        if name in _FLAG_MAP:
            return (eng["flags"] & _FLAG_MAP[name]) != 0
        if name == "spark_status":
            return eng["spark_state"].clone()
        if name in ("target_voltage", "ON_time", "OFF_time"):
            return eng[{"target_voltage": "target_voltage", "ON_time": "on_time", "OFF_time": "off_time"}[name]].clone()
        if name == "wire_average_temperature":
            return eng["wire_avg_T"].clone()
        return None

    # reference signal name -> row name of the device-side signal log (_native.SIGNALS)
    _SIGLOG = {"time": "time", "voltage": "voltage", "current": "current", "wire_position": "wire_position",
               "wire_velocity": "wire_velocity", "workpiece_position": "workpiece_position",
               "target_delta": "target_delta", "debris_density": "debris_density",
               "debris_concentration": "debris_density", "flow_rate": "flow_rate",
               "dielectric_flow_rate": "flow_rate", "is_short_circuit": "is_short_circuit",
               "spark_status": "spark_state", "wire_average_temperature": "wire_average_temperature",
               "debris_volume": "debris_volume", "time_in_critical_temp": "time_in_critical_temp"}

    def step_kwargs(self) -> Dict[str, int]:
        """Keyword arguments for `BatchedWireEDMEnv.step` that make the kernel sample at this
        logger's frequency: every µs for "every_step", every `value` µs for "interval" (when it
        divides the control step), nothing for "control_step"."""
        freq = self.config["log_frequency"]
        k = 1 if freq["type"] == "every_step" else int(freq.get("value", 0)) if freq["type"] == "interval" else 0
        if k <= 0 or (self.env is not None and self.env.servo_interval % k):
            return {}
        kw = {}
        if any(n in self._SIGLOG for n in self.config["signals_to_log"]):
            kw["log_signals_every"] = k
        if "wire_temperature" in self.config["signals_to_log"]:
            kw["log_temperature_every"] = k
        return kw

    def _collect_reference_style(self, state, info: Dict[str, Any]) -> None:
        """`collect(state, info)` exactly as the reference's drivers call it after every 1-µs
        `WireEDMEnv.step()` (utils/logger.py:118-160): every_step / control_step / interval counted in
        calls; values read attribute by attribute from the state view, ndarrays copied."""
        self.step_counter += 1
        freq = self.config["log_frequency"]
        if freq["type"] == "every_step":
            should_log = True
        elif freq["type"] == "control_step":
            should_log = bool(info.get("control_step", False))
        else:
            should_log = self.step_counter % int(freq["value"]) == 0
        if not should_log:
            return
        for name in self.config["signals_to_log"]:
            if not hasattr(state, name):
                print(f"Warning: No accessor found for signal '{name}'. Skipping.")
                continue
            value = getattr(state, name)
            self.log_data[name].append(value.copy() if isinstance(value, np.ndarray) else value)

    def collect(self, info: Optional[Dict[str, Any]] = None, n_us: Optional[int] = None):
        """Call after every `env.step()`.  `n_us` defaults to the env's control step.  The reference's
        call `collect(env.state, info)` (drop-in `WireEDMEnv`, one µs per step) is accepted too."""
        if self.env is None:
            raise ValueError("BatchedSimulationLogger needs env_reference")
        if info is not None and not isinstance(info, dict) and isinstance(n_us, dict):
            return self._collect_reference_style(info, n_us)
        import torch

        n_us = int(self.env.servo_interval if n_us is None else n_us)
        before = self.step_counter
        self.step_counter += n_us
        freq = self.config["log_frequency"]
        tlog = (info or {}).get("wire_temperature_log")
        slog = (info or {}).get("signal_log")
        if freq["type"] == "interval":
            k = int(freq["value"])
            should_log = self.step_counter // k > before // k
        else:                                      # "control_step"; "every_step" at control-step granularity
            should_log = True
        for name in self.config["signals_to_log"]:
            if name == "wire_temperature" and tlog is not None:
                for r in range(tlog.shape[1]):     # kernel-sampled rows, already at the requested interval
                    self.log_data[name].append(tlog[:, r])
                continue
            if slog is not None and name in self._SIGLOG:
                rows = slog[self._SIGLOG[name]]    # kernel-sampled rows [n_samples, n_envs]
                if name == "dielectric_flow_rate":
                    rows = rows * self._flow_scale()
                elif name == "is_short_circuit":
                    rows = rows != 0
                elif name in ("time", "spark_status", "time_in_critical_temp"):
                    rows = rows.to(dtype=torch.int32)
                for r in range(rows.shape[0]):
                    self.log_data[name].append(rows[r])
                continue
            if not should_log:
                continue
            value = self._signal(name)
            if value is None:
                print(f"Warning: No accessor found for signal '{name}'. Skipping.")
                continue
            self.log_data[name].append(value)

    def finalize(self):
        if self.config["backend"]["type"] != "numpy":
            return
        if not self.log_data:
            print("No data collected, skipping .npz file creation.")
            return
        import torch

        numpy_data = {}
        for name, samples in self.log_data.items():
            if not isinstance(samples[0], torch.Tensor):       # reference-style samples (drop-in env)
                numpy_data[name] = np.array(samples)
                continue
            arr = torch.stack(list(samples)).cpu().numpy()
            if arr.ndim >= 2 and arr.shape[1] == 1:
                arr = arr[:, 0]                    # one env: the reference's (n_timesteps[, n_segments])
            numpy_data[name] = arr
        output_path = pathlib.Path(self.config["backend"]["filepath"])
        output_path.parent.mkdir(parents=True, exist_ok=True)
        (np.savez_compressed if self.config["backend"].get("compress", False) else np.savez)(output_path, **numpy_data)
        print(f"Logged data saved to {output_path}")

    def get_data(self):
        if self.config["backend"]["type"] == "memory":
            return self.log_data
        if self.config["backend"]["type"] == "numpy":
            return self.config["backend"].get("filepath")
        return None

    def reset(self):
        self.log_data = defaultdict(list)
        self.step_counter = 0


SimulationLogger = BatchedSimulationLogger          # reference name (wedm.utils.logger.SimulationLogger)
