"""Run the UNMODIFIED reference (`/root/reference/src/wedm`) with injected random draws.

TEST INFRASTRUCTURE ONLY.  This module imports the reference package, so it only
works in the build container (the GPU box has no `/root/reference`).  It is used by
`oracle/make_golden.py` to generate the committed fixtures under `tests/golden/` and by
`oracle/validate_oracle.py` to pin the C restatement (`oracle/wedm_oracle.c`).

Draw injection: the reference reads `self.env.np_random` at call time
(`modules/ignition.py:233,239,261,327`, `modules/material.py:127`), so replacing
`env.np_random` after `reset()` with a tape object makes every stochastic decision a
pure function of `draws[t, slot]`:

    slot 0  u1  debris-short roll      ignition.py:233  (1st .random() of the µs)
    slot 1  u2  random-short roll      ignition.py:239  (2nd .random())
    slot 2  u3  ignition roll          ignition.py:327  (3rd .random())
    slot 3  u4  spark location         ignition.py:261  (.uniform(0,h) = 0 + (h-0)*u4)
    slot 4  z5  crater volume          material.py:127  (.normal(m,s) = m + s*z5)

`t` is `env.state.time` at the moment of the call (µs since reset).
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
from typing import Callable, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_SRC = "/root/reference/src"

N_SLOTS = 5


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "wedm"))


def import_reference():
    """Import `wedm` from /root/reference behind the gymnasium stub."""
    if not reference_available():
        raise RuntimeError("reference sources not present (only available in the build container)")
    stub = os.path.join(_HERE, "gymnasium_stub")
    for p in (REFERENCE_SRC, stub):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.setdefault("NUMBA_NUM_THREADS", "1")
    import wedm  # noqa: F401

    return wedm


class TapeRNG:
    """Stands in for `numpy.random.Generator`; serves draws from a [T,5] fp64 tape."""

    def __init__(self, env, draws: np.ndarray):
        assert draws.ndim == 2 and draws.shape[1] == N_SLOTS
        self.env = env
        self.draws = draws
        self._t = -1
        self._n_random = 0
        self.consumed = np.zeros(draws.shape, dtype=bool)

    def _row(self) -> int:
        t = int(self.env.state.time)
        if t != self._t:
            self._t = t
            self._n_random = 0
        return t

    def random(self):
        t = self._row()
        slot = self._n_random
        self._n_random += 1
        assert slot < 3, "more than three .random() calls in one µs"
        self.consumed[t, slot] = True
        return float(self.draws[t, slot])

    def uniform(self, low, high):
        t = self._row()
        self.consumed[t, 3] = True
        return low + (high - low) * float(self.draws[t, 3])

    def normal(self, loc, scale):
        t = self._row()
        self.consumed[t, 4] = True
        return loc + scale * float(self.draws[t, 4])


def make_env(config_kwargs=None, mechanics_control_mode="position", ignition=None, wire=None,
             material=None, dielectric=None, mechanics=None):
    """Construct a reference WireEDMEnv (stdout of WireModule.__init__ silenced, wire.py:254-257)."""
    wedm = import_reference()
    cfg = wedm.EnvironmentConfig(**(config_kwargs or {}))
    kw = {}
    if ignition:
        kw["ignition_params"] = wedm.IgnitionModuleParameters(**ignition)
    if wire:
        kw["wire_params"] = wedm.WireModuleParameters(**wire)
    if material:
        kw["material_params"] = wedm.MaterialModuleParameters(**material)
    if dielectric:
        kw["dielectric_params"] = wedm.DielectricModuleParameters(**dielectric)
    if mechanics:
        kw["mechanics_params"] = wedm.MechanicsModuleParameters(**mechanics)
    with contextlib.redirect_stdout(io.StringIO()):
        env = wedm.WireEDMEnv(mechanics_control_mode=mechanics_control_mode, config=cfg, **kw)
    return env


# Per-µs scalar record.  Names match the C oracle / CUDA state field names.
F64_FIELDS = ("wp", "wire", "v", "prev_accel", "voltage", "current", "debris_volume",
              "debris_density", "flow", "last_crater_volume", "spark_y", "tmax")
I32_FIELDS = ("time", "t_servo", "t_spark_ign", "t_spark_end", "spark_state", "spark_dur",
              "is_short", "rand_short_rem", "deb_short_rem", "ion_rem", "crit_count",
              "broken", "target_reached", "terminated")


def _snapshot(env, terminated: bool):
    s = env.state
    y = s.spark_status[1]
    ion = s.ionized_channel
    f = (
        s.workpiece_position, s.wire_position, s.wire_velocity, env.mechanics.prev_accel,
        0.0 if s.voltage is None else float(s.voltage),
        0.0 if s.current is None else float(s.current),
        s.debris_volume, s.debris_density, s.flow_rate, s.last_crater_volume,
        -1.0 if y is None else float(y),
        float(np.max(s.wire_temperature)) if len(s.wire_temperature) else 0.0,
    )
    i = (
        s.time, s.time_since_servo, s.time_since_spark_ignition, s.time_since_spark_end,
        int(s.spark_status[0]), int(s.spark_status[2]), int(bool(s.is_short_circuit)),
        env.ignition.random_short_remaining, env.ignition.debris_short_remaining,
        0 if ion is None else int(ion[1]), s.time_in_critical_temp,
        int(bool(s.is_wire_broken)), int(bool(s.is_target_distance_reached)), int(terminated),
    )
    return f, i


def make_action(servo, target_voltage, current_mode, on_time, off_time):
    """Action dict in the reference's format (wire_edm.py:84-98,162-170)."""
    return {
        "servo": np.array([servo], dtype=np.float32),
        "generator_control": {
            "target_voltage": np.array([target_voltage], dtype=np.float32),
            "current_mode": np.array([current_mode], dtype=np.int32),
            "ON_time": np.array([on_time], dtype=np.float32),
            "OFF_time": np.array([off_time], dtype=np.float32),
        },
    }


def run_reference(env, draws: np.ndarray, n_us: int,
                  controller: Callable[[object, int], dict],
                  per_us: bool = True, t_every: int = 1000,
                  initial: Optional[dict] = None, do_reset: bool = True):
    """Reset `env` (unless do_reset=False: the caller did, e.g. a second episode on used modules), inject `draws`, run up to `n_us` step() calls (stops at termination).

    controller(env, ctrl_index) -> action dict; called before the first step and after
    every step whose info["control_step"] is True (the loop of README.md:101-106 /
    experiments/run_simulation.py:245-297).  Returns dict of arrays:
      f64[n,len(F64_FIELDS)], i32[n,len(I32_FIELDS)] per µs (or per control step when
      per_us=False), T[k,N] snapshots every `t_every` µs, actions[c,5] that were passed.
    """
    if do_reset:
        env.reset(seed=0)
    if initial:
        for k, v in initial.items():
            setattr(env.state, k, v)
    tape = TapeRNG(env, draws)
    env.np_random = tape
    f_rows, i_rows, t_rows, t_times, actions = [], [], [], [], []
    ctrl = 0
    action = controller(env, ctrl)
    gc = action["generator_control"]
    actions.append([float(action["servo"][0]), float(gc["target_voltage"][0]), float(gc["current_mode"][0]),
                    float(gc["ON_time"][0]), float(gc["OFF_time"][0])])
    n_done = 0
    for k in range(n_us):
        obs, rew, terminated, truncated, info = env.step(action)
        n_done += 1
        if per_us or (k + 1) % t_every == 0 or terminated:
            f, i = _snapshot(env, terminated)
            f_rows.append(f)
            i_rows.append(i)
        if (k + 1) % t_every == 0 or terminated:
            t_rows.append(env.state.wire_temperature.copy())
            t_times.append(k + 1)
        if terminated:
            break
        if info.get("control_step"):
            ctrl += 1
            action = controller(env, ctrl)
            gc = action["generator_control"]
            actions.append([float(action["servo"][0]), float(gc["target_voltage"][0]),
                            float(gc["current_mode"][0]), float(gc["ON_time"][0]), float(gc["OFF_time"][0])])
    return {
        "f64": np.asarray(f_rows, dtype=np.float64).reshape(-1, len(F64_FIELDS)),
        "i32": np.asarray(i_rows, dtype=np.int64).reshape(-1, len(I32_FIELDS)),
        "T": np.asarray(t_rows, dtype=np.float32),
        "T_times": np.asarray(t_times, dtype=np.int64),
        "actions": np.asarray(actions, dtype=np.float64),
        "n_done": n_done,
        "consumed": tape.consumed[:n_done],
    }
