"""Time the UNMODIFIED reference's own `WireEDMEnv.step` on this machine's host cores.

MEASUREMENT INFRASTRUCTURE ONLY (bench.py's `cpu_baseline` / `--impl reference` legs).  The reference is
imported from `/root/reference/src` in the build container, else from `oracle/_ref/` — the copy
`oracle/fetch_ref.sh` pip-installs so that it travels to the GPU box (git-ignored) — behind the
gymnasium stub of `oracle/gymnasium_stub/` (gymnasium itself is not installed; the package only
touches `gym.Env`, `spaces.Box`, `spaces.Dict`).  Nothing under `sparc_b200/` imports this module.

What is timed: `time.perf_counter` around N `env.step(action)` calls after a warm-up
call (Numba JIT), `NUMBA_NUM_THREADS=1`, README config (`README.md:45-50`), the quickstart constant
action (`examples/quickstart.py:25-33`); once on ONE core (the reference is single-threaded), once with
one env per process on ALL host cores.  cfg1 (`BASELINE.json` configs[0]): `reset(seed=0)` and 1000
`step(action_space.sample())` calls, µs of wall time per call.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import time

_HERE = os.path.dirname(os.path.abspath(__file__))
README_CFG = dict(workpiece_height=15.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)


def reference_root():
    """Directory that holds the `wedm` package, or None."""
    for root in ("/root/reference/src", os.path.join(_HERE, "_ref")):
        if os.path.isdir(os.path.join(root, "wedm")):
            return root
    return None


def available():
    """(ok, why): the reference and its only compiled dependency (numba) can be imported here."""
    if reference_root() is None:
        return False, "reference package not found (run oracle/fetch_ref.sh in the build container)"
    try:
        import numba  # noqa: F401
    except Exception as e:  # pragma: no cover - depends on the box
        return False, f"numba is not importable on this box ({e.__class__.__name__})"
    return True, ""


def _import_reference():
    os.environ.setdefault("NUMBA_NUM_THREADS", "1")
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    for p in (reference_root(), os.path.join(_HERE, "gymnasium_stub")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import wedm  # noqa: F401

    return wedm


def _make_env(mode="position"):
    wedm = _import_reference()
    from wedm.core.env_config import EnvironmentConfig

    with contextlib.redirect_stdout(io.StringIO()):      # WireModule.__init__ prints two lines (wire.py:254-257)
        env = wedm.WireEDMEnv(mechanics_control_mode=mode, config=EnvironmentConfig(**README_CFG))
    return env


def _quickstart_action():
    import numpy as np

    return {"servo": np.array([0.1], dtype=np.float32),
            "generator_control": {"target_voltage": np.array([80.0], dtype=np.float32),
                                  "current_mode": np.array([5], dtype=np.int32),
                                  "ON_time": np.array([3.0], dtype=np.float32),
                                  "OFF_time": np.array([80.0], dtype=np.float32)}}


def _worker(args):
    """One process: one env, `seconds` of env.step() calls; returns (calls, elapsed)."""
    seed, seconds = args
    env = _make_env()
    env.reset(seed=seed)
    action = _quickstart_action()
    env.step(action)                                     # JIT warm-up
    calls, t0 = 0, time.perf_counter()
    deadline = t0 + seconds
    while True:
        for _ in range(2000):
            _, _, term, _, _ = env.step(action)
            calls += 1
            if term:
                env.reset(seed=seed + calls)
        now = time.perf_counter()
        if now >= deadline:
            return calls, now - t0


def time_reference(seconds: float = 10.0, cores=None):
    """Reference env.step throughput: dict(single_core, all_cores, cores, sample) in env-us/s (= calls/s)."""
    import multiprocessing as mp

    cores = cores or os.cpu_count() or 1
    calls, dt = _worker((0, seconds * 0.4))
    single = calls / dt
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        res = pool.map(_worker, [(100 + k, seconds * 0.6) for k in range(cores)])
    total = sum(c / t for c, t in res)
    return dict(single_core=single, all_cores=total, cores=cores, kind="reference",
                sample=f"reference WireEDMEnv.step, README config, quickstart action: 1 env x {dt:.1f} s on one core; "
                       f"{cores} processes x 1 env x {seconds * 0.6:.1f} s (NUMBA_NUM_THREADS=1)")


def time_reference_cfg1(n_steps: int = 1000):
    """BASELINE configs[0]: reset(seed=0), 1000 x step(action_space.sample()); wall µs per call."""
    env = _make_env()
    env.reset(seed=0)
    env.step(env.action_space.sample())                  # JIT warm-up (not counted)
    env.reset(seed=0)
    actions = [env.action_space.sample() for _ in range(n_steps)]
    t0 = time.perf_counter()
    for a in actions:
        env.step(a)
    dt = time.perf_counter() - t0
    return dict(us_per_call=1e6 * dt / n_steps, calls=n_steps, time_after=int(env.state.time))


if __name__ == "__main__":
    ok, why = available()
    print(ok, why)
    if ok:
        print(time_reference_cfg1())
        print(time_reference(float(sys.argv[1]) if len(sys.argv) > 1 else 6.0))
