"""Reference statistics for the on-device-RNG parity tests (build container only).

TEST INFRASTRUCTURE ONLY.  Runs the UNMODIFIED reference with its own NumPy generator
(`reset(seed=k)`) on all host cores and records, per env: sparks, short-circuit µs, removal, steps, break /
target flags (SURVEY §8d "statistical parity": 256 envs x 100 ms for the metric's config):
  cfg2     README config, quickstart action                                      256 envs x 100 ms
  cfg3     position mode, gap controller servo = clip(0.1 (gap - 25), +-1) evaluated from the state at
           every control-step boundary (stabilising sign, run_simulation.py:28-33), I5 ON 3 / OFF 80   128 envs x 100 ms
  shorts   debris-choked gap (oracle/cases.py "debris_shorts" parameters): short-circuit fraction     96 envs x 30 ms
  target   target distance 0.02 um beyond the initial gap: P(target reached within 40 ms)             160 envs x 40 ms
  hot      initial_gap 12 um, I17 / ON 5 / OFF 6 (wire breaks)                                          64 envs x <= 30 ms
Output: tests/golden/stats_reference.json (committed).   python -m oracle.make_stats_golden
"""
from __future__ import annotations

import json
import multiprocessing as mp
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from oracle import cases as CS          # noqa: E402
from oracle import ref_runner as RR     # noqa: E402


SHORTS = CS.CASES["debris_shorts"]
TARGET_CFG = dict(CS.README_CFG, target_cutting_distance=50.02)
JOBS = {"cfg2": (256, 100000, 1000), "cfg3": (128, 100000, 3000), "shorts": (96, 30000, 7000),
        "target": (160, 40000, 9000), "hot": (64, 30000, 5000)}


def _one(args):
    kind, seed, n_us = args
    controller = None
    if kind == "cfg2":
        env = RR.make_env(CS.README_CFG, "position")
        action = RR.make_action(*CS.QUICKSTART_ACTION)
    elif kind == "cfg3":
        env = RR.make_env(CS.README_CFG, "position")
        action = RR.make_action(0.0, 80.0, 5, 3.0, 80.0)

        def controller(e):
            gap = e.state.workpiece_position - e.state.wire_position
            return RR.make_action(float(np.float32(np.clip((gap - 25.0) * 0.1, -1.0, 1.0))), 80.0, 5, 3.0, 80.0)
    elif kind == "shorts":
        env = RR.make_env(SHORTS["config"], "position", ignition=SHORTS["ignition"], dielectric=SHORTS["dielectric"])
        action = RR.make_action(*SHORTS["actions"][0])
    elif kind == "target":
        env = RR.make_env(TARGET_CFG, "position")
        action = RR.make_action(*CS.QUICKSTART_ACTION)
    else:
        env = RR.make_env(dict(CS.README_CFG, initial_gap=12.0), "position")
        action = RR.make_action(0.0, 100.0, 17, 5.0, 6.0)
    env.reset(seed=seed)
    sparks, short_us, prev = 0, 0, 0
    t_end = n_us
    for k in range(n_us):
        if controller is not None and k % 1000 == 0:        # the action a latch at this boundary would take
            action = controller(env)
        _, _, term, _, info = env.step(action)
        st = env.state.spark_status[0]
        if st == 1 and prev != 1:
            sparks += 1
        prev = st
        short_us += int(env.state.is_short_circuit)
        if term:
            t_end = k + 1
            break
    return dict(seed=seed, sparks=sparks, short_us=short_us, steps=t_end,
                removed_um=env.state.workpiece_position - env.config.initial_gap,
                gap=env.state.workpiece_position - env.state.wire_position,
                broken=bool(env.state.is_wire_broken), target=bool(env.state.is_target_distance_reached),
                tmax=float(env.state.wire_temperature.max()))


def main():
    jobs = [(kind, base + k, n_us) for kind, (n, n_us, base) in JOBS.items() for k in range(n)]
    with mp.Pool(os.cpu_count()) as pool:
        res = pool.map(_one, jobs, chunksize=1)
    out = {"n_us": {k: v[1] for k, v in JOBS.items()},
           "generator": "oracle/make_stats_golden.py (unmodified reference, numpy default_rng seeds)"}
    pos = 0
    for kind, (n, n_us, _) in JOBS.items():
        out[kind] = res[pos:pos + n]
        pos += n
    path = os.path.join(_ROOT, "tests", "golden", "stats_reference.json")
    with open(path, "w") as f:
        json.dump(out, f)
    for kind in JOBS:
        r = out[kind]
        print(kind, "sparks/env-ms %.3f" % np.mean([x["sparks"] / (x["steps"] / 1000.0) for x in r]),
              "short frac %.4f" % (np.sum([x["short_us"] for x in r]) / np.sum([x["steps"] for x in r])),
              "removed um %.4f" % np.mean([x["removed_um"] for x in r]), "P(broken) %.3f" % np.mean([x["broken"] for x in r]),
              "P(target) %.3f" % np.mean([x["target"] for x in r]), "mean steps %.0f" % np.mean([x["steps"] for x in r]))


if __name__ == "__main__":
    main()
