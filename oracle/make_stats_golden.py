"""Reference statistics for the on-device-RNG parity test (build container only).

TEST INFRASTRUCTURE ONLY.  Runs the UNMODIFIED reference with its own NumPy generator
(`reset(seed=k)`) on all host cores and records per-env spark counts, removal and break times:
  cfg2   README config, quickstart action, 64 envs x 50 ms
  hot    initial_gap 12 um, I17 / ON 5 / OFF 6 (wire breaks), 48 envs x <= 30 ms
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


def _one(args):
    kind, seed, n_us = args
    if kind == "cfg2":
        env = RR.make_env(CS.README_CFG, "position")
        action = RR.make_action(*CS.QUICKSTART_ACTION)
    else:
        env = RR.make_env(dict(CS.README_CFG, initial_gap=12.0), "position")
        action = RR.make_action(0.0, 100.0, 17, 5.0, 6.0)
    env.reset(seed=seed)
    sparks, short_us, prev = 0, 0, 0
    t_end = n_us
    for k in range(n_us):
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
                broken=bool(env.state.is_wire_broken), tmax=float(env.state.wire_temperature.max()))


def main():
    jobs = [("cfg2", 1000 + k, 50000) for k in range(64)] + [("hot", 5000 + k, 30000) for k in range(48)]
    with mp.Pool(os.cpu_count()) as pool:
        res = pool.map(_one, jobs, chunksize=1)
    out = {"cfg2": res[:64], "hot": res[64:], "n_us": {"cfg2": 50000, "hot": 30000},
           "generator": "oracle/make_stats_golden.py (unmodified reference, numpy default_rng seeds)"}
    path = os.path.join(_ROOT, "tests", "golden", "stats_reference.json")
    with open(path, "w") as f:
        json.dump(out, f)
    c = out["cfg2"]
    print("cfg2: sparks/env-ms", np.mean([r["sparks"] / 50 for r in c]), "removed um/env", np.mean([r["removed_um"] for r in c]))
    h = out["hot"]
    print("hot: broken frac", np.mean([r["broken"] for r in h]), "mean steps", np.mean([r["steps"] for r in h]))


if __name__ == "__main__":
    main()
