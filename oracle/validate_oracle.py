"""Pin the C oracle against the RUNNING reference (build container only).

TEST INFRASTRUCTURE ONLY.  For every case in oracle/cases.py (CASES and the GEOMETRY_CASES sweep): run the unmodified reference
with the injected tape, run `oracle/wedm_oracle.c` on the same tape and recorded actions,
compare every per-µs field.  With `--write-golden` the reference trajectories are written
to tests/golden/<case>.npz (the committed fixtures; this script is their generator).

    python -m oracle.validate_oracle [--write-golden] [case ...]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from oracle import cases as CS          # noqa: E402
from oracle import oracle as O          # noqa: E402
from oracle import ref_runner as RR     # noqa: E402

GOLDEN_DIR = os.path.join(_ROOT, "tests", "golden")


def reference_controller(case):
    acts = case["actions"]
    if isinstance(acts, tuple) and acts[0] == "random":
        rows = CS.random_actions(acts[1], acts[2])
        return lambda env, c: RR.make_action(*rows[min(c, len(rows) - 1)])
    if isinstance(acts, tuple) and acts[0] == "gap_controller":
        desired = acts[1]

        def ctl(env, c):
            # stabilising sign of experiments/run_simulation.py:28-33
            gap = env.state.workpiece_position - env.state.wire_position
            servo = float(np.clip(0.1 * (gap - desired), -1.0, 1.0))
            return RR.make_action(servo, 80.0, 5, 3.0, 80.0)
        return ctl
    rows = acts
    return lambda env, c: RR.make_action(*rows[min(c, len(rows) - 1)])


def run_case_reference(name, per_us=True):
    case = CS.ALL_CASES[name]
    env = RR.make_env(case.get("config"), case.get("mode", "position"),
                      ignition=case.get("ignition"), wire=case.get("wire"),
                      material=case.get("material"), dielectric=case.get("dielectric"),
                      mechanics=case.get("mechanics"))
    tape = CS.case_tape(case)
    initial = {}
    if "unwind_v" in case:
        initial["wire_unwinding_velocity"] = case["unwind_v"]
    if "initial_gap" in case.get("config", {}):
        pass  # reset() applies config.initial_gap itself
    out = RR.run_reference(env, tape, case["n_us"], reference_controller(case), per_us=per_us,
                           t_every=1000, initial=initial)
    out["geometry"] = np.array([env.wire.n_segments, env.wire.zone_start, env.wire.zone_end,
                                env.wire.contact_bottom_idx, env.wire.contact_top_idx])
    out["avg_T"] = np.array([np.nan if env.state.wire_average_temperature is None
                             else env.state.wire_average_temperature])
    return case, tape, out


def run_case_oracle(case, tape, actions, rec_every=1):
    e = O.OracleEnv(**CS.oracle_overrides(case))
    if "unwind_v" in case:
        e.set_unwind_velocity(case["unwind_v"])
    res = e.run(case["n_us"], actions, draws=tape, rec_every=rec_every, t_every=1000)
    res["geometry"] = np.array(list(e.geometry().values()))
    return e, res


def compare(name, ref, orc, verbose=True):
    ok = True
    n = min(len(ref["f64"]), len(orc["f64"]))
    if len(ref["f64"]) != len(orc["f64"]):
        print(f"  [{name}] LENGTH MISMATCH ref {len(ref['f64'])} oracle {len(orc['f64'])}")
        ok = False
    if not np.array_equal(ref["geometry"], orc["geometry"]):
        print(f"  [{name}] geometry mismatch {ref['geometry']} vs {orc['geometry']}")
        ok = False
    ii = np.asarray(ref["i32"][:n]) != np.asarray(orc["i32"][:n])
    if ii.any():
        r, c = np.argwhere(ii)[0]
        print(f"  [{name}] first int mismatch row {r} field {O.I32_FIELDS[c]}: "
              f"ref {ref['i32'][r, c]} oracle {orc['i32'][r, c]}")
        ok = False
    worst = {}
    for c, fname in enumerate(O.F64_FIELDS):
        a, b = ref["f64"][:n, c], orc["f64"][:n, c]
        err = np.abs(a - b)
        rel = err / np.maximum(np.abs(a), 1e-300)
        worst[fname] = (float(err.max()) if n else 0.0, float(np.max(np.minimum(rel, err))) if n else 0.0,
                        int((a != b).sum()))
    nT = min(len(ref["T"]), len(orc["T"]))
    dT = np.abs(ref["T"][:nT].astype(np.float64) - orc["T"][:nT].astype(np.float64))
    nT_diff = int((ref["T"][:nT] != orc["T"][:nT]).sum())
    if verbose:
        print(f"  [{name}] rows {n}, T snapshots {nT}: max|dT| {dT.max() if nT else 0:.3e} K, "
              f"{nT_diff} of {dT.size} T values differ bitwise")
        for fname, (e, _, nd) in worst.items():
            if nd:
                print(f"      {fname:20s} max abs err {e:.3e}  ({nd} rows differ bitwise)")
    return ok, worst, (float(dT.max()) if nT else 0.0)


RESET_CASE, RESET_N1, RESET_N2 = "debris_shorts", 3200, 4000


def run_reference_reset_case():
    """Second episode on USED modules: run RESET_N1 µs of the debris_shorts case (short-circuit countdowns, debris,
    hysteresis memory, convection all active), call the reference's reset(), run RESET_N2 µs with a second tape.
    Returns (case, tape1, tape2, ref1, ref2)."""
    case = CS.CASES[RESET_CASE]
    env = RR.make_env(case.get("config"), case.get("mode", "position"), ignition=case.get("ignition"),
                      wire=case.get("wire"), material=case.get("material"), dielectric=case.get("dielectric"),
                      mechanics=case.get("mechanics"))
    tape1 = CS.make_tape(4101, RESET_N1)
    tape2 = CS.make_tape(4102, RESET_N2)
    ctl = reference_controller(case)
    ref1 = RR.run_reference(env, tape1, RESET_N1, ctl, per_us=True, t_every=1000)
    env.reset()                                   # the reference's reset: new EDMState, modules untouched
    ref2 = RR.run_reference(env, tape2, RESET_N2, ctl, per_us=True, t_every=1000, do_reset=False)
    geo = np.array([env.wire.n_segments, env.wire.zone_start, env.wire.zone_end, env.wire.contact_bottom_idx,
                    env.wire.contact_top_idx])
    ref1["geometry"] = ref2["geometry"] = geo
    return case, tape1, tape2, ref1, ref2


def check_reference_reset(write_golden=False):
    case, tape1, tape2, ref1, ref2 = run_reference_reset_case()
    e = O.OracleEnv(**CS.oracle_overrides(case))
    o1 = e.run(RESET_N1, ref1["actions"], draws=tape1, rec_every=1, t_every=1000)
    e.reset(reference_semantics=True)
    o2 = e.run(RESET_N2, ref2["actions"], draws=tape2, rec_every=1, t_every=1000)
    o1["geometry"] = o2["geometry"] = np.array(list(e.geometry().values()))
    ok1, _, _ = compare("reference_reset/episode1", ref1, o1)
    ok2, _, _ = compare("reference_reset/episode2", ref2, o2)
    carried = dict(short_us_first_100=int(ref2["i32"][:100, 6].sum()), debris_volume_t0=float(ref2["f64"][0, 6]))
    print(f"{'OK  ' if ok1 and ok2 else 'FAIL'} reference_reset: episode 2 starts with {carried}")
    if write_golden:
        n = len(ref2["f64"])
        np.savez_compressed(os.path.join(GOLDEN_DIR, "reference_reset.npz"), f64=ref2["f64"], i32=ref2["i32"].astype(np.int32),
                            T=ref2["T"], T_times=ref2["T_times"], actions1=ref1["actions"], actions2=ref2["actions"],
                            n1=np.array([RESET_N1]), n2=np.array([n]))
    return ok1 and ok2


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("cases", nargs="*")
    ap.add_argument("--write-golden", action="store_true")
    args = ap.parse_args()
    names = [c for c in args.cases if c != "reference_reset"] or ([] if args.cases else list(CS.ALL_CASES))
    all_ok = True
    for name in names:
        case, tape, ref = run_case_reference(name)
        e, orc = run_case_oracle(case, tape, ref["actions"])
        ok, worst, dT = compare(name, ref, orc)
        sparks = int((np.diff(np.concatenate([[0], (ref["i32"][:, 4] == 1).astype(int)])) == 1).sum())
        print(f"{'OK  ' if ok else 'FAIL'} {name}: {ref['n_done']} µs, {sparks} sparks, "
              f"final state {ref['i32'][-1, 4]}, broken {ref['i32'][-1, 11]}, target {ref['i32'][-1, 12]}, "
              f"tmax {ref['f64'][:, 11].max():.2f} K, short µs {int(ref['i32'][:, 6].sum())}, "
              f"debris-short starts {int((ref['i32'][:, 8] == case.get('ignition', {}).get('debris_short_duration', 50)).sum())}, "
              f"random-short starts {int((ref['i32'][:, 7] == 100).sum())}, min gap "
              f"{(ref['f64'][:, 0] - ref['f64'][:, 1]).min():.2f}")
        all_ok &= ok
        if args.write_golden:
            os.makedirs(GOLDEN_DIR, exist_ok=True)
            # keep fixtures small: per-µs ints (compress well) + per-µs f64 for the first
            # 3000 µs, then every 50th µs; T snapshots every 1000 µs.
            n = len(ref["f64"])
            dense = 0 if name in CS.GEOMETRY_CASES else min(n, 3000)      # geometry sweep: every 50th µs only
            keep = np.unique(np.concatenate([np.arange(dense), np.arange(49, n, 50), np.arange(0, n, 50), [n - 1]]))
            np.savez_compressed(
                os.path.join(GOLDEN_DIR, f"{name}.npz"),
                f64=ref["f64"][keep], f64_rows=keep, i32=ref["i32"].astype(np.int32),
                T=ref["T"], T_times=ref["T_times"], actions=ref["actions"],
                geometry=ref["geometry"], n_done=np.array([ref["n_done"]]), avg_T=ref["avg_T"],
                consumed=np.packbits(ref["consumed"], axis=0))
    if not args.cases or "reference_reset" in args.cases:
        all_ok &= check_reference_reset(args.write_golden)
    print("ALL OK" if all_ok else "SOME FAILED")
    return 0 if all_ok else 1


if __name__ == "__main__":
    sys.exit(main())
