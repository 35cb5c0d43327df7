"""Helpers shared by the GPU parity tests: build an engine for an oracle/golden case, replay the
action list with the reference's latch timing, read the state back in the oracle's row layout."""
from __future__ import annotations

import numpy as np

from oracle import cases as CS
from oracle import oracle as O

F64_FIELDS = O.F64_FIELDS
I32_FIELDS = O.I32_FIELDS

# tolerances: discrete fields exact; fp64 scalars 1e-12 relative (they are
# bit-identical in practice: same IEEE operations, -fmad=false); temperature field float32.
F64_RTOL = 1e-12
T_ATOL = 2e-3
T_RTOL = 1e-5


def case_params(case):
    """sparc_b200 parameter objects for an oracle case dict."""
    from sparc_b200 import params as P

    cfg = P.EnvironmentConfig(**case.get("config", {}))
    kw = {}
    for grp, cls in (("ignition", P.IgnitionModuleParameters), ("wire", P.WireModuleParameters),
                     ("material", P.MaterialModuleParameters), ("dielectric", P.DielectricModuleParameters),
                     ("mechanics", P.MechanicsModuleParameters)):
        if case.get(grp):
            kw[grp] = cls(**case[grp])
    return P.derive_params(cfg, case.get("mode", "position"), **kw)


def make_engine(case, n_envs=1, seed=0, device="cuda:0"):
    from sparc_b200.engine import Engine

    eng = Engine(n_envs, case_params(case), device=device, seed=seed)
    if "unwind_v" in case:
        eng["unwind_v"].fill_(case["unwind_v"])
    return eng


def stage_action(eng, row):
    eng.set_actions(float(row[0]), float(row[1]), int(row[2]), float(row[3]), float(row[4]))


def action_for_window(actions, t0, n, servo_interval=1000):
    """Action row the reference would latch inside entry times [t0, t0+n): latch k (entry time
    k*servo_interval, k>=1) uses actions[k-1] (driver loop of README.md:101-106)."""
    lo, hi = max(t0, 1), t0 + n - 1
    k = -(-lo // servo_interval)       # first multiple >= lo
    if k * servo_interval > hi:
        return actions[0]
    return actions[min(k - 1, len(actions) - 1)]


def read_rows(eng, envs=(0,)):
    """(f64[len(envs),12], i32[len(envs),14]) in the oracle's record layout."""
    import torch
    from sparc_b200 import _native as N

    idx = torch.as_tensor(list(envs), device=eng.device)
    g = lambda name: eng[name][idx].cpu().numpy()
    flags = g("flags")
    tmax = eng.T[idx].max(dim=1).values.cpu().numpy().astype(np.float64)
    f = np.stack([g("wp"), g("wire"), g("v"), g("prev_accel"), g("voltage"), g("current"),
                  g("debris_volume"), g("debris_density"), g("flow"), g("last_crater_volume"),
                  g("spark_y"), tmax], axis=1)
    i = np.stack([g("time"), g("t_servo"), g("t_spark_ign"), g("t_spark_end"), g("spark_state"),
                  g("spark_dur"), flags & N.F_SHORT, g("rand_short_rem"), g("deb_short_rem"), g("ion_rem"),
                  g("crit_count"), (flags & N.F_BROKEN) // N.F_BROKEN, (flags & N.F_TARGET) // N.F_TARGET,
                  ((flags & N.F_DONE) != 0).astype(np.int32)], axis=1).astype(np.int64)
    return f, i


def run_oracle_case(case, actions, rec_every=1, t_every=1000):
    tape = CS.case_tape(case)
    e = O.OracleEnv(**CS.oracle_overrides(case))
    if "unwind_v" in case:
        e.set_unwind_velocity(case["unwind_v"])
    res = e.run(case["n_us"], actions, draws=tape, rec_every=rec_every, t_every=t_every)
    return e, tape, res


def assert_rows_match(f_gpu, i_gpu, f_ref, i_ref, where=""):
    if not np.array_equal(i_gpu, i_ref):
        bad = np.argwhere(i_gpu != i_ref)[0]
        raise AssertionError(f"{where}: discrete field {I32_FIELDS[bad[-1]]} differs: gpu {i_gpu[tuple(bad)]} "
                             f"ref {i_ref[tuple(bad)]}")
    # tmax (last column) is a float32 quantity: temperature tolerance
    np.testing.assert_allclose(f_gpu[..., :-1], f_ref[..., :-1], rtol=F64_RTOL, atol=1e-300, err_msg=where)
    np.testing.assert_allclose(f_gpu[..., -1], f_ref[..., -1], rtol=T_RTOL, atol=T_ATOL, err_msg=where + " tmax")
