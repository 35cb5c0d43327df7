"""Parity of the CUDA path (through the C-ABI) against the committed reference trajectories and
against the C oracle, with the reference's random draws injected (BASELINE cfg 2 protocol).

Discrete events must match exactly; fp64 scalars to 1e-12; the float32 temperature field within
|dT| <= 2e-3 K + 1e-5*|T| (measured errors are printed: <= 2.1e-4 K with the four-instruction
stencil of the coarse geometries, <= 1.5e-4 K with the difference form, 7e-4 K above 1000 K).
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import cases as CS
import parity_util as PU

pytestmark = pytest.mark.gpu


def _golden(name):
    return np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))


@pytest.mark.parametrize("name", sorted(CS.ALL_CASES))
def test_trajectory_matches_reference_golden(name):
    """µs-by-µs (1-µs launches) for the first 3000 µs, then at every 50th µs, against the
    fixtures generated from the unmodified reference."""
    import torch

    g = _golden(name)
    case = CS.ALL_CASES[name]
    tape = CS.case_tape(case)
    actions = g["actions"]
    eng = PU.make_engine(case, n_envs=3)
    # env 0 and 2 replay the case; env 1 gets a different tape (must not disturb its neighbours)
    tapes = np.stack([tape, np.roll(tape, 17, axis=0), tape])
    draws = torch.from_numpy(tapes).to(eng.device)
    rows = g["f64_rows"]
    n_done = int(g["n_done"][0])
    done = 0
    max_t_err = 0.0
    for r in rows:
        n = int(r) + 1 - done
        PU.stage_action(eng, PU.action_for_window(actions, done, n))
        eng.step(n, draws[:, done:done + n].contiguous())
        done += n
        f, i = PU.read_rows(eng, envs=(0, 2))
        k = int(np.searchsorted(rows, r))
        for env_row in range(2):
            PU.assert_rows_match(f[env_row], i[env_row], g["f64"][k], g["i32"][r].astype(np.int64),
                                 where=f"{name} after {done} µs env {env_row * 2}")
        max_t_err = max(max_t_err, abs(f[0, -1] - g["f64"][k, -1]))
    assert done == n_done
    print(f"\n[{name}] {n_done} µs matched; max |tmax_gpu - tmax_ref| = {max_t_err:.3e} K")


@pytest.mark.parametrize("name", sorted(CS.ALL_CASES))
def test_control_step_launches_match_oracle(name):
    """1000-µs launches (the production path) against the oracle: every scalar and the whole
    temperature field at each control-step boundary."""
    import torch

    g = _golden(name)
    case = CS.ALL_CASES[name]
    actions = g["actions"]
    orc_env, tape, orc = PU.run_oracle_case(case, actions, rec_every=1000, t_every=1000)
    eng = PU.make_engine(case, n_envs=5)
    draws = torch.from_numpy(np.stack([tape] * 5)).to(eng.device)
    n_us = case["n_us"]
    si = 1000
    worst = 0.0
    for w, t0 in enumerate(range(0, n_us, si)):
        n = min(si, n_us - t0)
        PU.stage_action(eng, PU.action_for_window(actions, t0, n))
        eng.step(n, draws[:, t0:t0 + n].contiguous())
        f, i = PU.read_rows(eng, envs=(0, 4))
        if w >= len(orc["f64"]):
            break
        for env_row in range(2):
            PU.assert_rows_match(f[env_row], i[env_row], orc["f64"][w], orc["i32"][w].astype(np.int64),
                                 where=f"{name} window {w} env {env_row * 4}")
        T_gpu = eng.T[0].cpu().numpy()
        T_ref = orc["T"][w]
        err = np.abs(T_gpu.astype(np.float64) - T_ref.astype(np.float64))
        worst = max(worst, float(err.max()))
        assert np.all(err <= PU.T_ATOL + PU.T_RTOL * np.abs(T_ref)), f"{name} window {w}: max dT {err.max()}"
        if i[0, -1]:
            break
    print(f"\n[{name}] max |T_gpu - T_oracle| over all snapshots = {worst:.3e} K")


def test_even_mode_sets_error_flag():
    """Crater data exists for odd modes only: the reference raises ValueError at the first spark
    (material.py:108-113); the kernel freezes the env and raises the mode-error flag."""
    import torch
    from sparc_b200 import _native as N

    case = dict(config=CS.README_CFG, mode="position")
    eng = PU.make_engine(case, n_envs=2)
    tape = CS.never_ignite_tape(3000, ignite_at=(1500,))
    draws = torch.from_numpy(np.stack([tape, tape])).to(eng.device)
    eng.set_actions(0.0, 80.0, np.array([6, 5]), 3.0, 80.0)
    for t0 in range(0, 3000, 1000):
        eng.step(1000, draws[:, t0:t0 + 1000].contiguous())
    flags = eng["flags"].cpu().numpy()
    assert flags[0] & N.F_MODE_ERROR and not (flags[1] & N.F_MODE_ERROR)
    assert eng["time"].cpu().numpy().tolist() == [1500, 3000]


def test_even_mode_with_crater_mapping_extension():
    """crater_mode_mapping="fill_missing": an even mode sparks with the documented record instead of
    freezing; the workpiece advance is that record's (material.py:152-174 arithmetic)."""
    import torch
    from sparc_b200 import _native as N

    case = dict(config=CS.README_CFG, mode="position", material={"crater_mode_mapping": "fill_missing"})
    eng = PU.make_engine(case, n_envs=2)
    prm = PU.case_params(case)
    tape = CS.never_ignite_tape(3000, ignite_at=(1500,))
    tape[:, 4] = 0.25                                         # crater volume = mean + 0.25 std
    draws = torch.from_numpy(np.stack([tape, tape])).to(eng.device)
    eng.set_actions(0.0, 80.0, np.array([6, 5]), 3.0, 80.0)
    for t0 in range(0, 3000, 1000):
        eng.step(1000, draws[:, t0:t0 + 1000].contiguous())
    flags = eng["flags"].cpu().numpy()
    assert not (flags[0] & N.F_MODE_ERROR) and not (flags[1] & N.F_MODE_ERROR)
    assert eng["time"].cpu().numpy().tolist() == [3000, 3000]
    wp = eng["wp"].cpu().numpy()
    for env, mode in ((0, 6), (1, 5)):
        vol = prm.crater_mean[mode] + 0.25 * prm.crater_std[mode]
        want = 50.0 + (vol / 1e9) / prm.kerf_times_height[mode] * 1000.0
        assert abs(wp[env] - want) <= 1e-12 * want, (mode, wp[env], want)
    assert prm.crater_mean[6] == prm.crater_mean[1]           # 68 A maps to the first record


@pytest.mark.parametrize("cfg_name,n_us", [("cfg2_default", 30000), ("hot_break", 20000), ("debris_shorts", 12000)])
def test_on_device_philox_matches_oracle_philox(cfg_name, n_us):
    """Production path (on-device Philox4x32-10 draws, quiet-µs fast path): the oracle implements
    the same counter-based stream, so trajectories must agree event-for-event."""
    from oracle import oracle as O

    case = CS.CASES[cfg_name]
    action = [float(np.float32(x)) for x in case["actions"][0]]
    seed, base, n_envs = 0x1234567887654321, 1000, 6
    from sparc_b200.engine import Engine

    eng = Engine(n_envs, PU.case_params(case), seed=seed, env_id_base=base)
    PU.stage_action(eng, action)
    orcs = [O.OracleEnv(**CS.oracle_overrides(case)) for _ in range(n_envs)]
    alive = [True] * n_envs
    worst = 0.0
    for t0 in range(0, n_us, 1000):
        eng.step(1000)
        f, i = PU.read_rows(eng, envs=range(n_envs))
        T = eng.T.cpu().numpy()
        for e in range(n_envs):
            if alive[e]:
                res = orcs[e].run(1000, [action], draws=None, seed=seed, env_id=base + e, rec_every=0)
                alive[e] = res["n_done"] == 1000
            fo, io = orcs[e].record(0)
            io = io.astype(np.int64)
            io[-1] = int(io[11] or io[12])
            PU.assert_rows_match(f[e], i[e], fo, io, where=f"{cfg_name} philox env {e} t0 {t0}")
            err = np.abs(T[e].astype(np.float64) - orcs[e].temperature.astype(np.float64))
            worst = max(worst, float(err.max()))
            assert np.all(err <= PU.T_ATOL + PU.T_RTOL * np.abs(orcs[e].temperature))
    sparks = int(eng["stat_sparks"].sum().item())
    assert sparks > 0
    print(f"\n[{cfg_name}] philox mode: {sparks} sparks over {n_envs} envs, max |dT| {worst:.3e} K")


@pytest.mark.parametrize("name,log_every", [("cfg2_default", 50), ("thick_2200", 100), ("thick_5500", 250),
                                            ("hot_break", 200)])
def test_logged_temperature_field_matches_oracle(name, log_every):
    """BASELINE cfg 5: sparc_step_logged samples the whole wire temperature field every `log_every` µs
    into the reference logger's (n_timesteps, n_segments) layout; compared with the oracle's
    snapshots at the same µs (rows after a wire break hold the final field)."""
    import torch

    g = _golden(name)
    case = CS.ALL_CASES[name]
    actions = g["actions"]
    n_us = min(case["n_us"], 4000) if name != "hot_break" else case["n_us"]
    orc_env, tape, orc = PU.run_oracle_case(dict(case, n_us=n_us), actions, rec_every=1000, t_every=log_every)
    eng = PU.make_engine(case, n_envs=3)
    draws = torch.from_numpy(np.stack([tape[:n_us]] * 3)).to(eng.device)
    rows_per_call = 1000 // log_every
    worst, seen = 0.0, 0
    T_ref_all = orc["T"]
    for t0 in range(0, n_us, 1000):
        n = min(1000, n_us - t0)
        PU.stage_action(eng, PU.action_for_window(actions, t0, n))
        log = eng.step_logged(n, log_every, n_log_envs=2, draws=draws[:, t0:t0 + n].contiguous())
        assert tuple(log.shape) == (2, n // log_every, eng.n_segments)
        L = log.cpu().numpy()
        for r in range(L.shape[1]):
            k = t0 // log_every + r
            ref = T_ref_all[min(k, len(T_ref_all) - 1)]      # after termination the oracle stops sampling
            for env_row in range(2):
                err = np.abs(L[env_row, r].astype(np.float64) - ref.astype(np.float64))
                worst = max(worst, float(err.max()))
                assert np.all(err <= PU.T_ATOL + PU.T_RTOL * np.abs(ref)), f"{name} row {k}: max dT {err.max()}"
            seen += 1
        if int(eng["flags"][0].item()) & 2:
            break
    assert seen >= rows_per_call
    print(f"\n[{name}] {seen} logged rows of {eng.n_segments} segments, max |dT| {worst:.3e} K")


def test_long_launch_equals_split_launches():
    """A 2500-µs call (three mechanics epochs of <= 1024 µs inside the library) gives bit-identical
    state and log rows to 1000 + 1000 + 500 µs calls, with injected draws and with on-device Philox."""
    import torch
    from sparc_b200.engine import Engine

    case = CS.CASES["short_servo"]
    prm = PU.case_params(case)
    tape = CS.case_tape(case)[:2500]
    skip = {"win_vsum", "win_sparks", "win_short_us", "win_steps", "mech_traj", "mech_ckpt", "mech_prog"}
    for injected in (True, False):
        engs = [Engine(4, prm, seed=11, env_id_base=5) for _ in range(2)]
        draws = torch.from_numpy(np.stack([tape] * 4)).to(engs[0].device) if injected else None
        for eng in engs:
            PU.stage_action(eng, case["actions"][0])
        log_a = engs[0].step_logged(2500, 100, draws=draws)
        parts, t0 = [], 0
        for n in (1000, 1000, 500):
            d = draws[:, t0:t0 + n].contiguous() if injected else None
            parts.append(engs[1].step_logged(n, 100, draws=d))
            t0 += n
        log_b = torch.cat(parts, dim=1)
        assert torch.equal(log_a, log_b)
        for name, t in engs[0].fields.items():
            if name not in skip:
                assert torch.equal(t, engs[1].fields[name]), name
        assert int(engs[0]["time"][0].item()) == 2500
        assert int(engs[0]["win_steps"][0].item()) == 2500       # the window spans the epochs of one call
        assert int(engs[1]["win_steps"][0].item()) == 500


@pytest.mark.parametrize("name", ["cfg2_default", "debris_shorts"])
def test_signal_log_matches_reference_golden(name):
    """Device-side signal logger (sparc_step_sampled): every µs of the first 3000 µs in three
    1000-µs calls, then every 50th µs, against the per-µs rows recorded from the unmodified
    reference — what SimulationLogger("every_step" / "interval") would have stored."""
    import torch
    from sparc_b200 import _native as N

    g = _golden(name)
    case = CS.ALL_CASES[name]
    tape = CS.case_tape(case)
    actions = g["actions"]
    rows = g["f64_rows"]
    n_done = int(g["n_done"][0])
    eng = PU.make_engine(case, n_envs=2)
    draws = torch.from_numpy(np.stack([tape, tape])).to(eng.device)
    S = {n: k for k, n in enumerate(N.SIGNALS)}
    f_of = {"workpiece_position": 0, "wire_position": 1, "wire_velocity": 2, "voltage": 4, "current": 5,
            "debris_volume": 6, "debris_density": 7, "flow_rate": 8}
    i_of = {"time": 0, "spark_state": 4, "is_short_circuit": 6, "time_in_critical_temp": 10}
    done, checked = 0, 0
    # golden rows: every µs below 3000, then µs 3000, 3050, ...: a 1-µs call re-aligns the sampling phase
    plan = [(1000, 1)] * 3 + [(1, 1)] + [(1000, 50)] * ((n_done - 3001) // 1000)
    for n, every in plan:
        PU.stage_action(eng, PU.action_for_window(actions, done, n))
        sig, _ = eng.step_sampled(n, every, draws[:, done:done + n].contiguous())
        sig = sig.cpu().numpy()
        assert sig.shape == (n // every, N.N_SIGNALS, 2)
        for r in range(sig.shape[0]):
            us = done + (r + 1) * every - 1                 # last µs stepped before this sample
            k = int(np.searchsorted(rows, us))
            assert rows[k] == us
            for env in range(2):
                for nm, col in f_of.items():
                    a, b = sig[r, S[nm], env], g["f64"][k, col]
                    assert a == b or abs(a - b) <= PU.F64_RTOL * abs(b), (name, us, nm, a, b)
                for nm, col in i_of.items():
                    assert int(sig[r, S[nm], env]) == int(g["i32"][us, col]), (name, us, nm)
                tm, tr = sig[r, S["max_wire_temperature"], env], g["f64"][k, 11]
                assert abs(tm - tr) <= PU.T_ATOL + PU.T_RTOL * abs(tr), (name, us, tm, tr)
            checked += 1
        done += n
        if int(eng["flags"][0].item()) & N.F_DONE:
            break
    assert checked >= 3000
    # the per-call window spans the whole call even though it was cut into launches at the samples
    assert int(eng["win_steps"][0].item()) == 1000 or int(eng["flags"][0].item()) & N.F_DONE


def test_strict_reference_reset_matches_reference_golden():
    """`strict_reference_reset=True`: reset() clears only what the reference's reset() clears (a new EDMState); the
    module-internal state — short-circuit countdown, debris volume, hysteresis and convection memory, prev_accel —
    survives.  Second episode of the unmodified reference (tests/golden/reference_reset.npz) µs by µs for the first
    150 µs (the carried countdown), then per 1000-µs launch."""
    import torch
    from sparc_b200 import (BatchedWireEDMEnv, DielectricModuleParameters, EnvironmentConfig, IgnitionModuleParameters)

    g = _golden("reference_reset")
    case = CS.CASES["debris_shorts"]
    n1, n2 = int(g["n1"][0]), int(g["n2"][0])
    env = BatchedWireEDMEnv(3, config=EnvironmentConfig(**case["config"]), ignition_params=IgnitionModuleParameters(**case["ignition"]),
                            dielectric_params=DielectricModuleParameters(**case["dielectric"]), latch_first_action=False,
                            strict_reference_reset=True)
    eng = env.engine
    t1 = torch.from_numpy(np.stack([CS.make_tape(4101, n1)] * 3)).to(eng.device)
    t2 = torch.from_numpy(np.stack([CS.make_tape(4102, 4000)] * 3)).to(eng.device)
    done = 0
    for n in (1000, 1000, 1000, n1 - 3000):
        PU.stage_action(eng, PU.action_for_window(g["actions1"], done, n))
        eng.step(n, t1[:, done:done + n].contiguous())
        done += n
    assert float(eng["debris_volume"][0]) > 1e-4 and int(eng["deb_short_rem"][0]) + int(eng["rand_short_rem"][0]) > 0
    env.reset()                                        # reference semantics: EDMState only
    assert int(eng["time"][0]) == 0 and float(eng["debris_volume"][0]) > 1e-4
    done, checked = 0, 0
    plan = [1] * 150 + [850] + [1000] * ((n2 - 1000) // 1000)
    for n in plan:
        PU.stage_action(eng, PU.action_for_window(g["actions2"], done, n))
        eng.step(n, t2[:, done:done + n].contiguous())
        done += n
        f, i = PU.read_rows(eng, envs=(0, 2))
        for row in range(2):
            PU.assert_rows_match(f[row], i[row], g["f64"][done - 1], g["i32"][done - 1].astype(np.int64),
                                 where=f"strict reset, episode 2 after {done} µs")
        checked += 1
        if done % 1000 == 0:
            T_ref = g["T"][done // 1000 - 1]
            err = np.abs(eng.T[0].cpu().numpy().astype(np.float64) - T_ref.astype(np.float64))
            assert np.all(err <= PU.T_ATOL + PU.T_RTOL * np.abs(T_ref))
    assert checked >= 150 and done == n2
