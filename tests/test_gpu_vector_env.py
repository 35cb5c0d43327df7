"""SURVEY §8(f) rank 2 on the GPU: on-device controllers in closed loop against the oracle driven by
the reference's controller arithmetic, and the Gymnasium VectorEnv-shaped adapter (auto-reset
modes, truncation, rewards)."""
import numpy as np
import pytest

import parity_util as PU
from oracle import cases as CS
from oracle import oracle as O

pytestmark = pytest.mark.gpu

README = dict(workpiece_height=15.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)


def _tape(n_envs, n_us, key):
    g = np.random.Generator(np.random.Philox(key=key))
    u = g.random((n_envs, n_us, 4))
    z = g.standard_normal((n_envs, n_us, 1))
    return np.ascontiguousarray(np.concatenate([u, z], axis=2))


THICK = dict(workpiece_height=50.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)


@pytest.mark.parametrize("in_kernel", [False, True])
@pytest.mark.parametrize("mode,cfg,seg_len,steps", [("position", README, 0.2, 40), ("velocity", README, 0.2, 12),
                                                    ("position", THICK, 0.1, 10)])
def test_gap_controller_closed_loop_matches_oracle(mode, cfg, seg_len, steps, in_kernel):
    """BASELINE cfg 3: the batched GapController closes the loop on the device; the oracle is driven
    by the reference's controller arithmetic (run_simulation.py:27-39) evaluated in numpy from its
    own state each control step.  Same injected draws => same trajectories, discrete fields exact.
    The moving servo leaves the dielectric hysteresis band every ~32 µs (in-place refresh path), the
    40-step run also leaves the debris-density band; the thick case runs two warps per env.
    in_kernel: the controller object itself is passed to step(): the same law is evaluated inside the
    pre-pass and step kernels (SparcController), no torch op between control steps."""
    import torch
    from sparc_b200 import BatchedWireEDMEnv, EnvironmentConfig, GapController

    from sparc_b200 import WireModuleParameters

    E, STEPS, SI = 6, steps, 1000
    gen = dict(target_voltage=80.0, current_mode=5, ON_time=3.0, OFF_time=80.0)
    if mode == "position":
        ctrl = GapController(25.0, "position", limit=1.0, generator=gen)        # cfg 3 settings
        ref_servo = lambda gap: np.float32(np.clip((gap - 25.0) * 0.1, -1.0, 1.0))
    else:
        ctrl = GapController(25.0, "velocity", generator=gen)                   # reference gains
        ref_servo = lambda gap: np.float32(np.clip((gap - 25.0) * 50.0, -1000.0, 1000.0))
    tape = _tape(E, STEPS * SI, key=777)
    env = BatchedWireEDMEnv(E, config=EnvironmentConfig(**cfg), mechanics_control_mode=mode,
                            wire_params=WireModuleParameters(segment_len=seg_len), latch_first_action=False)
    env.reset()
    case = {"config": cfg, "mode": mode, "wire": {"segment_len": seg_len}}
    oracles = [O.OracleEnv(**CS.oracle_overrides(case)) for _ in range(E)]
    tape_dev = torch.from_numpy(tape).cuda()
    for k in range(STEPS):
        env.step(ctrl if in_kernel else ctrl(env), draws=tape_dev[:, k * SI:(k + 1) * SI].contiguous())
        f_ref, i_ref = [], []
        for e, orc in enumerate(oracles):
            f, _ = orc.record()
            servo = ref_servo(f[0] - f[1])
            orc.run(SI, [[float(servo), 80.0, 5, 3.0, 80.0]], draws=tape[e, k * SI:(k + 1) * SI], rec_every=0)
            f, i = orc.record()
            f_ref.append(f)
            i_ref.append(i)
        f_gpu, i_gpu = PU.read_rows(env.engine, range(E))
        PU.assert_rows_match(f_gpu, i_gpu, np.stack(f_ref), np.stack(i_ref).astype(np.int64),
                             where=f"{mode} control step {k}")
    assert float(env.state_tensor("stat_sparks").sum()) > 0


def test_voltage_controller_closed_loop():
    """PI voltage controller: the integral state and the command follow the reference's update
    (run_simulation.py:74-98) restated in numpy on the mean voltages the env reports."""
    import torch
    from sparc_b200 import BatchedWireEDMEnv, EnvironmentConfig, VoltageController

    E = 256
    env = BatchedWireEDMEnv(E, config=EnvironmentConfig(**README), seed=3)
    env.reset()
    ctrl = VoltageController(30.0, "position")
    integ = np.zeros(E)
    for k in range(15):
        mv = env.mean_voltage().cpu().numpy()
        act = ctrl(env)
        err = 30.0 - mv
        integ = np.clip(integ + err, -100.0, 100.0)
        want = np.clip(-(0.05 * err + 0.1 * integ * 0.001), -5.0, 5.0).astype(np.float32)
        np.testing.assert_array_equal(act["servo"].cpu().numpy(), want)
        np.testing.assert_array_equal(ctrl.integral.cpu().numpy(), integ)
        obs, *_ = env.step(act)
        # the observation's mean_voltage is the float32 view of the same average
        np.testing.assert_allclose(obs["mean_voltage"].cpu().numpy(), env.mean_voltage().cpu().numpy(), rtol=1e-6)
    assert torch.all(env.state_tensor("time") == 15000)
    mv = env.mean_voltage()
    assert float(mv.min()) >= 0.0 and float(mv.max()) <= 80.0 + 1e-9
    ctrl.reset(torch.arange(E, device=env.device) < 10)
    assert float(ctrl.integral[:10].abs().sum()) == 0.0


def _short_episode_env(n, **kw):
    from sparc_b200 import EnvironmentConfig, WireEDMVectorEnv

    cfg = dict(README, target_cutting_distance=50.0005)      # the first spark reaches the target
    return WireEDMVectorEnv(n, config=EnvironmentConfig(**cfg), seed=9, **kw)


def test_vector_env_spaces_and_next_step_autoreset():
    n = 128
    venv = _short_episode_env(n, reward="progress")
    assert venv.num_envs == n and venv.metadata["autoreset_mode"] == "next_step"
    assert venv.single_observation_space.shape == (7,) and venv.observation_space.shape == (n, 7)
    assert set(venv.single_action_space.spaces) == {"servo", "generator_control"}
    obs, info = venv.reset(seed=9)
    assert isinstance(obs, np.ndarray) and obs.shape == (n, 7) and obs.dtype == np.float32 and info == {}
    assert np.all(obs[:, 0] == 50.0)
    act = venv.action_space.sample()
    act["servo"][:] = 0.05
    act["generator_control"]["current_mode"][:] = 5
    act["generator_control"]["target_voltage"][:] = 80.0
    act["generator_control"]["ON_time"][:] = 3.0
    act["generator_control"]["OFF_time"][:] = 80.0
    prev_term = np.zeros(n, dtype=bool)
    seen_term = 0
    for k in range(12):
        obs, rew, term, trunc, info = venv.step(act)
        assert obs.shape == (n, 7) and rew.shape == (n,) and term.dtype == bool and not trunc.any()
        # an env that was terminal in the previous call comes back reset: gap 50, time 0, reward 0
        if prev_term.any():
            assert not term[prev_term].any()
            assert np.all(obs[prev_term, 0] == 50.0) and np.all(info["time"][prev_term] == 0)
            assert np.all(rew[prev_term] == 0.0)
        assert np.all(info["target_reached"] == term)        # the only way to end in this config
        assert np.all(rew[term] > 0.0)                        # the spark's workpiece advance
        seen_term += int(term.sum())
        prev_term = term.copy()
    assert seen_term > n // 4


def test_vector_env_same_step_truncation_and_disabled():
    import torch

    n = 64
    venv = _short_episode_env(n, autoreset_mode="same_step", max_episode_steps=3, to_numpy=False)
    obs, _ = venv.reset()
    assert isinstance(obs, torch.Tensor) and obs.is_cuda
    act = {"servo": 0.0, "generator_control": {"target_voltage": 80.0, "current_mode": 5, "ON_time": 3.0, "OFF_time": 80.0}}
    ends = 0
    for k in range(9):
        obs, rew, term, trunc, info = venv.step(act)
        done = term | trunc
        assert not (term & trunc).any()
        if done.any():
            assert torch.all(obs[done, 0] == 50.0)                       # reset observation
            assert torch.equal(info["_final_obs"], done)
            assert torch.all(info["time"][term] > 0)                       # info describes the final state
            ends += int(done.sum())
        assert torch.all(rew == 0.0)                                       # reference placeholder reward
    assert ends >= n * 2                                                   # every env ends at least every 3 steps
    ven2 = _short_episode_env(8, autoreset_mode="disabled")
    ven2.reset()
    for _ in range(10):
        obs, rew, term, trunc, info = ven2.step(act)
    assert term.all() and np.all(info["time"] < 10000)                    # frozen once terminal, never reset
    with pytest.raises(ValueError):
        _short_episode_env(4, autoreset_mode="sometimes")


def test_grouped_env_equals_separate_homogeneous_batches():
    """Heterogeneous batch = groups of homogeneous ones on concurrent streams: every group evolves
    exactly as a stand-alone BatchedWireEDMEnv with the same global env ids (different geometry,
    segment count, control mode and kernel variant per group)."""
    import torch
    from sparc_b200 import (BatchedWireEDMEnv, EnvironmentConfig, GroupedWireEDMEnv, WireModuleParameters)

    thick = dict(workpiece_height=50.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)
    groups = [
        dict(n_envs=96, config=EnvironmentConfig(**README), mechanics_control_mode="position"),
        dict(n_envs=40, config=EnvironmentConfig(**thick), mechanics_control_mode="velocity"),
        dict(n_envs=24, config=EnvironmentConfig(**thick), mechanics_control_mode="position",
             wire_params=WireModuleParameters(segment_len=0.1)),
    ]
    genv = GroupedWireEDMEnv(groups, seed=21, env_id_base=1000)
    assert genv.n_envs == 160 and [e.n_segments for e in genv.envs] == [375, 550, 1100]
    assert genv.group_of(0) == 0 and genv.group_of(96) == 1 and genv.group_of(159) == 2
    solo, base = [], 1000
    for g in groups:
        kw = dict(g)
        n = kw.pop("n_envs")
        solo.append(BatchedWireEDMEnv(n, seed=21, env_id_base=base, **kw))
        base += n
    obs, _ = genv.reset()                       # a reset starts a new episode (new Philox streams) ...
    for env in solo:
        env.reset()                             # ... so the stand-alone batches are reset once too
    assert obs["gap"].shape == (160,)
    servo = torch.linspace(-0.5, 0.5, 160, device=genv.device)
    servo[96:136] *= 400.0                                              # velocity-mode group: µm/s
    mode = torch.full((160,), 5, dtype=torch.int32, device=genv.device)
    mode[::3] = 13
    act = {"servo": servo, "generator_control": {"target_voltage": 80.0, "current_mode": mode, "ON_time": 3.0,
                                                 "OFF_time": 40.0}}
    for _ in range(5):
        obs, rew, term, trunc, info = genv.step(act)
        for sl, env in zip(genv.group_slices(), solo):
            env.step({"servo": servo[sl], "generator_control": {"target_voltage": 80.0, "current_mode": mode[sl],
                                                                 "ON_time": 3.0, "OFF_time": 40.0}})
    assert obs["gap"].shape == (160,) and term.shape == (160,) and info["time"].shape == (160,)
    for genv_e, env in zip(genv.envs, solo):
        for name in ("wp", "wire", "v", "debris_volume", "time", "spark_state", "stat_sparks", "flags"):
            assert torch.equal(genv_e.state_tensor(name), env.state_tensor(name)), name
        assert torch.equal(genv_e.wire_temperature, env.wire_temperature)
    assert torch.equal(genv.state_tensor("wp"), torch.cat([e.state_tensor("wp") for e in solo]))
    st = genv.episode_stats()
    assert st["env_us"] == 160 * 5000 and st["sparks"] > 0
