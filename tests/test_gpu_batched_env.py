"""The vectorised surface on the GPU: API shapes, controllers on device, auto-reset,
GPU-count invariance of the Philox streams, statistics against the running reference."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR

pytestmark = pytest.mark.gpu

README = dict(workpiece_height=15.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)


def _env(n, **kw):
    from sparc_b200 import BatchedWireEDMEnv, EnvironmentConfig

    cfg = kw.pop("config", README)
    return BatchedWireEDMEnv(n, config=EnvironmentConfig(**cfg), **kw)


def test_batched_api_shapes_and_gap_controller():
    """BASELINE cfg 3 flavour: position mode, proportional gap controller evaluated on device."""
    import torch

    env = _env(512, seed=5)
    obs, info = env.reset()
    assert set(obs) == {"gap", "wire_velocity", "mean_voltage", "sparks", "short_us", "max_temperature", "debris_density"}
    assert obs["gap"].shape == (512,) and torch.all(obs["gap"] == 50.0)
    for _ in range(60):
        servo = torch.clamp(0.1 * (obs["gap"] - 25.0), -1.0, 1.0)        # run_simulation.py:28-33 sign
        act = {"servo": servo, "generator_control": {"target_voltage": 80.0, "current_mode": 5, "ON_time": 3.0, "OFF_time": 80.0}}
        obs, rew, term, trunc, info = env.step(act)
    assert rew.shape == (512,) and term.dtype == torch.bool and not trunc.any()
    assert int(info["time"][0]) == 60000 and not term.any()
    g = obs["gap"].mean().item()
    assert 25.0 < g < 45.0, g                      # the stabilising controller closes the gap
    st = env.episode_stats()
    assert st["env_us"] == 512 * 60000 and st["sparks"] > 512 * 10


def test_streams_do_not_depend_on_sharding():
    """Global env ids key the Philox streams: 2 shards of 64 == one batch of 128 (cfg 4 invariant)."""
    import torch

    a = _env(128, seed=11)
    b0 = _env(64, seed=11, env_id_base=0)
    b1 = _env(64, seed=11, env_id_base=64)
    act = a.constant_action(servo=0.1)
    for _ in range(6):
        a.step(act); b0.step(act); b1.step(act)
    for name in ("wp", "wire", "v", "debris_volume", "time", "spark_state", "stat_sparks"):
        whole = a.state_tensor(name)
        parts = torch.cat([b0.state_tensor(name), b1.state_tensor(name)])
        assert torch.equal(whole, parts), name
    assert torch.equal(a.wire_temperature, torch.cat([b0.wire_temperature, b1.wire_temperature]))
    assert a.state_tensor("stat_sparks").sum() > 0


def test_auto_reset_and_first_latch():
    import torch

    env = _env(64, config=dict(README, initial_gap=10.0, target_cutting_distance=10.004), seed=3, auto_reset=True)
    act = env.constant_action(servo=0.0, OFF_time=20.0)
    seen = torch.zeros(64, dtype=torch.bool, device=env.device)
    for _ in range(40):
        obs, rew, term, trunc, info = env.step(act)
        seen |= info["final_terminated"]
        # a reset env is fresh again: time 0, latch armed
        if info["final_terminated"].any():
            i = int(info["final_terminated"].nonzero()[0])
            assert int(env.state_tensor("time")[i]) == 0 and int(env.state_tensor("episode")[i]) >= 1
    assert seen.float().mean() > 0.9
    # latch_first_action: the very first control step already runs on the action's generator settings
    env2 = _env(4, seed=1)
    env2.step(env2.constant_action(target_voltage=123.0))
    assert torch.all(env2.state_tensor("target_voltage") == 123.0)
    env3 = _env(4, seed=1, latch_first_action=False)        # reference quirk: first ms on defaults
    env3.step(env3.constant_action(target_voltage=123.0))
    assert torch.all(env3.state_tensor("target_voltage") == 0.0)


def test_statistics_match_running_reference():
    """On-device RNG: spark rate and removal per env must agree with the unmodified reference
    (tests/golden/stats_reference.json) within a 99.9 % confidence interval (|z| < 3.29)."""
    ref = json.load(open(os.path.join(GOLDEN_DIR, "stats_reference.json")))
    n = 4096
    env = _env(n, seed=2026, latch_first_action=False)      # same protocol as the reference runs
    act = env.constant_action(servo=np.float32(0.1))
    for _ in range(50):
        env.step(act)
    sparks = env.state_tensor("stat_sparks").double().cpu().numpy()
    removed = (env.state_tensor("wp").double().cpu().numpy() - 50.0)
    for name, gpu, key in (("sparks", sparks, "sparks"), ("removed_um", removed, "removed_um")):
        r = np.array([x[key] for x in ref["cfg2"]], dtype=np.float64)
        z = (gpu.mean() - r.mean()) / np.sqrt(gpu.var(ddof=1) / len(gpu) + r.var(ddof=1) / len(r))
        print(f"\n{name}: gpu {gpu.mean():.4f} ref {r.mean():.4f} z {z:+.2f}")
        assert abs(z) < 3.29, (name, gpu.mean(), r.mean(), z)
    # wire-break statistics: aggressive generator, every env breaks; compare time to break
    hot = _env(n, config=dict(README, initial_gap=12.0), seed=77, latch_first_action=False)
    act = hot.constant_action(servo=0.0, target_voltage=100.0, current_mode=17, ON_time=5.0, OFF_time=6.0)
    for _ in range(30):
        hot.step(act)
    steps = hot.state_tensor("stat_steps").double().cpu().numpy()
    broken = ((hot.state_tensor("flags") & 2) != 0).double().cpu().numpy()
    r_steps = np.array([x["steps"] for x in ref["hot"]], dtype=np.float64)
    r_broken = np.array([x["broken"] for x in ref["hot"]], dtype=np.float64)
    z = (steps.mean() - r_steps.mean()) / np.sqrt(steps.var(ddof=1) / n + r_steps.var(ddof=1) / len(r_steps))
    print(f"\nbreak: gpu P {broken.mean():.3f} mean t {steps.mean():.0f} us; ref P {r_broken.mean():.3f} mean t {r_steps.mean():.0f} us; z {z:+.2f}")
    assert abs(broken.mean() - r_broken.mean()) < 0.05
    assert abs(z) < 3.29


def test_batched_logger_npz_layout(tmp_path):
    """SimulationLogger equivalent: reference config dict and signal names; `.npz` arrays are
    [n_samples, n_envs(, n_segments)]; kernel-sampled temperature rows at a sub-control-step interval."""
    import numpy as np
    from sparc_b200 import BatchedWireEDMEnv, EnvironmentConfig
    from sparc_b200.utils.logger import SimulationLogger

    cfg = EnvironmentConfig(workpiece_height=15.0, wire_diameter=0.25, target_cutting_distance=800.0)
    venv = BatchedWireEDMEnv(8, config=cfg, seed=3)
    venv.reset()
    path = str(tmp_path / "log.npz")
    logger = SimulationLogger({"signals_to_log": ["workpiece_position", "wire_position", "voltage", "is_short_circuit",
                                                  "wire_temperature"],
                               "log_frequency": {"type": "interval", "value": 250},
                               "backend": {"type": "numpy", "filepath": path}}, env_reference=venv)
    act = venv.constant_action(servo=0.1)
    for _ in range(3):
        _, _, _, _, info = venv.step(act, log_temperature_every=250)
        logger.collect(info)
    logger.finalize()
    d = np.load(path)
    assert d["wire_temperature"].shape == (12, 8, venv.n_segments)          # 3 steps x 4 rows
    assert d["workpiece_position"].shape == (3, 8) and d["is_short_circuit"].dtype == np.bool_
    np.testing.assert_array_equal(d["wire_temperature"][-1], venv.wire_temperature.cpu().numpy())
    np.testing.assert_array_equal(d["wire_position"][-1], venv.state_tensor("wire").cpu().numpy())
    with pytest.raises(ValueError):
        SimulationLogger({"signals_to_log": [], "log_frequency": {"type": "control_step"}, "backend": {"type": "memory"}})


def test_logger_every_step_uses_the_device_signal_log(tmp_path):
    """SimulationLogger "every_step": one sample per µs of scalars and of the temperature field,
    produced inside the step call (reference: collect() after every step(), utils/logger.py:118-160)."""
    import torch
    from sparc_b200.utils import BatchedSimulationLogger

    env = _env(3, seed=2)
    env.reset()
    path = str(tmp_path / "every_step.npz")
    lg = BatchedSimulationLogger({"signals_to_log": ["time", "voltage", "wire_position", "is_short_circuit",
                                                     "debris_concentration", "wire_temperature"],
                                  "log_frequency": {"type": "every_step"},
                                  "backend": {"type": "numpy", "filepath": path}}, env_reference=env)
    kw = lg.step_kwargs()
    assert kw == {"log_signals_every": 1, "log_temperature_every": 1}
    act = env.constant_action(servo=0.1)
    for _ in range(2):
        _, _, _, _, info = env.step(act, **kw)
        lg.collect(info)
    lg.finalize()
    d = np.load(path)
    assert d["time"].shape == (2000, 3) and d["time"].dtype == np.int32
    assert np.array_equal(d["time"][:, 0], np.arange(1, 2001))
    assert d["wire_temperature"].shape == (2000, 3, env.n_segments)
    assert d["is_short_circuit"].dtype == bool and d["voltage"].max() <= 80.0
    # the last sample is the live state
    assert np.array_equal(d["wire_position"][-1], env.state_tensor("wire").cpu().numpy())
    assert np.array_equal(d["wire_temperature"][-1], env.wire_temperature.cpu().numpy())
    # interval logging below one control step
    lg2 = BatchedSimulationLogger({"signals_to_log": ["voltage"], "log_frequency": {"type": "interval", "value": 250},
                                   "backend": {"type": "memory"}}, env_reference=env)
    assert lg2.step_kwargs() == {"log_signals_every": 250}
    _, _, _, _, info = env.step(act, **lg2.step_kwargs())
    lg2.collect(info)
    assert len(lg2.get_data()["voltage"]) == 4 and lg2.get_data()["voltage"][0].shape == (3,)


@pytest.mark.parametrize("seg_len", [0.2, 0.05])
def test_fused_observation_equals_observe_kernel(seg_len):
    """The observation written by the step kernel's epilogue (sparc_step_observe, sparc_step_host) is
    bit-identical to sparc_observe on the stored state: one warp per env (N=375) and several warps
    per env (N=2200), running and terminated envs, and the host-buffer path."""
    import torch
    from sparc_b200 import BatchedWireEDMEnv, EnvironmentConfig, WireModuleParameters

    cfg = dict(README, target_cutting_distance=50.002) if seg_len == 0.2 else \
        dict(workpiece_height=50.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=50.0005)
    env = BatchedWireEDMEnv(96, config=EnvironmentConfig(**cfg), seed=4,
                            wire_params=WireModuleParameters(segment_len=seg_len))
    env.reset()
    act = env.constant_action(servo=0.3, current_mode=13, OFF_time=30.0)
    saw_done = False
    for k in range(6):
        obs, _, term, _, _ = env.step(act)
        fused = torch.stack([obs[key] for key in obs]).clone()
        ref = env.engine.observe()[:7].clone()
        assert torch.equal(fused, ref), k
        saw_done = saw_done or bool(term.any())
    assert saw_done and not bool(term.all())
    # host-buffer path: observation from the epilogue, copied to pinned host memory
    ha = env.engine.host_actions
    ha["servo"].fill_(0.3); ha["target_voltage"].fill_(80.0); ha["current_mode"].fill_(13)
    ha["on_time"].fill_(3.0); ha["off_time"].fill_(30.0)
    oh = env.step_host().clone()
    assert torch.equal(oh, env.engine.observe().cpu())


def test_reset_with_seed_replays_the_same_episode():
    """Gymnasium contract (the reference re-seeds np_random in reset(seed=s), wire_edm.py:106-107):
    two rollouts after reset(seed=s) on ONE instance are bit-identical; an unseeded reset starts a
    new episode with a different stream."""
    import torch
    from sparc_b200 import BatchedWireEDMEnv, EnvironmentConfig

    env = BatchedWireEDMEnv(256, config=EnvironmentConfig(**README), seed=1)
    act = env.constant_action(0.1)

    def rollout():
        for _ in range(12):
            env.step(act)
        return {k: env.state_tensor(k).clone() for k in ("wp", "wire", "debris_volume", "stat_sparks", "time", "T")}

    env.reset(seed=5)
    a = rollout()
    env.reset(seed=5)
    b = rollout()
    for k in a:
        assert torch.equal(a[k], b[k]), k
    assert int(a["stat_sparks"].sum()) > 0
    env.reset()                                   # new episode, same seed: a different stream
    c = rollout()
    assert not torch.equal(a["wp"], c["wp"])
    env.reset(seed=5, mask=torch.arange(256, device=env.device) < 128)
    assert int(env.state_tensor("episode")[:128].abs().sum()) == 0


def test_reference_style_logger_on_the_dropin_env():
    """`logger.collect(env.state, info)` after every 1-us step(), exactly as the reference's drivers call
    it (utils/logger.py:118-160, run_simulation.py:245-297), works on the drop-in env; `dielectric_flow_rate`
    carries the reference's scaling flow * base_flow_rate / 1e9 (dielectric.py:158)."""
    from sparc_b200 import EnvironmentConfig, WireEDMEnv
    from sparc_b200.utils.logger import SimulationLogger

    env = WireEDMEnv(config=EnvironmentConfig(**README))
    env.reset(seed=3)
    logger = SimulationLogger({"signals_to_log": ["time", "voltage", "dielectric_flow_rate", "wire_temperature"],
                               "log_frequency": {"type": "interval", "value": 5}, "backend": {"type": "memory"}}, env)
    act = env.action_space.sample()
    for _ in range(40):
        _, _, _, _, info = env.step(act)
        logger.collect(env.state, info)
    data = logger.get_data()
    assert [int(t) for t in data["time"]] == list(range(5, 45, 5))
    assert len(data["wire_temperature"]) == 8 and data["wire_temperature"][0].shape == (375,)
    assert abs(data["dielectric_flow_rate"][-1] - env.state.flow_rate * 100.0 / 1e9) < 1e-18
