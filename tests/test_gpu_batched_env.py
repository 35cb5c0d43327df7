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


def _welch(gpu, ref):
    gpu, ref = np.asarray(gpu, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    se = np.sqrt(gpu.var(ddof=1) / len(gpu) + ref.var(ddof=1) / len(ref))
    return (gpu.mean() - ref.mean()) / se if se > 0 else 0.0


def _prop_z(k1, n1, k2, n2):
    """Two-proportion z statistic (pooled)."""
    p = (k1 + k2) / (n1 + n2)
    se = np.sqrt(p * (1 - p) * (1 / n1 + 1 / n2))
    return (k1 / n1 - k2 / n2) / se if se > 0 else 0.0


Z999 = 3.29          # two-sided 99.9 % ("accept if the GPU mean is within the reference's 99.9 % CI")


def test_statistics_match_running_reference():
    """On-device RNG (SURVEY 8d statistical parity): sparks per env, removal, short-circuit fraction, P(target
    reached) and time to break must agree with the UNMODIFIED reference run with its own NumPy generator
    (tests/golden/stats_reference.json: cfg2 256 envs x 100 ms, closed-loop cfg3 128 x 100 ms, debris shorts
    96 x 30 ms, target 160 x 40 ms, hot 64 x <= 30 ms) — Welch z for means, pooled z for proportions, |z| < 3.29."""
    import torch
    from sparc_b200 import (BatchedWireEDMEnv, DielectricModuleParameters, EnvironmentConfig, GapController,
                            IgnitionModuleParameters)
    from oracle import cases as CS

    ref = json.load(open(os.path.join(GOLDEN_DIR, "stats_reference.json")))
    n = 4096

    def check_means(tag, env, r, keys=("sparks", "removed_um")):
        wp0 = env.state_tensor("prm_initial_gap").double().cpu().numpy()
        gpu = {"sparks": env.state_tensor("stat_sparks").double().cpu().numpy(),
               "removed_um": env.state_tensor("wp").double().cpu().numpy() - wp0,
               "steps": env.state_tensor("stat_steps").double().cpu().numpy(),
               "short_us": env.state_tensor("stat_short_us").double().cpu().numpy()}
        for key in keys:
            z = _welch(gpu[key], [x[key] for x in r])
            print(f"\n{tag} {key}: gpu {gpu[key].mean():.5f} ref {np.mean([x[key] for x in r]):.5f} z {z:+.2f}")
            assert abs(z) < Z999, (tag, key, z)
        return gpu

    # ---- cfg2: README config, quickstart action, 100 ms
    env = _env(n, seed=2026, latch_first_action=False)      # same protocol as the reference runs
    act = env.constant_action(servo=np.float32(0.1))
    for _ in range(ref["n_us"]["cfg2"] // 1000):
        env.step(act)
    g = check_means("cfg2", env, ref["cfg2"])
    assert g["short_us"].sum() == 0 and sum(x["short_us"] for x in ref["cfg2"]) == 0

    # ---- cfg3: closed loop, stabilising sign, controller inside the kernels (gap -> 25 um, ~2x the spark rate)
    env = _env(n, seed=31, latch_first_action=False)
    ctrl = GapController(25.0, "position", limit=1.0, generator=dict(target_voltage=80.0, current_mode=5, ON_time=3.0, OFF_time=80.0))
    for _ in range(ref["n_us"]["cfg3"] // 1000):
        env.step(ctrl)
    check_means("cfg3", env, ref["cfg3"])
    gap = (env.state_tensor("wp") - env.state_tensor("wire")).double().cpu().numpy()
    z = _welch(gap, [x["gap"] for x in ref["cfg3"]])
    print(f"\ncfg3 final gap: gpu {gap.mean():.3f} ref {np.mean([x['gap'] for x in ref['cfg3']]):.3f} z {z:+.2f}")
    assert abs(z) < Z999

    # ---- shorts: debris-choked gap, short-circuit fraction ~0.8
    sc = CS.CASES["debris_shorts"]
    env = BatchedWireEDMEnv(n, config=EnvironmentConfig(**sc["config"]), ignition_params=IgnitionModuleParameters(**sc["ignition"]),
                            dielectric_params=DielectricModuleParameters(**sc["dielectric"]), seed=5, latch_first_action=False)
    env.reset()
    a = sc["actions"][0]
    act = env.constant_action(servo=a[0], target_voltage=a[1], current_mode=a[2], ON_time=a[3], OFF_time=a[4])
    for _ in range(ref["n_us"]["shorts"] // 1000):
        env.step(act)
    g = check_means("shorts", env, ref["shorts"], keys=("sparks", "removed_um", "short_us"))
    frac_gpu = g["short_us"].sum() / g["steps"].sum()
    frac_ref = sum(x["short_us"] for x in ref["shorts"]) / sum(x["steps"] for x in ref["shorts"])
    print(f"\nshort-circuit fraction: gpu {frac_gpu:.4f} ref {frac_ref:.4f}")
    assert frac_ref > 0.5 and abs(frac_gpu - frac_ref) < 0.02

    # ---- target: P(target reached within 40 ms) and time to reach it
    env = _env(n, config=dict(README, target_cutting_distance=50.02), seed=9, latch_first_action=False)
    act = env.constant_action(servo=np.float32(0.1))
    for _ in range(ref["n_us"]["target"] // 1000):
        env.step(act)
    reached = int(((env.state_tensor("flags") & 4) != 0).sum().item())
    r_reached = sum(bool(x["target"]) for x in ref["target"])
    z = _prop_z(reached, n, r_reached, len(ref["target"]))
    print(f"\nP(target): gpu {reached / n:.3f} ref {r_reached / len(ref['target']):.3f} z {z:+.2f}")
    assert abs(z) < Z999 and 0.2 < reached / n < 0.95
    check_means("target", env, ref["target"], keys=("steps",))

    # ---- hot: aggressive generator, every env breaks; compare time to break
    hot = _env(n, config=dict(README, initial_gap=12.0), seed=77, latch_first_action=False)
    act = hot.constant_action(servo=0.0, target_voltage=100.0, current_mode=17, ON_time=5.0, OFF_time=6.0)
    for _ in range(ref["n_us"]["hot"] // 1000):
        hot.step(act)
    broken = ((hot.state_tensor("flags") & 2) != 0).double().cpu().numpy()
    r_broken = np.array([x["broken"] for x in ref["hot"]], dtype=np.float64)
    assert abs(broken.mean() - r_broken.mean()) < 0.05
    check_means("hot", hot, ref["hot"], keys=("steps", "sparks"))


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
