"""The reference's own test-suite contract (tests/test_state.py:10-28,
tests/test_env_integration.py:11-65), re-stated against the drop-in classes."""
import numpy as np
import pytest


def test_state_initialization():
    from sparc_b200 import EDMState

    s = EDMState()
    assert s.time == 0
    assert s.workpiece_position == 0.0
    assert s.wire_position == 0.0


def test_state_attributes():
    from sparc_b200.core.state import EDMState

    s = EDMState()
    s.time = 1000
    s.workpiece_position = 50.0
    assert s.time == 1000 and s.workpiece_position == 50.0
    assert s.spark_status == [0, None, 0] and s.wire_unwinding_velocity == 0.2 and s.target_position == 500.0


def test_import_paths_match_reference_layout():
    import sparc_b200
    from sparc_b200.core import EDMModule, EnvironmentConfig, get_material_db  # noqa: F401
    from sparc_b200.modules.ignition import IgnitionModuleParameters
    from sparc_b200.modules.wire import WireModule, WireModuleParameters  # noqa: F401
    from sparc_b200.modules import MaterialRemovalModule, DielectricModule, MechanicsModule  # noqa: F401

    assert IgnitionModuleParameters().default_current_mode == "I5"
    assert WireModuleParameters().segment_len == 0.2
    assert set(EDMModule.registry) >= {"IgnitionModule", "WireModule", "MaterialRemovalModule",
                                       "DielectricModule", "MechanicsModule"}
    assert get_material_db().get_wire_material("brass").breaking_temperature == 1500
    assert "WireEDMEnv" in sparc_b200.__all__


def test_invalid_mechanics_mode_raises_without_gpu():
    """Validation happens before any device work (wire_edm.py:39-42)."""
    from sparc_b200 import WireEDMEnv

    with pytest.raises(ValueError, match="mechanics_control_mode"):
        WireEDMEnv(mechanics_control_mode="invalid_mode")


@pytest.mark.gpu
def test_env_initialization():
    from sparc_b200 import WireEDMEnv

    env = WireEDMEnv()
    assert env is not None and env.mechanics_control_mode == "position"
    env = WireEDMEnv(mechanics_control_mode="velocity")
    assert env.mechanics_control_mode == "velocity"
    assert set(env.modules) == {"ignition", "material", "dielectric", "wire", "mechanics"}
    assert env.wire.n_segments == 400 and env.workpiece_height == 20.0


@pytest.mark.gpu
def test_env_reset_and_step():
    from sparc_b200 import WireEDMEnv

    env = WireEDMEnv()
    obs, info = env.reset()
    assert obs == {} and info == {}
    assert env.state.time == 0
    assert env.state.workpiece_position == env.config.initial_gap
    env.action_space["generator_control"]["current_mode"]  # structure as in the reference
    action = env.action_space.sample()
    obs, reward, terminated, truncated, info = env.step(action)
    assert env.state.time > 0
    assert isinstance(reward, float) and isinstance(terminated, bool) and isinstance(truncated, bool)
    assert isinstance(info, dict)
    assert set(info) == {"wire_broken", "target_reached", "spark_state", "time", "control_step"}
    assert obs is None and info["control_step"] is False and info["time"] == 1


@pytest.mark.gpu
def test_env_with_custom_config_and_state_view():
    from sparc_b200 import EnvironmentConfig, WireEDMEnv

    config = EnvironmentConfig(workpiece_height=15.0, wire_diameter=0.25, initial_gap=30.0)
    env = WireEDMEnv(config=config)
    assert env.config.workpiece_height == 15.0 and env.config.wire_diameter == 0.25
    env.reset(seed=3)
    assert env.state.workpiece_position == 30.0
    assert env.state.wire_temperature.shape == (375,) and env.state.wire_temperature.dtype == np.float32
    assert env.state.spark_status == [0, None, 0] and env.state.current_mode is None
    env.state.workpiece_position = 42.5          # scripts poke the state (run_simulation.py:199-203)
    assert env.state.workpiece_position == 42.5
    a = {"servo": np.array([0.1], np.float32),
         "generator_control": {"target_voltage": np.array([80.0], np.float32), "current_mode": np.array([5], np.int32),
                               "ON_time": np.array([3.0], np.float32), "OFF_time": np.array([80.0], np.float32)}}
    for _ in range(3):
        env.step(a)
    obs, rew, term, trunc, info = env.step_control(a, 997)   # 997 µs in one launch
    assert info["time"] == 1000 and obs == {} and not term
    obs, rew, term, trunc, info = env.step(a)                # the µs that latches the action
    assert info["control_step"] is True and obs == {}
    assert env.state.current_mode == "I5" and env.state.target_voltage == 80.0


@pytest.mark.gpu
def test_even_mode_raises_value_error_like_reference():
    from sparc_b200 import EnvironmentConfig, WireEDMEnv

    env = WireEDMEnv(config=EnvironmentConfig(workpiece_height=15.0, wire_diameter=0.25, initial_gap=12.0))
    env.reset(seed=1)
    a = {"servo": np.array([0.0], np.float32),
         "generator_control": {"target_voltage": np.array([80.0], np.float32), "current_mode": np.array([6], np.int32),
                               "ON_time": np.array([3.0], np.float32), "OFF_time": np.array([10.0], np.float32)}}
    with pytest.raises(ValueError, match="Current mode I6 is not available in crater data"):
        for _ in range(50):
            env.step_control(a)
