"""Controller arithmetic on CPU tensors against a scalar restatement of the reference's closures
(experiments/run_simulation.py:24-109): float64 error, gain, clip, then float32."""
import numpy as np
import pytest
import torch

from sparc_b200.controllers import GapController, VoltageController


class _FakeEnv:
    def __init__(self, wp, wire, mv):
        self._t = {"wp": torch.tensor(wp, dtype=torch.float64), "wire": torch.tensor(wire, dtype=torch.float64)}
        self._mv = torch.tensor(mv, dtype=torch.float64)

    def state_tensor(self, k):
        return self._t[k]

    def mean_voltage(self):
        return self._mv


@pytest.mark.parametrize("mode", ["position", "velocity"])
def test_gap_controller_matches_reference_arithmetic(mode):
    rng = np.random.default_rng(0)
    wp = rng.uniform(40, 900, 500)
    wire = wp - rng.uniform(-5, 120, 500)
    ctrl = GapController(5.0, mode)
    act = ctrl(_FakeEnv(wp, wire, np.zeros(500)))
    for k in range(500):
        error = (wp[k] - wire[k]) - 5.0
        if mode == "position":
            delta = error * 0.1
        else:
            delta = np.clip(error * 50.0, -1000.0, 1000.0)
        assert act["servo"][k].item() == np.array([delta], dtype=np.float32)[0]
    gc = act["generator_control"]
    assert gc["current_mode"].dtype == torch.int32 and int(gc["current_mode"][0]) == 7
    assert float(gc["ON_time"][0]) == 2.0 and float(gc["OFF_time"][0]) == 33.0 and float(gc["target_voltage"][0]) == 80.0


@pytest.mark.parametrize("mode", ["position", "velocity"])
def test_voltage_controller_matches_reference_arithmetic(mode):
    rng = np.random.default_rng(1)
    ctrl = VoltageController(30.0, mode)
    integral = np.zeros(16)
    for step in range(40):
        mv = rng.uniform(0, 80, 16)
        act = ctrl(_FakeEnv(np.zeros(16), np.zeros(16), mv))
        for k in range(16):
            error = 30.0 - mv[k]
            integral[k] = np.clip(integral[k] + error, -100.0, 100.0)
            out = -(0.05 * error + 0.1 * integral[k] * 0.001)
            delta = np.clip(out, -5.0, 5.0) if mode == "position" else np.clip(out * 100.0, -1000.0, 1000.0)
            assert act["servo"][k].item() == np.array([delta], dtype=np.float32)[0]
    ctrl.reset()
    assert float(ctrl.integral.abs().sum()) == 0.0


def test_bad_control_mode():
    with pytest.raises(ValueError):
        GapController(5.0, "force")
    with pytest.raises(ValueError):
        VoltageController(30.0, "force")
