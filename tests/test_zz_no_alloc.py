"""`sparc_step_host_static`: the allocation-free variant of `sparc_step_host` (scratch in a file-scope
`static __device__` array inside the library; no counterpart in the reference).
Called through the C-ABI and compared with the existing entry point on two calls in a row.

This file is named to be collected LAST: a failure here hides nothing.
"""
import ctypes as C

import numpy as np
import pytest

import parity_util as PU
from oracle import cases as CS

pytestmark = pytest.mark.gpu

SEED = 0xA110C0DE


def _engines(n_envs):
    from sparc_b200.engine import Engine

    prm = PU.case_params(CS.CASES["cfg2_default"])
    a = Engine(n_envs, prm, seed=SEED, env_id_base=11)
    b = Engine(n_envs, PU.case_params(CS.CASES["cfg2_default"]), seed=SEED, env_id_base=11)
    for eng in (a, b):                               # arm the latch: the first call latches its action
        eng["t_servo"].fill_(1000)
    return a, b


def _fill(host_actions, k, n_envs):
    rng = np.random.default_rng(100 + k)
    host_actions["servo"].numpy()[:] = rng.uniform(-0.2, 0.5, n_envs).astype(np.float32)
    host_actions["target_voltage"].numpy()[:] = 80.0
    host_actions["current_mode"].numpy()[:] = rng.choice([3, 5, 7], n_envs).astype(np.int32)
    host_actions["on_time"].numpy()[:] = 3.0
    host_actions["off_time"].numpy()[:] = 40.0 + 10.0 * k


@pytest.mark.parametrize("n_envs", [8, 4096])
def test_static_scratch_step_matches_step_host_on_two_calls(n_envs):
    import torch
    from sparc_b200 import _native as N

    L = N.lib()
    assert L.sparc_static_max_envs() == 16384
    ref, sta = _engines(n_envs)
    obs_static = torch.zeros(N.N_OBS, n_envs, dtype=torch.float32).pin_memory()
    acts = sta._actions_struct(sta.host_actions)
    for k in range(2):                               # two calls in a row: the array keeps the first call's contents
        _fill(ref.host_actions, k, n_envs)
        _fill(sta.host_actions, k, n_envs)
        want = ref.step_host(1000).numpy().copy()
        with torch.cuda.device(sta.device):
            N.check(L.sparc_step_host_static(sta.blob.data_ptr(), C.byref(sta.params), C.byref(acts), n_envs, 1000,
                                             sta.seed, sta.env_id_base, obs_static.data_ptr(), sta._stream()))
        got = obs_static.numpy()
        # rows 8..10 carry int32 bit patterns: compare them (and the flags) exactly
        assert np.array_equal(got.view(np.int32)[8:11], want.view(np.int32)[8:11]), f"call {k}: integer rows differ"
        assert np.array_equal(got[7], want[7]) and np.array_equal(got[3:5], want[3:5]), f"call {k}: flags / counts differ"
        np.testing.assert_allclose(got[:3], want[:3], rtol=1e-6, atol=0.0, err_msg=f"call {k}")       # float32 views of fp64 state
        np.testing.assert_allclose(got[5], want[5], rtol=PU.T_RTOL, atol=PU.T_ATOL, err_msg=f"call {k} max temperature")
        np.testing.assert_allclose(got[6], want[6], rtol=1e-6, atol=0.0, err_msg=f"call {k} debris density")
        # and the state the two entry points leave behind, in the parity suite's tolerances
        envs = list(range(0, n_envs, max(1, n_envs // 64)))
        f_a, i_a = PU.read_rows(ref, envs)
        f_b, i_b = PU.read_rows(sta, envs)
        PU.assert_rows_match(f_b, i_b, f_a, i_a, where=f"static vs step_host, call {k}")
        Ta, Tb = ref.T.cpu().numpy(), sta.T.cpu().numpy()
        assert np.all(np.abs(Ta - Tb) <= PU.T_ATOL + PU.T_RTOL * np.abs(Ta))
    assert float(obs_static[3].sum()) > 0 or n_envs < 100       # sparks were simulated at the large size
    assert int(sta["time"][0]) == 2000


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from sparc_b200 import _native as N

    L = N.lib()
    _, eng = _engines(4)
    acts = eng._actions_struct(eng.host_actions)
    too_many = L.sparc_static_max_envs() + 1
    rc = L.sparc_step_host_static(eng.blob.data_ptr(), C.byref(eng.params), C.byref(acts), too_many, 1000, eng.seed, 0,
                                  eng.obs_host.data_ptr(), eng._stream())
    assert rc == -3                                   # SPARC_E_UNSUPPORTED, nothing launched
    msg = L.sparc_last_error().decode()
    assert "16384" in msg and "static scratch" in msg
    assert int(eng["time"][0]) == 0
