"""The C oracle (oracle/wedm_oracle.c) against the committed reference trajectories.

tests/golden/*.npz were produced by running the UNMODIFIED reference with injected draws
(generator: oracle/validate_oracle.py --write-golden).  The oracle reproduced every field
bit-for-bit when they were generated; this test keeps it that way.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import cases as CS
from oracle import oracle as O


def _load(name):
    return np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))


@pytest.mark.parametrize("name", sorted(CS.ALL_CASES))
def test_oracle_matches_reference_golden(name):
    g = _load(name)
    case = CS.ALL_CASES[name]
    tape = CS.case_tape(case)
    e = O.OracleEnv(**CS.oracle_overrides(case))
    if "unwind_v" in case:
        e.set_unwind_velocity(case["unwind_v"])
    res = e.run(case["n_us"], g["actions"], draws=tape, rec_every=1, t_every=1000)
    assert np.array_equal(np.array(list(e.geometry().values())), g["geometry"])
    assert res["n_done"] == int(g["n_done"][0])
    # discrete fields: exact, every µs
    assert np.array_equal(res["i32"], g["i32"])
    # fp64 scalars: the oracle is bit-exact against the reference on this platform; allow
    # 1e-12 relative so a different libm (exp/pow last-bit) cannot make this flaky.
    rows = g["f64_rows"]
    np.testing.assert_allclose(res["f64"][rows], g["f64"], rtol=1e-12, atol=1e-300)
    # temperature field snapshots (float32): bit-exact when generated; 1 ulp slack
    assert np.array_equal(res["T_times"], g["T_times"])
    np.testing.assert_allclose(res["T"], g["T"], rtol=2e-7, atol=0)


def test_py_floordiv_geometry_quirk():
    """30.0 // 0.2 == 149 and 15.0 // 0.2 == 74 (wire.py:151-154; SURVEY §8 index quirk)."""
    e = O.OracleEnv(workpiece_height=15.0, wire_diameter=0.25)
    assert e.geometry() == dict(n=375, zone_start=149, zone_end=223, contact_bottom=100, contact_top=275)
    e = O.OracleEnv()
    assert e.geometry() == dict(n=400, zone_start=149, zone_end=248, contact_bottom=100, contact_top=300)
    e = O.OracleEnv(workpiece_height=50.0, wire_diameter=0.25, segment_len=0.05)
    g = e.geometry()
    assert (g["n"], g["zone_start"], g["zone_end"]) == (2200, 599, 1598)


def test_even_mode_is_an_error():
    """area_corrected.json has odd modes only; the reference raises at the first spark."""
    e = O.OracleEnv(workpiece_height=15.0, wire_diameter=0.25)
    tape = CS.never_ignite_tape(3000, ignite_at=(1500,))
    res = e.run(3000, [[0.0, 80.0, 6, 3.0, 80.0]], draws=tape, rec_every=0)
    assert res["n_done"] == 1501  # the step that sparks is the one that fails


def test_philox_known_answer():
    """Philox4x32-10 known-answer vectors from the Random123 distribution (kat_vectors)."""
    out = O.philox4x32_10([0, 0, 0, 0], [0, 0])
    assert [hex(x) for x in out] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    out = O.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2)
    assert [hex(x) for x in out] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    out = O.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0])
    assert [hex(x) for x in out] == ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_oracle_reference_style_reset_matches_reference_golden():
    """strict_reference_reset: a second episode on USED modules.  The golden holds the UNMODIFIED
    reference's second episode after RESET_N1 µs of the debris_shorts case and a plain `env.reset()`
    (oracle/validate_oracle.py::check_reference_reset); the oracle's `reset(reference_semantics=True)` must
    reproduce it bit for bit — the short-circuit countdown and the debris volume carry over."""
    import numpy as np
    from oracle import cases as CS
    from oracle import oracle as O

    g = np.load(os.path.join(GOLDEN_DIR, "reference_reset.npz"))
    case = CS.CASES["debris_shorts"]
    n1, n2 = int(g["n1"][0]), int(g["n2"][0])
    e = O.OracleEnv(**CS.oracle_overrides(case))
    e.run(n1, g["actions1"], draws=CS.make_tape(4101, n1), rec_every=0)
    e.reset(reference_semantics=True)
    res = e.run(n2, g["actions2"], draws=CS.make_tape(4102, 4000), rec_every=1, t_every=1000)
    assert np.array_equal(res["i32"], g["i32"].astype(res["i32"].dtype))
    assert np.array_equal(res["f64"], g["f64"])
    assert np.array_equal(res["T"], g["T"])
    assert g["i32"][:100, 6].sum() > 50 and g["f64"][0, 6] > 1e-4      # carried short circuit and debris
    # a fresh-env reset gives a different second episode
    e2 = O.OracleEnv(**CS.oracle_overrides(case))
    e2.run(n1, g["actions1"], draws=CS.make_tape(4101, n1), rec_every=0)
    e2.reset()
    res2 = e2.run(n2, g["actions2"], draws=CS.make_tape(4102, 4000), rec_every=1)
    assert not np.array_equal(res2["i32"][:100], g["i32"][:100])
