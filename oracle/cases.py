"""Parity cases shared by the golden generator, the oracle validation and the tests.

TEST INFRASTRUCTURE ONLY.  Each case is: reference constructor arguments, an action list
(one row per control window, replayed the way the reference driver loop latches them —
README.md:101-106), an injected-draw tape recipe and a horizon.

Action rows are (servo, target_voltage, current_mode, ON_time, OFF_time)
(wire_edm.py:84-98,162-170).
"""
from __future__ import annotations

import numpy as np

README_CFG = dict(workpiece_height=15.0, wire_diameter=0.25, wire_material="brass",
                  target_cutting_distance=800.0)          # README.md:45-50
QUICKSTART_ACTION = [0.1, 80.0, 5, 3.0, 80.0]              # examples/quickstart.py:25-33


def make_tape(key: int, n_us: int, forced_ignitions=()) -> np.ndarray:
    """draws[n_us,5]: u1..u4 uniform [0,1), z5 standard normal (SURVEY §8d cfg 2 recipe).

    forced_ignitions: µs indices whose u3 is set to 0.0 (ignites whenever the env is idle and
    not shorted on that µs: `u3 < lambda` always holds, ignition.py:327)."""
    g = np.random.Generator(np.random.Philox(key=key))
    d = np.empty((n_us, 5), dtype=np.float64)
    d[:, :4] = g.random((n_us, 4))
    d[:, 4] = g.standard_normal(n_us)
    for t in forced_ignitions:
        d[t, 2] = 0.0
    return d


def never_ignite_tape(n_us: int, ignite_at=(), z=0.0) -> np.ndarray:
    """Deterministic known-answer tape: u3 = 1.0 (never ignites) except at `ignite_at`."""
    d = np.empty((n_us, 5), dtype=np.float64)
    d[:, 0] = 0.5
    d[:, 1] = 0.5
    d[:, 2] = 1.0
    d[:, 3] = 0.5
    d[:, 4] = z
    for t in ignite_at:
        d[t, 2] = 0.0
    return d


def random_actions(key: int, n: int) -> list:
    """Random actions restricted to odd current modes (even modes raise, material.py:108-113);
    some values are exactly 0.0 to exercise the falsy defaults (ignition.py:329-343)."""
    g = np.random.default_rng(key)
    rows = []
    for _ in range(n):
        servo = float(np.float32(g.uniform(-1, 1)))
        tv = float(np.float32(g.uniform(0, 200))) if g.random() > 0.2 else 0.0
        mode = int(g.choice([1, 3, 5, 7, 9, 11, 13, 15, 17]))
        on = float(np.float32(g.uniform(0, 5))) if g.random() > 0.2 else 0.0
        off = float(np.float32(g.uniform(0, 100))) if g.random() > 0.2 else 0.0
        rows.append([servo, tv, mode, on, off])
    return rows


# name -> dict(config, mode, params per module, actions, tape=(kind, args), n_us, initial)
CASES = {
    # BASELINE cfg 2: README config, quickstart constant action, 20 control steps
    "cfg2_default": dict(config=README_CFG, mode="position", actions=[QUICKSTART_ACTION],
                         tape=("philox", 1234), n_us=20000),
    # BASELINE cfg 1 flavour: actions change every control step, falsy zeros, odd modes
    "random_actions": dict(config=README_CFG, mode="position", actions=("random", 7, 12),
                           tape=("philox", 99), n_us=12000),
    # cfg 3 flavour: position mode, stabilising P gap controller recorded from the reference
    "gap_controller": dict(config=README_CFG, mode="position", actions=("gap_controller", 25.0),
                           tape=("philox", 5), n_us=15000),
    # velocity mode with saturating servo commands (accel/jerk/speed clips, mechanics.py:87-108)
    "velocity_saturate": dict(config=README_CFG, mode="velocity",
                              actions=[[100.0, 80.0, 5, 3.0, 80.0], [-4.0e4, 80.0, 5, 3.0, 80.0],
                                       [3.5e4, 60.0, 3, 2.5, 40.0], [0.0, 80.0, 5, 3.0, 80.0]],
                              tape=("philox", 11), n_us=6000),
    # aggressive generator: I17, long ON, short OFF, small gap -> heating, wire break
    "hot_break": dict(config=dict(README_CFG, initial_gap=12.0), mode="position",
                      actions=[[0.0, 100.0, 17, 5.0, 6.0]], tape=("philox", 3), n_us=30000),
    # small initial gap, wire driven into the workpiece: hard shorts (gap<2) and collision
    "collision": dict(config=dict(README_CFG, initial_gap=6.0), mode="velocity",
                      actions=[[2.0e4, 80.0, 5, 3.0, 80.0]], tape=("philox", 21), n_us=40000),
    # debris-choked gap: sigmoid debris shorts + random shorts (ignition.py:115-146,221-242)
    "debris_shorts": dict(config=dict(README_CFG, initial_gap=8.0), mode="position",
                          ignition=dict(random_short_max_probability=0.002, base_critical_density=0.01,
                                        gap_coefficient=0.0005, sigmoid_steepness=300.0),
                          dielectric=dict(debris_removal_efficiency=0.0002),
                          actions=[[0.0, 80.0, 9, 3.0, 20.0]], tape=("philox", 8), n_us=15000),
    # dataclass-default geometry (20 mm / 0.2 mm wire, N=400) with zone mean enabled
    "dataclass_default_zone_mean": dict(config=dict(), mode="position",
                                        wire=dict(compute_zone_mean=True, zone_mean_interval=100),
                                        actions=[[0.05, 90.0, 7, 2.0, 50.0]], tape=("philox", 4), n_us=8000),
    # BASELINE cfg 5 flavour: 50 mm workpiece, 0.1 mm segments (N=1100), I13, OFF 30
    "thick_fine": dict(config=dict(workpiece_height=50.0, wire_diameter=0.25, target_cutting_distance=800.0),
                       mode="position", wire=dict(segment_len=0.1),
                       actions=[[0.0, 80.0, 13, 3.0, 30.0]], tape=("philox", 15), n_us=6000),
    # BASELINE cfg 5: 50 mm workpiece, 0.05 mm segments (N=2200, 4 warps per env) and 0.02 mm segments
    # (N=5500, 8 warps per env), I13 ON 3 / OFF 30, a few forced sparks so the field is not flat
    "thick_2200": dict(config=dict(workpiece_height=50.0, wire_diameter=0.25, target_cutting_distance=800.0),
                       mode="position", wire=dict(segment_len=0.05),
                       actions=[[0.0, 80.0, 13, 3.0, 30.0]], tape=("philox_forced", (31, (40, 400, 1300, 2100, 3300))),
                       n_us=4000),
    "thick_5500": dict(config=dict(workpiece_height=50.0, wire_diameter=0.25, target_cutting_distance=800.0),
                       mode="position", wire=dict(segment_len=0.02, compute_zone_mean=True, zone_mean_interval=250),
                       actions=[[0.0, 80.0, 13, 3.0, 30.0]], tape=("philox_forced", (32, (25, 500, 1200, 2600))),
                       n_us=3000),
    # short servo interval: an action latch every 37 µs (look-ahead blocks are cut at every latch)
    "short_servo": dict(config=dict(README_CFG, servo_interval=37, initial_gap=20.0), mode="position",
                        actions=[[0.3, 90.0, 7, 2.0, 25.0]], tape=("philox", 77), n_us=6000),
    # deterministic KAT (i): single forced spark, no unwinding (single_spark_animation.py:214-249)
    "kat_single_spark": dict(config=README_CFG, mode="position", actions=[[0.0, 80.0, 5, 3.0, 80.0]],
                             tape=("never", (50,)), n_us=3000, unwind_v=0.0),
    # deterministic KAT (iii): mechanics step response only, no sparks at all
    "kat_mechanics_position": dict(config=README_CFG, mode="position", actions=[[1.0, 80.0, 5, 3.0, 80.0], [-1.0, 80.0, 5, 3.0, 80.0]],
                                   tape=("never", ()), n_us=4000),
    # target reached: tiny target distance
    "target_reached": dict(config=dict(README_CFG, initial_gap=10.0, target_cutting_distance=10.004),
                           mode="position", actions=[[0.0, 80.0, 5, 3.0, 20.0]], tape=("philox", 2), n_us=20000),
}


def case_tape(case: dict) -> np.ndarray:
    kind, arg = case["tape"]
    n = case["n_us"]
    if kind == "philox":
        return make_tape(arg, n)
    if kind == "philox_forced":
        return make_tape(arg[0], n, forced_ignitions=arg[1])
    if kind == "never":
        return never_ignite_tape(n, ignite_at=arg)
    raise ValueError(kind)


def oracle_overrides(case: dict) -> dict:
    """Flatten a case into oracle.make_params(**overrides)."""
    ov = {}
    for k, v in case.get("config", {}).items():
        if k == "wire_material":
            continue
        ov[k] = v
    ov["mechanics_control_mode"] = case.get("mode", "position")
    for grp in ("ignition", "wire", "material", "dielectric", "mechanics"):
        for k, v in case.get(grp, {}).items():
            ov[k] = int(v) if isinstance(v, bool) else v
    return ov


# ------------------------------------------------------------------------------------------------
# Geometry sweep: the register layout of the wire depends on (n_segments, zone, contacts); these
# cases move every one of them — tiny wires, zones that touch either end, contacts clipped to
# the wire, buffers shorter than the contact offsets, segment counts on both sides of every
# cells-per-lane / warps-per-env boundary.  Each runs 3000 µs with a few forced sparks.
# Reference geometry expressions: wire.py:143-156, 208-252.
# ------------------------------------------------------------------------------------------------
def _geometry_case(h, seg, bb, bt, ob, ot, key, dia=0.25, zone_mean=False, action=None, gap=30.0):
    n_us = 3000
    forced = tuple(int(t) for t in np.random.default_rng(key).choice(np.arange(20, n_us - 300), 7, replace=False))
    wire = dict(segment_len=seg, buffer_len_bottom=bb, buffer_len_top=bt, contact_offset_bottom=ob,
                contact_offset_top=ot)
    if zone_mean:
        wire.update(compute_zone_mean=True, zone_mean_interval=64)
    return dict(config=dict(workpiece_height=h, wire_diameter=dia, target_cutting_distance=800.0, initial_gap=gap),
                mode="position", wire=wire, actions=[action or [0.02, 90.0, 11, 2.5, 25.0]],
                tape=("philox_forced", (key, forced)), n_us=n_us)


def geometry_cases() -> dict:
    out = {
        # N = 5: four cells in one lane, zone of two segments
        "geo_tiny": _geometry_case(0.4, 0.2, 0.3, 0.3, 0.2, 0.2, 901),
        # N = 33..34: one cell per lane and a remainder
        "geo_lanes_plus_one": _geometry_case(3.0, 0.2, 1.9, 1.9, 1.0, 1.0, 902, zone_mean=True),
        # zone starts at segment 0 (no bottom buffer), contact offset longer than the buffer
        "geo_zone_at_start": _geometry_case(12.0, 0.2, 0.0, 25.0, 10.0, 5.0, 903),
        # zone runs to the last segment (no top buffer)
        "geo_zone_at_end": _geometry_case(18.0, 0.25, 22.0, 0.0, 6.0, 10.0, 904, dia=0.2),
        # the workpiece covers the whole wire
        "geo_all_zone": _geometry_case(30.0, 0.1, 0.0, 0.0, 10.0, 10.0, 905, zone_mean=True),
        # workpiece thinner than one segment: empty zone in index terms
        "geo_thin_plate": _geometry_case(0.15, 0.2, 20.0, 31.0, 10.0, 10.0, 906),
        # N = 769 / 1153 / 2305: first sizes past one, two and four warps per env
        "geo_769": _geometry_case(36.9, 0.1, 20.0, 20.0, 10.0, 10.0, 907),
        "geo_1153": _geometry_case(55.3, 0.1, 30.0, 30.0, 12.0, 3.0, 908, dia=0.3),
        "geo_2305": _geometry_case(45.25, 0.05, 35.0, 35.0, 10.0, 40.0, 909),
        # asymmetric buffers, odd segment length (floor vs divide differ), thin wire
        "geo_odd_pitch": _geometry_case(17.3, 0.07, 11.0, 43.0, 3.3, 17.0, 910, dia=0.1, zone_mean=True,
                                        action=[0.0, 70.0, 7, 1.5, 15.0]),
        # N = 6650: near the largest supported size, short zone in the middle
        "geo_max": _geometry_case(13.0, 0.02, 60.0, 60.0, 10.0, 10.0, 911),
    }
    g = np.random.default_rng(20260207)
    for k in range(6):
        seg = float(np.round(g.choice([0.03, 0.05, 0.08, 0.13, 0.2, 0.37, 0.5]), 3))
        h = float(np.round(g.uniform(0.5, 60.0), 2))
        bb, bt = (float(np.round(x, 2)) for x in g.uniform(0.0, 40.0, 2))
        ob, ot = (float(np.round(x, 2)) for x in g.uniform(0.0, 20.0, 2))
        out[f"geo_random_{k}"] = _geometry_case(h, seg, bb, bt, ob, ot, 920 + k, dia=float(g.choice([0.15, 0.2, 0.25])),
                                                zone_mean=bool(k % 2))
    return out


GEOMETRY_CASES = geometry_cases()
ALL_CASES = {**CASES, **GEOMETRY_CASES}
