"""The two forms of the heat stencil (sparc_b200/csrc/step_kernel.cuh: stencil_apply, "stencil coefficients"),
emulated in NumPy with the kernel's rounding recipe, against the reference's arithmetic (wire.py:58-123 as
restated in oracle/wedm_oracle.c: float32 arrays, float64 scalars, every accumulation rounded to float32).

This pins the NUMERICAL ARGUMENT of DESIGN §4.1 on the CPU — which form is used when, that the rounded weights
cancel exactly, how far each form drifts from the reference — not the CUDA code (the `-m gpu` parity tests do
that).  A float32 FMA is emulated as round32(float64(a)*float64(b) + float64(c)): the product of two float32 is exact
in float64, the sum is rounded twice (53 then 24 bits), which differs from a true FMA in < 1e-7 of the cases.
"""
import numpy as np
import pytest

f32 = np.float32


def fma(a, b, c):
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


def setup(segment_len, unwind_v=0.2, h_base=14000.0, h_zone=21000.0):
    from sparc_b200 import params as P

    cfg = P.EnvironmentConfig(workpiece_height=15.0, wire_diameter=0.25, target_cutting_distance=800.0)
    prm = P.derive_params(cfg, "position", wire=P.WireModuleParameters(segment_len=segment_len))
    n = prm.n_segments
    h = np.full(n, h_base, f32)
    h[prm.actual_zone_start:prm.actual_zone_end] = h_zone
    adv = prm.density_times_cp * abs(unwind_v) * prm.S
    return prm, n, h, adv


def reference_step(T, prm, h, adv):
    """One quiet µs in the reference's evaluation order and precision (oracle/wedm_oracle.c:thermal_update)."""
    k, f, A, Td = prm.k_cond_coeff, prm.temp_update_factor, prm.A, prm.dielectric_temperature
    T = T.copy()
    T64 = T.astype(np.float64)
    q = np.zeros_like(T)
    q[1:-1] = (k * ((T64[:-2] - 2.0 * T64[1:-1]) + T64[2:])).astype(f32)
    q[-1] = f32(k * np.float64(f32(T[-2] - T[-1])))
    q = (q.astype(np.float64) - (h.astype(np.float64) * A) * (T64 - Td)).astype(f32)
    d = (T[:-1] - T[1:]).astype(f32)
    q[1:] = (q[1:].astype(np.float64) + adv * d.astype(np.float64)).astype(f32)
    out = (T64 + q.astype(np.float64) * f).astype(f32)
    out[0] = T[0]
    return out


def kernel_weights(prm, h, adv):
    """The kernel's rounding recipe.  Returns (diff_form, cR, cL, cC_or_cE, c0)."""
    f, k, A, Td = prm.temp_update_factor, prm.k_cond_coeff, prm.A, prm.dielectric_temperature
    cR = f32(f * k)
    cL = f32(f * (k + adv))
    csum = f32(cR + cL)
    hA = h.astype(np.float64) * A
    cE = (-(f * hA)).astype(f32)
    if not (csum <= f32(0.015625)):
        return True, cR, cL, cE, (f * hA * Td).astype(f32)
    if abs(cL) >= abs(cR):
        cR = f32(csum - cL)
    else:
        cL = f32(csum - cR)
    cC = (cE - csum).astype(f32)
    c0 = (-(cC.astype(np.float64) + np.float64(csum)) * Td).astype(f32)
    return False, cR, cL, cC, c0


def kernel_step(T, w):
    diff, cR, cL, ce, c0 = w
    n = len(T)
    nxt = np.concatenate([T[1:], T[-1:]])          # zero-flux end: its right neighbour is the cell itself
    prv = np.concatenate([T[:1], T[:-1]])
    q = fma(ce, T, c0)
    if diff:
        q = fma(np.full(n, cR, f32), (nxt - T).astype(f32), q)
        q = fma(np.full(n, -cL, f32), (T - prv).astype(f32), q)
    else:
        q = fma(np.full(n, cR, f32), nxt, q)
        q = fma(np.full(n, cL, f32), prv, q)
    out = (T + q).astype(f32)
    out[0] = T[0]                                   # spool cell (Dirichlet)
    return out


def initial_field(n, peak=600.0):
    x = np.arange(n, dtype=np.float64)
    return (293.15 + peak * np.exp(-((x - 0.5 * n) / (0.016 * n)) ** 2) + 80.0 * np.exp(-((x - 0.32 * n) / (0.07 * n)) ** 2)).astype(f32)


def test_form_choice_follows_the_segmentation():
    # README / dataclass default 0.2 mm and 0.1 mm segments: four-instruction form; 0.05 mm and finer: differences
    for seg, want_diff in ((0.2, False), (0.1, False), (0.05, True), (0.02, True)):
        prm, n, h, adv = setup(seg)
        assert kernel_weights(prm, h, adv)[0] is want_diff, seg
    # a very fast wire pushes the advection weight over the threshold on its own
    prm, n, h, adv = setup(0.2, unwind_v=5.0)
    assert kernel_weights(prm, h, adv)[0] is True


def test_rounded_weights_cancel_exactly():
    for seg, v in ((0.2, 0.2), (0.2, 0.0), (0.1, 0.2), (0.2, 1.0), (0.13, 0.37)):
        prm, n, h, adv = setup(seg, unwind_v=v)
        diff, cR, cL, cC, c0 = kernel_weights(prm, h, adv)
        assert not diff
        csum = np.float64(cR) + np.float64(cL)
        assert csum == np.float64(f32(csum)), "cR + cL must be a float32"
        f, A = prm.temp_update_factor, prm.A
        cE_true = -(f * (h.astype(np.float64) * A))
        cE_eff = cC.astype(np.float64) + csum                       # exact in float64
        assert np.all(np.abs(cE_eff - cE_true) <= 3e-6 * np.abs(cE_true))
        # conduction weights moved by at most one part in 2^23
        assert abs(np.float64(cR) - f * prm.k_cond_coeff) <= 2.0 ** -23 * csum
        assert abs(np.float64(cL) - f * (prm.k_cond_coeff + adv)) <= 2.0 ** -23 * csum
        # a uniform field at any temperature is a fixed point (q stays below half an ulp of T)
        for T0 in (293.15, 350.0, 900.0):
            hz = np.zeros_like(h)
            w = kernel_weights(prm, hz, adv)
            T = np.full(n, T0, f32)
            assert np.array_equal(kernel_step(T, w), T)


@pytest.mark.parametrize("seg,bound", [(0.2, 4e-4), (0.1, 6e-4)])
def test_four_instruction_form_stays_close_to_the_reference(seg, bound):
    prm, n, h, adv = setup(seg)
    w = kernel_weights(prm, h, adv)
    assert not w[0]
    w_diff = (True, f32(prm.temp_update_factor * prm.k_cond_coeff), f32(prm.temp_update_factor * (prm.k_cond_coeff + adv)),
              (-(prm.temp_update_factor * (h.astype(np.float64) * prm.A))).astype(f32),
              (prm.temp_update_factor * (h.astype(np.float64) * prm.A) * prm.dielectric_temperature).astype(f32))
    Tr = initial_field(n)
    T4, Td = Tr.copy(), Tr.copy()
    worst4 = worstd = 0.0
    for t in range(6000):
        Tr, T4, Td = reference_step(Tr, prm, h, adv), kernel_step(T4, w), kernel_step(Td, w_diff)
        if t % 500 == 499:
            worst4 = max(worst4, float(np.abs(T4.astype(np.float64) - Tr).max()))
            worstd = max(worstd, float(np.abs(Td.astype(np.float64) - Tr).max()))
    print(f"\n[seg {seg}] max |T - T_ref| over 6 ms: four-instruction form {worst4:.2e} K, difference form {worstd:.2e} K")
    assert worstd <= 1.3e-4                         # a couple of float32 ulp at these temperatures
    assert worst4 <= bound                          # GPU tolerance: 2e-3 K + 1e-5 T


def test_four_instruction_form_would_not_do_for_fine_segments():
    """Why the form is chosen per env: at 0.02 mm segments the products are ~0.1 T and the error leaves the tolerance."""
    prm, n, h, adv = setup(0.02)
    f, k = prm.temp_update_factor, prm.k_cond_coeff
    cR, cL = f32(f * k), f32(f * (k + adv))
    csum = f32(cR + cL)
    cR = f32(csum - cL)
    hA = h.astype(np.float64) * prm.A
    cC = ((-(f * hA)).astype(f32) - csum).astype(f32)
    c0 = (-(cC.astype(np.float64) + np.float64(csum)) * prm.dielectric_temperature).astype(f32)
    forced = (False, cR, cL, cC, c0)
    shipped = kernel_weights(prm, h, adv)
    assert shipped[0]
    Tr = initial_field(n, peak=400.0)
    T4, Td = Tr.copy(), Tr.copy()
    for _ in range(1500):
        Tr, T4, Td = reference_step(Tr, prm, h, adv), kernel_step(T4, forced), kernel_step(Td, shipped)
    e4 = float(np.abs(T4.astype(np.float64) - Tr).max())
    ed = float(np.abs(Td.astype(np.float64) - Tr).max())
    print(f"\n[seg 0.02] after 1.5 ms: four-instruction form {e4:.2e} K, difference form (shipped) {ed:.2e} K")
    assert ed <= 1.3e-4 and e4 > 5 * ed
