"""CPU tests of properties the domain offers independently of any stored vector (ideas carried over from the
reference's stale unit tests, dynamics/awfl/unit/recon_regular/recon_regular.cpp:6-8,111-154, and its PAM_DEBUG mass
check, dynamics/awfl/Dycore.h:224-251): reconstruction order of accuracy, overshoot control, mass conservation,
tracer positivity. They pin the ORACLE; the same properties are asserted on the CUDA path in test_gpu_parity.py."""
import numpy as np
import pytest

import helpers as H
from pam_b200 import synthetic as syn


def _oracle(ord):
    case = syn.make_case(12, 1, 10, 1, grid="equal")
    return H.make_oracle(case, ord), case


def _cell_avgs(f_int, n):
    """exact cell averages of a function with antiderivative f_int on [0,1] split into n cells"""
    x = np.linspace(0.0, 1.0, n + 1)
    return (f_int(x[1:]) - f_int(x[:-1])) * n


@pytest.mark.parametrize("ord", [3, 5, 7, 9])
def test_reconstruction_converges_at_the_design_order(ord):
    o, _ = _oracle(ord)
    f = lambda x: np.cos(2 * np.pi * x - np.pi / 10)
    F = lambda x: np.sin(2 * np.pi * x - np.pi / 10) / (2 * np.pi)
    errs = []
    grids = (16, 32) if ord >= 7 else (32, 64)
    for n in grids:
        avg = _cell_avgs(F, n)
        hs = (ord - 1) // 2
        ext = np.concatenate([avg[-hs:], avg, avg[:hs]])
        st = np.stack([ext[i:i + ord] for i in range(n)])
        right = o.reconstruct(st, 1)                         # value at the right edge of each cell
        exact = f(np.linspace(0.0, 1.0, n + 1)[1:])
        errs.append(np.max(np.abs(right - exact)))
    rate = np.log2(errs[0] / errs[1])
    assert rate >= ord - 1.1, (errs, rate)     # WENO-limited: at least ord-1 on smooth data (the unlimited scheme gives ord)


@pytest.mark.parametrize("ord", [3, 5, 7, 9])
def test_weno_limits_overshoot_at_a_step(ord):
    o, _ = _oracle(ord)
    t = o.tables()
    n = 32
    avg = np.where(np.arange(n) < n // 2, 0.0, 1.0).astype(float)
    avg[n // 2 - 1] = 0.85 * 0 + 0.0
    hs = (ord - 1) // 2
    ext = np.concatenate([np.zeros(hs), avg, np.ones(hs)])
    st = np.stack([ext[i:i + ord] for i in range(n)])
    weno = np.concatenate([o.reconstruct(st, 0), o.reconstruct(st, 1)])
    c2g = t["coefs_to_gll"]; s2c = t["sten_to_coefs"]
    plain = np.concatenate([st @ s2c @ c2g[:, 0], st @ s2c @ c2g[:, 1]])     # unlimited high-order polynomial
    over_w = max(weno.max() - 1.0, -weno.min(), 0.0)
    over_p = max(plain.max() - 1.0, -plain.min(), 0.0)
    assert over_p > 1e-3
    assert over_w < 0.25 * over_p, (over_w, over_p)


def test_constant_stencil_is_reproduced_exactly():
    for ord in (3, 5, 7, 9):
        o, _ = _oracle(ord)
        st = np.full((4, ord), 3.25)
        for ind in (0, 1):
            assert np.allclose(o.reconstruct(st, ind), 3.25, rtol=1e-14, atol=0)


@pytest.mark.parametrize("nt", [1, 10])
def test_mass_conservation_and_positivity(nt):
    tr = syn.TRACERS_NT1 if nt == 1 else syn.TRACERS_P3_SHOC
    case = syn.make_case(12, 1, 12, 2, tracers=tr, grid="tanh")
    o = H.make_oracle(case, 5, crm_dt=2.0)
    o.declare_hydrostatic()
    m0 = o.compute_mass()
    for _ in range(5):
        o.time_step()
    m1 = o.compute_mass()
    # total mass row = nt; reference's own check is rel & abs > 1e-10 (Dycore.h:238-239): here to round-off
    assert np.max(np.abs(m1[nt] - m0[nt]) / np.abs(m0[nt])) < 1e-13
    f = o.get_fields()
    for (name, pos, adds) in case["tracers"]:
        if pos:
            assert f[name].min() >= 0.0, name
    for v in f.values():
        assert np.isfinite(v).all()
