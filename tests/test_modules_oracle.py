"""CPU tests of the oracle for the coupler modules either side of the dycore: the plain-C restatement
(oracle/modules_oracle.c) against the compiled reference (oracle/_ref: the reference's own modules::sponge_layer,
compute/apply_gcm_forcing_tendencies, perturb_temperature, serial backend) -- bit for bit -- and against the committed
golden vectors the compiled reference produced (tests/golden/modules_*.npz, tools/make_golden.py)."""
import os

import numpy as np
import pytest

import helpers as H
from oracle import modules_oraclelib as M
from oracle import reflib

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {"2d": dict(nz=12, ny=1, nx=10, nens=5, seed=7, overshoot=False),
         "3d": dict(nz=8, ny=4, nx=6, nens=3, seed=11, overshoot=False),
         "overshoot": dict(nz=12, ny=1, nx=10, nens=5, seed=7, overshoot=True)}


def run_oracle_sequence(dm, case, opts, ids):
    """the driver's order (driver.cpp:76, :239-250): perturb, compute forcing, apply forcing, [dycore], sponge"""
    tr = [t[0] for t in case["tracers"]]
    M.perturb_temperature(dm, ids, 0.1)
    M.compute_gcm_forcing_tendencies(dm, opts["gcm_physics_dt"])
    M.apply_gcm_forcing_tendencies(dm, opts["crm_dt"], opts["gcm_physics_dt"])
    M.sponge_layer(dm, tr, opts["crm_dt"], opts["sponge_time_scale"], opts["sponge_num_layers"])


def run_reference_sequence(case, dm, opts, ids):
    r = reflib.RefDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                         tracers=case["tracers"], consts=case["consts"], crm_dt=opts["crm_dt"], ord=5)
    names, tends = H.modules_fields(case)
    for n in names + list(H.MODULE_GCM):
        r.set_field(n, dm[n])
    r.set_option("gcm_physics_dt", float(opts["gcm_physics_dt"]))
    r.set_option("latvap", 2.5e6); r.set_option("latice", 3.34e5)       # read, not used (gcm_forcing.h:27-28)
    r.set_option("sponge_time_scale", float(opts["sponge_time_scale"]))
    r.set_option("sponge_num_layers", int(opts["sponge_num_layers"]))
    r.perturb_temperature(ids, 0.1)
    r.compute_gcm_forcing_tendencies()
    r.apply_gcm_forcing_tendencies()
    r.sponge_layer()
    out = {n: r.get_field(n) for n in names}
    for n in tends:
        out[n] = r.get_field(n, (case["nz"], case["nens"]))
    return out


@pytest.mark.parametrize("name", sorted(CASES))
def test_restatement_is_bit_identical_to_compiled_reference(name):
    if not reflib.available(5):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    case, dm, opts = H.make_modules_case(**CASES[name])
    ids = np.arange(case["nens"], dtype=np.int32) * 3 + 1
    ref = run_reference_sequence(case, dm, opts, ids)
    run_oracle_sequence(dm, case, opts, ids)
    for n, v in ref.items():
        assert np.array_equal(dm[n].reshape(v.shape), v), n


@pytest.mark.parametrize("name", sorted(CASES))
def test_restatement_matches_golden_vectors(name):
    g = np.load(os.path.join(GOLDEN, "modules_%s.npz" % name))
    case, dm, opts = H.make_modules_case(**CASES[name])
    ids = g["ids"]
    run_oracle_sequence(dm, case, opts, ids)
    names, tends = H.modules_fields(case)
    for n in names + tends:
        assert np.array_equal(dm[n], g[n]), n


def test_hole_filling_is_exercised_and_conserves_water():
    """the cases must actually reach fill_holes: level stage in every case, whole-CRM stage in 'overshoot'"""
    for name, want_global in (("2d", False), ("overshoot", True)):
        case, dm, opts = H.make_modules_case(**CASES[name])
        M.compute_gcm_forcing_tendencies(dm, opts["gcm_physics_dt"])
        before = {n: dm[n].copy() for n in ("water_vapor", "cloud_water", "ice", "density_dry")}
        M.apply_gcm_forcing_tendencies(dm, opts["crm_dt"], opts["gcm_physics_dt"])
        for n in ("water_vapor", "cloud_water", "ice"):
            assert dm[n].min() >= 0.0
        # un-filled result of the forcing alone (gcm_forcing.h:367-390) for cloud water: negative somewhere
        rd_new = before["density_dry"] + dm["gcm_forcing_tend_rho_d"][:, None, None, :] * opts["crm_dt"]
        qv = before["water_vapor"] / (before["density_dry"] + before["water_vapor"]) + dm["gcm_forcing_tend_qv"][:, None, None, :] * opts["crm_dt"]
        rv = qv * rd_new / (1 - qv)
        ql = before["cloud_water"] / (before["density_dry"] + before["water_vapor"]) + dm["gcm_forcing_tend_ql"][:, None, None, :] * opts["crm_dt"]
        rl = ql * (rd_new + rv)
        assert (rl < 0).any()
        dz = dm["vertical_cell_dz"][:, None, None, :]
        neg = np.where(rl < 0, -rl * dz, 0.0).sum(axis=(1, 2))
        pos = np.where(rl > 0, rl * dz, 0.0).sum(axis=(1, 2))
        assert bool((neg > pos).any()) == want_global
        if not want_global:      # level filling conserves the level's mass: sum(after) = sum(unfilled)
            after = (dm["cloud_water"] * dz).sum(axis=(1, 2))
            np.testing.assert_allclose(after, (rl * dz).sum(axis=(1, 2)), rtol=1e-12, atol=1e-18)


def test_perturbation_keeps_level_means_and_bounds():
    case, dm, opts = H.make_modules_case(**CASES["2d"])
    t0 = dm["temp"].copy()
    M.perturb_temperature(dm, np.arange(case["nens"], dtype=np.int32), 0.1)
    nl = case["nz"] // 4
    assert np.array_equal(dm["temp"][nl:], t0[nl:])
    assert not np.array_equal(dm["temp"][:nl], t0[:nl])
    np.testing.assert_allclose(dm["temp"][:nl].mean(axis=(1, 2)), t0[:nl].mean(axis=(1, 2)), rtol=1e-14)
    assert np.abs(dm["temp"][:nl] - t0[:nl]).max() < 0.2      # +-0.1 K, then the rescale to the old level mean
