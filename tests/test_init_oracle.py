"""CPU tests of the oracle for the dycore's idealized initial states: the plain-C restatement
(oracle/init_oracle.c) against the compiled reference (oracle/_ref: the reference's own Dycore::init_supercell +
convert_dynamics_to_coupler; the thermal branch re-driven around the reference's point functions) -- bit for bit -- and
against the committed golden vectors (tests/golden/init_*.npz, tools/make_golden_init.py)."""
import os

import numpy as np
import pytest

from oracle import modules_oraclelib as M
from oracle import reflib
from pam_b200 import synthetic as syn

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {"c1_supercell": dict(kind="supercell", nz=50, ny=1, nx=64, nens=1, grid="equal", tracers="nt1", default_consts=True),
         "sc_3d_p3": dict(kind="supercell", nz=8, ny=3, nx=5, nens=2, grid="tanh", tracers="p3", default_consts=False),
         "thermal_2d": dict(kind="thermal", nz=12, ny=1, nx=10, nens=3, grid="equal", tracers="nt1", default_consts=True),
         "thermal_3d": dict(kind="thermal", nz=8, ny=4, nx=6, nens=2, grid="tanh", tracers="nt1", default_consts=False)}
DEFAULT_CONSTS = dict(R_d=287., cp_d=1003., R_v=461., cp_v=1859., p0=1.e5, grav=9.81)     # Dycore.h:871-876


def build(kw):
    tr = syn.TRACERS_NT1 if kw["tracers"] == "nt1" else syn.TRACERS_P3_SHOC
    case = syn.make_case(kw["nz"], kw["ny"], kw["nx"], kw["nens"], tracers=tr, grid=kw["grid"])
    if kw["default_consts"]:
        case["consts"] = dict(DEFAULT_CONSTS)
    return case


def oracle_init(case, kind):
    gamma, C0 = M.derived_constants(case["consts"])
    return M.init_idealized(kind, case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                            case["tracers"], case["consts"], gamma, C0)


def reference_init(case, kind):
    r = reflib.RefDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                         tracers=case["tracers"], consts=case["consts"], ord=5)
    getattr(r, "init_" + kind)()
    out = r.get_fields()
    r.close()
    return out


def test_gll_tables_equal_the_reference():
    """tools/gen_tables.py derives the 9-point GLL rule from first principles; the reference tabulates 40-digit literals"""
    p, w = M.gll9()
    assert abs(w.sum() - 1.0) < 1e-15 and p[0] == -0.5 and p[8] == 0.5 and p[4] == 0.0
    assert np.array_equal(p, -p[::-1]) and np.array_equal(w, w[::-1])
    g = np.load(os.path.join(GOLDEN, "init_c1_supercell.npz"))
    assert np.array_equal(p, g["gll_pts"]) and np.array_equal(w, g["gll_wts"])


@pytest.mark.parametrize("name", sorted(CASES))
def test_restatement_is_bit_identical_to_compiled_reference(name):
    if not reflib.available(5):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    case = build(CASES[name])
    ref = reference_init(case, CASES[name]["kind"])
    got = oracle_init(case, CASES[name]["kind"])
    for k, v in ref.items():
        assert np.array_equal(got[k], v), k


@pytest.mark.parametrize("name", sorted(CASES))
def test_restatement_matches_golden_vectors(name):
    g = np.load(os.path.join(GOLDEN, "init_%s.npz" % name))
    case = build(CASES[name])
    got = oracle_init(case, CASES[name]["kind"])
    for k, v in got.items():
        assert np.array_equal(v, g[k]), k


def test_initial_states_are_physical():
    sc = oracle_init(build(CASES["c1_supercell"]), "supercell")
    assert 212.9 < sc["temp"].min() and sc["temp"].max() < 300.0
    assert np.all(np.diff(sc["density_dry"][:, 0, 0, 0]) < 0)                 # density decreases with height
    assert abs(sc["uvel"][0].mean() + 15.0) < 1.5 and abs(sc["uvel"][-1].mean() - 15.0) < 1e-12
    assert sc["water_vapor"].min() > 0 and (sc["water_vapor"] / sc["density_dry"]).max() <= 0.014 * (1 + 1e-12)
    th = oracle_init(build(CASES["thermal_2d"]), "thermal")
    assert np.all(th["uvel"] == 0) and np.all(th["water_vapor"] == 0)
    # the bubble: warmer than the level mean somewhere near z = 2 km, nowhere colder
    anom = th["temp"] - th["temp"].min(axis=(1, 2), keepdims=True)
    assert anom.max() > 0.5 and anom.min() == 0.0
