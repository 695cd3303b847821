"""CPU tests (no GPU): the plain-C oracle (oracle/awfl_oracle.c) against

  1. the committed golden vectors tests/golden/*.npz, generated from the UNMODIFIED compiled reference by
     tools/make_golden.py (the reference tree does not travel to the GPU box; the vectors do), and
  2. the compiled reference itself (oracle/_ref) whenever it is present.

Arithmetic without libm (tables, reconstruct, vertical matrices) must agree bit for bit; everything downstream of pow()
must agree to 1e-13 relative L2 (the two are bit-identical on one host; the slack only covers a different libm pow
variant being selected on another CPU).
"""
import glob
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from conftest import rel_l2, field_floors
import helpers as H
import make_golden as MG
from oracle import oraclelib, reflib

GOLDEN = sorted(p for p in glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz")) if not os.path.basename(p).startswith(("modules_", "init_")))
TOL = 1e-13


def _load(path):
    z = np.load(path)
    meta = json.loads(bytes(z["meta"]).decode())
    return z, meta


def test_golden_fixtures_exist():
    assert len(GOLDEN) >= 8
    names = {os.path.basename(p)[:-4] for p in GOLDEN}
    assert names == set(MG.CASES)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_golden(path):
    z, meta = _load(path)
    case = MG.build_case(meta["case"])
    ord = meta["ord"]
    o = H.make_oracle(case, ord, crm_dt=meta["crm_dt"])
    # tables and options: bit for bit
    for k, v in o.tables().items():
        assert np.array_equal(np.asarray(v), z["tab_" + k]), k
    for k, v in o.vert_tables().items():
        assert np.array_equal(v, z[k]), k
    for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav", "cv_d", "gamma_d", "kappa_d", "C0"):
        assert o.option(k) == float(z["opt_" + k]), k
    st = MG.stencils(ord)
    assert np.array_equal(o.reconstruct(st, 0), z["recon_h0"])
    assert np.array_equal(o.reconstruct(st, 1), z["recon_h1"])
    lv = min(3, case["nz"] + 1)
    assert np.array_equal(o.reconstruct_vert(lv, 0, st, 0), z["recon_v0"])
    assert np.array_equal(o.reconstruct_vert(lv, 0, st, 1), z["recon_v1"])
    # hydrostatic balance, CFL, tendencies, steps
    o.declare_hydrostatic()
    assert rel_l2(o.get_field("variable_gravity"), z["variable_gravity"]) <= TOL
    assert abs(o.compute_time_step() / float(z["dt_cfl"]) - 1.0) <= 1e-15
    o.set_variable_gravity(z["variable_gravity"])
    state, tracers = MG.conserved(case, o)
    stt, ttt, _, _ = o.compute_tendencies(state, tracers, tracers, 0.7)
    for l in range(5):
        scale = max(np.linalg.norm(z["state_tend"][l]), 1e-6 * np.linalg.norm(z["state_tend"][0]) * 350.0)
        assert np.linalg.norm(stt[l] - z["state_tend"][l]) / scale <= TOL, l
    for t in range(len(case["tracers"])):
        scale = max(np.linalg.norm(z["tracers_tend"][t]), 1e-12 * np.linalg.norm(tracers[t]), 1e-300)
        assert np.linalg.norm(ttt[t] - z["tracers_tend"][t]) / scale <= 1e-12, t
    m0 = o.compute_mass()
    assert np.max(np.abs(m0 - z["mass0"]) / (np.abs(z["mass0"]) + 1e-300)) <= TOL
    assert o.time_step() == int(z["ncycles1"])
    ref = {k: z["s1_" + k] for k in o.get_fields()}
    fl = field_floors(ref)
    for k, v in o.get_fields().items():
        assert rel_l2(v, ref[k], fl[k]) <= TOL, (k, rel_l2(v, ref[k], fl[k]))
    o.time_step(); o.time_step()
    ref = {k: z["s3_" + k] for k in o.get_fields()}
    for k, v in o.get_fields().items():
        assert rel_l2(v, ref[k], fl[k]) <= 10 * TOL, (k, rel_l2(v, ref[k], fl[k]))
    m3 = o.compute_mass()
    assert np.max(np.abs(m3 - z["mass3"]) / (np.abs(z["mass3"]) + 1e-300)) <= 10 * TOL


@pytest.mark.skipif(not reflib.available(5), reason="oracle/_ref (compiled reference) not built on this host")
@pytest.mark.parametrize("name", ["c2d_ord5", "c3d_ord5", "c2d_ord5_nt10", "c2d_ord9"])
def test_oracle_is_bit_identical_to_compiled_reference(name):
    kw, ord, crm_dt = MG.CASES[name]
    if not reflib.available(ord):
        pytest.skip("reference library of this order not built")
    case = MG.build_case(kw)
    o = H.make_oracle(case, ord, crm_dt=crm_dt)
    r = reflib.RefDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                         tracers=case["tracers"], consts=case["consts"], crm_dt=crm_dt, ord=ord)
    r.set_fields(case["fields"])
    o.declare_hydrostatic(); r.declare_hydrostatic()
    assert np.array_equal(o.get_field("variable_gravity"), r.get_field("variable_gravity", (case["nz"], case["nens"])))
    for _ in range(2):
        assert o.time_step() == r.time_step()
    fo, fr = o.get_fields(), r.get_fields()
    for k in fo:
        assert np.array_equal(fo[k], fr[k]), k
    r.close()


@pytest.mark.skipif(not reflib.available(5), reason="oracle/_ref (compiled reference) not built on this host")
def test_golden_vectors_reproduce_from_reference():
    """The committed vectors are what the reference produces today (guards against a stale fixture)."""
    z, meta = _load(os.path.join(ROOT, "tests", "golden", "c2d_ord5.npz"))
    case = MG.build_case(meta["case"])
    r = reflib.RefDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                         tracers=case["tracers"], consts=case["consts"], crm_dt=meta["crm_dt"], ord=meta["ord"])
    r.set_fields(case["fields"])
    r.declare_hydrostatic()
    r.time_step()
    for k, v in r.get_fields().items():
        assert rel_l2(v, z["s1_" + k], 1e-300) <= TOL, k
    r.close()


@pytest.mark.skipif(not reflib.available(5), reason="oracle/_ref (compiled reference) not built on this host")
@pytest.mark.parametrize("name", ["c2d_ord5", "c3d_ord5", "c2d_ord5_nt10", "c2d_ord3"])
def test_oracle_haloed_methods_are_bit_identical_to_the_reference_methods(name):
    """The reference's other public Dycore methods take its HALOED arrays (Dycore.h:37, :262, :608, :1281, :1336). The
    oracle's restatements of them -- the checkers of the C ABI's awfl_*_haloed / awfl_convert_* / awfl_halo_exchange --
    against the compiled reference's own methods on the same arrays, ghosts included."""
    kw, ord, crm_dt = MG.CASES[name]
    if not reflib.available(ord):
        pytest.skip("reference library of this order not built")
    case = MG.build_case(kw)
    o = H.make_oracle(case, ord, crm_dt=crm_dt)
    r = reflib.RefDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                         tracers=case["tracers"], consts=case["consts"], crm_dt=crm_dt, ord=ord)
    r.set_fields(case["fields"])
    o.declare_hydrostatic(); r.declare_hydrostatic()
    nt, hsh, hs = len(case["tracers"]), o.halo_shape(), o.hs
    assert hsh == r.halo_shape()
    rng = np.random.default_rng(3)
    junk = lambda n: rng.standard_normal((n,) + hsh)      # ghosts start as junk: every one the method owns is overwritten
    so, to = junk(5), junk(nt)
    sr, tr = so.copy(), to.copy()
    o.convert_coupler_to_dynamics(so, to); r.convert_coupler_to_dynamics(sr, tr)
    assert np.array_equal(so, sr) and np.array_equal(to, tr)
    assert np.array_equal(o.compute_mass_dynamics(so, to), r.compute_mass_dynamics(sr, tr))
    tto = np.ascontiguousarray(np.maximum(np.stack([case["fields"][t[0]] for t in case["tracers"]]), 0.0))
    ttr = tto.copy()
    sto = o.compute_tendencies_haloed(so, to, tto, 0.7)
    str_ = r.compute_tendencies_haloed(sr, tr, ttr, 0.7)
    assert np.array_equal(sto, str_) and np.array_equal(tto, ttr)
    assert np.array_equal(so, sr) and np.array_equal(to, tr)            # round-tripped interior AND ghosts
    # halo_exchange proper works on the divided arrays (Dycore.h:316-323); its ghosts come from junk here
    do, dto, po = junk(5), junk(nt), junk(1)[0]
    do[:, hs:-hs, hs:-hs, hs:-hs] = so[:, hs:-hs, hs:-hs, hs:-hs]
    dto[:, hs:-hs, hs:-hs, hs:-hs] = to[:, hs:-hs, hs:-hs, hs:-hs]
    do[1:] /= do[0]; dto /= do[0]
    dr, dtr, pr = do.copy(), dto.copy(), po.copy()
    o.halo_exchange(do, dto, po); r.halo_exchange(dr, dtr, pr)
    assert np.array_equal(do, dr) and np.array_equal(dto, dtr) and np.array_equal(po, pr)
    assert np.isfinite(do[:, :, hs:-hs, hs:-hs]).all() and np.isfinite(po[:hs, hs:-hs, hs:-hs]).all()
    o.convert_dynamics_to_coupler(so, to); r.convert_dynamics_to_coupler(sr, tr)
    fo, fr = o.get_fields(), r.get_fields()
    for k in fo:
        assert np.array_equal(fo[k], fr[k]), k
    r.close()
