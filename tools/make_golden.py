#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the UNMODIFIED compiled reference (oracle/_ref, built by oracle/ref_build/Makefile from
/root/reference). Run in the build container only (the reference tree does not exist on the GPU box):

    python tools/make_golden.py

Every fixture stores the generator arguments of its inputs (pam_b200.synthetic.make_case is deterministic) and the
reference's outputs: constant tables, vertical matrices, reconstruct() on seeded stencils, variable_gravity, one
compute_tendencies() call, the CFL time step, the fields after one and three time steps and compute_mass(). The
"intended-semantics" entry points of the harness are used (the bottom-boundary read/write hazard of the
raw reference is neutralised by pre-filling the ghost theta the racing thread would have written).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import reflib                      # noqa: E402
from pam_b200 import synthetic as syn          # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

# name -> (make_case kwargs, ord, crm_dt).  Small enough that the whole set stays ~1 MB compressed.
CASES = {
    "c2d_ord5": (dict(nz=12, ny=1, nx=10, nens=3, grid="tanh"), 5, 2.0),
    "c2d_ord5_perens": (dict(nz=12, ny=1, nx=10, nens=3, grid="tanh", per_ens_grid=True), 5, 2.0),
    "c3d_ord5": (dict(nz=10, ny=5, nx=6, nens=2, grid="tanh"), 5, 2.0),
    "c2d_ord5_nt10": (dict(nz=12, ny=1, nx=10, nens=2, grid="tanh", tracers="P3_SHOC"), 5, 2.0),
    "c2d_ord3": (dict(nz=12, ny=1, nx=10, nens=2, grid="tanh"), 3, 2.0),
    "c2d_ord7": (dict(nz=12, ny=1, nx=10, nens=2, grid="tanh"), 7, 2.0),
    "c2d_ord9": (dict(nz=12, ny=1, nx=10, nens=2, grid="equal"), 9, 2.0),
    "c1_like": (dict(nz=10, ny=1, nx=16, nens=1, grid="equal", consts="DEFAULT"), 5, 3.0),
}


def build_case(kw):
    kw = dict(kw)
    tr = kw.pop("tracers", "NT1")
    kw["tracers"] = syn.TRACERS_NT1 if tr == "NT1" else syn.TRACERS_P3_SHOC
    cs = kw.pop("consts", "P3")
    kw["consts"] = syn.P3_CONSTS if cs == "P3" else syn.DEFAULT_CONSTS
    return syn.make_case(**kw)


def stencils(ord, n=64, seed=7):
    rng = np.random.default_rng(seed)
    st = rng.standard_normal((n, ord)) * np.exp(rng.uniform(-3, 3, (n, 1))) + rng.uniform(-5, 5, (n, 1))
    st[:4] = 1.0
    st[4:8] = np.where(np.arange(ord) < ord // 2, 0.0, 1.0)
    st[8:12] *= 1e8
    return st


def conserved(case, r):
    f = case["fields"]
    R_d, R_v, C0, gamma = r.option("R_d"), r.option("R_v"), r.option("C0"), r.option("gamma_d")
    rho = f["density_dry"].copy()
    for (n, pos, adds) in case["tracers"]:
        if adds:
            rho = rho + f[n]
    press = f["density_dry"] * R_d * f["temp"] + f["water_vapor"] * R_v * f["temp"]
    theta = np.power(press / C0, 1.0 / gamma) / rho
    state = np.stack([rho, rho * f["uvel"], rho * f["vvel"], rho * f["wvel"], rho * theta])
    tracers = np.stack([np.maximum(f[n], 0.0) for (n, _, _) in case["tracers"]])
    return state, tracers


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, (kw, ord, crm_dt) in CASES.items():
        if not reflib.available(ord):
            raise SystemExit("oracle/_ref/libpam_ref_ord%d.so missing: run __graft_entry__.build() in the build container" % ord)
        case = build_case(kw)
        r = reflib.RefDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                             tracers=case["tracers"], consts=case["consts"], crm_dt=crm_dt, ord=ord)
        out = {"meta": np.frombuffer(json.dumps(dict(case=kw, ord=ord, crm_dt=crm_dt)).encode(), dtype=np.uint8)}
        for k, v in r.tables().items():
            out["tab_" + k] = np.asarray(v)
        for k, v in r.vert_tables().items():
            out[k] = v
        for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav", "cv_d", "gamma_d", "kappa_d", "C0"):
            out["opt_" + k] = np.float64(r.option(k))
        st = stencils(ord)
        out["recon_h0"], out["recon_h1"] = r.reconstruct(st, 0), r.reconstruct(st, 1)
        lv = min(3, case["nz"] + 1)
        out["recon_v0"], out["recon_v1"] = r.reconstruct_vert(lv, 0, st, 0), r.reconstruct_vert(lv, 0, st, 1)
        r.set_fields(case["fields"])
        r.declare_hydrostatic()
        out["variable_gravity"] = r.get_field("variable_gravity", (case["nz"], case["nens"]))
        out["dt_cfl"] = np.float64(r.compute_time_step())
        state, tracers = conserved(case, r)
        stt, ttt, _, _ = r.compute_tendencies(state, tracers, tracers, 0.7)
        out["state_tend"], out["tracers_tend"] = stt, ttt
        out["mass0"] = r.compute_mass()
        out["ncycles1"] = np.int64(r.time_step())
        for k, v in r.get_fields().items():
            out["s1_" + k] = v
        r.time_step(); r.time_step()
        for k, v in r.get_fields().items():
            out["s3_" + k] = v
        out["mass3"] = r.compute_mass()
        path = os.path.join(OUT, name + ".npz")
        np.savez_compressed(path, **out)
        print("%-18s %7.1f kB" % (name, os.path.getsize(path) / 1e3))
        r.close()


if __name__ == "__main__":
    main()
