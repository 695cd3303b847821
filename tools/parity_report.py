#!/usr/bin/env python3
"""Prints the parity margins of the CUDA path against the plain-C oracle (relative L2 per field) for a few small
cases after 1 and N steps. Test-side tool (uses oracle/): run on the GPU box, e.g. `python tools/parity_report.py 20`."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
from conftest import rel_l2, field_floors
from pam_b200 import synthetic as syn

nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
CASES = [
    ("2d ord5 tanh", dict(nz=16, ny=1, nx=20, nens=40, grid="tanh"), 5),
    ("2d ord5 per-ens grid", dict(nz=16, ny=1, nx=20, nens=40, grid="tanh", per_ens_grid=True), 5),
    ("3d ord5", dict(nz=10, ny=6, nx=8, nens=33, grid="tanh"), 5),
    ("2d ord5 nt10", dict(nz=16, ny=1, nx=20, nens=32, grid="tanh", tracers=syn.TRACERS_P3_SHOC), 5),
    ("2d ord3", dict(nz=16, ny=1, nx=20, nens=40, grid="tanh"), 3),
    ("2d ord7", dict(nz=16, ny=1, nx=20, nens=40, grid="tanh"), 7),
    ("2d ord9", dict(nz=16, ny=1, nx=20, nens=40, grid="tanh"), 9),
]
for name, kw, ord in CASES:
    case = syn.make_case(**kw)
    o = H.make_oracle(case, ord, crm_dt=2.0)
    c, d = H.make_gpu(case, ord, crm_dt=2.0)
    o.declare_hydrostatic(); d.declare_current_profile_as_hydrostatic(c)
    out = []
    for step in range(1, nsteps + 1):
        o.time_step(); d.timeStep(c)
        if step in (1, nsteps):
            ref, got = o.get_fields(), H.gpu_fields(c)
            fl = field_floors(ref)
            worst = max((rel_l2(got[k], ref[k], fl[k]), k) for k in ref)
            out.append("step %d: worst %.2e (%s)" % (step, worst[0], worst[1]))
    print("%-24s %s" % (name, "; ".join(out)), flush=True)
    d.finalize()

# ---- against the reference's OWN kernels on this GPU (oracle/_ref/libpam_ref_ord5_cuda.so: the unmodified reference through
# its YAKL CUDA backend): its step with the bottom-ghost race neutralised ("intended", the oracle's semantics) and its
# unmodified Dycore::timeStep ("raw": whatever the racing threads and the pool's stale memory give) ----
from oracle import pamcuda
if pamcuda.available("ref", 5):
    print("\nagainst the reference's YAKL-CUDA build on the same GPU (one step, worst field):")
    for name, kw, ord in [c for c in CASES if c[2] == 5]:
        case = syn.make_case(**kw)
        c, d = H.make_gpu(case, ord, crm_dt=2.0)
        d.declare_current_profile_as_hydrostatic(c)
        gv = c.get_data_manager_device_readonly().get("variable_gravity").cpu().numpy()
        res = {}
        for mode in ("raw", "intended"):      # raw FIRST: its haloed arrays then come from pool memory no intended run has pre-filled
            r = pamcuda.PamCudaDycore("ref", case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                                      tracers=case["tracers"], consts=case["consts"], crm_dt=2.0, ord=ord)
            r.set_fields(case["fields"])
            r.set_field("variable_gravity", gv)
            if mode == "intended":
                r.time_step_public()
            else:
                r.time_step()
            res[mode] = r.get_fields()
            r.close()
        d.timeStep(c)
        got = H.gpu_fields(c)
        fl = field_floors(res["intended"])
        w_int = max((rel_l2(got[k], res["intended"][k], fl[k]), k) for k in got)
        w_raw = max((rel_l2(got[k], res["raw"][k], fl[k]), k) for k in got)
        w_rr = max((rel_l2(res["raw"][k], res["intended"][k], fl[k]), k) for k in got)
        print("%-24s this repo vs intended %.2e (%s); vs raw timeStep %.2e (%s); reference raw vs its own intended %.2e (%s)"
              % (name, w_int[0], w_int[1], w_raw[0], w_raw[1], w_rr[0], w_rr[1]), flush=True)
        d.finalize()
