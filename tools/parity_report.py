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
