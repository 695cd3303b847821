"""Which fields differ, and by how much, between the 2-D state marches without the v reconstruction and the general ones
(AWFL_GENERAL_2D=1)? Run on a GPU box from the repo root: python tools/diag_general2d.py"""
import os, sys, numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import helpers as H
from pam_b200 import synthetic as syn
for ord in (3, 5, 7, 9):
    for grid, per in (("equal", False), ("tanh", False), ("tanh", True)):
        case = syn.make_case(14, 1, 10, 40, tracers=syn.TRACERS_NT1, grid=grid, per_ens_grid=per, seed=12345)
        out = {}
        for general in ("0", "1"):
            os.environ["AWFL_GENERAL_2D"] = general
            c, d = H.make_gpu(case, ord, crm_dt=2.0)
            d.declare_current_profile_as_hydrostatic(c)
            for _ in range(int(os.environ.get("DIAG_STEPS", "2"))):
                d.timeStep(c)
            out[general] = H.gpu_fields(c)
        msg = []
        for k in out["0"]:
            a, b = out["0"][k], out["1"][k]
            if not np.array_equal(a, b):
                nd = int((a != b).sum())
                msg.append("%s: %d/%d differ, max rel %.2e, l2 rel %.2e" % (k, nd, a.size, np.abs(a - b).max() / np.abs(b).max(),
                                                                           np.linalg.norm(a - b) / np.linalg.norm(b)))
        print(ord, grid, per, msg or "identical", flush=True)
