#!/usr/bin/env python3
"""Small cases through every kernel of the time step, for compute-sanitizer (racecheck / initcheck / memcheck): 2-D and 3-D,
one and ten tracers (fused and two-kernel FCT), orders 3 / 5 / 9, shared and member-dependent vertical grids, the constant-bank
and the global table homes, the host-buffer entry point with member slabs, the haloed-array entry points. Each case is checked
against the oracle as well, so a sanitizer run is also a parity run.
   compute-sanitizer --tool racecheck python tools/sanitize_case.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import helpers as H
from conftest import rel_l2, field_floors
from pam_b200 import synthetic as syn

CASES = [("2d ord5", dict(nz=10, ny=1, nx=12, nens=40), 5, {}),
         ("3d ord5", dict(nz=8, ny=5, nx=6, nens=33), 5, {}),
         ("2d nt10 ord5", dict(nz=10, ny=1, nx=12, nens=32, tracers=syn.TRACERS_P3_SHOC), 5, {}),
         ("3d nt10 ord5", dict(nz=8, ny=5, nx=6, nens=8, tracers=syn.TRACERS_P3_SHOC), 5, {}),
         ("2d per-member grid", dict(nz=10, ny=1, nx=12, nens=33, per_ens_grid=True), 5, {}),
         ("2d global table", dict(nz=10, ny=1, nx=12, nens=33), 5, {"AWFL_NO_CVTAB": "1"}),
         ("2d ord3", dict(nz=10, ny=1, nx=12, nens=33), 3, {}),
         ("2d ord9", dict(nz=12, ny=1, nx=12, nens=33), 9, {})]
worst_all = 0.0
for name, kw, ord, env in CASES:
    os.environ.update(env)
    case = syn.make_case(grid="tanh", **kw)
    o = H.make_oracle(case, ord, crm_dt=2.0)
    c, d = H.make_gpu(case, ord, crm_dt=2.0)
    o.declare_hydrostatic(); d.declare_current_profile_as_hydrostatic(c)
    names = list(H.STATE_FIELDS) + c.get_tracer_names()
    host = [np.ascontiguousarray(case["fields"][n]).copy() for n in names]
    n = d.timeStep(c)
    assert n == o.time_step()
    d.set_host_slabs(2)
    d.time_step_host(host, 2.0, ncycles=n)
    ref, got = o.get_fields(), H.gpu_fields(c)
    fl = field_floors(ref)
    worst = max(rel_l2(got[k], ref[k], fl[k]) for k in ref)
    for a, k in zip(host, names):
        assert np.array_equal(a, got[k]), k
    worst_all = max(worst_all, worst)
    d.compute_mass(c); d.compute_time_step(c)
    # the haloed-array entry points
    hs, nt = d.hs, len(case["tracers"])
    hsh = (case["nz"] + 2 * hs, case["ny"] + 2 * hs, case["nx"] + 2 * hs, case["nens"])
    st = torch.zeros((5,) + hsh, dtype=torch.float64, device=c.device); tr = torch.zeros((nt,) + hsh, dtype=torch.float64, device=c.device)
    stt = torch.zeros((5, case["nz"], case["ny"], case["nx"], case["nens"]), dtype=torch.float64, device=c.device)
    trt = torch.zeros((nt, case["nz"], case["ny"], case["nx"], case["nens"]), dtype=torch.float64, device=c.device)
    pr = torch.zeros(hsh, dtype=torch.float64, device=c.device)
    f, keep = d._fields(c)
    import ctypes as C
    lib = d.lib
    assert lib.awfl_convert_coupler_to_dynamics(d.h, C.byref(f), st.data_ptr(), tr.data_ptr()) == 0
    assert lib.awfl_compute_tendencies_haloed(d.h, st.data_ptr(), stt.data_ptr(), tr.data_ptr(), trt.data_ptr(), 0.5) == 0
    assert lib.awfl_halo_exchange(d.h, st.data_ptr(), tr.data_ptr(), pr.data_ptr()) == 0
    assert lib.awfl_convert_dynamics_to_coupler(d.h, st.data_ptr(), tr.data_ptr(), C.byref(f)) == 0
    torch.cuda.synchronize()
    print("%-22s ok: %d sub-cycles, worst rel-L2 vs oracle %.2e" % (name, n, worst), flush=True)
    d.finalize()
    for k in env:
        os.environ.pop(k, None)
assert worst_all <= 1e-12
print("sanitize_case: all cases done, worst %.2e" % worst_all)
