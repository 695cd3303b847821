#!/usr/bin/env python3
"""Times the reference's OWN GPU path -- the unmodified dynamics/awfl/Dycore.h over the YAKL CUDA backend, compiled for
sm_100a into oracle/_ref/libpam_ref_ord<N>_cuda.so (oracle/ref_build/pam_cuda_harness.cpp) -- on a bench.py workload, on the
same B200 as this repo's kernels. Dycore::timeStep through coupler.run_module, CUDA events on the default stream.
Prints one JSON line.   python tools/gpu_reference_bench.py --workload C2 [--nens N] [--steps K]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def gpu_reference_rate(w, crm_dt=1.0, steps=5, warm=2, nens=None, pool_mb=None):
    import bench
    from oracle import pamcuda
    nens = nens or w["nens"]
    case = bench.build_case(w, nens=nens)
    ncell = w["nz"] * w["ny"] * w["nx"] * nens
    nt = len(case["tracers"])
    hs = (w["ord"] + 1) // 2
    halo = (w["nz"] + 2 * hs) * (w["ny"] + 2 * hs) * (w["nx"] + 2 * hs) * nens
    # the reference's temporaries per timeStep: 2 haloed state + 2 haloed tracer copies, pressure, 6 flux arrays, tendencies
    need = 8 * (2 * 5 * halo + 2 * nt * halo + halo + 4 * (5 + nt) * ncell + 4 * (5 + nt) * ncell) + 8 * ncell * (5 + nt) * 2
    pool_mb = pool_mb or int(need / 2**20 * 1.3) + 2048
    r = pamcuda.PamCudaDycore("ref", w["nz"], w["ny"], w["nx"], nens, case["xlen"], case["ylen"], case["zint"], tracers=case["tracers"],
                              consts=case["consts"], crm_dt=crm_dt, ord=w["ord"], pool_mb=pool_mb)
    r.set_fields(case["fields"])
    r.declare_hydrostatic(); r.declare_hydrostatic()       # twice: SURVEY.md F2/F4
    dt = r.compute_time_step()
    ncycles = int(np.ceil(crm_dt / dt))
    r.time_steps_ms(warm)
    ms = r.time_steps_ms(steps)
    ok = bool(np.isfinite(r.get_field("temp")).all())
    r.close()
    upd = float(ncell) * ncycles * steps
    return {"value": upd / (ms * 1e-3), "unit": "cell-updates/s", "ms_per_step": ms / steps, "steps": steps, "ncycles": ncycles,
            "nens": nens, "finite": ok, "pool_mb": pool_mb,
            "what": "the UNMODIFIED reference dycore (dynamics/awfl/Dycore.h) through its YAKL CUDA backend, nvcc -O3 sm_100a, "
                    "Dycore::timeStep via coupler.run_module on the same GPU (oracle/_ref/libpam_ref_ord%d_cuda.so)" % w["ord"]}


if __name__ == "__main__":
    import bench
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--nens", type=int, default=0)
    ap.add_argument("--ord", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    w = dict(bench.WORKLOADS[a.workload])
    if a.ord:
        w["ord"] = a.ord
    out = gpu_reference_rate(w, steps=a.steps, nens=a.nens or None)
    out["workload"] = a.workload
    print(json.dumps(out))
