#!/usr/bin/env python3
"""Measure the coupler-module kernels (sponge layer, GCM forcing compute/apply, perturb_temperature) and the
idealized initial states on one B200: CUDA events around repeated calls on an MMF-shaped coupler (2-D 64x58, P3+SHOC
tracers), algorithmic bytes / time against the measured HBM peak. Prints one JSON line.

    python tools/bench_modules.py [--nens 4096] [--reps 20]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nens", type=int, default=4096)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--ny", type=int, default=1)
    ap.add_argument("--nx", type=int, default=64)
    args = ap.parse_args()
    import torch
    from pam_b200 import PamCoupler, Dycore, modules
    from pam_b200 import synthetic as syn
    if not torch.cuda.is_available():
        raise RuntimeError("needs a CUDA device")
    nz, ny, nx, nens = 58, args.ny, args.nx, args.nens
    base = syn.make_case(nz, ny, nx, min(nens, 68), tracers=syn.TRACERS_P3_SHOC, grid="equal")
    c = PamCoupler("cuda:0")
    c.set_option("crm_dt", 10.0)
    c.set_option("gcm_physics_dt", 1200.0)
    for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav"):
        c.set_option(k, base["consts"][k])
    c.allocate_coupler_state(nz, ny, nx, nens)
    c.set_grid(base["xlen"], base["ylen"], base["zint"])
    for name, pos, adds in base["tracers"]:
        c.add_tracer(name, name, pos, adds)
    d = Dycore(5)
    d.init(c)
    dm = c.get_data_manager_device_readwrite()
    reps = (nens + base["nens"] - 1) // base["nens"]
    for k, v in base["fields"].items():
        dm.get(k).copy_(torch.from_numpy(np.ascontiguousarray(np.tile(v, (1, 1, 1, reps))[..., :nens])))
    for g, f in zip(modules.GCM_NAMES, modules.CRM_NAMES):
        dm.get(g).copy_(0.97 * dm.get(f).mean(dim=(1, 2)))
    ncell = nz * ny * nx * nens
    nfields = 5 + len(base["tracers"])
    ids = torch.arange(nens, dtype=torch.int32)

    def timed(fn, reps=args.reps):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    out = {"workload": "%dx%dx58, nens=%d, P3+SHOC tracers (15 fields)" % (nx, ny, nens), "hbm_peak_gbs": hbm, "kernels": {}}

    def rec(name, ms, bytes_alg, note):
        out["kernels"][name] = {"ms": ms, "algorithmic_bytes": bytes_alg, "gbs": bytes_alg / ms / 1e6, "frac_hbm": bytes_alg / ms / 1e6 / hbm,
                                "note": note}

    modules.compute_gcm_forcing_tendencies(c)
    rec("compute_gcm_forcing_tendencies", timed(lambda: modules.compute_gcm_forcing_tendencies(c)), 8 * 10 * ncell,
        "reads the 10 forced fields once")
    rec("apply_gcm_forcing_tendencies", timed(lambda: modules.apply_gcm_forcing_tendencies(c)), 8 * 20 * ncell,
        "reads and writes the 10 forced fields once (+ the levels that need hole filling a second time); 2 launches")
    rec("sponge_layer", timed(lambda: modules.sponge_layer(c)), 8 * 3 * nfields * 5 * (ncell // nz),
        "top 5 levels of 15 fields: read (mean), read + write (relaxation)")
    rec("perturb_temperature", timed(lambda: modules.perturb_temperature(c, ids)), 8 * 4 * (nz // 4) * (ncell // nz),
        "lowest nz/4 levels of temp: read + write (perturb), read + write (rescale); 21 smallprng rounds per cell")
    if ny > 1:
        print(json.dumps(out))
        return
    rec("init_supercell", timed(lambda: d.init_idealized_state_and_tracers(c, "supercell"), reps=3), 8 * nfields * ncell,
        "init-time: writes 15 fields; 729-point quadrature per cell + per-member hydrostatic integration (synchronises)")
    rec("init_thermal", timed(lambda: d.init_idealized_state_and_tracers(c, "thermal"), reps=3), 8 * nfields * ncell,
        "init-time: writes 15 fields; 729-point quadrature with cos/sqrt per cell (FP64-bound, synchronises)")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
