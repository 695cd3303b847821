"""Shared by the GPU parity tests: build the same configuration on the CUDA path (through the C ABI) and on the oracle."""
import numpy as np

from oracle import oraclelib
from pam_b200 import synthetic as syn

STATE_FIELDS = ("density_dry", "uvel", "vvel", "wvel", "temp")


def make_oracle(case, ord=5, crm_dt=1.0):
    o = oraclelib.OracleDycore(case["nz"], case["ny"], case["nx"], case["nens"], case["xlen"], case["ylen"], case["zint"],
                               tracers=case["tracers"], consts=case["consts"], crm_dt=crm_dt, ord=ord)
    o.set_fields(case["fields"])
    return o


def make_gpu(case, ord=5, crm_dt=1.0, device="cuda:0"):
    import torch
    from pam_b200 import PamCoupler, Dycore
    c = PamCoupler(device)
    c.set_option("crm_dt", float(crm_dt))
    if case["consts"] is not None:
        for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav"):
            c.set_option(k, case["consts"][k])
    c.allocate_coupler_state(case["nz"], case["ny"], case["nx"], case["nens"])
    c.set_grid(case["xlen"], case["ylen"], case["zint"])
    for name, pos, adds in case["tracers"]:
        c.add_tracer(name, name, pos, adds)
    d = Dycore(ord)
    d.init(c)
    dm = c.get_data_manager_device_readwrite()
    for k, v in case["fields"].items():
        dm.get(k).copy_(torch.from_numpy(v))
    return c, d


def gpu_fields(c):
    dm = c.get_data_manager_device_readonly()
    names = list(STATE_FIELDS) + c.get_tracer_names()
    return {n: dm.get(n, readonly=True).cpu().numpy() for n in names}


def dycore_state_from_fields(o, fields):
    """Conserved (rho, rho u, rho v, rho w, rho theta) + tracers from coupler fields, the way
    Dycore::convert_coupler_to_dynamics does (Dycore.h:1370-1387); numpy, test-side only."""
    R_d, R_v, C0, gamma = o.option("R_d"), o.option("R_v"), o.option("C0"), o.option("gamma_d")
    names = o.tracer_names
    rho_d, T = fields["density_dry"], fields["temp"]
    rho_v = fields["water_vapor"]
    press = rho_d * R_d * T + rho_v * R_v * T
    rho = rho_d.copy()
    for (n, pos, adds) in zip(names, [True] * len(names), [None] * len(names)):
        pass
    return R_d, R_v, C0, gamma, press, rho
