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


# ---------------------------------------------------------------------------------------------------------------------
# Coupler modules either side of the dycore: a numpy "DataManager" with the P3+SHOC tracer set and GCM
# columns, shared by the oracle tests (CPU) and the CUDA parity tests.
# ---------------------------------------------------------------------------------------------------------------------
MODULE_GCM = ("gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice",
              "gcm_num_liq", "gcm_num_ice", "gcm_num_rain")
MODULE_TEND = ("rho_d", "uvel", "vvel", "temp", "qtot", "qv", "ql", "qi", "rho_v", "rho_l", "rho_i", "nc", "ni", "nr")


def make_modules_case(nz=12, ny=1, nx=10, nens=5, seed=7, overshoot=False):
    """Fields of syn.make_case (P3+SHOC tracers: exact zeros over a third of the cells) + GCM columns built from the CRM
    column means: winds / temperature shifted, water species scaled DOWN so that the forcing removes water and cells that
    hold none go negative (level hole filling). overshoot=True sets crm_dt = 3 x gcm_physics_dt -- unphysical, but it makes
    whole levels owe more water than they hold, which is what sends fill_holes to its whole-CRM stage."""
    case = syn.make_case(nz, ny, nx, nens, tracers=syn.TRACERS_P3_SHOC, grid="tanh", seed=seed)
    dm = {k: np.ascontiguousarray(v).copy() for k, v in case["fields"].items()}
    zint = np.asarray(case["zint"], dtype=np.float64)
    if zint.ndim == 1:
        zint = np.repeat(zint[:, None], nens, axis=1)
    dm["vertical_interface_height"] = np.ascontiguousarray(zint)
    dm["vertical_cell_dz"] = np.ascontiguousarray(zint[1:] - zint[:-1])
    dm["vertical_midpoint_height"] = np.ascontiguousarray(0.5 * (zint[1:] + zint[:-1]))
    rng = np.random.default_rng(seed)
    mean = lambda n: dm[n].mean(axis=(1, 2))
    dm["gcm_density_dry"] = mean("density_dry") * (1.0 + 1e-3 * rng.standard_normal((nz, nens)))
    dm["gcm_uvel"] = mean("uvel") + 0.5 * rng.standard_normal((nz, nens))
    dm["gcm_vvel"] = mean("vvel") + 0.5 * rng.standard_normal((nz, nens))
    dm["gcm_temp"] = mean("temp") + 0.3 * rng.standard_normal((nz, nens))
    dm["gcm_water_vapor"] = mean("water_vapor") * rng.uniform(0.3, 1.1, (nz, nens))
    dm["gcm_cloud_water"] = mean("cloud_water") * rng.uniform(0.0, 0.8, (nz, nens))
    dm["gcm_cloud_ice"] = mean("ice") * rng.uniform(0.0, 1.2, (nz, nens))
    dm["gcm_num_liq"] = mean("cloud_water_num") * rng.uniform(0.0, 1.5, (nz, nens))
    dm["gcm_num_ice"] = mean("ice_num") * rng.uniform(0.0, 1.5, (nz, nens))
    dm["gcm_num_rain"] = mean("rain_num") * rng.uniform(0.0, 1.5, (nz, nens))
    dm["gcm_wvel"] = np.zeros((nz, nens))
    for k in list(dm):
        dm[k] = np.ascontiguousarray(dm[k], dtype=np.float64)
    gcm_dt = 20.0
    opts = dict(crm_dt=(3.0 * gcm_dt if overshoot else 5.0), gcm_physics_dt=gcm_dt, sponge_time_scale=60.0, sponge_num_layers=min(5, nz - 1))
    return case, dm, opts


def modules_fields(case):
    """names of every array the module tests compare"""
    names = list(STATE_FIELDS) + [t[0] for t in case["tracers"]]
    return names, ["gcm_forcing_tend_" + n for n in MODULE_TEND]
