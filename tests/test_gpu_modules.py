"""GPU parity tests (through the C ABI of include/pam_modules_b200.h) of the coupler modules either side of the dycore
(SURVEY.md 8f): the sm_100a kernels against the CPU oracle (oracle/modules_oracle.c, pinned bit for bit to the compiled
reference) on the same seeded inputs and against the committed golden vectors of the reference itself.

Tolerance: FP64; the kernels sum every column in the reference's serial (j,i) order, so the only differences are FMA
contraction and CUDA's cos() vs glibc's -- relative L2 <= 1e-13 per field (measured ~1e-16). perturb_temperature is
integer-exact in its random numbers."""
import os

import numpy as np
import pytest
import torch

import helpers as H
from oracle import modules_oraclelib as M
from test_modules_oracle import CASES, GOLDEN, run_oracle_sequence

pytestmark = pytest.mark.gpu
TOL = 1e-13


def make_gpu_coupler(case, dm, opts):
    from pam_b200 import PamCoupler
    c = PamCoupler("cuda:0")
    for k, v in opts.items():
        c.set_option(k, v)
    c.allocate_coupler_state(case["nz"], case["ny"], case["nx"], case["nens"])
    c.set_grid(case["xlen"], case["ylen"], case["zint"])
    for name, pos, adds in case["tracers"]:
        c.add_tracer(name, name, pos, adds)
    d = c.get_data_manager_device_readwrite()
    names, _ = H.modules_fields(case)
    for n in names + list(H.MODULE_GCM):
        d.get(n).copy_(torch.from_numpy(dm[n]))
    return c


def rel(a, b, floor=1e-300):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), floor))


def tend_floors(dm, opts):
    """A forcing tendency is (GCM column - CRM column mean) / gcm_physics_dt: a difference of nearly equal numbers, so its
    round-off is relative to the OPERANDS, not to the (much smaller) result. Norm floor per tendency = |GCM operand| / dt."""
    r = 1.0 / opts["gcm_physics_dt"]
    tot = dm["gcm_density_dry"] + dm["gcm_water_vapor"]
    op = dict(rho_d=dm["gcm_density_dry"], uvel=dm["gcm_uvel"], vvel=dm["gcm_vvel"], temp=dm["gcm_temp"],
              qv=dm["gcm_water_vapor"] / tot, ql=dm["gcm_cloud_water"] / tot, qi=dm["gcm_cloud_ice"] / tot,
              rho_v=dm["gcm_water_vapor"], rho_l=dm["gcm_cloud_water"], rho_i=dm["gcm_cloud_ice"],
              nc=dm["gcm_num_liq"], ni=dm["gcm_num_ice"], nr=dm["gcm_num_rain"])
    op["qtot"] = op["qv"] + op["ql"] + op["qi"]
    return {"gcm_forcing_tend_" + k: float(np.linalg.norm(v)) * r for k, v in op.items()}


def run_gpu_sequence(c, ids):
    from pam_b200 import modules
    modules.perturb_temperature(c, ids, 0.1)
    modules.compute_gcm_forcing_tendencies(c)
    c.run_module("apply_gcm_forcing_tendencies", modules.apply_gcm_forcing_tendencies)
    c.run_module("sponge_layer", modules.sponge_layer)
    torch.cuda.synchronize()


@pytest.mark.parametrize("name", sorted(CASES))
def test_module_sequence_matches_oracle_and_golden(name):
    from pam_b200 import modules
    case, dm, opts = H.make_modules_case(**CASES[name])
    c = make_gpu_coupler(case, dm, opts)
    g = np.load(os.path.join(GOLDEN, "modules_%s.npz" % name))
    ids = g["ids"]
    run_gpu_sequence(c, ids)
    run_oracle_sequence(dm, case, opts, ids)
    d = c.get_data_manager_device_readonly()
    names, tends = H.modules_fields(case)
    fl = tend_floors(dm, opts)
    for n in names + tends:
        got = d.get(n, readonly=True).cpu().numpy()
        assert rel(got, dm[n], fl.get(n, 1e-300)) <= TOL, (n, rel(got, dm[n], fl.get(n, 1e-300)))
        assert rel(got, g[n], fl.get(n, 1e-300)) <= TOL, (n, "golden", rel(got, g[n], fl.get(n, 1e-300)))
    for n in ("water_vapor", "cloud_water", "ice", "cloud_water_num", "ice_num", "rain_num"):
        assert float(d.get(n, readonly=True).min()) >= 0.0
    assert modules.launch_count(c) == 5          # perturb, compute, apply + whole-CRM fill, sponge: one launch each


def test_each_module_alone_matches_oracle():
    from pam_b200 import modules
    case, dm, opts = H.make_modules_case(nz=16, ny=3, nx=7, nens=40, seed=3)
    tr = [t[0] for t in case["tracers"]]
    names, tends = H.modules_fields(case)

    fl = tend_floors(dm, opts)

    def check(c, which):
        d = c.get_data_manager_device_readonly()
        for n in which:
            got = d.get(n, readonly=True).cpu().numpy()
            assert rel(got, dm[n], fl.get(n, 1e-300)) <= TOL, (n, rel(got, dm[n], fl.get(n, 1e-300)))

    c = make_gpu_coupler(case, dm, opts)
    modules.sponge_layer(c)
    M.sponge_layer(dm, tr, opts["crm_dt"], opts["sponge_time_scale"], opts["sponge_num_layers"])
    check(c, names)
    ids = (np.arange(case["nens"]) * 7 + 2).astype(np.int32)
    modules.perturb_temperature(c, ids, 0.25)
    M.perturb_temperature(dm, ids, 0.25)
    d = c.get_data_manager_device_readonly()
    assert np.array_equal(d.get("temp", readonly=True).cpu().numpy()[case["nz"] // 4:], dm["temp"][case["nz"] // 4:])
    check(c, ["temp"])
    modules.compute_gcm_forcing_tendencies(c)
    M.compute_gcm_forcing_tendencies(dm, opts["gcm_physics_dt"])
    check(c, tends)
    modules.apply_gcm_forcing_tendencies(c)
    M.apply_gcm_forcing_tendencies(dm, opts["crm_dt"], opts["gcm_physics_dt"])
    check(c, names + tends)


def test_module_errors_are_reported():
    from pam_b200 import PamCoupler, modules
    c = PamCoupler("cuda:0")
    c.set_option("crm_dt", 1.0)
    c.allocate_coupler_state(8, 1, 8, 2)
    c.set_grid(8000.0, 1000.0, np.arange(9) * 100.0)
    c.add_tracer("water_vapor", "", True, True)
    with pytest.raises(RuntimeError):
        modules.apply_gcm_forcing_tendencies(c)       # compute_* has not registered the tendencies
    with pytest.raises(RuntimeError):
        modules.compute_gcm_forcing_tendencies(c)     # option gcm_physics_dt / P3 tracers missing
    with pytest.raises(RuntimeError):
        modules.perturb_temperature(c, [1, 2, 3])     # size of id array must be nens
    c.set_option("sponge_num_layers", 50)
    with pytest.raises(RuntimeError):
        modules.sponge_layer(c)


@pytest.mark.parametrize("cfg", [dict(nz=10, ny=12, nx=12, nens=35, seed=5), dict(nz=8, ny=32, nx=32, nens=33, seed=9),
                                 dict(nz=8, ny=16, nx=24, nens=40, seed=13, overshoot=True)])
def test_module_sequence_with_segmented_columns_matches_oracle(cfg):
    """3-D CRMs: a (level, member) column has ny*nx cells and is summed by S = ny*nx/64 (<= 16) threads, each over a
    contiguous run of cells, the partial sums combined in segment order. Same numbers as the oracle's serial sum up to
    the association of that last step: relative L2 <= 5e-13 per field (2, 16 and 6 segments; a ragged last member tile;
    hole filling at the level stage and -- overshoot -- the whole-CRM stage, whose x - neg * (x dz / pos) / dz cancels
    most of x and so amplifies the 1e-16 difference of the sums: 1.3e-13 measured on cloud_water)."""
    case, dm, opts = H.make_modules_case(**cfg)
    c = make_gpu_coupler(case, dm, opts)
    ids = (np.arange(case["nens"]) * 3 + 1).astype(np.int32)
    run_gpu_sequence(c, ids)
    run_oracle_sequence(dm, case, opts, ids)
    d = c.get_data_manager_device_readonly()
    names, tends = H.modules_fields(case)
    fl = tend_floors(dm, opts)
    for n in names + tends:
        got = d.get(n, readonly=True).cpu().numpy()
        assert rel(got, dm[n], fl.get(n, 1e-300)) <= 5 * TOL, (n, rel(got, dm[n], fl.get(n, 1e-300)))
    for n in ("water_vapor", "cloud_water", "ice", "cloud_water_num", "ice_num", "rain_num"):
        assert float(d.get(n, readonly=True).min()) >= 0.0
