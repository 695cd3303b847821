"""Headless standalone driver: the step loop of the reference's standalone/mmf_simplified/driver.cpp:79-297 without
MPI, NetCDF, P3 or SHOC. Same YAML keys and defaults, same order of operations:

    read config -> vertical grid -> coupler.allocate_coupler_state / set_grid -> dycore.init
    [idealized: initData inside dycore.init | else: supercell column -> GCM columns -> broadcast -> perturb_temperature]
    for each GCM step:    compute_gcm_forcing_tendencies                     (apply_gcm_forcing)
      for each CRM step:  apply_gcm_forcing_tendencies -> dycore.timeStep -> sponge_layer  (apply_sponge)
                          output every out_freq seconds

What differs, and why:
  * physics: "none" for microphysics and SGS (the reference's P3/SHOC need Fortran + an un-vendored SCREAM checkout);
    the water_vapor tracer the dycore needs is registered the way the "none" microphysics does
    (physics/micro/none/Microphysics.h).
  * output: snapshots of the coupler fields go to <out_prefix>_0.npz instead of NetCDF (output.h:8-123): arrays
    "etime" (nout) and one (nout, nz, ny, nx, nens) array per field; the vertical grid as zint.
  * vcoords: "uniform" (driver.cpp:149-168, needs crm_nz and zlen) is supported as is; a NetCDF file name cannot be read
    here, so the reference's generated equal grids are recognised by their name, vcoords_equal_<nz>_<ztop>km.nc
    (utils/generate_vertical_levels.py, function=equal), and a .npy / .txt file of interface heights is accepted.

  * hydrostatic balance: the reference's standalone driver never calls declare_current_profile_as_hydrostatic, so its
    "variable_gravity" stays whatever the allocator returned; this driver declares the initial profile
    hydrostatic once before the loop, which is what the MMF driver does at the start of every GCM step.

    python -m pam_b200.driver input.yaml
"""
import math
import re
import sys
import time

import numpy as np
import torch
import yaml

from . import modules
from .coupler import PamCoupler, endrun
from .dycore import Dycore
from . import synthetic


def vertical_interfaces(config):
    """zint (crm_nz+1,) from the vcoords key (driver.cpp:146-184)."""
    vc = str(config["vcoords"])
    if vc == "uniform":
        if "crm_nz" not in config:
            endrun("To use uniform vertical grid you need to specify crm_nz")
        nz, zlen = int(config["crm_nz"]), float(config.get("zlen", -1))            # driver.cpp:101: as<real>(-1)
        if zlen <= 0:
            endrun("ERROR: a uniform vertical grid needs zlen (with the AWFL dycore the reference's set_domain_sizes is a no-op, "
                   "pamc_init.h:43-53, and zlen stays -1)")
        dz = zlen / (nz - 1)                                   # driver.cpp:156 (sic: nz-1, first and last cells are half cells)
        z = np.array([0.0 if k == 0 else (zlen if k == nz else k * dz - dz / 2) for k in range(nz + 1)])
        return z
    m = re.match(r".*vcoords_equal_(\d+)_(\d+(?:\.\d+)?)km\.nc$", vc)
    if m:
        nz, ztop = int(m.group(1)), 1000.0 * float(m.group(2))
        return np.arange(nz + 1, dtype=np.float64) * (ztop / nz)
    if vc.endswith(".npy"):
        return np.asarray(np.load(vc), dtype=np.float64).reshape(-1)
    if vc.endswith(".txt"):
        return np.loadtxt(vc, dtype=np.float64).reshape(-1)
    endrun("ERROR: vcoords must be 'uniform', vcoords_equal_<nz>_<ztop>km.nc, or a .npy/.txt file of interface heights "
           "(NetCDF is not available in this build)")


def initialize_from_supercell_column(zint, coupler):
    """driver.cpp:14-77: supercell column -> GCM columns (+ reference state) -> broadcast -> perturb_temperature(seeds=0)."""
    dm = coupler.get_data_manager_device_readwrite()
    nz, nens = coupler.get_nz(), coupler.get_nens()
    col = synthetic.supercell_column(zint, coupler.get_option("R_d"), coupler.get_option("R_v"), coupler.get_option("grav"))
    for name, key in (("gcm_density_dry", "density_dry"), ("gcm_uvel", "uvel"), ("gcm_vvel", "vvel"), ("gcm_wvel", "wvel"),
                      ("gcm_temp", "temp"), ("gcm_water_vapor", "water_vapor")):
        t = torch.from_numpy(np.repeat(np.asarray(col[key])[:, None], nens, axis=1)).to(coupler.device)
        dm.get(name, 2).copy_(t)
    for name, key in (("ref_density_dry", "density_dry"), ("ref_density_vapor", "water_vapor"), ("ref_temp", "temp")):
        if not dm.entry_exists(name):
            dm.register_and_allocate(name, "", (nz, nens))
        dm.get(name, 2).copy_(torch.from_numpy(np.repeat(np.asarray(col[key])[:, None], nens, axis=1)).to(coupler.device))
    for name in ("ref_density_liq", "ref_density_ice"):
        if not dm.entry_exists(name):
            dm.register_and_allocate(name, "", (nz, nens))
    modules.broadcast_initial_gcm_column(coupler)
    modules.perturb_temperature(coupler, np.zeros(nens, dtype=np.int32))


class Output:
    """output.h:8-123 without NetCDF: snapshots of every coupler field, written as one .npz at the end."""

    def __init__(self, coupler, prefix):
        self.c, self.prefix, self.times, self.snaps = coupler, prefix, [], {}

    def write(self, etime):
        dm = self.c.get_data_manager_device_readonly()
        self.times.append(float(etime))
        for n in ["density_dry", "uvel", "vvel", "wvel", "temp"] + self.c.get_tracer_names():
            self.snaps.setdefault(n, []).append(dm.get(n, readonly=True).cpu().numpy().copy())

    def close(self):
        if not self.times:
            return None
        path = "%s_0.npz" % self.prefix
        zint = self.c.get_data_manager_device_readonly().get("vertical_interface_height", readonly=True).cpu().numpy()
        np.savez(path, etime=np.array(self.times), zint=zint, **{k: np.stack(v) for k, v in self.snaps.items()})
        return path


def run(config_path, device="cuda:0", ord=5, quiet=False):
    """driver.cpp main(). Returns a dict (simulated time, wall time, steps, output path, the coupler and the dycore)."""
    with open(config_path) as fh:
        config = yaml.safe_load(fh)
    if not config:
        endrun("ERROR: Invalid YAML input file")
    idealized = bool(config.get("idealized", False))
    verbose = bool(config.get("verbose", idealized))
    apply_sponge = bool(config.get("apply_sponge", not idealized))
    apply_gcm_forcing = bool(config.get("apply_gcm_forcing", not idealized))
    sim_time = float(config["sim_time"])
    crm_nx, crm_ny, nens = int(config["crm_nx"]), int(config["crm_ny"]), int(config["nens"])
    xlen, ylen = float(config.get("xlen", -1)), float(config.get("ylen", -1))
    dt_gcm = float(config.get("dt_gcm", sim_time))
    dt_crm_phys = float(config["dt_crm_phys"])
    out_freq = float(config["out_freq"])
    out_prefix = str(config["out_prefix"])
    if xlen <= 0 or ylen <= 0:
        # driver.cpp:130-133 asks set_domain_sizes for missing lengths, which only the PAM-C dycore answers (pamc_init.h:43-53:
        # the body sits under #ifdef PAMC_DYCORE); with AWFL the reference goes on with xlen = -1. Stop instead.
        endrun("ERROR: xlen and ylen must be given (the AWFL build of the reference has no default domain size)")
    nsteps_gcm = int(math.ceil(sim_time / dt_gcm))                       # driver.cpp:120-123
    dt_gcm = sim_time / nsteps_gcm
    nsteps_crm_phys = int(math.ceil(dt_gcm / dt_crm_phys))
    dt_crm_phys = dt_gcm / nsteps_crm_phys

    coupler = PamCoupler(device)
    coupler.set_option("gcm_physics_dt", dt_gcm)
    coupler.set_option("crm_dt", dt_crm_phys)
    if idealized:
        coupler.set_option("standalone_input_file", config_path)         # driver.cpp:131-134
    zint = vertical_interfaces(config)
    crm_nz = zint.size - 1
    coupler.allocate_coupler_state(crm_nz, crm_ny, crm_nx, nens)
    coupler.set_grid(xlen, ylen, zint)
    # physics/micro/none: registers the water vapour tracer and the constants it owns
    coupler.add_tracer("water_vapor", "Water Vapor Density", True, True)
    dycore = Dycore(ord)
    dycore.init(coupler, verbose)
    if not quiet:
        print("Dycore: %s\nMicro : none\nSGS   : none\n" % dycore.dycore_name())
        print("crm_nx:   %d\ncrm_ny:   %d\ncrm_nz:   %d\nxlen (m): %g\nylen (m): %g" % (crm_nx, crm_ny, crm_nz, xlen, ylen))
        print("Vertical interface heights: " + "  ".join("%g" % z for z in zint) + "\n")
    if not idealized:
        initialize_from_supercell_column(zint, coupler)
    # The reference's standalone driver never calls declare_current_profile_as_hydrostatic, so its AWFL run integrates with
    # whatever the pool block behind "variable_gravity" last held. Here the initial profile is declared
    # hydrostatic once, like the MMF driver does; `b200_declare_hydrostatic: false` in the YAML file reproduces the
    # reference's call sequence exactly (variable_gravity then stays 0).
    if bool(config.get("b200_declare_hydrostatic", True)):
        dycore.declare_current_profile_as_hydrostatic(coupler)

    out = Output(coupler, out_prefix)
    etime_gcm, num_out, steps = 0.0, 0, 0
    if out_freq >= 0.0:
        out.write(etime_gcm)
    torch.cuda.synchronize(coupler.device)
    t0 = time.perf_counter()
    for step_gcm in range(nsteps_gcm):
        if apply_gcm_forcing:
            modules.compute_gcm_forcing_tendencies(coupler)
        for step_crm in range(nsteps_crm_phys):
            if apply_gcm_forcing:
                coupler.run_module("apply_gcm_forcing_tendencies", modules.apply_gcm_forcing_tendencies)
            coupler.run_module("dycore", lambda c: dycore.timeStep(c))
            if apply_sponge:
                coupler.run_module("sponge_layer", modules.sponge_layer)
            steps += 1
            etime_gcm = step_gcm * dt_gcm + (step_crm + 1) * dt_crm_phys
            if out_freq >= 0.0 and (out_freq == 0.0 or etime_gcm / out_freq >= num_out + 1):
                out.write(etime_gcm)
                maxw = float(coupler.get_data_manager_device_readonly().get("wvel", readonly=True).abs().max())
                if not quiet:
                    print("Etime , dtphys, maxw: %g , %g , %10.6g" % (etime_gcm, dt_crm_phys, maxw))
                num_out += 1
    torch.cuda.synchronize(coupler.device)
    runtime = time.perf_counter() - t0
    path = out.close()
    if not quiet:
        print("Simulation Time: %g\nRun Time: %g" % (etime_gcm, runtime))
    return dict(sim_time=etime_gcm, run_time=runtime, steps=steps, output=path, coupler=coupler, dycore=dycore,
                dt_crm_phys=dt_crm_phys, dt_gcm=dt_gcm)


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 1:
        endrun("ERROR: Must pass the input YAML filename as a parameter")      # driver.cpp:88
    r = run(argv[0])
    r["dycore"].finalize(r["coupler"])


if __name__ == "__main__":
    main()
