"""The C++ host shim (pam_b200/cxx/Dycore.h + pam_coupler_b200.h), i.e. the reference's own `class Dycore` interface
re-homed on the C ABI, driven by tests/cxx/shim_driver.cpp the way standalone/mmf_simplified/driver.cpp drives the
reference. CPU: it compiles with plain g++ (no CUDA headers), links against libawfl_b200.so and refuses loudly to run
without a device. GPU: the same case stepped through the C++ shim and through the Python mirror gives bit-identical
fields (same kernels behind the same C ABI), and matches the oracle within the stated tolerance."""
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cxx", "shim_driver.cpp")
STATE = ("density_dry", "uvel", "vvel", "wvel", "temp")


def _build(tmp_path, extra=()):
    from pam_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    exe = os.path.join(str(tmp_path), "shim_driver")
    libdir = os.path.dirname(_lib.LIB_PATH)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-O1", "-Wall", "-Werror", *extra, "-I", os.path.join(ROOT, "pam_b200", "cxx"), SRC, "-o", exe,
           "-L", libdir, "-lawfl_b200", "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_shim_compiles_without_cuda_headers_and_refuses_to_run_without_a_device(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe, "selftest"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "selftest ok" in r.stdout
    for order in (3, 9):      # MW_ORD selects the order at compile time like the reference (Dycore.h:17-21)
        _build(tmp_path, ("-DMW_ORD=%d" % order,))


def _write_case(path, case, crm_dt, default_consts=False):
    c = case["consts"]
    with open(path, "wb") as f:
        f.write(struct.pack("6i", case["nz"], case["ny"], case["nx"], case["nens"], len(case["tracers"]), int(default_consts)))
        f.write(struct.pack("9d", case["xlen"], case["ylen"], crm_dt, c["R_d"], c["cp_d"], c["R_v"], c["cp_v"], c["p0"], c["grav"]))
        f.write(np.ascontiguousarray(case["zint"], dtype=np.float64).tobytes())
        for (name, pos, adds) in case["tracers"]:
            f.write(name.encode().ljust(32, b"\0"))
            f.write(struct.pack("2i", int(pos), int(adds)))
        for n in list(STATE) + [t[0] for t in case["tracers"]]:
            f.write(np.ascontiguousarray(case["fields"][n], dtype=np.float64).tobytes())


def _read_result(path, case):
    nz, nens, nt = case["nz"], case["nens"], len(case["tracers"])
    shape = (nz, case["ny"], case["nx"], nens)
    a = np.fromfile(path, dtype=np.float64)
    o = 0
    dt = a[o]; o += 1
    gv = a[o:o + nz * nens].reshape(nz, nens); o += nz * nens
    mass = a[o:o + (nt + 2) * nens].reshape(nt + 2, nens); o += (nt + 2) * nens
    fields = {}
    n = int(np.prod(shape))
    for name in list(STATE) + [t[0] for t in case["tracers"]]:
        fields[name] = a[o:o + n].reshape(shape); o += n
    assert o == a.size
    return dt, gv, mass, fields


@pytest.mark.gpu
@pytest.mark.parametrize("nt", [1, 10])
def test_cxx_shim_matches_python_mirror_and_oracle(tmp_path, nt):
    import helpers as H
    from conftest import rel_l2, field_floors
    from pam_b200 import synthetic as syn
    tr = syn.TRACERS_NT1 if nt == 1 else syn.TRACERS_P3_SHOC
    case = syn.make_case(14, 1, 12, 35, tracers=tr, grid="tanh")
    exe = _build(tmp_path)
    inp, out = os.path.join(str(tmp_path), "case.bin"), os.path.join(str(tmp_path), "out.bin")
    _write_case(inp, case, 2.0)
    r = subprocess.run([exe, "run", inp, out, "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    dt, gv, mass, got = _read_result(out, case)
    # the Python mirror drives the same C ABI: bit-identical
    c, d = H.make_gpu(case, 5, crm_dt=2.0)
    d.declare_current_profile_as_hydrostatic(c)
    assert d.compute_time_step(c) == dt
    assert np.array_equal(c.get_data_manager_device_readonly().get("variable_gravity").cpu().numpy(), gv)
    d.timeStep(c); d.timeStep(c)
    py = H.gpu_fields(c)
    for k in py:
        assert np.array_equal(py[k], got[k]), k
    assert np.array_equal(d.compute_mass(c), mass)
    # and the oracle within the stated tolerance (two steps: 2e-12)
    o = H.make_oracle(case, 5, crm_dt=2.0)
    o.declare_hydrostatic()
    o.time_step(); o.time_step()
    ref = o.get_fields()
    fl = field_floors(ref)
    for k in ref:
        assert rel_l2(got[k], ref[k], fl[k]) <= 2e-12, k


@pytest.mark.gpu
def test_cxx_modules_shim_matches_python_mirror(tmp_path):
    """MMF-style loop (forcing -> dycore -> sponge, driver.cpp:239-250) through pam_b200/cxx/modules_b200.h and through
    pam_b200/modules.py: same kernels behind the same C ABI, so the fields agree bit for bit."""
    import torch
    import helpers as H
    from pam_b200 import modules
    from pam_b200 import synthetic as syn
    case = syn.make_case(14, 1, 12, 35, tracers=syn.TRACERS_P3_SHOC, grid="tanh")
    exe = _build(tmp_path)
    inp, out = os.path.join(str(tmp_path), "case.bin"), os.path.join(str(tmp_path), "out.bin")
    _write_case(inp, case, 2.0)
    r = subprocess.run([exe, "run", inp, out, "2", "mmf"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    dt, gv, mass, got = _read_result(out, case)
    c, d = H.make_gpu(case, 5, crm_dt=2.0)
    d.declare_current_profile_as_hydrostatic(c)
    c.set_option("gcm_physics_dt", 20.0)
    dm = c.get_data_manager_device_readwrite()
    for gname, cname in zip(modules.GCM_NAMES, modules.CRM_NAMES):
        dm.get(gname).copy_(0.9 * dm.get(cname)[:, 0, 0, :])
    modules.compute_gcm_forcing_tendencies(c)
    for _ in range(2):
        c.run_module("apply_gcm_forcing_tendencies", modules.apply_gcm_forcing_tendencies)
        c.run_module("dycore", lambda cc: d.timeStep(cc))
        c.run_module("sponge_layer", modules.sponge_layer)
    py = H.gpu_fields(c)
    for k in py:
        assert np.array_equal(py[k], got[k]), k
