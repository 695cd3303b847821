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


@pytest.mark.parametrize("order", [3, 5, 7, 9])
def test_static_reconstruct_of_the_shim_matches_the_oracle(tmp_path, order):
    """Dycore::reconstruct (Dycore.h:591-604) is a static, host- and device-callable function of caller-supplied matrices;
    the shim keeps it with the reference's signature. CPU test: stencils through the g++-built driver against the oracle's
    reconstruct with the uniform-grid tables (same algorithm, same association: equal to the last bits)."""
    from oracle import oraclelib
    import helpers as H
    from pam_b200 import synthetic as syn
    exe = _build(tmp_path, ("-DMW_ORD=%d" % order,))
    rng = np.random.default_rng(11)
    n = 2000
    st = rng.standard_normal((n, order)) * np.exp(rng.uniform(-3, 3, (n, 1))) + rng.uniform(-5, 5, (n, 1))
    st[:32] = 1.0
    st[32:64] = np.where(np.arange(order) < order // 2, 0.0, 1.0)
    st[64:96] *= 1e8
    inp, out = os.path.join(str(tmp_path), "st.bin"), os.path.join(str(tmp_path), "rec.bin")
    with open(inp, "wb") as f:
        f.write(struct.pack("i", n))
        f.write(np.ascontiguousarray(st).tobytes())
    r = subprocess.run([exe, "reconstruct", inp, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    got = np.fromfile(out, dtype=np.float64).reshape(2, n)
    case = syn.make_case(12, 1, 12, 2, tracers=syn.TRACERS_NT1, grid="equal")
    o = H.make_oracle(case, order)
    scale = np.abs(st).max(axis=1)
    for ind in (0, 1):
        ref = o.reconstruct(st, ind)
        assert np.max(np.abs(got[ind] - ref) / scale) <= 4e-16, ind


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(nt=1), dict(nt=10), dict(nt=1, ny=5)])
def test_cxx_shim_dynamics_array_methods_match_the_oracle(tmp_path, cfg):
    """The reference's other public methods (Dycore.h:37, :262, :608, :1096, :1281, :1336) through the C++ shim on
    caller-owned HALOED arrays, each dumped and compared with the oracle's restatement of the same method (itself
    bit-identical to the compiled reference, tests/test_oracle_golden.py): conversion in (bit-exact up to pow),
    compute_mass, one compute_tendencies incl. the divided-and-multiplied-back interior and the ghosts it leaves,
    conversion out, halo_exchange on its own, and init_supercell."""
    import helpers as H
    from conftest import rel_l2
    from pam_b200 import synthetic as syn
    from test_init_oracle import oracle_init
    nt, ny = cfg["nt"], cfg.get("ny", 1)
    tr = syn.TRACERS_NT1 if nt == 1 else syn.TRACERS_P3_SHOC
    case = syn.make_case(12, ny, 10 if ny == 1 else 7, 35, tracers=tr, grid="tanh")
    exe = _build(tmp_path)
    inp, out = os.path.join(str(tmp_path), "case.bin"), os.path.join(str(tmp_path), "out.bin")
    _write_case(inp, case, 2.0)
    r = subprocess.run([exe, "dynamics", inp, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    a = np.fromfile(out, dtype=np.float64)
    o = H.make_oracle(case, 5, crm_dt=2.0)
    o.declare_hydrostatic()
    hsh, hs, n4 = o.halo_shape(), o.hs, (case["nz"], ny, case["nx"], case["nens"])
    pos = [0]

    def take(shape):
        n = int(np.prod(shape))
        v = a[pos[0]:pos[0] + n].reshape(shape)
        pos[0] += n
        return v

    names = list(STATE) + [t[0] for t in case["tracers"]]
    near = lambda x, y, tol: np.linalg.norm(x - y) <= tol * max(np.linalg.norm(y), 1e-300)
    # convert_coupler_to_dynamics
    so, to = np.zeros((5,) + hsh), np.zeros((nt,) + hsh)
    o.convert_coupler_to_dynamics(so, to)
    s_in, t_in = take((5,) + hsh), take((nt,) + hsh)
    assert np.array_equal(s_in[:4], so[:4]) and near(s_in[4], so[4], 1e-15) and np.array_equal(t_in, to)
    # compute_mass(coupler, state, tracers)
    assert np.max(np.abs(take((nt + 2, case["nens"])) - o.compute_mass_dynamics(so, to)) / (np.abs(o.compute_mass_dynamics(so, to)) + 1e-300)) < 1e-13
    # compute_tendencies: tendencies, and what it leaves in state / tracers (round-tripped interior + ghosts)
    tto = np.ascontiguousarray(np.maximum(np.stack([case["fields"][n] for n in names[5:]]), 0.0))
    sto = o.compute_tendencies_haloed(so, to, tto, 0.7)
    s_rt, t_rt, s_tend, t_tend = take((5,) + hsh), take((nt,) + hsh), take((5,) + n4), take((nt,) + n4)
    for l in range(5):
        scale = max(np.linalg.norm(sto[l]), 1e-6 * np.linalg.norm(sto[0]) * 350.0)
        assert np.linalg.norm(s_tend[l] - sto[l]) / scale < 1e-11, l
    for t in range(nt):
        scale = max(np.linalg.norm(tto[t]), 1e-12 * np.linalg.norm(to[t]), 1e-300)
        assert np.linalg.norm(t_tend[t] - tto[t]) / scale < 1e-10, t
    I = (slice(None), slice(hs, -hs), slice(hs, -hs), slice(hs, -hs))
    assert near(s_rt[I], so[I], 1e-15) and near(t_rt[I], to[I], 1e-15)
    for l in range(5):
        assert near(s_rt[l], so[l], 1e-13), l          # ghosts: pow() of the wall extrapolation
    assert near(t_rt, to, 1e-13)
    # convert_dynamics_to_coupler of the round-tripped arrays
    o.convert_dynamics_to_coupler(so, to)
    ref = o.get_fields()
    un = float(np.linalg.norm(ref["uvel"]))
    for n in names:
        got = take(n4)
        floor = 1e-3 * un if n in ("uvel", "vvel", "wvel") else 1e-300
        assert rel_l2(got, ref[n], floor) <= 1e-13, n
    # halo_exchange on its own (interior divided by density first, as compute_tendencies does)
    do, dto, po = so.copy(), to.copy(), np.zeros(hsh)
    nz0 = do[0] != 0.0
    for l in range(1, 5):
        do[l][nz0] /= do[0][nz0]
    for t in range(nt):
        dto[t][nz0] /= do[0][nz0]
    o.halo_exchange(do, dto, po)
    s_h, t_h, p_h = take((5,) + hsh), take((nt,) + hsh), take(hsh)
    for l in range(5):
        assert near(s_h[l], do[l], 1e-13), l
    assert near(t_h, dto, 1e-13)
    G = np.ones(hsh, dtype=bool); G[hs:-hs] = False       # pressure: halo_exchange only defines the wall ghosts
    assert near(p_h[G], po[G], 1e-13)
    # init_supercell -> convert_dynamics_to_coupler against the init oracle
    ref = oracle_init(case, "supercell")
    un = float(np.linalg.norm(ref["temp"])) * 1e-3
    for n in names:
        got = take(n4)
        floor = un if n in ("uvel", "vvel", "wvel") else 1e-300
        assert rel_l2(got, ref[n], floor) <= 1e-13, n
    assert pos[0] == a.size
