"""CPU tests of the drop-in boundary: libawfl_b200.so loads, exports every symbol include/awfl_b200.h declares, fails
loudly without a GPU (no CPU fallback), and the product package never touches the oracle."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "awfl_b200.h")


MODULES_HEADER = os.path.join(ROOT, "include", "pam_modules_b200.h")


def _declared_functions(header=HEADER, prefix="awfl_"):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(%s[a-z0-9_]+)\s*\(" % prefix, src)
    return sorted(set(names))


def _lib():
    from pam_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib, _lib.load()


def test_header_declares_the_reference_entry_points():
    fns = _declared_functions()
    # one C entry point per public method of the reference's class Dycore that sits on the path
    for need in ("awfl_create", "awfl_destroy", "awfl_time_step", "awfl_compute_time_step", "awfl_compute_tendencies",
                 "awfl_declare_hydrostatic", "awfl_compute_mass", "awfl_dycore_name", "awfl_last_error"):
        assert need in fns
    text = open(HEADER).read()
    assert "Dycore.h:107-255" in text and "Dycore.h:262-586" in text     # reference citations
    assert "torch" not in text.lower().replace("no torch", "")           # plain pointers and sizes only


def test_library_exports_every_declared_symbol():
    mod, lib = _lib()
    for fn in _declared_functions():
        assert hasattr(lib, fn), "libawfl_b200.so does not export %s" % fn
    for fn in mod.EXPORTS:
        assert fn in _declared_functions(), "%s bound in _lib.py but not declared in the header" % fn
    assert lib.awfl_dycore_name().decode() == "SSPRK3+WENO+FV A-grid"    # Dycore.h:1544


def test_library_exports_the_coupler_module_entry_points():
    """include/pam_modules_b200.h: sponge layer, GCM forcing, temperature perturbation."""
    mod, lib = _lib()
    fns = _declared_functions(MODULES_HEADER, "pamm_")
    for need in ("pamm_create", "pamm_destroy", "pamm_last_error", "pamm_sponge_layer", "pamm_compute_gcm_forcing_tendencies",
                 "pamm_apply_gcm_forcing_tendencies", "pamm_perturb_temperature"):
        assert need in fns
    for fn in fns:
        assert hasattr(lib, fn), "libawfl_b200.so does not export %s" % fn
    text = open(MODULES_HEADER).read()
    assert "sponge_layer.h:9-96" in text and "gcm_forcing.h:18-205" in text and "perturb_temperature.h:10-63" in text
    import torch
    if not torch.cuda.is_available():
        h = C.c_void_p()
        lib.pamm_create.argtypes = [C.c_int] * 5 + [C.c_void_p, C.POINTER(C.c_void_p)]
        lib.pamm_last_error.restype = C.c_char_p
        lib.pamm_last_error.argtypes = [C.c_void_p]
        assert lib.pamm_create(4, 1, 4, 2, 0, None, C.byref(h)) != 0        # no device: fails, does not fall back
        assert b"no CPU fallback" in lib.pamm_last_error(None)


def test_library_contains_sm100a_kernels():
    mod, _ = _lib()
    out = subprocess.run(["cuobjdump", "-lelf", mod.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


def test_create_rejects_bad_arguments_with_a_message():
    mod, lib = _lib()
    cfg = mod.AwflConfig()
    h = C.c_void_p()
    assert lib.awfl_create(C.byref(cfg), C.byref(h)) != 0
    assert b"bad dimensions" in lib.awfl_last_error(None)
    cfg.nz, cfg.ny, cfg.nx, cfg.nens, cfg.num_tracers, cfg.ord = 8, 1, 8, 2, 1, 4
    assert lib.awfl_create(C.byref(cfg), C.byref(h)) != 0
    assert b"ord must be" in lib.awfl_last_error(None)


def test_no_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    mod, lib = _lib()
    cfg = mod.AwflConfig()
    cfg.nz, cfg.ny, cfg.nx, cfg.nens, cfg.num_tracers, cfg.ord = 8, 1, 8, 2, 1, 5
    cfg.xlen, cfg.ylen = 16000.0, 2000.0
    cfg.R_d, cfg.cp_d, cfg.R_v, cfg.cp_v, cfg.p0, cfg.grav = 287.0, 1003.0, 461.0, 1859.0, 1e5, 9.81
    pos = (C.c_ubyte * 1)(1); adds = (C.c_ubyte * 1)(1)
    dz = np.full((8, 2), 100.0)
    cfg.tracer_positive = C.cast(pos, C.POINTER(C.c_ubyte)); cfg.tracer_adds_mass = C.cast(adds, C.POINTER(C.c_ubyte))
    cfg.dz_host = dz.ctypes.data_as(mod.c_double_p)
    h = C.c_void_p()
    assert lib.awfl_create(C.byref(cfg), C.byref(h)) != 0          # must fail, not fall back
    msg = lib.awfl_last_error(None).decode()
    assert "no CPU fallback" in msg or "cuda" in msg.lower()
    # and the Python mirror refuses a CPU coupler outright
    from pam_b200 import PamCoupler, Dycore
    c = PamCoupler("cpu")
    c.set_option("crm_dt", 1.0)
    c.allocate_coupler_state(8, 1, 8, 2)
    c.set_grid(16000.0, 2000.0, np.arange(9) * 100.0)
    c.add_tracer("water_vapor", "", True, True)
    with pytest.raises(RuntimeError):
        Dycore(5).init(c)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "pam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), f
                assert "oraclelib" not in txt and "reflib" not in txt and "awfl_oracle" not in txt, f
    # importing the package must not pull the oracle in either
    code = "import sys; sys.path.insert(0, %r); import pam_b200; assert not [m for m in sys.modules if m.startswith('oracle')]" % ROOT
    assert subprocess.run([sys.executable, "-c", code]).returncode == 0
