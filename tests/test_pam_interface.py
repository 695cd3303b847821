"""The GCM-facing host registry (include/pam_interface_b200.h): the extern "C" symbols of the reference's
pam_core/pam_interface/pam_interface_extern_c.cpp:5-274, exercised through ctypes the way pam_fortran_interface.F90 binds
them (values by value for set_option, everything else by reference)."""
import os
import re
import subprocess
import sys
import threading

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pam_interface_b200.h")

# the 34 names pam_fortran_interface.F90 binds (bind(C,name=...)), written out: /root/reference does not travel
FORTRAN_BOUND = ["pam_interface_finalize", "pam_interface_set_option_string", "pam_interface_get_option_stringlen",
                 "pam_interface_get_option_string", "pam_interface_option_exists", "pam_interface_remove_option",
                 "pam_interface_make_readonly", "pam_interface_destroy_array", "pam_interface_array_exists",
                 "pam_interface_register_dimension"] + [
    "pam_interface_%s_%s" % (op, t) for t in ("bool", "int", "float", "double")
    for op in ("mirror_array_readonly", "mirror_array_readwrite", "create_array", "get_array", "set_option", "get_option")]


@pytest.fixture()
def pi():
    from pam_b200 import interface
    interface.pam_finalize()
    yield interface
    interface.pam_finalize()


def test_every_symbol_the_fortran_interface_binds_is_exported():
    from pam_b200 import _lib
    lib = _lib.load()
    assert len(FORTRAN_BOUND) == 34
    text = open(HEADER).read()
    assert "pam_interface_extern_c.cpp:5-274" in text
    for fn in FORTRAN_BOUND:
        assert hasattr(lib, fn), "libawfl_b200.so does not export %s" % fn
    for fn in ("pam_interface_b200_array_bytes", "pam_interface_b200_copy_to_device", "pam_interface_b200_copy_from_device"):
        assert hasattr(lib, fn) and fn in text
    ref = "/root/reference/pam_core/pam_interface/pam_fortran_interface.F90"
    if os.path.exists(ref):                      # in the build container: the list above is what the reference binds
        bound = re.findall(r"bind\(C,name='([a-z_]+)'\)", open(ref).read())
        assert sorted(bound) == sorted(FORTRAN_BOUND)


def test_options_round_trip(pi):
    pi.pam_set_option("enable_physics_tend_stats", True)
    pi.pam_set_option("ncrms", 384)
    pi.pam_set_option("crm_dt", 10.0)
    pi.pam_set_option("cfl", 0.75, kind="float")
    pi.pam_set_option("standalone_input_file", "input_euler3d.yaml")
    assert pi.pam_get_option("enable_physics_tend_stats", "bool") is True
    assert pi.pam_get_option("ncrms", "int") == 384
    assert pi.pam_get_option("crm_dt", "double") == 10.0
    assert pi.pam_get_option("cfl", "float") == np.float32(0.75)
    assert pi.pam_get_option("standalone_input_file", "string") == "input_euler3d.yaml"
    pi.pam_set_option("ncrms", 4)                                   # overwrite (Options::set_option)
    assert pi.pam_get_option("ncrms", "int") == 4
    assert pi.pam_option_exists("crm_dt") and not pi.pam_option_exists("nope")
    pi.pam_remove_option("crm_dt")
    assert not pi.pam_option_exists("crm_dt")
    pi.pam_remove_option("crm_dt")                                  # removing a missing key is not an error
    with pytest.raises(KeyError):
        pi.pam_get_option("crm_dt", "double")


@pytest.mark.parametrize("dtype", [np.bool_, np.int32, np.float32, np.float64])
def test_arrays_create_mirror_get_destroy(pi, dtype):
    rng = np.random.default_rng(3)
    for ndims in range(1, 8):
        shape = tuple(int(s) for s in rng.integers(1, 4, ndims))
        host = (rng.random(shape) * 7).astype(dtype)
        pi.pam_mirror_array_readwrite("mirrored", host, "a GCM array")
        view = pi.pam_get_array("mirrored", dtype, ndims)
        assert view.shape == shape and np.array_equal(view, host)
        assert view.ctypes.data == host.ctypes.data                 # mirrored: the GCM's own memory, not a copy
        view.flat[0] = dtype(1)
        assert host.flat[0] == dtype(1)
        before = host.copy()
        pi.pam_create_array("owned", dtype, shape, "a PAM-owned array")
        own = pi.pam_get_array("owned", dtype, ndims)
        assert own.shape == shape and not own.any()                 # zero-initialised
        own[...] = host
        assert np.array_equal(pi.pam_get_array("owned", dtype, ndims), host)
        pi.pam_destroy_array("owned")
        pi.pam_destroy_array("mirrored")
        assert not pi.pam_array_exists("owned") and not pi.pam_array_exists("mirrored")
        assert np.array_equal(host, before)                         # destroying a mirror leaves the GCM memory alone


def test_readonly_dimension_duplicate_finalize(pi):
    a = np.arange(12, dtype=np.float64).reshape(3, 4)
    pi.pam_mirror_array_readonly("gcm_temp_in", a)
    assert pi.pam_array_exists("gcm_temp_in")
    with pytest.raises(KeyError):
        pi.pam_mirror_array_readonly("gcm_temp_in", a)              # duplicate name
    pi.pam_create_array("state", np.float64, (2, 2))
    pi.pam_make_readonly("state")
    with pytest.raises(KeyError):
        pi.pam_make_readonly("missing")
    pi.pam_register_dimension("z", 58)
    pi.pam_register_dimension("z", 58)                              # same length again is fine
    pi.pam_destroy_array("missing")                                 # unknown name: ignored
    pi.pam_finalize()
    assert not pi.pam_array_exists("gcm_temp_in") and not pi.pam_array_exists("state")


def test_one_registry_per_host_thread(pi):
    """pam_interface.h:16-32: the coupler is looked up by std::this_thread::get_id()."""
    pi.pam_set_option("who", 1)
    seen = {}

    def other():
        seen["exists_before"] = pi.pam_option_exists("who")
        pi.pam_set_option("who", 2)
        seen["value"] = pi.pam_get_option("who", "int")

    t = threading.Thread(target=other)
    t.start()
    t.join()
    assert seen == {"exists_before": False, "value": 2}
    assert pi.pam_get_option("who", "int") == 1


@pytest.mark.parametrize("snippet,msg", [
    ("i.pam_set_option('a', 1); i._fn('pam_interface_get_option_double', [C.c_char_p, C.POINTER(C.c_double)])(b'a', C.byref(C.c_double()))",
     "Requesting option with the wrong type"),
    ("i._fn('pam_interface_get_option_int', [C.c_char_p, C.POINTER(C.c_int)])(b'zz', C.byref(C.c_int()))", "option key does not exist"),
    ("i.pam_create_array('x', 'float64', (2, 3)); i.pam_get_array('x', 'float32', 2)", "Requested type is different"),
    ("i.pam_create_array('x', 'float64', (2, 3)); i.pam_get_array('x', 'float64', 3)", "Requested dimensions is different"),
    ("i.pam_register_dimension('z', 4); i.pam_register_dimension('z', 5)", "dimension already exists"),
])
def test_errors_end_the_run_like_the_reference(snippet, msg):
    """endrun (pam_const.h:249-252): message on stderr, then an exception nobody catches."""
    code = "import ctypes as C\nfrom pam_b200 import interface as i\n" + snippet.replace("; ", "\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and msg in r.stderr, r.stderr[-400:]


@pytest.mark.gpu
def test_registered_array_to_device_and_back(pi):
    import torch
    host = np.random.default_rng(0).random((5, 3, 4, 8))
    pi.pam_mirror_array_readwrite("crm_state_temp", host)
    dev = torch.empty(host.shape, dtype=torch.float64, device="cuda:0")
    pi.pam_copy_to_device("crm_state_temp", dev)
    assert np.array_equal(dev.cpu().numpy(), host)
    dev.mul_(2.0)
    pi.pam_copy_from_device("crm_state_temp", dev)
    assert np.array_equal(host, dev.cpu().numpy())
    with pytest.raises(ValueError):
        pi.pam_copy_to_device("crm_state_temp", dev[:2].contiguous())
