"""TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/_ref/libpam_{ref,dropin}_ord{N}_cuda.so: the coupler-level driver
oracle/ref_build/pam_cuda_harness.cpp over the REAL pam::PamCoupler + YAKL DataManager, built with nvcc -DYAKL_ARCH_CUDA
for sm_100a -- once with the reference's own dynamics/awfl/Dycore.h ("ref": the reference's YAKL kernels on the GPU),
once with this repo's pam_b200/cxx/Dycore.h over libawfl_b200.so ("dropin").

Only tests/, tools/ and bench.py's `gpu_reference` leg may import this module. The product path (pam_b200/) never does.
"""
import atexit
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_dp = C.POINTER(C.c_double)
STATE_FIELDS = ("density_dry", "uvel", "vvel", "wvel", "temp")


def _p(a):
    return a.ctypes.data_as(_dp)


def lib_path(kind, ord=5):
    assert kind in ("ref", "dropin")
    return os.path.join(_REF_DIR, "libpam_%s_ord%d_cuda.so" % (kind, ord))


def available(kind, ord=5):
    return os.path.exists(lib_path(kind, ord))


EXPORTS = ("pc_ord", "pc_is_dropin", "pc_shutdown", "pc_create", "pc_destroy", "pc_dycore_name", "pc_set_crm_dt", "pc_get_option",
           "pc_entry_exists", "pc_field_size", "pc_set_field", "pc_get_field", "pc_get_flags", "pc_compute_time_step",
           "pc_declare_hydrostatic", "pc_time_step", "pc_time_steps_ms", "pc_time_step_public", "pc_compute_mass")
_libs = {}


def load(kind, ord=5, pool_mb=None):
    """pool_mb: size of YAKL's device pool, honoured by the FIRST library that initialises YAKL in this process."""
    key = (kind, ord)
    if key in _libs:
        return _libs[key]
    if pool_mb:
        os.environ["PAM_REF_POOL_MB"] = str(int(pool_mb))
    lib = C.CDLL(lib_path(kind, ord))
    lib.pc_create.restype = C.c_void_p
    lib.pc_create.argtypes = [C.c_int] * 4 + [C.c_double] * 2 + [_dp, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                                                C.POINTER(C.c_int), _dp, C.c_double]
    lib.pc_destroy.argtypes = [C.c_void_p]
    lib.pc_destroy.restype = None
    lib.pc_dycore_name.argtypes = [C.c_void_p]
    lib.pc_dycore_name.restype = C.c_char_p
    lib.pc_set_crm_dt.argtypes = [C.c_void_p, C.c_double]
    lib.pc_set_crm_dt.restype = None
    lib.pc_get_option.argtypes = [C.c_void_p, C.c_char_p]
    lib.pc_get_option.restype = C.c_double
    lib.pc_entry_exists.argtypes = [C.c_void_p, C.c_char_p]
    lib.pc_field_size.argtypes = [C.c_void_p, C.c_char_p]
    lib.pc_field_size.restype = C.c_long
    lib.pc_set_field.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_long]
    lib.pc_get_field.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_long]
    lib.pc_get_flags.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int), C.c_int]
    lib.pc_compute_time_step.argtypes = [C.c_void_p, C.c_double]
    lib.pc_compute_time_step.restype = C.c_double
    lib.pc_declare_hydrostatic.argtypes = [C.c_void_p, C.c_int]
    lib.pc_time_step.argtypes = [C.c_void_p]
    lib.pc_time_steps_ms.argtypes = [C.c_void_p, C.c_int]
    lib.pc_time_steps_ms.restype = C.c_double
    lib.pc_time_step_public.argtypes = [C.c_void_p, C.c_int]
    lib.pc_compute_mass.argtypes = [C.c_void_p, _dp]
    lib.pc_shutdown.restype = None
    atexit.register(lib.pc_shutdown)
    _libs[key] = lib
    return lib


class PamCudaDycore:
    """pam::PamCoupler + Dycore on the GPU for one configuration ('ref': the reference's kernels, 'dropin': this repo's)."""

    def __init__(self, kind, nz, ny, nx, nens, xlen, ylen, zint, tracers=(("water_vapor", True, True),), consts=None, crm_dt=1.0,
                 ord=5, pool_mb=None):
        self.lib = load(kind, ord, pool_mb)
        assert self.lib.pc_ord() == ord and self.lib.pc_is_dropin() == (1 if kind == "dropin" else 0)
        self.kind, self.ord = kind, ord
        self.nz, self.ny, self.nx, self.nens = nz, ny, nx, nens
        self.tracer_names = [t[0] for t in tracers]
        self.nt = len(tracers)
        zint = np.ascontiguousarray(zint, dtype=np.float64)
        names = (C.c_char_p * max(self.nt, 1))(*[t[0].encode() for t in tracers])
        pos = (C.c_int * max(self.nt, 1))(*[int(t[1]) for t in tracers])
        adds = (C.c_int * max(self.nt, 1))(*[int(t[2]) for t in tracers])
        cptr = None
        if consts is not None:
            self._consts = np.array([consts[k] for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav")], dtype=np.float64)
            cptr = _p(self._consts)
        self.h = self.lib.pc_create(nz, ny, nx, nens, float(xlen), float(ylen), _p(zint), 1 if zint.ndim == 2 else 0, self.nt, names,
                                    pos, adds, cptr, float(crm_dt))
        if not self.h:
            raise RuntimeError("pc_create failed")
        self.shape4 = (nz, ny, nx, nens)

    def close(self):
        if getattr(self, "h", None):
            self.lib.pc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def dycore_name(self):
        return self.lib.pc_dycore_name(self.h).decode()

    def option(self, name):
        return self.lib.pc_get_option(self.h, name.encode())

    def entry_exists(self, name):
        return bool(self.lib.pc_entry_exists(self.h, name.encode()))

    def set_field(self, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        if self.lib.pc_set_field(self.h, name.encode(), _p(a), a.size):
            raise ValueError("size mismatch or copy failure for %s" % name)

    def get_field(self, name, shape=None):
        n = self.lib.pc_field_size(self.h, name.encode())
        out = np.empty(n, dtype=np.float64)
        assert self.lib.pc_get_field(self.h, name.encode(), _p(out), n) == 0
        if shape is None and n == int(np.prod(self.shape4)):
            shape = self.shape4
        return out.reshape(shape) if shape is not None else out

    def get_flags(self, name):
        out = (C.c_int * max(self.nt, 1))()
        assert self.lib.pc_get_flags(self.h, name.encode(), out, self.nt) == 0
        return [bool(x) for x in out[:self.nt]]

    def set_fields(self, fields):
        for k, v in fields.items():
            self.set_field(k, v)

    def get_fields(self):
        return {k: self.get_field(k) for k in list(STATE_FIELDS) + self.tracer_names}

    def compute_time_step(self, cfl=0.8):
        return self.lib.pc_compute_time_step(self.h, cfl)

    def declare_hydrostatic(self, use_gcm=False):
        assert self.lib.pc_declare_hydrostatic(self.h, int(use_gcm)) == 0

    def time_step(self):
        """coupler.run_module("dycore", [&](auto &c){ dycore.timeStep(c); }) -- standalone/mmf_simplified/driver.cpp:248"""
        assert self.lib.pc_time_step(self.h) == 0

    def time_steps_ms(self, nsteps):
        return self.lib.pc_time_steps_ms(self.h, int(nsteps))

    def time_step_public(self, ncycles=0):
        """SSPRK3 restated around the class's public methods on harness-owned haloed YAKL arrays (F2 neutralised)."""
        n = self.lib.pc_time_step_public(self.h, int(ncycles))
        assert n > 0
        return n

    def compute_mass(self):
        out = np.empty((self.nt + 2, self.nens))
        assert self.lib.pc_compute_mass(self.h, _p(out)) == 0
        return out
