"""TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/_lib/libawfl_oracle_ord{N}.so, the plain-C restatement
of the reference AWFL time step (oracle/awfl_oracle.c). Same method names as oracle/reflib.RefDycore so tests
can swap one for the other.

Only tests/, tools that generate tests/golden/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
import this module. The product path (pam_b200/) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_DIR = os.path.join(_HERE, "_lib")
_dp = C.POINTER(C.c_double)
_dpp = C.POINTER(_dp)
STATE_FIELDS = ("density_dry", "uvel", "vvel", "wvel", "temp")


def _p(a):
    return a.ctypes.data_as(_dp)


def _lib_path(ord):
    return os.path.join(_LIB_DIR, "libawfl_oracle_ord%d.so" % ord)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "all"])


def available(ord=5):
    return os.path.exists(_lib_path(ord))


_libs = {}


def load(ord=5):
    if ord in _libs:
        return _libs[ord]
    if not available(ord):
        build()
    lib = C.CDLL(_lib_path(ord))
    lib.orc_create.restype = C.c_void_p
    lib.orc_create.argtypes = [C.c_int] * 4 + [C.c_double] * 2 + [_dp, C.c_int, C.c_int, C.POINTER(C.c_int),
                                                                 C.POINTER(C.c_int), C.c_int, _dp]
    lib.orc_destroy.argtypes = [C.c_void_p]
    lib.orc_get_const.restype = C.c_double
    lib.orc_get_const.argtypes = [C.c_void_p, C.c_int]
    lib.orc_get_tables.argtypes = [_dp] * 5
    lib.orc_get_vert_tables.argtypes = [C.c_void_p, _dp, _dp]
    lib.orc_get_dz.argtypes = [C.c_void_p, _dp]
    lib.orc_get_grav_var.argtypes = [C.c_void_p, _dp]
    lib.orc_set_grav_var.argtypes = [C.c_void_p, _dp]
    lib.orc_coefs_to_tv.restype = C.c_double
    lib.orc_coefs_to_tv.argtypes = [C.c_int, _dp]
    lib.orc_reconstruct.argtypes = [_dp, C.c_int, C.c_int, _dp]
    lib.orc_reconstruct_vert.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, C.c_int, C.c_int, _dp]
    lib.orc_compute_tendencies.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp]
    lib.orc_compute_time_step.restype = C.c_double
    lib.orc_compute_time_step.argtypes = [C.c_void_p, _dpp, C.c_double]
    lib.orc_time_step.argtypes = [C.c_void_p, _dpp, C.c_double, C.c_int]
    lib.orc_declare_hydrostatic.argtypes = [C.c_void_p, _dpp, _dpp]
    lib.orc_compute_mass.argtypes = [C.c_void_p, _dpp, _dp]
    lib.orc_halo_shape.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    lib.orc_halo_shape.restype = None
    lib.orc_halo_exchange.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.orc_compute_tendencies_haloed.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double]
    lib.orc_convert_coupler_to_dynamics.argtypes = [C.c_void_p, _dpp, _dp, _dp]
    lib.orc_convert_dynamics_to_coupler.argtypes = [C.c_void_p, _dp, _dp, _dpp]
    lib.orc_compute_mass_dynamics.argtypes = [C.c_void_p, _dp, _dp, _dp]
    _libs[ord] = lib
    return lib


_CONST_NAMES = ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav", "cv_d", "gamma_d", "kappa_d", "C0")


class OracleDycore:
    """Plain-C oracle for one configuration; holds the coupler fields as numpy arrays."""

    def __init__(self, nz, ny, nx, nens, xlen, ylen, zint, tracers=(("water_vapor", True, True),),
                 consts=None, crm_dt=1.0, ord=5):
        self.lib = load(ord)
        assert self.lib.orc_ord() == ord
        self.ord, self.hs = ord, (ord + 1) // 2
        self.nz, self.ny, self.nx, self.nens = nz, ny, nx, nens
        self.tracer_names = [t[0] for t in tracers]
        self.nt = len(tracers)
        self.crm_dt = float(crm_dt)
        zint = np.ascontiguousarray(zint, dtype=np.float64)
        per_ens = 1 if zint.ndim == 2 else 0
        pos = (C.c_int * max(self.nt, 1))(*[int(t[1]) for t in tracers])
        adds = (C.c_int * max(self.nt, 1))(*[int(t[2]) for t in tracers])
        idwv = self.tracer_names.index("water_vapor") if "water_vapor" in self.tracer_names else 0
        cptr = None
        if consts is not None:
            self._consts = np.array([consts[k] for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav")], dtype=np.float64)
            cptr = _p(self._consts)
        self.h = self.lib.orc_create(nz, ny, nx, nens, float(xlen), float(ylen), _p(zint), per_ens, self.nt, pos, adds,
                                     idwv, cptr)
        self.idWV = idwv
        self.shape4 = (nz, ny, nx, nens)
        self.fields = {n: np.zeros(self.shape4) for n in list(STATE_FIELDS) + self.tracer_names}
        self.gcm = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _fptrs(self):
        names = list(STATE_FIELDS) + self.tracer_names
        arr = (_dp * len(names))(*[_p(self.fields[n]) for n in names])
        return arr

    def option(self, name):
        return self.lib.orc_get_const(self.h, _CONST_NAMES.index(name))

    def set_crm_dt(self, dt):
        self.crm_dt = float(dt)

    def set_field(self, name, arr):
        if name.startswith("gcm_"):
            self.gcm = self.gcm or {}
            self.gcm[name] = np.ascontiguousarray(arr, dtype=np.float64).reshape(self.nz, self.nens).copy()
            return
        self.fields[name] = np.ascontiguousarray(arr, dtype=np.float64).reshape(self.shape4).copy()

    def get_field(self, name, shape=None):
        if name == "variable_gravity":
            g = np.empty((self.nz, self.nens))
            self.lib.orc_get_grav_var(self.h, _p(g))
            return g
        if name == "vertical_cell_dz":
            g = np.empty((self.nz, self.nens))
            self.lib.orc_get_dz(self.h, _p(g))
            return g
        return self.fields[name].copy()

    def set_fields(self, fields):
        for k, v in fields.items():
            self.set_field(k, v)

    def get_fields(self):
        return {k: v.copy() for k, v in self.fields.items()}

    def set_variable_gravity(self, g):
        g = np.ascontiguousarray(g, dtype=np.float64).reshape(self.nz, self.nens)
        self.lib.orc_set_grav_var(self.h, _p(g))

    def tables(self):
        o, hs = self.ord, self.hs
        s2c = np.empty((o, o)); c2g = np.empty((o, 2)); wrl = np.empty((hs, hs, hs)); idl = np.empty(hs + 1)
        sig = np.empty(1)
        self.lib.orc_get_tables(_p(s2c), _p(c2g), _p(wrl), _p(idl), _p(sig))
        return dict(sten_to_coefs=s2c, coefs_to_gll=c2g, weno_recon_lower=wrl, idl=idl, sigma=float(sig[0]))

    def vert_tables(self):
        o, hs = self.ord, self.hs
        a = np.empty((self.nz + 2, o, o, self.nens)); b = np.empty((self.nz + 2, hs, hs, hs, self.nens))
        self.lib.orc_get_vert_tables(self.h, _p(a), _p(b))
        return dict(vert_sten_to_coefs=a, vert_weno_recon_lower=b)

    def coefs_to_tv(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        return self.lib.orc_coefs_to_tv(a.size, _p(a))

    def reconstruct(self, stencils, ind):
        st = np.ascontiguousarray(stencils, dtype=np.float64).reshape(-1, self.ord)
        out = np.empty(st.shape[0])
        self.lib.orc_reconstruct(_p(st), st.shape[0], ind, _p(out))
        return out

    def reconstruct_vert(self, level, iens, stencils, ind):
        st = np.ascontiguousarray(stencils, dtype=np.float64).reshape(-1, self.ord)
        out = np.empty(st.shape[0])
        self.lib.orc_reconstruct_vert(self.h, level, iens, _p(st), st.shape[0], ind, _p(out))
        return out

    def compute_time_step(self, cfl=0.8):
        return self.lib.orc_compute_time_step(self.h, self._fptrs(), cfl)

    def time_step(self, ncycles=0):
        n = self.lib.orc_time_step(self.h, self._fptrs(), self.crm_dt, ncycles)
        assert n > 0
        return n

    def declare_hydrostatic(self, use_gcm=False):
        g = None
        if use_gcm:
            names = ("gcm_density_dry", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice")
            g = (_dp * 5)(*[_p(self.gcm[n]) for n in names])
        assert self.lib.orc_declare_hydrostatic(self.h, self._fptrs(), g) == 0

    def compute_tendencies(self, state, tracers, fct_start, dt):
        n4 = self.shape4
        state = np.ascontiguousarray(state, dtype=np.float64); assert state.shape == (5,) + n4
        nt1 = max(self.nt, 1)
        t_in = np.zeros((nt1,) + n4); f_in = np.zeros((nt1,) + n4)
        if self.nt:
            t_in[:] = np.asarray(tracers).reshape((self.nt,) + n4)
            f_in[:] = np.asarray(fct_start).reshape((self.nt,) + n4)
        st = np.empty_like(state); tt = np.empty((nt1,) + n4); srt = np.empty_like(state); trt = np.empty((nt1,) + n4)
        rc = self.lib.orc_compute_tendencies(self.h, _p(state), _p(t_in), _p(f_in), float(dt), _p(st), _p(tt), _p(srt), _p(trt))
        assert rc == 0
        return st, tt[:self.nt], srt, trt[:self.nt]

    def compute_mass(self):
        out = np.empty((self.nt + 2, self.nens))
        assert self.lib.orc_compute_mass(self.h, self._fptrs(), _p(out)) == 0
        return out

    # ---- the reference's public methods on its HALOED arrays (Dycore.h:37,262,608,1281,1336); arrays are modified in place ----
    def halo_shape(self):
        hs = self.hs
        return (self.nz + 2 * hs, self.ny + 2 * hs, self.nx + 2 * hs, self.nens)

    def halo_exchange(self, state, tracers, pressure):
        assert self.lib.orc_halo_exchange(self.h, _p(state), _p(tracers), _p(pressure)) == 0

    def compute_tendencies_haloed(self, state, tracers, tracers_tend, dt):
        """state / tracers (haloed) and tracers_tend (FCT start mass in, tendencies out) are modified in place; returns state_tend."""
        st = np.empty((5,) + self.shape4)
        assert self.lib.orc_compute_tendencies_haloed(self.h, _p(state), _p(st), _p(tracers), _p(tracers_tend), float(dt)) == 0
        return st

    def convert_coupler_to_dynamics(self, state, tracers):
        assert self.lib.orc_convert_coupler_to_dynamics(self.h, self._fptrs(), _p(state), _p(tracers)) == 0

    def convert_dynamics_to_coupler(self, state, tracers):
        assert self.lib.orc_convert_dynamics_to_coupler(self.h, _p(state), _p(tracers), self._fptrs()) == 0

    def compute_mass_dynamics(self, state, tracers):
        out = np.empty((self.nt + 2, self.nens))
        assert self.lib.orc_compute_mass_dynamics(self.h, _p(state), _p(tracers), _p(out)) == 0
        return out
