"""TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/_ref/libpam_ref_ord{N}[_omp].so, the UNMODIFIED
reference AWFL dycore (E3SM-Project/PAM @ 58b3e77) behind oracle/ref_build/ref_harness.cpp.

Only tests/, tools that generate tests/golden/, __graft_entry__.smoke() and bench.py's CPU-baseline legs
may import this module. The product path (pam_b200/) never does.
"""
import atexit
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_dp = C.POINTER(C.c_double)


def _p(a):
    return a.ctypes.data_as(_dp)


def available(ord=5, omp=False):
    return os.path.exists(_lib_path(ord, omp))


def _lib_path(ord, omp):
    return os.path.join(_REF_DIR, "libpam_ref_ord%d%s.so" % (ord, "_omp" if omp else ""))


_libs = {}


def load(ord=5, omp=False):
    key = (ord, omp)
    if key in _libs:
        return _libs[key]
    lib = C.CDLL(_lib_path(ord, omp))
    lib.ref_create.restype = C.c_void_p
    lib.ref_create.argtypes = [C.c_int] * 4 + [C.c_double] * 2 + [_dp, C.c_int, C.c_int,
                                                                 C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                                                 C.POINTER(C.c_int), _dp, C.c_double]
    lib.ref_destroy.argtypes = [C.c_void_p]
    lib.ref_set_crm_dt.argtypes = [C.c_void_p, C.c_double]
    lib.ref_get_option.restype = C.c_double
    lib.ref_get_option.argtypes = [C.c_void_p, C.c_char_p]
    lib.ref_get_idWV.argtypes = [C.c_void_p]
    lib.ref_field_size.restype = C.c_long
    lib.ref_field_size.argtypes = [C.c_void_p, C.c_char_p]
    lib.ref_set_field.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_long]
    lib.ref_get_field.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_long]
    lib.ref_get_tables.argtypes = [_dp] * 5
    lib.ref_coefs_to_tv.restype = C.c_double
    lib.ref_coefs_to_tv.argtypes = [C.c_int, _dp]
    lib.ref_reconstruct.argtypes = [_dp, C.c_int, C.c_int, _dp]
    lib.ref_reconstruct_vert.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, C.c_int, C.c_int, _dp]
    lib.ref_compute_time_step.restype = C.c_double
    lib.ref_compute_time_step.argtypes = [C.c_void_p, C.c_double]
    lib.ref_timestep_raw.argtypes = [C.c_void_p]
    lib.ref_declare_hydrostatic_raw.argtypes = [C.c_void_p, C.c_int]
    lib.ref_compute_tendencies.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp]
    lib.ref_timestep_intended.argtypes = [C.c_void_p, C.c_int]
    lib.ref_declare_hydrostatic_intended.argtypes = [C.c_void_p, C.c_int]
    lib.ref_compute_mass.argtypes = [C.c_void_p, _dp]
    lib.ref_time_steps_raw.restype = C.c_double
    lib.ref_time_steps_raw.argtypes = [C.c_void_p, C.c_int]
    lib.ref_set_option_real.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
    lib.ref_set_option_real.restype = None
    lib.ref_set_option_int.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    lib.ref_set_option_int.restype = None
    lib.ref_entry_exists.argtypes = [C.c_void_p, C.c_char_p]
    lib.ref_sponge_layer.argtypes = [C.c_void_p]
    lib.ref_compute_gcm_forcing_tendencies.argtypes = [C.c_void_p]
    lib.ref_apply_gcm_forcing_tendencies.argtypes = [C.c_void_p]
    lib.ref_perturb_temperature.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_double]
    lib.ref_gll9.argtypes = [_dp, _dp]
    lib.ref_gll9.restype = None
    lib.ref_init_supercell.argtypes = [C.c_void_p]
    lib.ref_init_thermal.argtypes = [C.c_void_p]
    lib.ref_halo_exchange.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.ref_compute_tendencies_haloed.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double]
    lib.ref_convert_coupler_to_dynamics.argtypes = [C.c_void_p, _dp, _dp]
    lib.ref_convert_dynamics_to_coupler.argtypes = [C.c_void_p, _dp, _dp]
    lib.ref_compute_mass_dynamics.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.ref_init_supercell_dynamics.argtypes = [C.c_void_p, _dp, _dp]
    lib.ref_shutdown.restype = None
    atexit.register(lib.ref_shutdown)
    _libs[key] = lib
    return lib


STATE_FIELDS = ("density_dry", "uvel", "vvel", "wvel", "temp")


class RefDycore:
    """The reference coupler + Dycore pair for one configuration."""

    def __init__(self, nz, ny, nx, nens, xlen, ylen, zint, tracers=(("water_vapor", True, True),),
                 consts=None, crm_dt=1.0, ord=5, omp=False):
        self.lib = load(ord, omp)
        assert self.lib.ref_ord() == ord
        self.ord, self.hs = ord, (ord + 1) // 2
        self.nz, self.ny, self.nx, self.nens = nz, ny, nx, nens
        self.tracer_names = [t[0] for t in tracers]
        self.nt = len(tracers)
        zint = np.ascontiguousarray(zint, dtype=np.float64)
        per_ens = 1 if zint.ndim == 2 else 0
        names = (C.c_char_p * max(self.nt, 1))(*[t[0].encode() for t in tracers])
        pos = (C.c_int * max(self.nt, 1))(*[int(t[1]) for t in tracers])
        adds = (C.c_int * max(self.nt, 1))(*[int(t[2]) for t in tracers])
        cptr = None
        if consts is not None:
            self._consts = np.array([consts[k] for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav")], dtype=np.float64)
            cptr = _p(self._consts)
        self.h = self.lib.ref_create(nz, ny, nx, nens, float(xlen), float(ylen), _p(zint), per_ens, self.nt,
                                     names, pos, adds, cptr, float(crm_dt))
        if not self.h:
            raise RuntimeError("ref_create failed")
        self.shape4 = (nz, ny, nx, nens)

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- options / fields ----
    def option(self, name):
        return self.lib.ref_get_option(self.h, name.encode())

    @property
    def idWV(self):
        return self.lib.ref_get_idWV(self.h)

    def set_crm_dt(self, dt):
        self.lib.ref_set_crm_dt(self.h, float(dt))

    def set_field(self, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        rc = self.lib.ref_set_field(self.h, name.encode(), _p(a), a.size)
        if rc:
            raise ValueError("size mismatch for %s" % name)

    def get_field(self, name, shape=None):
        n = self.lib.ref_field_size(self.h, name.encode())
        out = np.empty(n, dtype=np.float64)
        rc = self.lib.ref_get_field(self.h, name.encode(), _p(out), n)
        assert rc == 0
        if shape is None and n == int(np.prod(self.shape4)):
            shape = self.shape4
        return out.reshape(shape) if shape is not None else out

    def set_fields(self, fields):
        for k, v in fields.items():
            self.set_field(k, v)

    def get_fields(self):
        return {k: self.get_field(k) for k in list(STATE_FIELDS) + self.tracer_names}

    # ---- coupler modules either side of the dycore: the reference's own functions ----
    def set_option(self, name, value):
        if isinstance(value, int):
            self.lib.ref_set_option_int(self.h, name.encode(), value)
        else:
            self.lib.ref_set_option_real(self.h, name.encode(), float(value))

    def entry_exists(self, name):
        return bool(self.lib.ref_entry_exists(self.h, name.encode()))

    def sponge_layer(self):
        assert self.lib.ref_sponge_layer(self.h) == 0

    def compute_gcm_forcing_tendencies(self):
        assert self.lib.ref_compute_gcm_forcing_tendencies(self.h) == 0

    def apply_gcm_forcing_tendencies(self):
        assert self.lib.ref_apply_gcm_forcing_tendencies(self.h) == 0

    def perturb_temperature(self, ids, magnitude=0.1):
        a = np.ascontiguousarray(ids, dtype=np.int32)
        assert a.size == self.nens
        assert self.lib.ref_perturb_temperature(self.h, a.ctypes.data_as(C.POINTER(C.c_int)), float(magnitude)) == 0

    # ---- idealized initial states (Dycore.h:1017-1086 thermal, :1096-1276 supercell) ----
    def gll9(self):
        p, w = np.empty(9), np.empty(9)
        self.lib.ref_gll9(_p(p), _p(w))
        return p, w

    def init_supercell(self):
        assert self.lib.ref_init_supercell(self.h) == 0

    def init_thermal(self):
        assert self.lib.ref_init_thermal(self.h) == 0

    # ---- tables ----
    def tables(self):
        o, hs = self.ord, self.hs
        s2c = np.empty((o, o)); c2g = np.empty((o, 2)); wrl = np.empty((hs, hs, hs)); idl = np.empty(hs + 1)
        sig = np.empty(1)
        self.lib.ref_get_tables(_p(s2c), _p(c2g), _p(wrl), _p(idl), _p(sig))
        return dict(sten_to_coefs=s2c, coefs_to_gll=c2g, weno_recon_lower=wrl, idl=idl, sigma=float(sig[0]))

    def coefs_to_tv(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        return self.lib.ref_coefs_to_tv(a.size, _p(a))

    def reconstruct(self, stencils, ind):
        st = np.ascontiguousarray(stencils, dtype=np.float64).reshape(-1, self.ord)
        out = np.empty(st.shape[0])
        self.lib.ref_reconstruct(_p(st), st.shape[0], ind, _p(out))
        return out

    def reconstruct_vert(self, level, iens, stencils, ind):
        st = np.ascontiguousarray(stencils, dtype=np.float64).reshape(-1, self.ord)
        out = np.empty(st.shape[0])
        self.lib.ref_reconstruct_vert(self.h, level, iens, _p(st), st.shape[0], ind, _p(out))
        return out

    def vert_tables(self):
        o, hs = self.ord, self.hs
        return dict(vert_sten_to_coefs=self.get_field("vert_sten_to_coefs", (self.nz + 2, o, o, self.nens)),
                    vert_weno_recon_lower=self.get_field("vert_weno_recon_lower", (self.nz + 2, hs, hs, hs, self.nens)))

    # ---- dycore entry points ----
    def compute_time_step(self, cfl=0.8):
        return self.lib.ref_compute_time_step(self.h, cfl)

    def time_step_raw(self):
        assert self.lib.ref_timestep_raw(self.h) == 0

    def time_step(self, ncycles=0):
        n = self.lib.ref_timestep_intended(self.h, ncycles)
        assert n > 0
        return n

    def declare_hydrostatic(self, use_gcm=False):
        assert self.lib.ref_declare_hydrostatic_intended(self.h, int(use_gcm)) == 0

    def declare_hydrostatic_raw(self, use_gcm=False):
        assert self.lib.ref_declare_hydrostatic_raw(self.h, int(use_gcm)) == 0

    def compute_tendencies(self, state, tracers, fct_start, dt):
        n4 = self.shape4
        state = np.ascontiguousarray(state, dtype=np.float64); assert state.shape == (5,) + n4
        tracers = np.ascontiguousarray(tracers, dtype=np.float64).reshape((self.nt,) + n4)
        fct_start = np.ascontiguousarray(fct_start, dtype=np.float64).reshape((self.nt,) + n4)
        st = np.empty_like(state); tt = np.empty((max(self.nt, 1),) + n4)
        srt = np.empty_like(state); trt = np.empty((max(self.nt, 1),) + n4)
        t_in = tracers if self.nt else np.zeros((1,) + n4)
        f_in = fct_start if self.nt else np.zeros((1,) + n4)
        rc = self.lib.ref_compute_tendencies(self.h, _p(state), _p(t_in), _p(f_in), float(dt),
                                             _p(st), _p(tt), _p(srt), _p(trt))
        assert rc == 0
        return st, tt[:self.nt], srt, trt[:self.nt]

    def compute_mass(self):
        out = np.empty((self.nt + 2, self.nens))
        assert self.lib.ref_compute_mass(self.h, _p(out)) == 0
        return out

    def time_steps_raw(self, nsteps):
        return self.lib.ref_time_steps_raw(self.h, nsteps)

    # ---- the reference's own public methods on caller-owned HALOED arrays (modified in place) ----
    def halo_shape(self):
        hs = self.hs
        return (self.nz + 2 * hs, self.ny + 2 * hs, self.nx + 2 * hs, self.nens)

    def halo_exchange(self, state, tracers, pressure):
        assert self.lib.ref_halo_exchange(self.h, _p(state), _p(tracers), _p(pressure)) == 0

    def compute_tendencies_haloed(self, state, tracers, tracers_tend, dt):
        st = np.empty((5,) + self.shape4)
        assert self.lib.ref_compute_tendencies_haloed(self.h, _p(state), _p(st), _p(tracers), _p(tracers_tend), float(dt)) == 0
        return st

    def convert_coupler_to_dynamics(self, state, tracers):
        assert self.lib.ref_convert_coupler_to_dynamics(self.h, _p(state), _p(tracers)) == 0

    def convert_dynamics_to_coupler(self, state, tracers):
        assert self.lib.ref_convert_dynamics_to_coupler(self.h, _p(state), _p(tracers)) == 0

    def compute_mass_dynamics(self, state, tracers):
        out = np.empty((self.nt + 2, self.nens))
        assert self.lib.ref_compute_mass_dynamics(self.h, _p(state), _p(tracers), _p(out)) == 0
        return out

    def init_supercell_dynamics(self):
        st = np.empty((5,) + self.halo_shape()); tr = np.empty((max(self.nt, 1),) + self.halo_shape())
        assert self.lib.ref_init_supercell_dynamics(self.h, _p(st), _p(tr)) == 0
        return st, tr[:self.nt]
