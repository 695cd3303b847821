"""TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/_lib/libpam_modules_oracle.so, the plain-C restatement
(oracle/modules_oracle.c) of the coupler modules either side of the dycore: sponge_layer, compute/apply GCM forcing
(+ fill_holes) and perturb_temperature. Works on a dict {name: numpy array} standing for the coupler's DataManager.

Only tests/, tools that generate tests/golden/ and __graft_entry__.smoke() may import this module.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_lib", "libpam_modules_oracle.so")
_dp = C.POINTER(C.c_double)

CRM_NAMES = ("density_dry", "uvel", "vvel", "temp", "water_vapor", "cloud_water", "ice", "cloud_water_num", "ice_num", "rain_num")
GCM_NAMES = ("gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice",
             "gcm_num_liq", "gcm_num_ice", "gcm_num_rain")
TEND_NAMES = ("rho_d", "uvel", "vvel", "temp", "qtot", "qv", "ql", "qi", "rho_v", "rho_l", "rho_i", "nc", "ni", "nr")


class _Crm(C.Structure):
    _fields_ = [(n, _dp) for n in CRM_NAMES]


class _Gcm(C.Structure):
    _fields_ = [(n, _dp) for n in GCM_NAMES]


class _Tend(C.Structure):
    _fields_ = [(n, _dp) for n in TEND_NAMES]


_lib = None


def load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_LIB)
        _lib.pamo_sponge_layer.argtypes = [C.c_int] * 4 + [C.POINTER(_dp), C.c_int, _dp, _dp, C.c_double, C.c_double, C.c_int]
        _lib.pamo_sponge_layer.restype = None
        _lib.pamo_compute_gcm_forcing_tendencies.argtypes = [C.c_int] * 4 + [C.POINTER(_Crm), C.POINTER(_Gcm), C.POINTER(_Tend), C.c_double]
        _lib.pamo_compute_gcm_forcing_tendencies.restype = None
        _lib.pamo_apply_gcm_forcing_tendencies.argtypes = [C.c_int] * 4 + [C.POINTER(_Crm), C.POINTER(_Gcm), C.POINTER(_Tend), _dp,
                                                                           C.c_double, C.c_double]
        _lib.pamo_apply_gcm_forcing_tendencies.restype = None
        _lib.pamo_perturb_temperature.argtypes = [C.c_int] * 4 + [_dp, C.POINTER(C.c_int), C.c_double]
        _lib.pamo_perturb_temperature.restype = None
        _lib.pamo_gll9.argtypes = [_dp, _dp]
        _lib.pamo_gll9.restype = None
        _lib.pamo_init_supercell.argtypes = [C.c_int] * 4 + [C.c_double] * 2 + [_dp] + [C.c_double] * 5 + [C.c_int, C.c_int,
                                                                                                         C.POINTER(C.c_int), C.POINTER(_dp)]
        _lib.pamo_init_supercell.restype = None
        _lib.pamo_init_thermal.argtypes = [C.c_int] * 4 + [C.c_double] * 2 + [_dp] + [C.c_double] * 7 + [C.c_int, C.c_int,
                                                                                                       C.POINTER(C.c_int), C.POINTER(_dp)]
        _lib.pamo_init_thermal.restype = None
    return _lib


def available():
    return os.path.exists(_LIB)


def _p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def _dims(dm):
    return dm["density_dry"].shape


def sponge_layer(dm, tracer_names, crm_dt, time_scale=60.0, num_layers=5):
    """modules::sponge_layer on dm (updated in place)."""
    nz, ny, nx, nens = _dims(dm)
    names = ["density_dry", "uvel", "vvel", "wvel", "temp"] + list(tracer_names)
    ptrs = (_dp * len(names))(*[_p(dm[n]) for n in names])
    load().pamo_sponge_layer(nz, ny, nx, nens, ptrs, len(names), _p(dm["vertical_interface_height"]), _p(dm["vertical_midpoint_height"]),
                             float(crm_dt), float(time_scale), int(num_layers))


def _bundles(dm):
    nz, ny, nx, nens = _dims(dm)
    for n in TEND_NAMES:
        dm.setdefault("gcm_forcing_tend_" + n, np.zeros((nz, nens)))
    crm = _Crm(*[_p(dm[n]) for n in CRM_NAMES])
    gcm = _Gcm(*[_p(dm[n]) for n in GCM_NAMES])
    tend = _Tend(*[_p(dm["gcm_forcing_tend_" + n]) for n in TEND_NAMES])
    return crm, gcm, tend


def compute_gcm_forcing_tendencies(dm, gcm_physics_dt):
    nz, ny, nx, nens = _dims(dm)
    crm, gcm, tend = _bundles(dm)
    load().pamo_compute_gcm_forcing_tendencies(nz, ny, nx, nens, C.byref(crm), C.byref(gcm), C.byref(tend), float(gcm_physics_dt))


def apply_gcm_forcing_tendencies(dm, crm_dt, gcm_physics_dt):
    nz, ny, nx, nens = _dims(dm)
    crm, gcm, tend = _bundles(dm)
    load().pamo_apply_gcm_forcing_tendencies(nz, ny, nx, nens, C.byref(crm), C.byref(gcm), C.byref(tend), _p(dm["vertical_cell_dz"]),
                                             float(crm_dt), float(gcm_physics_dt))


def perturb_temperature(dm, ids, magnitude=0.1):
    nz, ny, nx, nens = _dims(dm)
    a = np.ascontiguousarray(ids, dtype=np.int32)
    load().pamo_perturb_temperature(nz, ny, nx, nens, _p(dm["temp"]), a.ctypes.data_as(C.POINTER(C.c_int)), float(magnitude))


def gll9():
    p, w = np.empty(9), np.empty(9)
    load().pamo_gll9(_p(p), _p(w))
    return p, w


def derived_constants(consts):
    """gamma_d, C0 the way Dycore::init derives them (Dycore.h:877-891)"""
    R_d, cp_d, p0 = consts["R_d"], consts["cp_d"], consts["p0"]
    cv_d = cp_d - R_d
    gamma = cp_d / cv_d
    kappa = R_d / cp_d
    C0 = pow(R_d * pow(p0, -kappa), gamma)
    return gamma, C0


def init_idealized(kind, nz, ny, nx, nens, xlen, ylen, zint, tracers, consts, gamma, C0):
    """Dycore::init's idealized data (kind = "supercell" | "thermal") -> {field name: array}; tracers as (name, positive, adds_mass)."""
    z = np.asarray(zint, dtype=np.float64)
    if z.ndim == 1:
        z = np.repeat(z[:, None], nens, axis=1)
    z = np.ascontiguousarray(z)
    names = ["density_dry", "uvel", "vvel", "wvel", "temp"] + [t[0] for t in tracers]
    out = {n: np.zeros((nz, ny, nx, nens)) for n in names}
    ptrs = (_dp * len(names))(*[_p(out[n]) for n in names])
    adds = (C.c_int * max(len(tracers), 1))(*[int(t[2]) for t in tracers])
    idWV = [t[0] for t in tracers].index("water_vapor")
    lib = load()
    if kind == "supercell":
        lib.pamo_init_supercell(nz, ny, nx, nens, float(xlen), float(ylen), _p(z), consts["R_d"], consts["R_v"], consts["grav"], gamma, C0,
                                idWV, len(tracers), adds, ptrs)
    elif kind == "thermal":
        lib.pamo_init_thermal(nz, ny, nx, nens, float(xlen), float(ylen), _p(z), consts["R_d"], consts["R_v"], consts["cp_d"], consts["p0"],
                              consts["grav"], gamma, C0, idWV, len(tracers), adds, ptrs)
    else:
        raise ValueError(kind)
    return out
