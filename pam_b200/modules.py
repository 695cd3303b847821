"""Host-side mirror of the coupler modules that run either side of the dycore in PAM's step loop
(standalone/mmf_simplified/driver.cpp:239-250): same names, argument meaning, options and DataManager entries as
pam_core/modules/sponge_layer.h, gcm_forcing.h and perturb_temperature.h, over the C ABI of
include/pam_modules_b200.h (sm_100a CUDA kernels inside libawfl_b200.so). No CPU fallback: a missing library or a
non-CUDA coupler raises.

    coupler.run_module("apply_gcm_forcing_tendencies", modules.apply_gcm_forcing_tendencies)
    coupler.run_module("dycore", lambda c: dycore.timeStep(c))
    coupler.run_module("sponge_layer", modules.sponge_layer)
"""
import ctypes as C

import torch

from . import _lib
from .coupler import endrun

CRM_NAMES = ("density_dry", "uvel", "vvel", "temp", "water_vapor", "cloud_water", "ice", "cloud_water_num", "ice_num", "rain_num")
GCM_NAMES = ("gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice",
             "gcm_num_liq", "gcm_num_ice", "gcm_num_rain")
TEND_NAMES = ("rho_d", "uvel", "vvel", "temp", "qtot", "qv", "ql", "qi", "rho_v", "rho_l", "rho_i", "nc", "ni", "nr")


class _Crm(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in CRM_NAMES]


class _Gcm(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in GCM_NAMES]


class _Tend(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in TEND_NAMES]


_bound = False


def _bind(lib):
    global _bound
    if _bound:
        return
    lib.pamm_create.argtypes = [C.c_int] * 5 + [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.pamm_destroy.argtypes = [C.c_void_p]
    lib.pamm_destroy.restype = None
    lib.pamm_last_error.argtypes = [C.c_void_p]
    lib.pamm_last_error.restype = C.c_char_p
    lib.pamm_launch_count.argtypes = [C.c_void_p]
    lib.pamm_launch_count.restype = C.c_long
    lib.pamm_sponge_layer.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int]
    lib.pamm_compute_gcm_forcing_tendencies.argtypes = [C.c_void_p, C.POINTER(_Crm), C.POINTER(_Gcm), C.POINTER(_Tend), C.c_double]
    lib.pamm_apply_gcm_forcing_tendencies.argtypes = [C.c_void_p, C.POINTER(_Crm), C.POINTER(_Gcm), C.POINTER(_Tend), C.c_void_p,
                                                      C.c_double, C.c_double]
    lib.pamm_broadcast_initial_gcm_column.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    lib.pamm_perturb_temperature.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
    _bound = True


class _Workspace:
    """pamm_workspace of one coupler (created on first use, kept on the coupler object)."""

    def __init__(self, coupler):
        self.lib = _lib.load()
        _bind(self.lib)
        if coupler.device.type != "cuda":
            endrun("ERROR: the coupler modules need a CUDA coupler (no CPU fallback)")
        h = C.c_void_p()
        dev = coupler.device.index if coupler.device.index is not None else torch.cuda.current_device()
        stream = torch.cuda.current_stream(coupler.device).cuda_stream
        if self.lib.pamm_create(coupler.get_nz(), coupler.get_ny(), coupler.get_nx(), coupler.get_nens(), dev, C.c_void_p(stream), C.byref(h)):
            endrun("pamm_create: " + self.lib.pamm_last_error(None).decode())
        self.h = h

    def rc(self, code, what):
        if code:
            endrun("%s: %s" % (what, self.lib.pamm_last_error(self.h).decode()))

    def __del__(self):
        try:
            if self.h:
                self.lib.pamm_destroy(self.h)
                self.h = None
        except Exception:
            pass


def _ws(coupler):
    w = getattr(coupler, "_pamm_ws", None)
    if w is None:
        w = coupler._pamm_ws = _Workspace(coupler)
    return w


def launch_count(coupler):
    w = _ws(coupler)
    return w.lib.pamm_launch_count(w.h)


def sponge_layer(coupler):
    """modules::sponge_layer (pam_core/modules/sponge_layer.h:9-96). Options sponge_num_layers (5), sponge_time_scale (60 s)."""
    w = _ws(coupler)
    num_layers = int(coupler.get_option("sponge_num_layers")) if coupler.option_exists("sponge_num_layers") else 5
    time_scale = float(coupler.get_option("sponge_time_scale")) if coupler.option_exists("sponge_time_scale") else 60.0
    dm = coupler.get_data_manager_device_readwrite()
    names = ["density_dry", "uvel", "vvel", "wvel", "temp"] + coupler.get_tracer_names()
    ptrs = (C.c_void_p * len(names))(*[dm.get(n, 4).data_ptr() for n in names])
    zint = dm.get("vertical_interface_height", 2, readonly=True)
    zmid = dm.get("vertical_midpoint_height", 2, readonly=True)
    w.rc(w.lib.pamm_sponge_layer(w.h, ptrs, len(names), zint.data_ptr(), zmid.data_ptr(), float(coupler.get_option("crm_dt")),
                                 time_scale, num_layers), "pamm_sponge_layer")


def _bundles(coupler, dm):
    nz, nens = coupler.get_nz(), coupler.get_nens()
    if not dm.entry_exists("gcm_forcing_tend_uvel"):                      # gcm_forcing.h:134-149
        for n in TEND_NAMES:
            dm.register_and_allocate("gcm_forcing_tend_" + n, "GCM forcing", (nz, nens))
    crm = _Crm(*[dm.get(n, 4).data_ptr() for n in CRM_NAMES])
    gcm = _Gcm(*[dm.get(n, 2, readonly=True).data_ptr() for n in GCM_NAMES])
    tend = _Tend(*[dm.get("gcm_forcing_tend_" + n, 2).data_ptr() for n in TEND_NAMES])
    return crm, gcm, tend


def compute_gcm_forcing_tendencies(coupler):
    """modules::compute_gcm_forcing_tendencies (pam_core/modules/gcm_forcing.h:18-205); option gcm_physics_dt."""
    w = _ws(coupler)
    dm = coupler.get_data_manager_device_readwrite()
    crm, gcm, tend = _bundles(coupler, dm)
    w.rc(w.lib.pamm_compute_gcm_forcing_tendencies(w.h, C.byref(crm), C.byref(gcm), C.byref(tend), float(coupler.get_option("gcm_physics_dt"))),
         "pamm_compute_gcm_forcing_tendencies")


def apply_gcm_forcing_tendencies(coupler):
    """modules::apply_gcm_forcing_tendencies + fill_holes (pam_core/modules/gcm_forcing.h:208-436); options crm_dt, gcm_physics_dt."""
    w = _ws(coupler)
    dm = coupler.get_data_manager_device_readwrite()
    if not dm.entry_exists("gcm_forcing_tend_uvel"):
        endrun("ERROR: Could not find entry: gcm_forcing_tend_rho_d")     # DataManager.h:553-560: compute_* must run first
    crm, gcm, tend = _bundles(coupler, dm)
    dz = dm.get("vertical_cell_dz", 2, readonly=True)
    w.rc(w.lib.pamm_apply_gcm_forcing_tendencies(w.h, C.byref(crm), C.byref(gcm), C.byref(tend), dz.data_ptr(),
                                                 float(coupler.get_option("crm_dt")), float(coupler.get_option("gcm_physics_dt"))),
         "pamm_apply_gcm_forcing_tendencies")


def broadcast_initial_gcm_column(coupler):
    """modules::broadcast_initial_gcm_column (pam_core/modules/broadcast_initial_gcm_column.h:7-41)."""
    w = _ws(coupler)
    dm = coupler.get_data_manager_device_readwrite()
    cols = ("gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_wvel", "gcm_temp", "gcm_water_vapor")
    flds = ("density_dry", "uvel", "vvel", "wvel", "temp", "water_vapor")
    cp = (C.c_void_p * 6)(*[dm.get(n, 2, readonly=True).data_ptr() for n in cols])
    fp = (C.c_void_p * 6)(*[dm.get(n, 4).data_ptr() for n in flds])
    w.rc(w.lib.pamm_broadcast_initial_gcm_column(w.h, cp, fp), "pamm_broadcast_initial_gcm_column")


def perturb_temperature(coupler, ids, magnitude=0.1):
    """modules::perturb_temperature (pam_core/modules/perturb_temperature.h:10-63). ids: one int per ensemble member
    (globally unique per batch of CRMs, as the reference asks)."""
    w = _ws(coupler)
    idt = torch.as_tensor(ids, dtype=torch.int32).to(coupler.device).contiguous()
    if idt.numel() != coupler.get_nens():
        endrun("ERROR: size of id array must be the same as nens")        # perturb_temperature.h:20
    dm = coupler.get_data_manager_device_readwrite()
    temp = dm.get("temp", 4)
    w.rc(w.lib.pamm_perturb_temperature(w.h, temp.data_ptr(), idt.data_ptr(), float(magnitude)), "pamm_perturb_temperature")
    coupler._pamm_keepalive = idt
