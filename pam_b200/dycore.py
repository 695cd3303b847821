"""`Dycore`: the reference's AWFL dycore class (dynamics/awfl/Dycore.h:13-1550) re-homed on libawfl_b200.so.
Same public method names and argument meaning; every compute method goes through the C ABI of include/awfl_b200.h
into hand-written sm_100a kernels. No torch ops, no numpy and no oracle on this path; if the CUDA library is missing
the import of this module's compute entry points raises (no CPU fallback).
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from .coupler import endrun

STATE_FIELDS = ("density_dry", "uvel", "vvel", "wvel", "temp")


class Dycore:
    """Drop-in for `class Dycore` of dynamics/awfl/Dycore.h. `ord` replaces the compile-time MW_ORD (Dycore.h:17-21)."""

    num_state = 5
    idR, idU, idV, idW, idT = 0, 1, 2, 3, 4

    def __init__(self, ord=5):
        if ord not in (3, 5, 7, 9):
            endrun("ERROR: ord must be 3, 5, 7 or 9")
        self.ord = ord
        self.hs = (ord + 1) // 2
        self.h = None
        self.lib = None

    # ------------------------------------------------------------------ init / finalize
    def init(self, coupler, verbose=False):
        """Dycore::init (Dycore.h:835-1091): options, tracer flags, vertical matrices, DataManager entries."""
        self.lib = _lib.load()
        if coupler.device.type != "cuda":
            endrun("ERROR: the AWFL dycore needs a CUDA device (no CPU fallback)")
        nz, ny, nx, nens = coupler.get_nz(), coupler.get_ny(), coupler.get_nx(), coupler.get_nens()
        coupler.set_option("balance_hydrostasis_with_gravity", True)                  # Dycore.h:866
        for k, v in (("R_d", 287.), ("cp_d", 1003.), ("R_v", 461.), ("cp_v", 1859.), ("p0", 1.e5), ("grav", 9.81)):
            if not coupler.option_exists(k):
                coupler.set_option(k, v)                                                # Dycore.h:871-876
        names = coupler.get_tracer_names()
        nt = len(names)
        pos = (C.c_ubyte * max(nt, 1))()
        adds = (C.c_ubyte * max(nt, 1))()
        idwv = -1
        for i, n in enumerate(names):
            _, _, p, a = coupler.get_tracer_info(n)
            pos[i], adds[i] = int(p), int(a)
            if n == "water_vapor":
                idwv = i
        if idwv < 0:
            endrun("ERROR: the AWFL dycore needs a tracer named water_vapor (Dycore.h:83,974)")
        dm = coupler.get_data_manager_device_readwrite()
        dz_host = np.ascontiguousarray(dm.get("vertical_cell_dz", readonly=True).cpu().numpy(), dtype=np.float64)
        cfg = _lib.AwflConfig()
        cfg.nz, cfg.ny, cfg.nx, cfg.nens, cfg.num_tracers, cfg.ord = nz, ny, nx, nens, nt, self.ord
        cfg.xlen, cfg.ylen = coupler.get_xlen(), coupler.get_ylen()
        for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav"):
            setattr(cfg, k, float(coupler.get_option(k)))
        for k in ("gamma_d", "kappa_d", "C0"):
            setattr(cfg, k, float(coupler.get_option(k)) if coupler.option_exists(k) else 0.0)
        cfg.idWV = idwv
        cfg.tracer_positive = C.cast(pos, C.POINTER(C.c_ubyte))
        cfg.tracer_adds_mass = C.cast(adds, C.POINTER(C.c_ubyte))
        cfg.dz_host = dz_host.ctypes.data_as(_lib.c_double_p)
        cfg.device = coupler.device.index or 0
        self._stream = torch.cuda.current_stream(coupler.device)
        cfg.stream = C.c_void_p(self._stream.cuda_stream)
        h = C.c_void_p()
        if self.lib.awfl_create(C.byref(cfg), C.byref(h)) != 0:
            endrun("awfl_create: " + self.lib.awfl_last_error(None).decode())
        self.h = h
        self.nt, self.idWV, self.shape4 = nt, idwv, (nz, ny, nx, nens)
        self.device = coupler.device
        # options the reference sets if absent (Dycore.h:883-891,974)
        for k in ("cv_d", "gamma_d", "kappa_d", "C0"):
            if not coupler.option_exists(k):
                coupler.set_option(k, self.lib.awfl_get_option(self.h, k.encode()))
        if not coupler.option_exists("cv_v"):
            coupler.set_option("cv_v", coupler.get_option("R_v") - coupler.get_option("cp_v"))   # Dycore.h:886 (sic)
        coupler.set_option("idWV", idwv)
        # dycore-owned DataManager entries other modules may read (Dycore.h:868,897-898,975-984)
        o, hs = self.ord, self.hs
        dm.register_existing("variable_gravity", self._view(self.lib.awfl_variable_gravity(self.h), (nz, nens)))
        dm.register_existing("vert_sten_to_coefs", self._view(self.lib.awfl_vert_sten_to_coefs(self.h), (nz + 2, o, o, nens)))
        dm.register_existing("vert_weno_recon_lower",
                             self._view(self.lib.awfl_vert_weno_recon_lower(self.h), (nz + 2, hs, hs, hs, nens)))
        dm.register_existing("tracer_positive", torch.tensor([bool(x) for x in pos[:nt]], device=self.device))
        dm.register_existing("tracer_adds_mass", torch.tensor([bool(x) for x in adds[:nt]], device=self.device))
        if not dm.entry_exists("hy_dens_cells"):
            dm.register_and_allocate("hy_dens_cells", "", (nz, nens))
            dm.register_and_allocate("hy_pressure_cells", "", (nz, nens))
        # idealized initial data, Dycore.h:985-1091: chosen by "initData" of the standalone driver's YAML file
        if coupler.option_exists("standalone_input_file"):
            import yaml
            with open(coupler.get_option("standalone_input_file")) as fh:
                config = yaml.safe_load(fh)
            data = str(config.get("initData"))
            if data in ("thermal", "supercell"):
                self.init_idealized_state_and_tracers(coupler, data)
            elif data != "external":
                endrun("ERROR: Invalid data_spec")                                     # Dycore.h:1001

    def init_idealized_state_and_tracers(self, coupler, data_spec):
        """The DATA_SPEC_THERMAL / DATA_SPEC_SUPERCELL branches of Dycore::init (Dycore.h:1007-1088; init_supercell
        :1096-1276): overwrite the coupler fields with the idealized state."""
        spec = {"thermal": 0, "supercell": 1}.get(data_spec)
        if spec is None:
            endrun("ERROR: Invalid data_spec")
        f, keep = self._fields(coupler)
        zint = coupler.get_data_manager_device_readonly().get("vertical_interface_height", 2, readonly=True)
        self._rc(self.lib.awfl_init_idealized(self.h, C.byref(f), zint.data_ptr(), spec), "awfl_init_idealized")

    def _view(self, ptr, shape):
        """Non-owning torch view of handle-owned device memory (DataManager `get` returns non-owning views too)."""
        n = int(np.prod(shape))

        class _Arr:
            __cuda_array_interface__ = dict(shape=(n,), typestr="<f8", data=(int(ptr), False), version=3)
        t = torch.as_tensor(_Arr(), device=self.device)
        t._awfl_owner = self   # keep the handle alive as long as the view is
        return t.view(*shape)

    def finalize(self, coupler=None):
        if self.h is not None:
            torch.cuda.synchronize(self.device)
            self.lib.awfl_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.finalize()
        except Exception:
            pass

    def dycore_name(self):
        return _lib.load().awfl_dycore_name().decode()

    # ------------------------------------------------------------------ helpers
    def _fields(self, coupler):
        dm = coupler.get_data_manager_device_readwrite()
        f = _lib.AwflFields()
        keep = []
        for n in STATE_FIELDS:
            t = dm.get(n, 4)
            self._check(t, n)
            keep.append(t)
            setattr(f, n, t.data_ptr())
        names = coupler.get_tracer_names()
        arr = (C.c_void_p * max(len(names), 1))()
        for i, n in enumerate(names):
            t = dm.get(n, 4)
            self._check(t, n)
            keep.append(t)
            arr[i] = t.data_ptr()
        f.tracers = C.cast(arr, C.POINTER(C.c_void_p))
        return f, (keep, arr)

    def _check(self, t, name):
        if t.dtype != torch.float64 or not t.is_contiguous() or tuple(t.shape) != self.shape4 or t.device != self.device:
            endrun("ERROR: field %s must be a contiguous float64 (nz,ny,nx,nens) tensor on %s" % (name, self.device))

    def _rc(self, rc, what):
        if rc != 0:
            endrun("%s: %s" % (what, self.lib.awfl_last_error(self.h).decode()))

    # ------------------------------------------------------------------ compute entry points
    def timeStep(self, coupler, ncycles=0):
        """Dycore::timeStep (Dycore.h:107-255). ncycles>0 fixes the sub-cycle count (no host sync). Returns ncycles."""
        f, keep = self._fields(coupler)
        n = C.c_int(0)
        debug = bool(os.environ.get("PAM_DEBUG"))          # the reference's compile-time PAM_DEBUG (Dycore.h:136-138,224-251)
        mass_init = self.compute_mass(coupler) if debug else None
        self._rc(self.lib.awfl_time_step(self.h, C.byref(f), float(coupler.get_option("crm_dt")), int(ncycles), C.byref(n)),
                 "awfl_time_step")
        if debug:
            mass_final = self.compute_mass(coupler)
            diff = mass_final - mass_init
            rel = np.abs(diff) / (np.abs(mass_init) + 1.e-20)
            for ivar, iens in zip(*np.nonzero((rel > 1.e-10) & (np.abs(diff) > 1.e-10))):
                print("WARNING: conservation violated variable,ensemble,rel_diff,mass_diff,init,final: %d , %d , %10.6e , %10.6e , %10.6e , %10.6e"
                      % (ivar, iens, rel[ivar, iens], diff[ivar, iens], mass_init[ivar, iens], mass_final[ivar, iens]))
        return n.value

    def compute_time_step(self, coupler, cfl=0.8):
        """Dycore::compute_time_step (Dycore.h:65-102)."""
        f, keep = self._fields(coupler)
        dt = C.c_double(0.0)
        self._rc(self.lib.awfl_compute_time_step(self.h, C.byref(f), float(cfl), C.byref(dt)), "awfl_compute_time_step")
        return dt.value

    def declare_current_profile_as_hydrostatic(self, coupler, use_gcm_data=False):
        """Dycore::declare_current_profile_as_hydrostatic (Dycore.h:1392-1504)."""
        f, keep = self._fields(coupler)
        g = None
        if use_gcm_data:
            dm = coupler.get_data_manager_device_readonly()
            g = _lib.AwflGcmColumns()
            ts = []
            for n in ("gcm_density_dry", "gcm_temp", "gcm_water_vapor", "gcm_cloud_water", "gcm_cloud_ice"):
                t = dm.get(n, 2, readonly=True)
                ts.append(t)
                setattr(g, n, t.data_ptr())
            g = C.byref(g)
        self._rc(self.lib.awfl_declare_hydrostatic(self.h, C.byref(f), g), "awfl_declare_hydrostatic")

    def compute_mass(self, coupler):
        """Dycore::compute_mass (Dycore.h:37-60) on the coupler's current fields -> numpy (num_tracers+2, nens)."""
        f, keep = self._fields(coupler)
        out = np.empty((self.nt + 2, self.shape4[3]))
        self._rc(self.lib.awfl_compute_mass(self.h, C.byref(f), out.ctypes.data_as(_lib.c_double_p)), "awfl_compute_mass")
        return out

    def compute_tendencies(self, state, tracers, fct_start, dt):
        """Dycore::compute_tendencies (Dycore.h:262-586) on un-haloed conserved device tensors
        state (5,nz,ny,nx,nens), tracers / fct_start (nt,...). Returns (state_tend, tracers_tend)."""
        for t in (state, tracers, fct_start):
            if t.dtype != torch.float64 or not t.is_contiguous() or t.device != self.device:
                endrun("ERROR: compute_tendencies needs contiguous float64 tensors on " + str(self.device))
        st = torch.empty_like(state)
        tt = torch.empty_like(tracers)
        self._rc(self.lib.awfl_compute_tendencies(self.h, state.data_ptr(), tracers.data_ptr(), fct_start.data_ptr(), float(dt),
                                                  st.data_ptr(), tt.data_ptr()), "awfl_compute_tendencies")
        return st, tt

    def reconstruct(self, stencils, ind, level=-1, iens=0):
        """Dycore::reconstruct (Dycore.h:591-604) on a (n,ord) device tensor of stencils."""
        s = stencils.contiguous()
        out = torch.empty(s.shape[0], dtype=torch.float64, device=self.device)
        self._rc(self.lib.awfl_reconstruct(self.h, s.data_ptr(), s.shape[0], int(ind), int(level), int(iens), out.data_ptr()),
                 "awfl_reconstruct")
        return out

    def time_step_host(self, host_arrays, crm_dt, ncycles=0):
        """End-to-end step on HOST numpy arrays (5+nt fields, updated in place): H2D, step, D2H inside the call."""
        arr = (C.c_void_p * len(host_arrays))(*[a.ctypes.data for a in host_arrays])
        n = C.c_int(0)
        self._rc(self.lib.awfl_time_step_host(self.h, arr, float(crm_dt), int(ncycles), C.byref(n)), "awfl_time_step_host")
        return n.value

    def set_host_slabs(self, nslabs):
        """Member slabs pipelined by time_step_host (copies of one slab overlap the kernels of another)."""
        self._rc(self.lib.awfl_set_host_slabs(self.h, int(nslabs)), "awfl_set_host_slabs")


    def host_slabs(self):
        """Slab count the next time_step_host will use (0 passed to set_host_slabs = chosen from the grid)."""
        return int(self.lib.awfl_host_slabs(self.h))

    def poison_workspace(self):
        """Debug aid: NaN-fill every scratch buffer of the handle (a step that reads scratch before writing it shows NaN)."""
        self._rc(self.lib.awfl_poison_workspace(self.h), "awfl_poison_workspace")

    def table_home(self):
        """1: vertical matrices in the constant bank, 3: unit-stride global table, 2: member-strided global table."""
        return int(self.lib.awfl_table_home(self.h))

    def fused_tracer(self):
        """Bit 0 / bit 1: the x / z state march carries the first tracer (2-D runs); 0 otherwise."""
        return int(self.lib.awfl_fused_tracer(self.h))

    def host_mispredictions(self):
        """Calls of time_step_host whose predicted sub-cycle count was wrong (redone un-overlapped)."""
        return self.lib.awfl_host_mispredictions(self.h)

    def profile_begin(self):
        self._rc(self.lib.awfl_profile_begin(self.h), "awfl_profile_begin")

    def profile_end(self):
        """-> {class: (milliseconds, launches)} for the kernels launched since profile_begin (CUDA events)."""
        n = len(_lib.KERNEL_CLASSES)
        ms = (C.c_double * n)()
        cnt = (C.c_long * n)()
        self._rc(self.lib.awfl_profile_end(self.h, ms, cnt), "awfl_profile_end")
        return {k: (ms[i], cnt[i]) for i, k in enumerate(_lib.KERNEL_CLASSES)}

    def launch_count(self):
        return self.lib.awfl_launch_count(self.h)

    def measure_fp64_peak(self, iters=4096):
        v = C.c_double(0.0)
        self._rc(self.lib.awfl_measure_fp64_peak(self.h, int(iters), C.byref(v)), "awfl_measure_fp64_peak")
        return v.value
