"""Host-side mirror of the reference's coupler surface that the AWFL dycore touches -- same method names and argument
meaning as pam::PamCoupler / DataManager / Options (pam_core/pam_coupler.h:59-356, pam_core/DataManager.h:90-312,
pam_core/Options.h:63-131) -- backed by torch CUDA tensors (torch is plumbing here: device memory and streams).

The C++ twin for a compiled caller is pam_b200/cxx/pam_coupler_b200.h.
"""
import numpy as np
import torch


def endrun(msg):
    """pam_core/pam_const.h:249-252: print and throw."""
    raise RuntimeError(msg)


class DataManager:
    """Named-field registry (pam_core/DataManager.h). Entries are torch tensors on the coupler's device."""

    def __init__(self, device):
        self.device = device
        self.entries = {}
        self.dirty = {}

    def register_and_allocate(self, name, desc, dims, dtype=torch.float64):
        if name in self.entries:
            endrun("ERROR: Duplicate entry name: " + name)          # DataManager.h:104-108
        self.entries[name] = torch.zeros(tuple(int(d) for d in dims), dtype=dtype, device=self.device)
        self.dirty[name] = False
        return self.entries[name]

    def register_existing(self, name, tensor):
        self.entries[name] = tensor
        self.dirty[name] = False

    def unregister_and_deallocate(self, name):
        self.entries.pop(name, None)
        self.dirty.pop(name, None)

    def entry_exists(self, name):
        return name in self.entries

    def get(self, name, ndim=None, readonly=False):
        if name not in self.entries:
            endrun("ERROR: Could not find entry: " + name)           # DataManager.h:553-560
        t = self.entries[name]
        if ndim is not None and t.dim() != ndim:
            endrun("ERROR: Requested dimensions is different from entry dimensions for " + name)   # DataManager.h:569-575
        if not readonly:
            self.dirty[name] = True                                  # DataManager.h:306
        return t

    def get_collapsed(self, name):
        return self.get(name).reshape(-1)

    def finalize(self):
        self.entries.clear()
        self.dirty.clear()


class PamCoupler:
    """pam::PamCoupler look-alike: options + device DataManager + tracer registry + grid."""

    def __init__(self, device="cuda:0"):
        self.device = torch.device(device)
        self.options = {}
        self.dm = DataManager(self.device)
        self.tracers = []          # dicts: name, desc, positive, adds_mass
        self.xlen = self.ylen = 0.0
        self.nz = self.ny = self.nx = self.nens = 0

    # ---- Options (pam_core/Options.h) ----
    def set_option(self, key, value):
        self.options[key] = value

    def get_option(self, key):
        if key not in self.options:
            endrun("ERROR: option key does not exist: " + key)       # Options.h:84-87
        return self.options[key]

    def option_exists(self, key):
        return key in self.options

    # ---- state / grid (pam_coupler.h:184-356) ----
    def allocate_coupler_state(self, nz, ny, nx, nens):
        self.nz, self.ny, self.nx, self.nens = int(nz), int(ny), int(nx), int(nens)
        for n in ("density_dry", "uvel", "vvel", "wvel", "temp"):
            self.dm.register_and_allocate(n, n, (nz, ny, nx, nens))
        self.dm.register_and_allocate("vertical_interface_height", "", (nz + 1, nens))
        self.dm.register_and_allocate("vertical_cell_dz", "", (nz, nens))
        self.dm.register_and_allocate("vertical_midpoint_height", "", (nz, nens))
        for n in ("gcm_density_dry", "gcm_uvel", "gcm_vvel", "gcm_wvel", "gcm_temp", "gcm_water_vapor",
                  "gcm_cloud_water", "gcm_cloud_ice", "gcm_num_liq", "gcm_num_ice", "gcm_num_rain"):   # pam_coupler.h:268-281
            self.dm.register_and_allocate(n, "", (nz, nens))

    def set_grid(self, xlen, ylen, zint):
        """zint: (nz+1,) or (nz+1,nens) interface heights (pam_coupler.h:184-202)."""
        self.xlen, self.ylen = float(xlen), float(ylen)
        z = np.asarray(zint, dtype=np.float64)
        if z.ndim == 1:
            z = np.repeat(z[:, None], self.nens, axis=1)
        if z.shape != (self.nz + 1, self.nens):
            endrun("ERROR: zint has the wrong shape")
        zt = torch.from_numpy(np.ascontiguousarray(z)).to(self.device)
        self.dm.get("vertical_interface_height").copy_(zt)
        self.dm.get("vertical_cell_dz").copy_(zt[1:] - zt[:-1])
        self.dm.get("vertical_midpoint_height").copy_(0.5 * (zt[1:] + zt[:-1]))

    def add_tracer(self, name, desc="", positive=True, adds_mass=False):
        self.dm.register_and_allocate(name, desc, (self.nz, self.ny, self.nx, self.nens))
        self.tracers.append(dict(name=name, desc=desc, positive=bool(positive), adds_mass=bool(adds_mass)))

    def get_tracer_names(self):
        return [t["name"] for t in self.tracers]

    def get_tracer_info(self, name):
        for t in self.tracers:
            if t["name"] == name:
                return t["desc"], True, t["positive"], t["adds_mass"]
        return "", False, False, False

    def get_num_tracers(self):
        return len(self.tracers)

    def get_nx(self): return self.nx
    def get_ny(self): return self.ny
    def get_nz(self): return self.nz
    def get_nens(self): return self.nens
    def get_xlen(self): return self.xlen
    def get_ylen(self): return self.ylen
    def get_dx(self): return self.xlen / self.nx
    def get_dy(self): return self.ylen / self.ny

    def get_data_manager_device_readwrite(self):
        return self.dm

    def get_data_manager_device_readonly(self):
        return self.dm

    def run_module(self, name, fn):
        """pam_coupler.h:139-160"""
        fn(self)

    def finalize(self):
        self.dm.finalize()
