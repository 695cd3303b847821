"""ctypes binding of libawfl_b200.so (include/awfl_b200.h). There is no fallback: if the CUDA library is missing
or fails to load, importing a compute entry point raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("AWFL_B200_LIB") or os.path.join(_HERE, "libawfl_b200.so")   # env override: kernel experiments

c_double_p = C.POINTER(C.c_double)


class AwflConfig(C.Structure):
    _fields_ = [("nz", C.c_int), ("ny", C.c_int), ("nx", C.c_int), ("nens", C.c_int), ("num_tracers", C.c_int),
                ("ord", C.c_int), ("xlen", C.c_double), ("ylen", C.c_double),
                ("R_d", C.c_double), ("cp_d", C.c_double), ("R_v", C.c_double), ("cp_v", C.c_double),
                ("p0", C.c_double), ("grav", C.c_double),
                ("gamma_d", C.c_double), ("kappa_d", C.c_double), ("C0", C.c_double),
                ("idWV", C.c_int), ("tracer_positive", C.POINTER(C.c_ubyte)), ("tracer_adds_mass", C.POINTER(C.c_ubyte)),
                ("dz_host", c_double_p), ("device", C.c_int), ("stream", C.c_void_p)]


class AwflFields(C.Structure):
    _fields_ = [("density_dry", C.c_void_p), ("uvel", C.c_void_p), ("vvel", C.c_void_p), ("wvel", C.c_void_p),
                ("temp", C.c_void_p), ("tracers", C.POINTER(C.c_void_p))]


class AwflGcmColumns(C.Structure):
    _fields_ = [("gcm_density_dry", C.c_void_p), ("gcm_temp", C.c_void_p), ("gcm_water_vapor", C.c_void_p),
                ("gcm_cloud_water", C.c_void_p), ("gcm_cloud_ice", C.c_void_p)]


EXPORTS = ("awfl_create", "awfl_destroy", "awfl_last_error", "awfl_dycore_name", "awfl_time_step",
           "awfl_time_step_host", "awfl_compute_time_step", "awfl_compute_tendencies", "awfl_declare_hydrostatic",
           "awfl_compute_mass", "awfl_variable_gravity", "awfl_vert_sten_to_coefs", "awfl_vert_weno_recon_lower",
           "awfl_get_option", "awfl_reconstruct", "awfl_launch_count", "awfl_measure_fp64_peak", "awfl_profile_begin",
           "awfl_profile_end", "awfl_set_variable_gravity", "awfl_export_tables", "awfl_copy_to_host", "awfl_copy_to_device",
           "awfl_current_device", "awfl_device_alloc", "awfl_device_free", "awfl_set_host_slabs", "awfl_host_slabs", "awfl_host_mispredictions", "awfl_init_idealized",
           "awfl_table_home", "awfl_convert_coupler_to_dynamics", "awfl_convert_dynamics_to_coupler", "awfl_halo_exchange",
           "awfl_compute_tendencies_haloed", "awfl_compute_mass_dynamics", "awfl_init_supercell", "awfl_poison_workspace",
           "awfl_noalloc_setup", "awfl_noalloc_max_doubles", "awfl_compute_mass_dynamics_noalloc", "awfl_fused_tracer")

KERNEL_CLASSES = ("in", "out", "state_x", "state_y", "state_z", "tracer_x", "tracer_y", "tracer_z", "fct", "tracer_upd", "other")

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("pam_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.awfl_create.argtypes = [C.POINTER(AwflConfig), C.POINTER(C.c_void_p)]
    lib.awfl_destroy.argtypes = [C.c_void_p]
    lib.awfl_destroy.restype = None
    lib.awfl_last_error.argtypes = [C.c_void_p]
    lib.awfl_last_error.restype = C.c_char_p
    lib.awfl_dycore_name.restype = C.c_char_p
    lib.awfl_time_step.argtypes = [C.c_void_p, C.POINTER(AwflFields), C.c_double, C.c_int, C.POINTER(C.c_int)]
    lib.awfl_time_step_host.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_double, C.c_int, C.POINTER(C.c_int)]
    lib.awfl_compute_time_step.argtypes = [C.c_void_p, C.POINTER(AwflFields), C.c_double, c_double_p]
    lib.awfl_compute_tendencies.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p]
    lib.awfl_declare_hydrostatic.argtypes = [C.c_void_p, C.POINTER(AwflFields), C.POINTER(AwflGcmColumns)]
    lib.awfl_compute_mass.argtypes = [C.c_void_p, C.POINTER(AwflFields), c_double_p]
    lib.awfl_variable_gravity.argtypes = [C.c_void_p]
    lib.awfl_variable_gravity.restype = C.c_void_p
    lib.awfl_vert_sten_to_coefs.argtypes = [C.c_void_p]
    lib.awfl_vert_sten_to_coefs.restype = C.c_void_p
    lib.awfl_vert_weno_recon_lower.argtypes = [C.c_void_p]
    lib.awfl_vert_weno_recon_lower.restype = C.c_void_p
    lib.awfl_get_option.argtypes = [C.c_void_p, C.c_char_p]
    lib.awfl_get_option.restype = C.c_double
    lib.awfl_reconstruct.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.awfl_launch_count.argtypes = [C.c_void_p]
    lib.awfl_launch_count.restype = C.c_long
    lib.awfl_measure_fp64_peak.argtypes = [C.c_void_p, C.c_int, c_double_p]
    lib.awfl_set_variable_gravity.argtypes = [C.c_void_p, C.c_void_p]
    lib.awfl_export_tables.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.awfl_profile_begin.argtypes = [C.c_void_p]
    lib.awfl_init_idealized.argtypes = [C.c_void_p, C.POINTER(AwflFields), C.c_void_p, C.c_int]
    lib.awfl_set_host_slabs.argtypes = [C.c_void_p, C.c_int]
    lib.awfl_host_slabs.argtypes = [C.c_void_p]
    lib.awfl_host_slabs.restype = C.c_int
    lib.awfl_host_mispredictions.argtypes = [C.c_void_p]
    lib.awfl_host_mispredictions.restype = C.c_long
    lib.awfl_table_home.argtypes = [C.c_void_p]
    lib.awfl_poison_workspace.argtypes = [C.c_void_p]
    lib.awfl_convert_coupler_to_dynamics.argtypes = [C.c_void_p, C.POINTER(AwflFields), C.c_void_p, C.c_void_p]
    lib.awfl_convert_dynamics_to_coupler.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(AwflFields)]
    lib.awfl_halo_exchange.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.awfl_compute_tendencies_haloed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
    lib.awfl_compute_mass_dynamics.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.awfl_fused_tracer.argtypes = [C.c_void_p]
    lib.awfl_noalloc_setup.argtypes = []
    lib.awfl_noalloc_max_doubles.argtypes = []
    lib.awfl_noalloc_max_doubles.restype = C.c_ulong
    lib.awfl_compute_mass_dynamics_noalloc.argtypes = [C.c_int] * 6 + [C.c_void_p] * 5
    lib.awfl_init_supercell.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.awfl_profile_end.argtypes = [C.c_void_p, c_double_p, C.POINTER(C.c_long)]
    _lib = lib
    return lib
