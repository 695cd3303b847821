"""Python binding of the GCM-facing host registry (include/pam_interface_b200.h), with the generic names of the reference's
Fortran module pam_fortran_interface.F90 (pam_mirror_array_readonly, pam_create_array, pam_get_array, pam_set_option, ...):
the type suffix is chosen from the numpy dtype / Python type, as the Fortran generic interfaces choose it from the
argument's type and rank. Host data only (pam_interface.h:9-12). Errors inside the library are C++ exceptions, as in the
reference: they end the process with the message on stderr, so this wrapper checks what it can before calling.
"""
import ctypes as C

import numpy as np

from . import _lib

_SUFFIX = {np.dtype(np.bool_): ("bool", C.c_bool), np.dtype(np.int32): ("int", C.c_int),
           np.dtype(np.float32): ("float", C.c_float), np.dtype(np.float64): ("double", C.c_double)}
_keep = {}     # mirrored numpy arrays must outlive their registry entries


def _fn(name, argtypes, restype=None):
    f = getattr(_lib.load(), name)
    f.argtypes, f.restype = argtypes, restype
    return f


def _dims(shape):
    return (C.c_int * len(shape))(*shape)


def _mirror(kind, name, desc, arr):
    if not isinstance(arr, np.ndarray) or arr.dtype not in _SUFFIX or not arr.flags.c_contiguous:
        raise TypeError("mirror needs a C-contiguous numpy array of bool, int32, float32 or float64")
    if not 1 <= arr.ndim <= 7:
        raise ValueError("arrays must have between 1 and 7 dimensions")
    if pam_array_exists(name):
        raise KeyError("Duplicate entry name: " + name)
    suf, ct = _SUFFIX[arr.dtype]
    _fn("pam_interface_mirror_array_%s_%s" % (kind, suf), [C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.c_int, C.c_void_p])(
        name.encode(), desc.encode(), _dims(arr.shape), arr.ndim, arr.ctypes.data)
    _keep[name] = arr


def pam_mirror_array_readonly(name, arr, desc=""):
    _mirror("readonly", name, desc, arr)


def pam_mirror_array_readwrite(name, arr, desc=""):
    _mirror("readwrite", name, desc, arr)


def pam_create_array(name, dtype, shape, desc=""):
    dt = np.dtype(dtype)
    if dt not in _SUFFIX:
        raise TypeError("unsupported element type %s" % dt)
    if pam_array_exists(name):
        raise KeyError("Duplicate entry name: " + name)
    _fn("pam_interface_create_array_%s" % _SUFFIX[dt][0], [C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.c_int])(
        name.encode(), desc.encode(), _dims(shape), len(shape))


def pam_get_array(name, dtype, ndims):
    """A numpy view (no copy) of a registered array."""
    dt = np.dtype(dtype)
    suf, ct = _SUFFIX[dt]
    if not pam_array_exists(name):
        raise KeyError("Could not find entry: " + name)
    ptr, dims = C.POINTER(ct)(), (C.c_int * ndims)()
    _fn("pam_interface_get_array_%s" % suf, [C.c_char_p, C.POINTER(C.POINTER(ct)), C.POINTER(C.c_int), C.c_int])(
        name.encode(), C.byref(ptr), dims, ndims)
    n = int(np.prod(list(dims)))
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct * n)).contents).view(dt).reshape(tuple(dims))


def pam_destroy_array(name):
    _fn("pam_interface_destroy_array", [C.c_char_p])(name.encode())
    _keep.pop(name, None)


def pam_array_exists(name):
    e = C.c_bool(False)
    _fn("pam_interface_array_exists", [C.c_char_p, C.POINTER(C.c_bool)])(name.encode(), C.byref(e))
    return bool(e.value)


def pam_make_readonly(name):
    if not pam_array_exists(name):
        raise KeyError("Could not find entry: " + name)
    _fn("pam_interface_make_readonly", [C.c_char_p])(name.encode())


def pam_register_dimension(name, length):
    _fn("pam_interface_register_dimension", [C.c_char_p, C.c_int])(name.encode(), int(length))


def pam_set_option(name, value, kind=None):
    """kind: 'bool', 'int', 'float', 'double' or 'string'; default from the Python type (float -> double)."""
    if kind is None:
        kind = ("bool" if isinstance(value, (bool, np.bool_)) else "int" if isinstance(value, (int, np.integer)) else
                "string" if isinstance(value, str) else "double")
    if kind == "string":
        _fn("pam_interface_set_option_string", [C.c_char_p, C.c_char_p])(name.encode(), value.encode())
    else:
        ct = {"bool": C.c_bool, "int": C.c_int, "float": C.c_float, "double": C.c_double}[kind]
        _fn("pam_interface_set_option_%s" % kind, [C.c_char_p, ct])(name.encode(), value)


def pam_get_option(name, kind):
    if not pam_option_exists(name):
        raise KeyError("option key does not exist: " + name)
    if kind == "string":
        n = C.c_int(0)
        _fn("pam_interface_get_option_stringlen", [C.c_char_p, C.POINTER(C.c_int)])(name.encode(), C.byref(n))
        buf = C.create_string_buffer(n.value + 1)
        _fn("pam_interface_get_option_string", [C.c_char_p, C.c_char_p])(name.encode(), buf)
        return buf.raw[:n.value].decode()
    ct = {"bool": C.c_bool, "int": C.c_int, "float": C.c_float, "double": C.c_double}[kind]
    v = ct()
    _fn("pam_interface_get_option_%s" % kind, [C.c_char_p, C.POINTER(ct)])(name.encode(), C.byref(v))
    return v.value


def pam_option_exists(name):
    e = C.c_bool(False)
    _fn("pam_interface_option_exists", [C.c_char_p, C.POINTER(C.c_bool)])(name.encode(), C.byref(e))
    return bool(e.value)


def pam_remove_option(name):
    _fn("pam_interface_remove_option", [C.c_char_p])(name.encode())


def pam_finalize():
    _fn("pam_interface_finalize", [])()
    _keep.clear()


def pam_copy_to_device(name, tensor):
    """Registered host array -> a torch CUDA tensor of the same byte size (B200 addition)."""
    _check_bytes(name, tensor)
    if _fn("pam_interface_b200_copy_to_device", [C.c_char_p, C.c_void_p], C.c_int)(name.encode(), tensor.data_ptr()):
        raise RuntimeError("copy to device failed for " + name)


def pam_copy_from_device(name, tensor):
    _check_bytes(name, tensor)
    if _fn("pam_interface_b200_copy_from_device", [C.c_char_p, C.c_void_p], C.c_int)(name.encode(), tensor.data_ptr()):
        raise RuntimeError("copy from device failed for " + name)


def _check_bytes(name, tensor):
    n = C.c_ulong(0)
    if _fn("pam_interface_b200_array_bytes", [C.c_char_p, C.POINTER(C.c_ulong)], C.c_int)(name.encode(), C.byref(n)):
        raise KeyError("Could not find entry: " + name)
    if not tensor.is_cuda or not tensor.is_contiguous() or tensor.numel() * tensor.element_size() != n.value:
        raise ValueError("needs a contiguous CUDA tensor of %d bytes" % n.value)
