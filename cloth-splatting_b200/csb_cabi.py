"""ctypes binding of libcloth_splatting_b200.so (the C-ABI in include/cloth_splatting_b200.h).

This is the only place the shared library is loaded.  There is no CPU fallback: if the library is
missing (and cannot be built) importing the drop-in packages raises.
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_size_t, c_uint32, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcloth_splatting_b200.so")

MAX_VIEWS = 16


class CsbView(Structure):
    _fields_ = [
        ("viewmatrix", c_void_p),
        ("projmatrix", c_void_p),
        ("campos", c_void_p),
        ("background", c_void_p),
        ("tan_fovx", c_float),
        ("tan_fovy", c_float),
    ]


def _load():
    if not os.path.exists(LIB_PATH):
        # build in-tree on first use (nvcc is part of the image); never fall back to a CPU path
        import importlib.util

        spec = importlib.util.spec_from_file_location("_csb_build", os.path.join(_HERE, "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build()
    lib = ctypes.CDLL(LIB_PATH)

    lib.csb_abi_version.restype = c_int
    lib.csb_error_string.restype = c_char_p
    lib.csb_error_string.argtypes = [c_int]

    lib.csb_rasterize_geom_bytes.restype = c_size_t
    lib.csb_rasterize_geom_bytes.argtypes = [c_int, c_int]
    lib.csb_rasterize_image_bytes.restype = c_size_t
    lib.csb_rasterize_image_bytes.argtypes = [c_int, c_int, c_int]
    lib.csb_rasterize_binning_bytes.restype = c_size_t
    lib.csb_rasterize_binning_bytes.argtypes = [c_int, c_int, c_int, c_int, c_size_t]

    lib.csb_rasterize_binning_capacity.restype = c_size_t
    lib.csb_rasterize_binning_capacity.argtypes = [c_int, c_int, c_int, c_int, c_size_t]

    lib.csb_rasterize_forward.restype = c_int
    lib.csb_rasterize_forward.argtypes = [
        c_int, POINTER(CsbView), c_int, c_int, c_int, c_int, c_int,  # V, views, P, D, M, W, H
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int,
        c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, c_size_t,
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
    ]
    lib.csb_rasterize_backward.restype = c_int
    lib.csb_rasterize_backward.argtypes = [
        c_int, POINTER(CsbView), c_int, c_int, c_int, c_int, c_int,
        c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
        c_void_p, c_void_p, c_void_p, c_size_t,
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
    ]
    lib.csb_mark_visible.restype = c_int
    lib.csb_mark_visible.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.csb_rasterize_state_ptr.restype = c_void_p
    lib.csb_rasterize_state_ptr.argtypes = [c_char_p, c_int, c_int, c_int, c_int, c_size_t, c_void_p, c_void_p,
                                            c_void_p]
    for name, res, args in _OPTIONAL:
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
    return lib


# entry points added after ABI v1 are registered here (name, restype, argtypes)
_OPTIONAL = [
    ("csb_dist_cuda2_bytes", c_size_t, [c_int]),
    ("csb_dist_cuda2", c_int, [c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    ("csb_launch_count", ctypes.c_ulonglong, []),
    ("csb_debug_render_variant", c_int, [c_int]),
    ("csb_debug_gnn_small", c_int, [c_int]),
    ("csb_debug_gnn_trace", c_int, [c_void_p, c_int]),
    ("csb_measure_fp32_peaks", c_int, [c_void_p, c_size_t, POINTER(ctypes.c_double), POINTER(ctypes.c_double),
                                       POINTER(ctypes.c_double), c_void_p]),
    ("csb_profile_enable", None, [c_int]),
    ("csb_profile_read", c_int, [c_char_p, POINTER(ctypes.c_double), POINTER(ctypes.c_ulonglong)]),
    ("csb_profile_names", c_int, [c_char_p, c_size_t]),
    ("csb_rasterize_backward_staged", c_int, [
        c_int, POINTER(CsbView), c_int, c_int, c_int, c_int, c_int,
        c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
        c_void_p, c_void_p, c_void_p, c_size_t,
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, POINTER(c_void_p),
        c_void_p]),
    ("csb_peer_header_bytes", c_size_t, []),
    ("csb_peer_alloc", c_int, [c_size_t, POINTER(c_void_p)]),
    ("csb_peer_free", c_int, [c_void_p]),
    ("csb_peer_export", c_int, [c_void_p, c_void_p]),
    ("csb_peer_import", c_int, [c_char_p, POINTER(c_void_p)]),
    ("csb_peer_close", c_int, [c_void_p]),
    ("csb_peer_all_reduce", c_int, [c_int, c_int, POINTER(c_void_p), c_size_t, ctypes.c_uint, c_void_p]),
    ("csb_peer_set_timeout", c_int, [ctypes.c_double]),
    ("csb_peer_timed_out", c_int, [c_void_p, POINTER(ctypes.c_uint)]),
]

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
    return _lib


def check(code, what):
    if code != 0:
        msg = lib().csb_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")


def ptr(t):
    """Device (or host) address of a torch tensor, None -> NULL, empty -> NULL."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def current_stream_ptr():
    import torch

    return torch.cuda.current_stream().cuda_stream


def profile_read():
    """{kernel name: (total device ms, launches)} of the launches recorded since csb_profile_enable(1)."""
    L = lib()
    buf = ctypes.create_string_buffer(4096)
    L.csb_profile_names(buf, 4096)
    out = {}
    for name in [n for n in buf.value.decode().split(",") if n]:
        ms, n = ctypes.c_double(0), ctypes.c_ulonglong(0)
        check(L.csb_profile_read(name.encode(), ctypes.byref(ms), ctypes.byref(n)), "csb_profile_read")
        out[name] = (ms.value, n.value)
    return out
