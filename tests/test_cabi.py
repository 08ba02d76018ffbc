"""The C-ABI library loads and exports every symbol include/*.h declares (no compute calls: CPU only)."""
import ctypes
import glob
import os
import re

import conftest


def declared_symbols():
    names = []
    for h in glob.glob(os.path.join(conftest.ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(csb_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_entry_points():
    names = declared_symbols()
    for must in ("csb_rasterize_forward", "csb_rasterize_backward", "csb_mark_visible", "csb_dist_cuda2",
                 "csb_abi_version"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import csb_cabi

    lib = csb_cabi.lib()
    raw = ctypes.CDLL(csb_cabi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(raw, name), f"{name} declared in include/ but not exported"
    assert lib.csb_abi_version() >= 1
    assert lib.csb_error_string(0) == b"ok"
    assert lib.csb_error_string(-2) == b"workspace too small"


def test_workspace_size_queries_are_consistent():
    import csb_cabi

    lib = csb_cabi.lib()
    assert lib.csb_rasterize_geom_bytes(1, 0) > 0
    g1, g4 = lib.csb_rasterize_geom_bytes(1, 10_000), lib.csb_rasterize_geom_bytes(4, 10_000)
    assert g4 > g1 > 10_000 * 60
    assert lib.csb_rasterize_image_bytes(1, 800, 800) >= 800 * 800 * 8
    for cap in (0, 1, 4096, 123_457, 10_000_000):
        nbytes = lib.csb_rasterize_binning_bytes(1, 1000, 800, 800, cap)
        back = lib.csb_rasterize_binning_capacity(1, 1000, 800, 800, nbytes)
        assert back >= cap
        assert lib.csb_rasterize_binning_bytes(1, 1000, 800, 800, back) <= nbytes
    assert lib.csb_dist_cuda2_bytes(1_000_000) > 1_000_000 * 32


def test_product_never_imports_the_oracle():
    """The product path must not route through oracle/ (or any CPU fallback)."""
    pkg = os.path.join(conftest.ROOT, "cloth-splatting_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(root, f), errors="ignore").read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", src, flags=re.M), f
                assert "liboracle" not in src and "libref_cuda" not in src, f


def test_static_scratch_entry_point_refuses_oversize_without_touching_the_gpu():
    """csb_image_loss_forward_static checks its fixed 1 MiB array before any CUDA call: an oversize request is refused on
    a machine without a GPU too (the GPU parity test is tests/test_zz_no_alloc.py)."""
    import csb_cabi

    lib = csb_cabi.lib()
    c_int, c_void_p, c_float = ctypes.c_int, ctypes.c_void_p, ctypes.c_float
    lib.csb_image_loss_forward_static.restype = c_int
    lib.csb_image_loss_forward_static.argtypes = [c_int] * 4 + [c_void_p] * 3 + [c_int, c_float] + [c_void_p] * 3 + [c_void_p]
    lib.csb_image_loss_static_max_bytes.restype = ctypes.c_size_t
    assert lib.csb_image_loss_static_max_bytes() == 1 << 20
    dummy = ctypes.create_string_buffer(16)                      # never dereferenced: the size check comes first
    p = ctypes.addressof(dummy)
    # 16 x 91 x 91 tiles of 32 x 32 = 132 496 > 131 072
    assert lib.csb_image_loss_forward_static(16, 1, 2900, 2900, p, p, None, 0, 0.2, None, None, p, None) == -4
    assert b"no-allocation" in lib.csb_error_string(-4)
    assert lib.csb_image_loss_forward_static(1, 3, 64, 64, None, p, None, 0, 0.2, None, None, p, None) == -1   # NULL image
    assert lib.csb_image_loss_forward_static(1, 3, 0, 64, p, p, None, 0, 0.2, None, None, p, None) == -1      # empty plane
