"""TEST INFRASTRUCTURE ONLY: ctypes driver of oracle/_ref/libref_cuda.so.

libref_cuda.so is the reference's own CUDA rasterizer + simple-knn, compiled unmodified from
/root/reference by oracle/Makefile (`make ref`).  It needs a GPU, so it is used by the `-m gpu`
parity tests as the bit-exactness arbiter, by tests/make_golden.py to pin the CPU oracle, and by
`bench.py --impl reference` as the timed reference arm.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref_cuda.so")

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.ref_ctx_create.restype = ctypes.c_void_p
        _lib.ref_ctx_destroy.argtypes = [ctypes.c_void_p]
        _lib.ref_state_ptr.restype = ctypes.c_void_p
        _lib.ref_state_ptr.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        _lib.ref_rasterize_forward.restype = ctypes.c_int
        _lib.ref_rasterize_backward.restype = ctypes.c_int
        _lib.ref_dist_cuda2.restype = ctypes.c_int
        _lib.ref_copy_state.restype = ctypes.c_int
        _lib.ref_mark_visible.restype = ctypes.c_int
    return _lib


def _p(t):
    return ctypes.c_void_p(None if t is None or t.numel() == 0 else t.data_ptr())


class RefRasterizer:
    """One reference context (owns the three growable buffers).  The reference launches on the
    legacy default stream, so callers synchronise around it."""

    def __init__(self):
        self.ctx = ctypes.c_void_p(lib().ref_ctx_create())

    def __del__(self):
        try:
            lib().ref_ctx_destroy(self.ctx)
        except Exception:
            pass

    def forward(self, means3D, opacities, viewmatrix, projmatrix, campos, bg, tanfovx, tanfovy, H, W, shs=None,
                colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, sh_degree=3,
                scale_modifier=1.0, sync=True):
        """sync=False: no host synchronisation is added around the call -- torch's default stream IS the legacy default
        stream the reference launches on, so stream order already holds (bench.py times the reference this way; the
        tests keep the conservative default)."""
        dev = means3D.device
        P = means3D.shape[0]
        M = shs.shape[1] if shs is not None else 0
        color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        depth = torch.zeros((1, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        if sync:
            torch.cuda.synchronize()
        R = lib().ref_rasterize_forward(
            self.ctx, ctypes.c_int(P), ctypes.c_int(sh_degree), ctypes.c_int(M), _p(bg), ctypes.c_int(W),
            ctypes.c_int(H), _p(means3D), _p(shs), _p(colors_precomp), _p(opacities), _p(scales),
            ctypes.c_float(scale_modifier), _p(rotations), _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix),
            _p(campos), ctypes.c_float(tanfovx), ctypes.c_float(tanfovy), ctypes.c_int(0), _p(color), _p(depth),
            _p(radii))
        if R < 0:
            raise RuntimeError(f"reference forward failed: cudaError {-R}")
        self.P, self.H, self.W, self.M, self.R = P, H, W, M, R
        self._args = dict(means3D=means3D, shs=shs, colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                          cov3D_precomp=cov3D_precomp, viewmatrix=viewmatrix, projmatrix=projmatrix, campos=campos,
                          bg=bg, tanfovx=tanfovx, tanfovy=tanfovy, sh_degree=sh_degree,
                          scale_modifier=scale_modifier, radii=radii)
        return R, color, radii, depth

    def backward(self, dL_dpix, sync=True):
        a = self._args
        dev = dL_dpix.device
        P, M = self.P, self.M
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacities=z(P, 1), colors=z(P, 3), means3D=z(P, 3),
                 cov3D=z(P, 6), shs=z(P, max(M, 1), 3), scales=z(P, 3), rotations=z(P, 4))
        if sync:
            torch.cuda.synchronize()
        rc = lib().ref_rasterize_backward(
            self.ctx, ctypes.c_int(P), ctypes.c_int(a["sh_degree"]), ctypes.c_int(M), ctypes.c_int(self.R),
            _p(a["bg"]), ctypes.c_int(self.W), ctypes.c_int(self.H), _p(a["means3D"]), _p(a["shs"]),
            _p(a["colors_precomp"]), _p(a["scales"]), ctypes.c_float(a["scale_modifier"]), _p(a["rotations"]),
            _p(a["cov3D_precomp"]), _p(a["viewmatrix"]), _p(a["projmatrix"]), _p(a["campos"]),
            ctypes.c_float(a["tanfovx"]), ctypes.c_float(a["tanfovy"]), _p(a["radii"]), _p(dL_dpix.contiguous()),
            _p(g["means2D"]), _p(g["conic"]), _p(g["opacities"]), _p(g["colors"]), _p(g["means3D"]), _p(g["cov3D"]),
            _p(g["shs"]), _p(g["scales"]), _p(g["rotations"]))
        if rc != 0:
            raise RuntimeError(f"reference backward failed: cudaError {-rc}")
        if sync:
            torch.cuda.synchronize()
        if M == 0:
            g["shs"] = None
        return g

    def state(self, name, dtype, shape):
        """Copy of one carved sub-array of the last forward call (see ref_wrapper.cu:ref_state_ptr)."""
        n = 1
        for s_ in shape:
            n *= s_
        out = torch.empty(shape, dtype=dtype, device="cuda")
        if n:
            torch.cuda.synchronize()
            rc = lib().ref_copy_state(self.ctx, name.encode(), ctypes.c_void_p(out.data_ptr()),
                                      ctypes.c_size_t(n * out.element_size()))
            if rc != 0:
                raise KeyError(f"{name}: {rc}")
        return out


def dist_cuda2(points):
    P = points.shape[0]
    out = torch.zeros(P, dtype=torch.float32, device=points.device)
    pts = points.contiguous()
    torch.cuda.synchronize()
    rc = lib().ref_dist_cuda2(ctypes.c_int(P), _p(pts), _p(out))
    if rc != 0:
        raise RuntimeError(f"reference distCUDA2 failed: cudaError {-rc}")
    return out
