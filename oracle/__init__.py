"""TEST INFRASTRUCTURE ONLY: numpy front-end of the CPU oracle (oracle/liboracle.so, plain C).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  Nothing under cloth-splatting_b200/ does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("rast_oracle.c", "rast_oracle_bwd.c", "knn_oracle.c")]
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC",
                               "-o", _LIB, *srcs, "-lm"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB)
        _lib.oracle_preprocess.restype = ctypes.c_int64
        _lib.oracle_bin.restype = ctypes.c_int
        _lib.oracle_higher_msb.restype = ctypes.c_uint32
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class RasterState:
    """All intermediates of one forward pass of one view (numpy arrays)."""


def rasterize_forward(means3D, opacities, viewmatrix, projmatrix, campos, bg, tanfovx, tanfovy, H, W, shs=None,
                      colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, sh_degree=3,
                      scale_modifier=1.0):
    """CPU restatement of Rasterizer::forward for one view; returns a RasterState."""
    L = lib()
    means3D = _f32(means3D)
    P = means3D.shape[0]
    shs, colors_precomp, scales, rotations, cov3D_precomp = map(_f32, (shs, colors_precomp, scales, rotations, cov3D_precomp))
    opacities = _f32(opacities).reshape(-1)
    vm, pm, cp, bg = _f32(viewmatrix).reshape(-1), _f32(projmatrix).reshape(-1), _f32(campos).reshape(-1), _f32(bg).reshape(-1)
    M = shs.shape[1] if shs is not None else 0
    s = RasterState()
    s.P, s.H, s.W, s.M, s.D = P, H, W, M, sh_degree
    s.radii = np.zeros(P, np.int32)
    s.means2D = np.zeros((P, 2), np.float32)
    s.depths = np.zeros(P, np.float32)
    s.cov3D = np.zeros((P, 6), np.float32) if cov3D_precomp is None else cov3D_precomp
    s.conic_opacity = np.zeros((P, 4), np.float32)
    s.rgb = np.zeros((P, 3), np.float32)
    s.clamped = np.zeros((P, 3), np.uint8)
    s.tiles_touched = np.zeros(P, np.uint32)
    R = L.oracle_preprocess(
        ctypes.c_int(P), ctypes.c_int(sh_degree), ctypes.c_int(M), ctypes.c_int(W), ctypes.c_int(H), _p(means3D),
        _p(shs), _p(colors_precomp), _p(opacities), _p(scales), ctypes.c_float(scale_modifier), _p(rotations),
        _p(cov3D_precomp), _p(vm), _p(pm), _p(cp), ctypes.c_float(tanfovx), ctypes.c_float(tanfovy), _p(s.radii),
        _p(s.means2D), _p(s.depths), _p(s.cov3D), _p(s.conic_opacity), _p(s.rgb), _p(s.clamped), _p(s.tiles_touched))
    s.num_rendered = int(R)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    s.keys = np.zeros(max(R, 1), np.uint64)[:R]
    s.point_list = np.zeros(max(R, 1), np.uint32)[:R]
    s.ranges = np.zeros((gx * gy, 2), np.uint32)
    rc = L.oracle_bin(ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(s.radii), _p(s.means2D), _p(s.depths),
                      ctypes.c_int64(R), _p(s.keys), _p(s.point_list), _p(s.ranges))
    if rc != 0:
        raise RuntimeError(f"oracle_bin failed: {rc}")
    s.final_T = np.zeros((H, W), np.float32)
    s.n_contrib = np.zeros((H, W), np.uint32)
    s.color = np.zeros((3, H, W), np.float32)
    s.depth = np.zeros((1, H, W), np.float32)
    s.counts = np.zeros(3, np.int64)
    L.oracle_render(ctypes.c_int(W), ctypes.c_int(H), _p(s.ranges), _p(s.point_list), _p(s.means2D), _p(s.rgb),
                    _p(s.depths), _p(s.conic_opacity), _p(bg), _p(s.final_T), _p(s.n_contrib), _p(s.color),
                    _p(s.depth), _p(s.counts))
    s._inputs = dict(means3D=means3D, shs=shs, colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                     cov3D_precomp=cov3D_precomp, vm=vm, pm=pm, cp=cp, bg=bg, tanfovx=tanfovx, tanfovy=tanfovy,
                     scale_modifier=scale_modifier)
    return s


def rasterize_backward(s, dL_dpix):
    """CPU restatement of Rasterizer::backward for the view of RasterState `s`; returns a dict of gradients."""
    L = lib()
    i = s._inputs
    P, H, W, M = s.P, s.H, s.W, s.M
    dL_dpix = _f32(dL_dpix).reshape(3, H, W)
    acc = np.zeros((P, 9), np.float64)
    L.oracle_render_backward(ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(s.ranges), _p(s.point_list),
                             _p(s.means2D), _p(s.rgb), _p(s.conic_opacity), _p(i["bg"]), _p(s.final_T),
                             _p(s.n_contrib), _p(dL_dpix), _p(acc))
    g = dict(
        means2D=np.zeros((P, 3), np.float32), colors=np.zeros((P, 3), np.float32),
        opacities=np.zeros((P, 1), np.float32), means3D=np.zeros((P, 3), np.float32),
        cov3D=np.zeros((P, 6), np.float32), shs=np.zeros((P, M, 3), np.float32) if M else None,
        scales=np.zeros((P, 3), np.float32), rotations=np.zeros((P, 4), np.float32), acc=acc)
    use_scales = i["cov3D_precomp"] is None
    L.oracle_gaussian_backward(
        ctypes.c_int(P), ctypes.c_int(s.D), ctypes.c_int(M), ctypes.c_int(W), ctypes.c_int(H), _p(i["means3D"]),
        _p(i["shs"]), _p(i["scales"]) if use_scales else None, ctypes.c_float(i["scale_modifier"]),
        _p(i["rotations"]) if use_scales else None, _p(s.cov3D), _p(i["vm"]), _p(i["pm"]), _p(i["cp"]),
        ctypes.c_float(i["tanfovx"]), ctypes.c_float(i["tanfovy"]), _p(s.radii), _p(s.clamped), _p(acc),
        _p(g["means2D"]), _p(g["colors"]), _p(g["opacities"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["shs"]),
        _p(g["scales"]), _p(g["rotations"]))
    return g


def dist2(points):
    """CPU restatement of simple_knn distCUDA2."""
    pts = _f32(points)
    out = np.zeros(pts.shape[0], np.float32)
    lib().oracle_dist2(ctypes.c_int(pts.shape[0]), _p(pts), _p(out))
    return out


def higher_msb(n):
    return int(lib().oracle_higher_msb(ctypes.c_uint32(n)))
