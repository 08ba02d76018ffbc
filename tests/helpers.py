"""Shared helpers of the parity tests."""
import ctypes

import numpy as np
import torch

import csb_synth as S

GRAD_NAMES = ["means2D", "colors", "opacities", "means3D", "cov3D", "shs", "scales", "rotations"]


def to_cuda(g):
    return {k: torch.from_numpy(v).cuda() for k, v in g.items() if isinstance(v, np.ndarray)}


def cam_tensors(cam, bg=(1.0, 1.0, 1.0)):
    return dict(
        viewmatrix=torch.from_numpy(cam.world_view_transform).cuda(),
        projmatrix=torch.from_numpy(cam.full_proj_transform).cuda(),
        campos=torch.from_numpy(cam.camera_center).cuda(),
        bg=torch.tensor(bg, dtype=torch.float32).cuda(),
    )


def view_tuple(cam, bg=(1.0, 1.0, 1.0)):
    t = cam_tensors(cam, bg)
    return (t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"], cam.tanfovx, cam.tanfovy)


def settings_for(cam, sh_degree=3, bg=(1.0, 1.0, 1.0), scale_modifier=1.0):
    from diff_gaussian_rasterization import GaussianRasterizationSettings

    t = cam_tensors(cam, bg)
    return GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=t["bg"], scale_modifier=scale_modifier, viewmatrix=t["viewmatrix"], projmatrix=t["projmatrix"],
        sh_degree=sh_degree, campos=t["campos"], prefiltered=False, debug=False)


def small_scene(P=3000, W=160, H=112, seed=11):
    """A small random cloud in front of one camera: finishes in well under a second on the CPU oracle."""
    g = S.random_gaussians(P, seed, box=((-1.2, 1.2), (-0.9, 0.9), (-0.5, 5.0)), log_scale=(0.01, 0.15))
    cam = S.front_camera(W, H, 1.0)
    return g, cam


def mesh_scene(res=208, view=0, n_views=4):
    """The C2 cloth (10 108 mesh-anchored Gaussians) seen by one ring camera at a reduced resolution."""
    g = S.mesh_gaussians(20, 14, 1)
    center = g["verts"].mean(0)
    cams = S.ring_cameras(n_views, res, res, 0.6911, center)
    return g, cams[view]


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64).reshape(a.shape)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def oracle_forward(g, cam, bg=(1.0, 1.0, 1.0), sh_degree=3, scale_modifier=1.0):
    import oracle

    return oracle.rasterize_forward(
        g["means3D"], g["opacities"], cam.world_view_transform, cam.full_proj_transform, cam.camera_center,
        np.asarray(bg, np.float32), cam.tanfovx, cam.tanfovy, cam.image_height, cam.image_width, shs=g.get("shs"),
        colors_precomp=g.get("colors_precomp"), scales=g.get("scales"), rotations=g.get("rotations"),
        cov3D_precomp=g.get("cov3D_precomp"), sh_degree=sh_degree, scale_modifier=scale_modifier)


def product_forward(gc, cams, sh_degree=3, bg=(1.0, 1.0, 1.0), flags=1, scale_modifier=1.0):
    """Forward through the package's C-ABI wrapper for one or more views.
    Returns (views, (R, color, depth, radii, geom, binning, img))."""
    import diff_gaussian_rasterization as dgr

    cams = cams if isinstance(cams, (list, tuple)) else [cams]
    views = [view_tuple(c, bg) for c in cams]
    e = torch.Tensor([])
    out = dgr._forward_views(
        views, gc["means3D"], gc.get("colors_precomp", e), gc["opacities"], gc.get("scales", e), gc.get("rotations", e),
        scale_modifier, gc.get("cov3D_precomp", e), cams[0].image_height, cams[0].image_width, gc.get("shs", e),
        sh_degree, False, flags=flags)
    return views, out


def product_backward(gc, cams, views, fwd, dl, sh_degree=3, scale_modifier=1.0):
    import diff_gaussian_rasterization as dgr

    cams = cams if isinstance(cams, (list, tuple)) else [cams]
    R, color, depth, radii, geom, binning, img = fwd
    e = torch.Tensor([])
    gb = dgr._backward_views(
        views, gc["means3D"], radii, gc.get("colors_precomp", e), gc.get("scales", e), gc.get("rotations", e),
        scale_modifier, gc.get("cov3D_precomp", e), cams[0].image_height, cams[0].image_width, dl, gc.get("shs", e),
        sh_degree, geom, binning, img)
    return dict(zip(GRAD_NAMES, gb))


def product_state(name, dtype, shape, V, P, W, H, fwd):
    """Copy of one carved workspace array of a forward call (csb_rasterize_state_ptr)."""
    import csb_cabi

    R, color, depth, radii, geom, binning, img = fwd
    lib = csb_cabi.lib()
    cap = lib.csb_rasterize_binning_capacity(V, P, W, H, binning.numel())
    ptr = lib.csb_rasterize_state_ptr(name.encode(), V, P, W, H, cap, geom.data_ptr(), img.data_ptr(),
                                      binning.data_ptr())
    assert ptr, name
    n = int(np.prod(shape))
    out = torch.empty(shape, dtype=dtype, device="cuda")
    if n:
        torch.cuda.synchronize()
        rt = ctypes.CDLL("libcudart.so.12")
        rc = rt.cudaMemcpy(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(ptr),
                           ctypes.c_size_t(n * out.element_size()), ctypes.c_int(3))
        assert rc == 0, rc
    return out


def same_as_golden(a, g, rtol=4e-6, atol=1e-7):
    """Comparison with a golden the reference's own function produced on the build container's host: integers, booleans and
    indices bit for bit; floating point to a few ulp, because torch's CPU kernels (exp / log / rsqrt, LAPACK) round the last
    bit differently on hosts with different vector units (the goldens were written on an AVX-512 machine)."""
    import numpy as np

    a, g = np.asarray(a), np.asarray(g)
    if a.shape != g.shape:
        return False
    if a.dtype.kind in "fc" or g.dtype.kind in "fc":
        return bool(np.allclose(a, g, rtol=rtol, atol=atol, equal_nan=True))
    return bool(np.array_equal(a, g))
