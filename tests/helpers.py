"""Shared helpers of the parity tests."""
import numpy as np
import torch

import csb_synth as S


def to_cuda(g):
    return {k: torch.from_numpy(v).cuda() for k, v in g.items() if isinstance(v, np.ndarray)}


def cam_tensors(cam, bg=(1.0, 1.0, 1.0)):
    return dict(
        viewmatrix=torch.from_numpy(cam.world_view_transform).cuda(),
        projmatrix=torch.from_numpy(cam.full_proj_transform).cuda(),
        campos=torch.from_numpy(cam.camera_center).cuda(),
        bg=torch.tensor(bg, dtype=torch.float32).cuda(),
    )


def settings_for(cam, sh_degree=3, bg=(1.0, 1.0, 1.0), scale_modifier=1.0):
    from diff_gaussian_rasterization import GaussianRasterizationSettings

    t = cam_tensors(cam, bg)
    return GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=t["bg"], scale_modifier=scale_modifier, viewmatrix=t["viewmatrix"], projmatrix=t["projmatrix"],
        sh_degree=sh_degree, campos=t["campos"], prefiltered=False, debug=False)


def small_scene(P=3000, W=160, H=112, seed=11, sh_degree=3):
    """A small random cloud in front of one camera: finishes in well under a second on the CPU oracle."""
    g = S.random_gaussians(P, seed, box=((-1.2, 1.2), (-0.9, 0.9), (-0.5, 5.0)), log_scale=(0.01, 0.15))
    cam = S.front_camera(W, H, 1.0)
    return g, cam


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
