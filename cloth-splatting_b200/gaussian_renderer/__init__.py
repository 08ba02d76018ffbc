"""`gaussian_renderer.render`: the refinement step's forward glue (gaussian_renderer/__init__.py:22-206 of
KTH-RPL/cloth-splatting) over the sm_100a kernels:

    vertices  = simulator(time)                       ResidualMeshSimulator (mesh prediction + residual MLP)
    means3D, rotations = pc.get_xyz_rotation(vertices)   csb_mesh_transform_*  (get_xyz + get_rotation fused: barycentric
                                                          placement, closed-form Kabsch of the face, quaternion compose)
    image, radii, depth = GaussianRasterizer(...)     csb_rasterize_*

Same signature and the same 14-field RenderResults.  The `shadow_scalars` branch of the reference is dead code
(always None, :92); `pipe.compute_cov3D_python` needs GaussianModel.get_covariance and is not part of the loop
(arguments/__init__.py default False): it raises here instead of silently taking another path."""
import math
from typing import NamedTuple, Optional

import numpy as np
import torch

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


class RenderResults(NamedTuple):
    render: torch.Tensor
    viewspace_points: torch.Tensor
    visibility_filter: torch.Tensor
    radii: torch.Tensor
    depth: torch.Tensor
    means3D_deform: torch.Tensor
    vertice_deform: torch.Tensor
    shadows_mean: Optional[torch.Tensor]
    shadows_std: Optional[torch.Tensor]
    projections: torch.Tensor
    rotations: torch.Tensor
    opacities: torch.Tensor
    shadows: Optional[torch.Tensor]
    vertice_projections: Optional[torch.Tensor]


def project_to_pixels(cam, points):
    """Pixel coordinates of 3D points under cam.full_proj_transform (gaussian_renderer/__init__.py:166-180)."""
    ph = torch.cat([points, torch.ones_like(points[:, 0:1])], dim=1).T
    proj = cam.full_proj_transform.to(ph.device).T.matmul(ph)
    proj = (proj / proj[3, :])[:2].T
    H, W = int(cam.image_height), int(cam.image_width)
    return torch.stack([((proj[:, 0] + 1.0) * W - 1.0) * 0.5, ((proj[:, 1] + 1.0) * H - 1.0) * 0.5], dim=1)


def render(viewpoint_camera, pc, simulator, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None,
           log_deform_path=None, no_shadow=False, render_static=False, project_vertices=False,
           project_gaussians=True, precomputed=None) -> RenderResults:
    """Render one view of the mesh-anchored Gaussians at viewpoint_camera.time.  bg_color must be on the GPU.
    Extensions for callers that replay the step as a CUDA graph (train_utils.train_step): a camera may carry `time_tensor`
    (a [1, 1] device tensor holding its time: the value then is an input of the graph, not a constant baked into it), and
    `project_gaussians=False` skips the pixel projections the refinement loop never reads (`projections` is then None);
    `precomputed=(vertice_deform, means3D_deform, rotations_deform)` hands in this view's slice of a step-wide batched
    simulator + transform pass (train_utils._render_views) instead of running both per view."""
    if getattr(pipe, "compute_cov3D_python", False):
        raise NotImplementedError("compute_cov3D_python is outside the accelerated path (SURVEY.md section 8)")
    dev = pc.mesh.pos.device
    # the reference builds this zero tensor with zeros_like(pc.get_xyz()): only shape, dtype and device of the rest-pose
    # positions are used, so they are not computed unless the static branch or the deformation log needs them
    need_base = render_static or log_deform_path is not None
    base_xyz = pc.get_xyz() if need_base else None
    screenspace_points = torch.zeros((pc.face_ids.shape[0], 3), dtype=pc.face_bary.dtype, device=dev, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform.to(dev),
        projmatrix=viewpoint_camera.full_proj_transform.to(dev), sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center.to(dev), prefiltered=False, debug=False)
    rasterizer = GaussianRasterizer(raster_settings=settings)

    if render_static:
        vertice_deform = pc.mesh.pos
        means3D_deform = base_xyz
        rotations_deform = pc.get_rotation()
    elif precomputed is not None:
        vertice_deform, means3D_deform, rotations_deform = precomputed
    else:
        time = getattr(viewpoint_camera, "time_tensor", None)
        if time is None:
            time = torch.full((pc.mesh.pos.shape[0], 1), float(viewpoint_camera.time), device=dev)  # no host-to-device copy
        if hasattr(simulator, "check_time"):
            simulator.check_time(viewpoint_camera.time)  # on the host: no synchronisation inside the step
            vertice_deform = simulator(time_vector=time, checked=True)
        else:
            vertice_deform = simulator(time_vector=time)
        if callable(getattr(pc, "get_xyz_rotation", None)):  # one fused launch (csb_mesh_transform_*), same arithmetic
            means3D_deform, rotations_deform = pc.get_xyz_rotation(vertice_deform)
        else:
            means3D_deform = pc.get_xyz(vertice_deform)
            rotations_deform = pc.get_rotation(vertice_deform)
    opacity = pc.get_opacity
    if log_deform_path is not None:
        np.savez(log_deform_path, means3D=base_xyz.detach().cpu().numpy(),
                 means3D_deform=means3D_deform.detach().cpu().numpy(),
                 vertice_deform=vertice_deform.detach().cpu().numpy(), rotations=rotations_deform.detach().cpu().numpy())
    shs, colors_precomp = (pc.get_features, None) if override_color is None else (None, override_color)
    rendered_image, radii, depth = rasterizer(means3D=means3D_deform, means2D=screenspace_points, shs=shs,
                                              colors_precomp=colors_precomp, opacities=opacity,
                                              scales=pc.get_scaling, rotations=rotations_deform, cov3D_precomp=None)
    return RenderResults(
        render=rendered_image, viewspace_points=screenspace_points, visibility_filter=radii > 0, radii=radii,
        depth=depth, means3D_deform=means3D_deform, vertice_deform=vertice_deform, shadows_mean=None, shadows_std=None,
        projections=project_to_pixels(viewpoint_camera, means3D_deform) if project_gaussians else None,
        rotations=rotations_deform,
        opacities=opacity, shadows=None,
        vertice_projections=project_to_pixels(viewpoint_camera, vertice_deform) if project_vertices else None)
