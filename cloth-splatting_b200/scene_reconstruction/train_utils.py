"""`scene_reconstruction.train_utils.image_losses` on the fused B200 loss kernels.

Mirrors scene_reconstruction/train_utils.py:50-74 of the reference: same arguments, same `(loss, loss_dict)` return,
same keys ('l1', 'ssim_loss').  The reference reads both entries with `.item()` (two host synchronisations per
step); `sync=False` keeps them as device scalars instead."""
import torch

from utils.image_utils import psnr
from utils.loss_utils import image_loss


def image_losses(image_tensor, gt_image_tensor, opt, mask_tensor=None, sync=True):
    lam = float(opt if isinstance(opt, (int, float)) else opt.lambda_dssim)
    loss, l1, ssim_loss = image_loss(image_tensor, gt_image_tensor, lam, mask_tensor)
    loss_dict = {"l1": l1.item() if sync else l1}
    if lam != 0:
        loss_dict["ssim_loss"] = ssim_loss.item() if sync else ssim_loss
    return loss, loss_dict


def densification(gaussians, iteration, visibility_filter, radii, viewspace_point_tensor_grad, opt, cameras_extent):
    """The densification schedule of the refinement loop (scene_reconstruction/train_utils.py:324-345): per-Gaussian
    statistics every iteration; clone / split every `densification_interval` and prune every `pruning_interval`
    iterations with thresholds that fall linearly from their `*_init` to their `*_after` value over
    `densify_until_iter` iterations; the screen-size criterion (20 px) only after the first opacity reset."""
    # same values as the reference's boolean-mask assignment, without the host synchronisation of mask indexing
    gaussians.max_radii2D = torch.where(visibility_filter, torch.max(gaussians.max_radii2D, radii), gaussians.max_radii2D)
    gaussians.add_densification_stats(viewspace_point_tensor_grad, visibility_filter)

    def ramp(init, after):
        return init - iteration * (init - after) / opt.densify_until_iter

    opacity_threshold = ramp(opt.opacity_threshold_fine_init, opt.opacity_threshold_fine_after)
    densify_threshold = ramp(opt.densify_grad_threshold_fine_init, opt.densify_grad_threshold_after)
    size_threshold = 20 if iteration > opt.opacity_reset_interval else None
    if iteration > opt.densify_from_iter and iteration % opt.densification_interval == 0:
        gaussians.densify(densify_threshold, opacity_threshold, cameras_extent, size_threshold)
    if iteration > opt.pruning_from_iter and iteration % opt.pruning_interval == 0:
        gaussians.prune(densify_threshold, opacity_threshold, cameras_extent, size_threshold)


def regularization(all_vertice_deform, gaussians, opt, static=False):
    """The mesh regularisers added to the photometric loss (scene_reconstruction/train_utils.py:77-102, 238), over the
    deformed vertices of the views of a step `[n_cams, N, 3]` (consecutive time steps): mean deformation magnitude
    between neighbouring views, L1 between rest and deformed edge lengths, and an L1 second-difference (momentum)
    term.  The first and the last need at least three views."""
    n_cams = all_vertice_deform.shape[0]
    loss = torch.zeros([], device=all_vertice_deform.device)
    if static:
        return loss
    if opt.lambda_deform_mag > 0. and n_cams >= 3:
        mag0 = torch.linalg.norm(all_vertice_deform[1] - all_vertice_deform[0], dim=-1).mean()
        mag1 = torch.linalg.norm(all_vertice_deform[2] - all_vertice_deform[1], dim=-1).mean()
        loss += opt.lambda_deform_mag * (0.5 * (mag0 + mag1))
    if opt.lambda_rigid > 0:
        ei = gaussians.mesh.edge_index
        deformed_norm = torch.linalg.norm(all_vertice_deform[:, ei[1]] - all_vertice_deform[:, ei[0]], dim=-1, keepdim=True)
        static_norm = gaussians.edge_norm.unsqueeze(0).expand(n_cams, -1, -1)
        loss += opt.lambda_rigid * torch.nn.functional.l1_loss(static_norm, deformed_norm)
    if opt.lambda_momentum > 0 and n_cams >= 3:
        second = all_vertice_deform[2] - 2 * all_vertice_deform[1] + all_vertice_deform[0]
        loss += opt.lambda_momentum * torch.linalg.norm(second, dim=-1, ord=1).mean().mean()
    return loss


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """Log-linear learning-rate decay with an optional eased-in start (utils/general_utils.py:32-65)."""
    import numpy as np

    def helper(step):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        delay = 1.0
        if lr_delay_steps > 0:
            delay = lr_delay_mult + (1 - lr_delay_mult) * np.sin(0.5 * np.pi * np.clip(step / lr_delay_steps, 0, 1))
        t = np.clip(step / max_steps, 0, 1)
        return delay * np.exp(np.log(lr_init) * (1 - t) + np.log(lr_final) * t)

    return helper


def train_step(iteration, viewpoint_cams, gaussians, simulator, meshnet_optimizer, pipeline_params, opt_params,
               cameras_extent, background, static=False, white_background=False, user_args=None):
    """One iteration of the refinement loop (scene_reconstruction/train_utils.py:240-321), same arguments and the same
    `(psnr, loss, loss_dict)` return: render every view of the step at its time (`gaussian_renderer.render`: simulator ->
    mesh-anchored Gaussians -> rasterizer), image losses over the stacked views + the mesh regularisers, one backward, the
    densification / opacity-reset / barycentric-cleanup schedule on the summed screen-space gradients, then the Adam
    steps of the Gaussians and (unless `static`) of the simulator."""
    from gaussian_renderer import render

    gaussians.update_learning_rate(iteration)
    if iteration % 1000 == 0:
        gaussians.oneupSHdegree()
    dev = background.device
    no_shadow = getattr(user_args, "no_shadow", False)
    pkgs = [render(cam, gaussians, simulator, pipeline_params, background, no_shadow=no_shadow, render_static=static)
            for cam in viewpoint_cams]
    image_tensor = torch.stack([p.render for p in pkgs])
    gt_image_tensor = torch.stack([cam.original_image.to(dev, non_blocking=True) for cam in viewpoint_cams])
    mask_tensor = None
    if viewpoint_cams[0].mask is not None:
        mask_tensor = torch.stack([cam.mask.to(dev, non_blocking=True) for cam in viewpoint_cams])
    all_vertice_deform = torch.stack([p.vertice_deform for p in pkgs])
    radii = torch.stack([p.radii for p in pkgs]).max(dim=0).values
    visibility_filter = torch.stack([p.visibility_filter for p in pkgs]).any(dim=0)

    psnr_ = psnr(image_tensor, gt_image_tensor).mean().double()
    loss, loss_dict = image_losses(image_tensor, gt_image_tensor, opt_params, mask_tensor)
    loss = loss + regularization(all_vertice_deform, gaussians, opt_params, static)
    loss.backward()

    viewspace_point_tensor_grad = torch.zeros_like(pkgs[0].viewspace_points)
    for p in pkgs:
        viewspace_point_tensor_grad = viewspace_point_tensor_grad + p.viewspace_points.grad
    with torch.no_grad():
        if iteration < opt_params.densify_until_iter:
            densification(gaussians, iteration, visibility_filter, radii, viewspace_point_tensor_grad, opt_params,
                          cameras_extent)
            if iteration % opt_params.opacity_reset_interval == 0 or (
                    white_background and iteration == opt_params.densify_from_iter):
                gaussians.reset_opacity()
        if iteration % opt_params.bary_cleanup == 0:
            gaussians.cleanup_barycentric_coordinates()
        if iteration < opt_params.iterations:
            gaussians.optimizer.step()
            if not static:
                meshnet_optimizer.step()
            gaussians.optimizer.zero_grad(set_to_none=True)
            meshnet_optimizer.zero_grad()
    return psnr_, loss, loss_dict
