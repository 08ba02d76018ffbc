"""`scene_reconstruction.train_utils.image_losses` on the fused B200 loss kernels.

Mirrors scene_reconstruction/train_utils.py:50-74 of the reference: same arguments, same `(loss, loss_dict)` return,
same keys ('l1', 'ssim_loss').  The reference reads both entries with `.item()` (two host synchronisations per
step); `sync=False` keeps them as device scalars instead."""
import torch

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
    gaussians.max_radii2D[visibility_filter] = torch.max(gaussians.max_radii2D[visibility_filter], radii[visibility_filter])
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
