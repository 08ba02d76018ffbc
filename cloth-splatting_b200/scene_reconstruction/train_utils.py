"""`scene_reconstruction.train_utils.image_losses` on the fused B200 loss kernels.

Mirrors scene_reconstruction/train_utils.py:50-74 of the reference: same arguments, same `(loss, loss_dict)` return,
same keys ('l1', 'ssim_loss').  The reference reads both entries with `.item()` (two host synchronisations per
step); `sync=False` keeps them as device scalars instead."""
from utils.loss_utils import image_loss


def image_losses(image_tensor, gt_image_tensor, opt, mask_tensor=None, sync=True):
    lam = float(opt if isinstance(opt, (int, float)) else opt.lambda_dssim)
    loss, l1, ssim_loss = image_loss(image_tensor, gt_image_tensor, lam, mask_tensor)
    loss_dict = {"l1": l1.item() if sync else l1}
    if lam != 0:
        loss_dict["ssim_loss"] = ssim_loss.item() if sync else ssim_loss
    return loss, loss_dict
