"""`utils.image_utils` of the reference (utils/image_utils.py:14-19): per-image mean squared error and PSNR, read by
`train_step` for logging.  Plain device torch, no gradient."""
import torch


def mse(img1, img2):
    return ((img1 - img2) ** 2).view(img1.shape[0], -1).mean(1, keepdim=True)


@torch.no_grad()
def psnr(img1, img2):
    return 20 * torch.log10(1.0 / torch.sqrt(mse(img1, img2)))
