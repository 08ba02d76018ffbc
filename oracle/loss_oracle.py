"""TEST INFRASTRUCTURE (never imported by the product): CPU restatement of the photometric loss.

Follows the reference operation by operation in torch on the CPU:
  l1_loss       utils/loss_utils.py:20-23
  gaussian / create_window / ssim / _ssim   utils/loss_utils.py:28-70   (2-D 11x11 window, conv2d with zero padding)
  image_losses  scene_reconstruction/train_utils.py:50-74               (loss = l1 + lambda_dssim * ssim_loss)
Gradients come from torch autograd over this restatement (what loss.backward() does in the reference).
Pinned by tests/golden/loss_*.npz, generated from the reference's own functions by tests/make_golden_loss.py."""
from math import exp

import torch
import torch.nn.functional as F


def gaussian(window_size, sigma):
    g = torch.tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)],
                     dtype=torch.float32)
    return g / g.sum()


def create_window(window_size, channel, dtype=torch.float32):
    w1 = gaussian(window_size, 1.5).unsqueeze(1)
    w2 = w1.mm(w1.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, window_size, window_size).contiguous().to(dtype)


def ssim_map(img1, img2, window_size=11):
    channel = img1.size(-3)
    window = create_window(window_size, channel, img1.dtype).to(img1.device)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))


def l1_loss(out, gt, mask=None):
    if mask is not None:
        return torch.abs((out - gt) * mask).mean()
    return torch.abs(out - gt).mean()


def image_losses(img, gt, lambda_dssim, mask=None):
    """-> (loss, l1, ssim_loss) as 0-d tensors (ssim_loss = 0 when lambda_dssim == 0, like the skipped branch)."""
    l1 = l1_loss(img, gt, mask)
    loss = l1
    ssim_loss = torch.zeros((), dtype=img.dtype)
    if lambda_dssim != 0:
        if mask is None:
            ssim_loss = 1.0 - ssim_map(img, gt).mean()
        else:
            ssim_loss = ((1.0 - ssim_map(img, gt)) * mask).mean()
        loss = loss + lambda_dssim * ssim_loss
    return loss, l1, ssim_loss


def image_losses_with_grad(img, gt, lambda_dssim, mask=None, dtype=torch.float32):
    """numpy in, numpy out: (loss, l1, ssim_loss, dL/dimg) evaluated in `dtype` on the CPU."""
    import numpy as np

    x = torch.from_numpy(np.ascontiguousarray(img)).to(dtype).requires_grad_(True)
    y = torch.from_numpy(np.ascontiguousarray(gt)).to(dtype)
    m = None if mask is None else torch.from_numpy(np.ascontiguousarray(mask)).to(dtype)
    loss, l1, ds = image_losses(x, y, lambda_dssim, m)
    loss.backward()
    return float(loss.detach()), float(l1.detach()), float(ds.detach()), x.grad.numpy()
