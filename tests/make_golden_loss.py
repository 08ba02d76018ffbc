"""Generates tests/golden/loss_*.npz from the REFERENCE's own loss functions, run in the build container on the CPU
(needs /root/reference):

    python tests/make_golden_loss.py

utils/loss_utils.py cannot be imported as a module (its first lines import `lpips`, which is not installed), so the
`def` blocks of l1_loss, gaussian, create_window, ssim and _ssim are compiled at run time from the reference file's
own lines -- nothing is copied into this repository.  image_losses (scene_reconstruction/train_utils.py:50-74) is
lifted the same way (its module imports wandb / lpips / the rasterizer).  Outputs: loss, l1, ssim_loss, the SSIM map
and dL/dimage from torch autograd, all fp32 on the CPU, plus an fp64 evaluation of the same functions that the tests
use to tell rounding noise from error."""
import ast
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"


def lift(path, names, glob):
    src = open(path).read()
    tree = ast.parse(src)
    lines = src.splitlines()
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            block = "\n".join(lines[node.lineno - 1:node.end_lineno])
            exec(compile(block, f"{path}:{node.lineno}", "exec"), glob)
    missing = [n for n in names if n not in glob]
    assert not missing, missing
    return glob


def reference_functions():
    import torch.nn.functional as F
    from math import exp
    from torch.autograd import Variable

    g = dict(torch=torch, F=F, exp=exp, Variable=Variable)
    lift(os.path.join(REF, "utils", "loss_utils.py"), ["l1_loss", "gaussian", "create_window", "ssim", "_ssim"], g)

    class OptimizationParams:  # annotation in image_losses' signature
        pass

    g["OptimizationParams"] = OptimizationParams
    lift(os.path.join(REF, "scene_reconstruction", "train_utils.py"), ["image_losses"], g)
    return g


def synth(seed, n_img, H, W, mask_channels):
    """Rendered image = ground truth + structured error + noise, in [0,1]; flat regions included (they stress the
    sigma cancellation)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.linspace(0, 1, H), np.linspace(0, 1, W), indexing="ij")
    gt = np.empty((n_img, 3, H, W), np.float32)
    for i in range(n_img):
        for c in range(3):
            f = 2 + i + c
            gt[i, c] = 0.5 + 0.4 * np.sin(2 * np.pi * f * xx + c) * np.cos(2 * np.pi * (f - 1) * yy + i)
    gt[:, :, : H // 4, : W // 3] = 1.0  # white background block
    img = gt + 0.05 * rng.standard_normal(gt.shape).astype(np.float32)
    img += 0.1 * np.sin(2 * np.pi * 3 * xx)[None, None].astype(np.float32) * (rng.random((n_img, 3, 1, 1)) > 0.5)
    img = np.clip(img, 0, 1).astype(np.float32)
    img[:, :, : H // 8, : W // 6] = 1.0  # exactly equal, flat: |x - y| = 0 and sigma = 0
    mask = None
    if mask_channels:
        mask = (rng.random((n_img, mask_channels, H, W)) > 0.3).astype(np.float32)
        mask[:, :, H // 2:, : W // 2] *= 0.5  # soft values too
    return img, gt, mask


CASES = [  # name, seed, n_img, H, W, mask_channels, lambda_dssim
    ("loss_2x3x40x52_l01", 11, 2, 40, 52, 0, 0.1),
    ("loss_1x3x67x33_mask1_l02", 12, 1, 67, 33, 1, 0.2),
    ("loss_2x3x32x64_mask3_l01", 13, 2, 32, 64, 3, 0.1),
    ("loss_1x3x9x7_l01", 14, 1, 9, 7, 0, 0.1),  # smaller than the window
]


def run(g, img, gt, mask, lam, dtype):
    class Opt:
        lambda_dssim = lam

    x = torch.from_numpy(img).to(dtype).requires_grad_(True)
    y = torch.from_numpy(gt).to(dtype)
    m = None if mask is None else torch.from_numpy(mask).to(dtype)
    loss, d = g["image_losses"](x, y, Opt(), m)
    loss.backward()
    smap = g["ssim"](x.detach(), y, return_map=True)
    return dict(loss=float(loss), l1=d["l1"], ssim_loss=d.get("ssim_loss", 0.0), grad=x.grad.numpy(),
                ssim_map=smap.numpy())


def main():
    torch.manual_seed(0)
    torch.set_num_threads(1)
    g = reference_functions()
    os.makedirs(os.path.join(HERE, "golden"), exist_ok=True)
    for name, seed, n, H, W, mc, lam in CASES:
        img, gt, mask = synth(seed, n, H, W, mc)
        r32 = run(g, img, gt, mask, lam, torch.float32)
        r64 = run(g, img, gt, mask, lam, torch.float64)
        out = dict(ref_commit="KTH-RPL/cloth-splatting f7acc90: utils/loss_utils.py:20-70, "
                              "scene_reconstruction/train_utils.py:50-74", img=img, gt=gt, lambda_dssim=np.float64(lam), loss=np.float64(r32["loss"]), l1=np.float64(r32["l1"]),
                   ssim_loss=np.float64(r32["ssim_loss"]), grad=r32["grad"].astype(np.float32),
                   ssim_map=r32["ssim_map"].astype(np.float32), loss64=np.float64(r64["loss"]),
                   grad64=r64["grad"].astype(np.float64), ssim_map64=r64["ssim_map"].astype(np.float64))
        if mask is not None:
            out["mask"] = mask
        np.savez_compressed(os.path.join(HERE, "golden", name + ".npz"), **out)
        gerr = np.linalg.norm(r32["grad"] - r64["grad"]) / np.linalg.norm(r64["grad"])
        print(name, "loss", r32["loss"], "fp32-vs-fp64: loss", abs(r32["loss"] - r64["loss"]), "grad rel", gerr)


if __name__ == "__main__":
    main()
