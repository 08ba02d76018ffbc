"""`utils.loss_utils`: L1 and SSIM of the refinement step on the B200 kernels (csb_image_loss_*).

Drop-in for the reference's utils/loss_utils.py:20-70 (`l1_loss`, `l2_loss`, `ssim`) plus `image_loss`, the fused
form of scene_reconstruction/train_utils.py:50-74 (`image_losses`): ONE forward launch produces the loss values and
the per-pixel SSIM derivatives, ONE backward launch produces dL/dimage -- instead of five depthwise conv2d, ~15
elementwise kernels and their autograd graph.  Gradients flow to the rendered image only (the ground truth is data).
No CPU fallback: tensors must live on a CUDA device.
"""
import ctypes

import torch

import csb_cabi as _abi


def _bind(lib):
    if getattr(lib, "_loss_bound", False):
        return
    c_int, c_void_p, c_float, c_size_t = ctypes.c_int, ctypes.c_void_p, ctypes.c_float, ctypes.c_size_t
    lib.csb_image_loss_workspace_bytes.restype = c_size_t
    lib.csb_image_loss_workspace_bytes.argtypes = [c_int] * 4
    lib.csb_image_loss_forward.restype = c_int
    lib.csb_image_loss_forward.argtypes = [c_int] * 4 + [c_void_p] * 3 + [c_int, c_float] + [c_void_p] * 4 + [
        c_size_t, c_void_p]
    lib.csb_image_loss_backward.restype = c_int
    lib.csb_image_loss_backward.argtypes = [c_int] * 4 + [c_void_p] * 3 + [c_int] + [c_void_p] * 3 + [
        c_float, c_float, c_void_p, c_void_p]
    lib._loss_bound = True


def _as_planes(img):
    """[..., C, H, W] -> (n_img, C, H, W) view description of a contiguous fp32 tensor."""
    if img.dim() < 3:
        raise ValueError("expected an image tensor [..., C, H, W]")
    C, H, W = img.shape[-3:]
    n = 1
    for d in img.shape[:-3]:
        n *= d
    return n, C, H, W


def _prep_mask(mask, img):
    if mask is None:
        return None, 1
    n, C, H, W = _as_planes(img)
    m = mask.detach().to(device=img.device, dtype=torch.float32)
    while m.dim() < img.dim():
        m = m.unsqueeze(0)
    mc = m.shape[-3]
    if mc not in (1, C):
        raise ValueError(f"mask has {mc} channels, image {C}")
    lead = img.shape[:-3]
    m = m.expand(*lead, mc, H, W).contiguous()
    return m, mc


def _forward(img, gt, mask, lambda_dssim, want_map, want_grad):
    lib = _abi.lib()
    _bind(lib)
    if not img.is_cuda:
        raise ValueError("image losses (B200) need CUDA tensors; there is no CPU path")
    if img.shape != gt.shape:
        raise ValueError(f"image {tuple(img.shape)} and ground truth {tuple(gt.shape)} differ in shape")
    x = img.detach().float().contiguous()
    y = gt.detach().to(device=x.device, dtype=torch.float32).contiguous()
    m, mc = _prep_mask(mask, x)
    n, C, H, W = _as_planes(x)
    if n * C * H * W == 0:
        raise ValueError("image losses of an empty tensor")
    losses = torch.empty(3, dtype=torch.float32, device=x.device)
    smap = torch.empty_like(x) if want_map else None
    dmaps = torch.empty((3,) + tuple(x.shape), dtype=torch.float32, device=x.device) if want_grad else None
    ws_bytes = lib.csb_image_loss_workspace_bytes(n, C, H, W)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.csb_image_loss_forward(n, C, H, W, _abi.ptr(x), _abi.ptr(y), _abi.ptr(m), mc, float(lambda_dssim),
                                        _abi.ptr(smap), _abi.ptr(dmaps), _abi.ptr(losses), _abi.ptr(ws), ws_bytes,
                                        _abi.current_stream_ptr())
    _abi.check(rc, "csb_image_loss_forward")
    return x, y, m, mc, losses, smap, dmaps


def _backward(x, y, m, mc, dmaps, grad_map, grad_scale, w_ssim, w_l1):
    lib = _abi.lib()
    n, C, H, W = _as_planes(x)
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        rc = lib.csb_image_loss_backward(n, C, H, W, _abi.ptr(x), _abi.ptr(y), _abi.ptr(m), mc, _abi.ptr(dmaps),
                                         _abi.ptr(grad_map), _abi.ptr(grad_scale), float(w_ssim), float(w_l1),
                                         _abi.ptr(out), _abi.current_stream_ptr())
    _abi.check(rc, "csb_image_loss_backward")
    return out


class _ImageLoss(torch.autograd.Function):
    """(loss, l1, ssim_loss) of image_losses(); only `loss` carries a gradient (the reference reads the other two
    with .item() for logging)."""

    @staticmethod
    def forward(ctx, img, gt, mask, lambda_dssim):
        need = img.requires_grad
        x, y, m, mc, losses, _, dmaps = _forward(img, gt, mask, lambda_dssim, False, need)
        ctx.lambda_dssim = float(lambda_dssim)
        ctx.mc = mc
        ctx.in_dtype, ctx.in_shape = img.dtype, img.shape
        if need:
            ctx.save_for_backward(x, y, m, dmaps) if m is not None else ctx.save_for_backward(x, y, dmaps)
        ctx.has_mask = m is not None
        loss, l1, ds = losses[0], losses[1], losses[2]
        ctx.mark_non_differentiable(l1, ds)
        return loss, l1, ds

    @staticmethod
    def backward(ctx, g_loss, _g_l1, _g_ds):
        if ctx.has_mask:
            x, y, m, dmaps = ctx.saved_tensors
        else:
            (x, y, dmaps), m = ctx.saved_tensors, None
        count = float(x.numel())
        scale = g_loss.detach().float().reshape(1).contiguous()
        g = _backward(x, y, m, ctx.mc, dmaps, None, scale, -ctx.lambda_dssim / count, 1.0 / count)
        return g.view(ctx.in_shape).to(ctx.in_dtype), None, None, None


class _SsimMap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, gt):
        need = img.requires_grad
        x, y, _, _, _, smap, dmaps = _forward(img, gt, None, 0.0, True, need)
        ctx.in_dtype, ctx.in_shape = img.dtype, img.shape
        if need:
            ctx.save_for_backward(x, y, dmaps)
        return smap.view(img.shape)

    @staticmethod
    def backward(ctx, g_map):
        x, y, dmaps = ctx.saved_tensors
        gm = g_map.detach().float().expand(ctx.in_shape).contiguous().view(x.shape)
        g = _backward(x, y, None, 1, dmaps, gm, None, 0.0, 0.0)
        return g.view(ctx.in_shape).to(ctx.in_dtype), None


def image_loss(image, gt, lambda_dssim, mask=None):
    """Fused image_losses() (train_utils.py:50-74): returns device scalars (loss, l1, ssim_loss) with
    loss = l1 + lambda_dssim * ssim_loss; no host synchronisation."""
    return _ImageLoss.apply(image, gt, mask, float(lambda_dssim))


def l1_loss(network_output, gt, mask=None):
    """utils/loss_utils.py:20-23."""
    return _ImageLoss.apply(network_output, gt, mask, 0.0)[0]


def l2_loss(network_output, gt):
    """utils/loss_utils.py:25-26 (not on the hot path: plain torch on the device)."""
    return ((network_output - gt) ** 2).mean()


def ssim(img1, img2, window_size=11, size_average=True, return_map=False):
    """utils/loss_utils.py:39-70.  The window is the reference's default (11 taps, sigma 1.5); other sizes are
    not part of the hot path and raise."""
    if window_size != 11:
        raise NotImplementedError("ssim (B200): only the reference's default window_size=11 is built")
    if img2.requires_grad:
        raise NotImplementedError("ssim (B200): gradients flow to img1 only (img2 is the ground truth)")
    smap = _SsimMap.apply(img1, img2)
    if return_map:
        return smap
    if size_average:
        return smap.mean()
    return smap.mean(1).mean(1).mean(1)
