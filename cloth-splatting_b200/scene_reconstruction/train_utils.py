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


_side_streams = {}


def _view_streams(like, n):
    """n side streams on the device of `like` (kept for the life of the process: a captured step must find the same
    streams at every capture); none on the CPU or when CSB_VIEW_STREAMS=0."""
    import os

    if n <= 0 or not like.is_cuda or os.environ.get("CSB_VIEW_STREAMS", "1") == "0":
        return []
    have = _side_streams.setdefault(like.device, [])
    while len(have) < n:
        have.append(torch.cuda.Stream(device=like.device))
    return have[:n]


def _render_views(views, gaussians, simulator, pipeline_params, background, static, user_args, project=True):
    """First half of an iteration (train_utils.py:256-272): every view rendered at its time."""
    from gaussian_renderer import render

    no_shadow = getattr(user_args, "no_shadow", False)
    extra = {} if project else {"project_gaussians": False}  # the eager path calls render() with the reference's arguments
    import contextlib

    share = gaussians.shared_activations() if hasattr(gaussians, "shared_activations") else contextlib.nullcontext()
    pre, batched_vertices = [None] * len(views), None
    if (not static and len(views) > 1 and hasattr(simulator, "forward_times") and hasattr(simulator, "check_time")
            and callable(getattr(gaussians, "get_xyz_rotation", None))):
        # one pass of the residual simulator and ONE fused transform launch for all views of the step
        for cam in views:
            simulator.check_time(cam.time)
        tt = [getattr(cam, "time_tensor", None) for cam in views]
        if all(t is not None for t in tt):
            times = torch.cat([t.reshape(1) for t in tt])  # device values: inputs of a captured step, not constants
        else:
            times = torch.tensor([float(cam.time) for cam in views], dtype=torch.float32, device=gaussians.mesh.pos.device)
        vertices = simulator.forward_times(times)
        xyz, rot = gaussians.get_xyz_rotation(vertices)
        # unbind: ONE stack in the backward pass per tensor instead of a zero-fill + add per view and tensor
        pre = list(zip(vertices.unbind(0), xyz.unbind(0), rot.unbind(0)))
        batched_vertices = vertices
    def one(v, cam):
        return render(cam, gaussians, simulator, pipeline_params, background, no_shadow=no_shadow, render_static=static,
                      **({} if pre[v] is None else {"precomputed": pre[v]}), **extra)

    with share:  # exp / sigmoid / cat of the view-independent parameters once per step instead of once per view
        side = _view_streams(background, len(views) - 1) if batched_vertices is not None else []
        if side:
            # Every view has its own time, hence its own Gaussian positions and rotations: its rasterization is its own
            # launch set, a chain of short kernels (preprocess, scan, fill, sort, forward -- and their mirror image in the
            # backward pass) that each leave most of the GPU idle.  The chains of the views share nothing but their
            # inputs, so views 1.. run on side streams beside view 0 (fork after everything they share has been enqueued,
            # join before the images are stacked); autograd runs each view's backward on the stream of its forward.
            # Inside the captured step the fork and the join become graph branches.
            gaussians.get_opacity, gaussians.get_scaling, gaussians.get_features
            main = torch.cuda.current_stream(background.device)
            for st in side:
                st.wait_stream(main)
            pkgs = [None] * len(views)
            for v in range(1, len(views)):
                with torch.cuda.stream(side[v - 1]):
                    pkgs[v] = one(v, views[v])
            pkgs[0] = one(0, views[0])
            for st in side:
                main.wait_stream(st)
        else:
            pkgs = [one(v, cam) for v, cam in enumerate(views)]
    image_tensor = torch.stack([p.render for p in pkgs])
    all_vertice_deform = batched_vertices if batched_vertices is not None else torch.stack([p.vertice_deform for p in pkgs])
    radii = torch.stack([p.radii for p in pkgs]).max(dim=0).values
    visibility_filter = torch.stack([p.visibility_filter for p in pkgs]).any(dim=0)
    return pkgs, image_tensor, all_vertice_deform, radii, visibility_filter


def _loss_and_step(rendered, gt_image_tensor, mask_tensor, iteration_stats, gaussians, meshnet_optimizer, opt_params,
                   static, take_steps, sync):
    """Second half (train_utils.py:274-321 minus the densification / opacity-reset / clean-up schedule): image losses over
    the stacked views + the mesh regularisers, one backward, the densification statistics, both Adam steps.  No host
    synchronisation when `sync` is False (the loss_dict entries then are device scalars)."""
    pkgs, image_tensor, all_vertice_deform, radii, visibility_filter = rendered
    psnr_ = psnr(image_tensor, gt_image_tensor).mean().double()
    if sync:  # the reference's call (the CPU tests put their own image_losses in its place)
        loss, loss_dict = image_losses(image_tensor, gt_image_tensor, opt_params, mask_tensor)
    else:
        loss, loss_dict = image_losses(image_tensor, gt_image_tensor, opt_params, mask_tensor, sync=False)
    loss = loss + regularization(all_vertice_deform, gaussians, opt_params, static)
    loss.backward()

    viewspace_point_tensor_grad = torch.zeros_like(pkgs[0].viewspace_points)
    for p in pkgs:
        viewspace_point_tensor_grad = viewspace_point_tensor_grad + p.viewspace_points.grad
    with torch.no_grad():
        if iteration_stats:
            # per-Gaussian statistics of `densification` (train_utils.py:324-331), in place: the tensors keep their addresses
            gaussians.max_radii2D.copy_(torch.where(visibility_filter, torch.max(gaussians.max_radii2D, radii),
                                                    gaussians.max_radii2D))
            gaussians.add_densification_stats(viewspace_point_tensor_grad, visibility_filter)
        if take_steps:
            gaussians.optimizer.step()
            if not static:
                meshnet_optimizer.step()
            gaussians.optimizer.zero_grad(set_to_none=True)
            meshnet_optimizer.zero_grad()
    return psnr_, loss, loss_dict, viewspace_point_tensor_grad, visibility_filter, radii


class ResidentPlanes:
    """Ground-truth images and masks kept on the device after their first use.  The reference uploads `cam.original_image`
    (3 x 800 x 800 floats = 7.7 MB) from the host in every iteration of every view (`viewpoint_cam.original_image.cuda()`,
    train_utils.py:262); a scene's views are a fixed set that fits HBM many times over (180 GB: ~23 000 such planes), so a
    view's planes cross PCIe once and every later iteration that draws the view reads them from HBM.  LRU over
    `CSB_GT_CACHE_MB` megabytes (default 8192; 0 switches it off); an entry is keyed on the host tensor OBJECT and its
    version counter, so an image edited in place is uploaded again."""

    def __init__(self):
        import collections
        import os

        self.budget = int(float(os.environ.get("CSB_GT_CACHE_MB", "8192")) * (1 << 20))
        self.used = 0
        self.entries = collections.OrderedDict()  # id(host tensor) -> (weakref, version, device tensor)
        self.hits = self.misses = 0

    def fetch(self, host, device, into=None):
        """The device copy of `host` (a CPU tensor); written into `into` when given (a static buffer of a captured step)."""
        import weakref

        if host.is_cuda or self.budget <= 0:
            src = host if host.is_cuda and into is not None else host.to(device, non_blocking=True)
            return src if into is None else into.copy_(src, non_blocking=True)
        key = id(host)
        ent = self.entries.get(key)
        if ent is not None and ent[0]() is host and ent[1] == host._version and ent[2].device == torch.device(device):
            self.entries.move_to_end(key)
            self.hits += 1
            dev_t = ent[2]
        else:
            self.misses += 1
            if ent is not None:
                self.used -= ent[2].numel() * ent[2].element_size()
                del self.entries[key]
            dev_t = host.to(device, non_blocking=True)
            nbytes = dev_t.numel() * dev_t.element_size()
            if nbytes <= self.budget:
                while self.used + nbytes > self.budget and self.entries:
                    _, (_, _, old) = self.entries.popitem(last=False)
                    self.used -= old.numel() * old.element_size()
                self.entries[key] = (weakref.ref(host), host._version, dev_t)
                self.used += nbytes
        return dev_t if into is None else into.copy_(dev_t, non_blocking=True)

    def clear(self):
        self.entries.clear()
        self.used = 0


resident_planes = ResidentPlanes()


class _StaticView:
    """A camera whose tensors are the static device buffers of a captured step."""

    def __init__(self, H, W, fovx, fovy, dev):
        self.image_height, self.image_width, self.FoVx, self.FoVy = H, W, fovx, fovy
        self.world_view_transform = torch.zeros((4, 4), device=dev)
        self.full_proj_transform = torch.zeros((4, 4), device=dev)
        self.camera_center = torch.zeros(3, device=dev)
        self.time_tensor = torch.zeros((1, 1), device=dev)
        self.time = 0.0
        self.mask = None

    def load(self, cam):
        self.world_view_transform.copy_(cam.world_view_transform, non_blocking=True)
        self.full_proj_transform.copy_(cam.full_proj_transform, non_blocking=True)
        self.camera_center.copy_(cam.camera_center, non_blocking=True)
        self.time = float(cam.time)
        self.time_tensor.fill_(self.time)


def _host_tensor(x):
    """`cam.original_image` / `cam.mask` as a tensor (the CPU stand-ins of the tests wrap theirs)."""
    return x if isinstance(x, torch.Tensor) else x.to("cpu")


class _StepGraph:
    """The frame-invariant part of `train_step` captured once as two CUDA graphs and replayed: ~380 kernel launches and the
    autograd bookkeeping of an iteration become two graph launches (launched from Python one op at a time the iteration
    took 6 ms for ~1 ms of device work).  Graph A renders the views; graph B computes the losses, runs the backward pass, the
    statistics and both Adam steps.  The ground-truth images (3 x 7.7 MB over PCIe) are uploaded on a copy stream while
    graph A renders; graph B waits for them.  Inputs that change from frame to frame (camera matrices, times, ground truth,
    masks) land in static buffers; everything else the graphs read (parameters, Adam moments and step counts, statistics)
    lives at fixed addresses and is updated in place.
    The graphs are dropped whenever the schedule changes tensors (densify / prune / opacity reset / clean-up / SH degree) or
    the signature (image size, field of view, number of views, learning rates) changes."""

    WARM = 2  # eager iterations on the side stream before the capture (allocator, cuBLAS workspaces, size hints)

    def __init__(self, sig, cams, gaussians, dev):
        self.sig = sig
        c0 = cams[0]
        H, W = int(c0.image_height), int(c0.image_width)
        self.views = [_StaticView(H, W, c.FoVx, c.FoVy, dev) for c in cams]
        self.gt = torch.zeros((len(cams), 3, H, W), device=dev)
        self.mask = None if c0.mask is None else torch.zeros((len(cams),) + tuple(c0.mask.shape), device=dev,
                                                             dtype=c0.mask.dtype)
        self.stream = torch.cuda.Stream(device=dev)
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.graph_a, self.graph_b, self.rendered, self.outputs, self.words, self.warm = None, None, None, None, None, 0
        self.mesh_adam = None

    def load_views(self, cams):
        for v, c in zip(self.views, cams):
            v.load(c)

    def load_images(self, cams):
        """On the copy stream: the planes cross PCIe while the views are rendered."""
        for i, c in enumerate(cams):
            resident_planes.fetch(_host_tensor(c.original_image), self.gt.device, into=self.gt[i])
            if self.mask is not None:
                resident_planes.fetch(_host_tensor(c.mask), self.mask.device, into=self.mask[i])


def _graph_signature(viewpoint_cams, gaussians, simulator, meshnet_optimizer, opt_params, static, stats):
    c0 = viewpoint_cams[0]
    for c in viewpoint_cams:
        if (int(c.image_height), int(c.image_width), float(c.FoVx), float(c.FoVy)) != (
                int(c0.image_height), int(c0.image_width), float(c0.FoVx), float(c0.FoVy)) or (c.mask is None) != (c0.mask is None):
            return None
    lrs = tuple(float(g["lr"]) for g in gaussians.optimizer.param_groups) + tuple(
        float(g["lr"]) for g in meshnet_optimizer.param_groups)
    return (len(viewpoint_cams), int(c0.image_height), int(c0.image_width), float(c0.FoVx), float(c0.FoVy),
            c0.mask is None, gaussians.num_gaussians, gaussians.active_sh_degree, bool(static), bool(stats), lrs,
            id(simulator), id(meshnet_optimizer), id(gaussians.optimizer), float(opt_params.lambda_dssim))


class _FusedTorchAdam:
    """The caller's torch.optim.Adam (the simulator's optimiser, train.py builds it) stepped by csb_adam_step_capturable
    on ITS OWN state tensors: one launch instead of the ~15 multi-tensor launches of torch's capturable path, same
    arithmetic.  `release()` writes the step count back the way torch keeps it."""

    def __init__(self, opt):
        import ctypes

        self.opt, self.ct = opt, ctypes
        self.items, self.steps = [], set()
        for g in opt.param_groups:
            if g.get("amsgrad") or g.get("maximize") or g.get("weight_decay", 0) != 0 or g.get("differentiable"):
                raise ValueError("unsupported torch.optim.Adam configuration")
            for p in g["params"]:
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                    raise ValueError("parameters must be contiguous CUDA fp32 tensors")
                st = opt.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                self.steps.add(int(float(st["step"])))
                self.items.append((p, st, g))
        if len(self.steps) > 1 or len({(g["betas"], g["eps"]) for _, _, g in self.items}) > 1:
            raise ValueError("parameters with different step counts / betas")
        dev = self.items[0][0].device
        self.step_dev = torch.full((1,), self.steps.pop() if self.steps else 0, dtype=torch.int64, device=dev)
        self.bias_dev = torch.zeros(2, dtype=torch.float64, device=dev)

    def step(self):
        import csb_cabi as _abi
        import csb_optim

        lib = _abi.lib()
        csb_optim._bind(lib)
        ct = self.ct
        items = [(p, st, g) for p, st, g in self.items if p.grad is not None]
        if not items:
            return
        n = len(items)
        g0 = items[0][2]
        P = (ct.c_void_p * n)(*[p.data_ptr() for p, _, _ in items])
        G = (ct.c_void_p * n)(*[p.grad.data_ptr() for p, _, _ in items])
        M = (ct.c_void_p * n)(*[s["exp_avg"].data_ptr() for _, s, _ in items])
        V = (ct.c_void_p * n)(*[s["exp_avg_sq"].data_ptr() for _, s, _ in items])
        N = (ct.c_int64 * n)(*[p.numel() for p, _, _ in items])
        LR = (ct.c_double * n)(*[float(g["lr"]) for _, _, g in items])
        with torch.cuda.device(items[0][0].device):
            rc = lib.csb_adam_step_capturable(n, P, G, M, V, N, LR, float(g0["betas"][0]), float(g0["betas"][1]),
                                              float(g0["eps"]), self.step_dev.data_ptr(), self.bias_dev.data_ptr(), 0,
                                              _abi.current_stream_ptr())
        _abi.check(rc, "csb_adam_step_capturable")
        torch.autograd.graph.increment_version([p for p, _, _ in items])

    def zero_grad(self, set_to_none=True):
        self.opt.zero_grad(set_to_none=set_to_none)

    def release(self):
        n = float(self.step_dev.item())
        for _, st, _ in self.items:
            st["step"] = torch.tensor(n)


def _make_capturable(st, gaussians, meshnet_optimizer):
    """Both optimisers with their step counts on the device; False if one of them cannot (then the step stays eager)."""
    from csb_optim import Adam as CsbAdam

    opt = gaussians.optimizer
    if not isinstance(opt, CsbAdam):
        return False
    opt.make_capturable()
    if isinstance(meshnet_optimizer, CsbAdam):
        meshnet_optimizer.make_capturable()
        st.mesh_adam = meshnet_optimizer
        return True
    if type(meshnet_optimizer) is torch.optim.Adam:
        try:
            st.mesh_adam = _FusedTorchAdam(meshnet_optimizer)
            return True
        except ValueError:
            return False
    return False


def drop_step_graph(gaussians):
    """Forget the captured iteration of this model (its tensors are about to change) and hand the step counts back."""
    st = gaussians.__dict__.pop("_csb_step_graph", None)
    if st is None:
        return
    torch.cuda.synchronize()
    if hasattr(gaussians.optimizer, "leave_capturable"):
        gaussians.optimizer.leave_capturable()
    if isinstance(st.mesh_adam, _FusedTorchAdam):
        st.mesh_adam.release()
    elif st.mesh_adam is not None and hasattr(st.mesh_adam, "leave_capturable"):
        st.mesh_adam.leave_capturable()


def _schedule_event(iteration, opt_params, white_background):
    """Does this iteration change tensors (not only their values)?  Then it runs eagerly and the graph is dropped."""
    if iteration % 1000 == 0 or iteration >= opt_params.iterations:
        return True
    if iteration % opt_params.bary_cleanup == 0:
        return True
    if iteration < opt_params.densify_until_iter:
        if iteration > opt_params.densify_from_iter and iteration % opt_params.densification_interval == 0:
            return True
        if iteration > opt_params.pruning_from_iter and iteration % opt_params.pruning_interval == 0:
            return True
        if iteration % opt_params.opacity_reset_interval == 0 or (
                white_background and iteration == opt_params.densify_from_iter):
            return True
    return False


def _graphed_step(iteration, viewpoint_cams, gaussians, simulator, meshnet_optimizer, pipeline_params, opt_params,
                  background, static, user_args):
    """One iteration through the captured graphs; None when this configuration cannot be captured."""
    import diff_gaussian_rasterization as dgr

    dev = background.device
    stats = iteration < opt_params.densify_until_iter
    sig = _graph_signature(viewpoint_cams, gaussians, simulator, meshnet_optimizer, opt_params, static, stats)
    if sig is None or not hasattr(gaussians, "max_radii2D"):
        return None
    st = gaussians.__dict__.get("_csb_step_graph")
    if st is None or st.sig != sig:
        drop_step_graph(gaussians)
        st = _StepGraph(sig, viewpoint_cams, gaussians, dev)
        if not _make_capturable(st, gaussians, meshnet_optimizer):
            return None
        gaussians.__dict__["_csb_step_graph"] = st
    for c in viewpoint_cams:
        if hasattr(simulator, "check_time") and not static:
            simulator.check_time(c.time)
    cur = torch.cuda.current_stream(dev)
    st.stream.wait_stream(cur)
    st.copy_stream.wait_stream(cur)
    with torch.cuda.stream(st.copy_stream):
        st.load_images(viewpoint_cams)
    with torch.cuda.stream(st.stream):
        st.load_views(viewpoint_cams)
        a_args = (st.views, gaussians, simulator, pipeline_params, background, static, user_args, False)

        def b_args(rendered):
            return (rendered, st.gt, st.mask, stats, gaussians, st.mesh_adam, opt_params, static, True, False)

        if st.graph_a is None and st.warm < _StepGraph.WARM:
            rendered = _render_views(*a_args)
            st.stream.wait_stream(st.copy_stream)
            out = _loss_and_step(*b_args(rendered))[:3]
            st.warm += 1
        else:
            if st.graph_a is None:
                st.words = dgr.reserve_count_words(len(st.views))
                gaussians.optimizer.zero_grad(set_to_none=True)
                meshnet_optimizer.zero_grad(set_to_none=True)
                st.stream.wait_stream(st.copy_stream)
                ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
                dgr._captured_forwards.clear()
                with torch.cuda.graph(ga, stream=st.stream):
                    st.rendered = _render_views(*a_args)
                dgr._reserved_words.clear()
                st.forwards = list(dgr._captured_forwards)
                with torch.cuda.graph(gb, stream=st.stream, pool=ga.pool()):
                    st.outputs = _loss_and_step(*b_args(st.rendered))[:3]
                st.graph_a, st.graph_b = ga, gb
            st.words.fill_(-1)
            st.graph_a.replay()
            st.stream.wait_stream(st.copy_stream)
            st.graph_b.replay()
            out = st.outputs
    cur.wait_stream(st.stream)
    psnr_, loss, loss_dict = out
    # the reference returns Python floats in loss_dict (two .item() per step): read them once, after everything is enqueued
    vals = torch.stack([psnr_.float(), loss.detach().float()] + [v.detach().float() for v in loss_dict.values()]).cpu()
    if st.graph_a is not None and st.words is not None:
        # the instance counts of the replayed forwards have arrived with the values above: a frame that outgrew the
        # workspace the graph was captured with rendered incompletely -- drop the graphs, the next frame re-captures larger
        if not all(f.check_replay() for f in st.forwards):
            drop_step_graph(gaussians)
    out_dict = {k: float(vals[2 + i]) for i, k in enumerate(loss_dict)}
    return psnr_.detach().clone(), loss.detach().clone(), out_dict


def train_step(iteration, viewpoint_cams, gaussians, simulator, meshnet_optimizer, pipeline_params, opt_params,
               cameras_extent, background, static=False, white_background=False, user_args=None):
    """One iteration of the refinement loop (scene_reconstruction/train_utils.py:240-321), same arguments and the same
    `(psnr, loss, loss_dict)` return: render every view of the step at its time (`gaussian_renderer.render`: simulator ->
    mesh-anchored Gaussians -> rasterizer), image losses over the stacked views + the mesh regularisers, one backward, the
    densification / opacity-reset / barycentric-cleanup schedule on the summed screen-space gradients, then the Adam
    steps of the Gaussians and (unless `static`) of the simulator.

    Iterations on which the schedule does nothing but accumulate statistics -- all but one in a hundred -- are replayed
    from a CUDA graph captured on the first such frames (`_StepGraph`; opt_params.cuda_graph = False turns it off);
    the others run op by op and drop the graph, because they change tensor shapes."""
    gaussians.update_learning_rate(iteration)
    if (getattr(opt_params, "cuda_graph", True) and background.is_cuda
            and not _schedule_event(iteration, opt_params, white_background)):
        out = _graphed_step(iteration, viewpoint_cams, gaussians, simulator, meshnet_optimizer, pipeline_params,
                            opt_params, background, static, user_args)
        if out is not None:
            return out
    drop_step_graph(gaussians)
    if iteration % 1000 == 0:
        gaussians.oneupSHdegree()
    dev = background.device
    gt_image_tensor = torch.stack([resident_planes.fetch(_host_tensor(cam.original_image), dev) for cam in viewpoint_cams])
    mask_tensor = None
    if viewpoint_cams[0].mask is not None:
        mask_tensor = torch.stack([resident_planes.fetch(_host_tensor(cam.mask), dev) for cam in viewpoint_cams])
    rendered = _render_views(viewpoint_cams, gaussians, simulator, pipeline_params, background, static, user_args)
    psnr_, loss, loss_dict, viewspace_point_tensor_grad, visibility_filter, radii = _loss_and_step(
        rendered, gt_image_tensor, mask_tensor, False, gaussians, meshnet_optimizer, opt_params, static, False, True)
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
