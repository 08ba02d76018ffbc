"""Drop-in replacement of the reference's `diff_gaussian_rasterization` package.

Same public surface as RAST/diff_gaussian_rasterization/__init__.py (RAST =
submodules/depth-diff-gaussian-rasterization of KTH-RPL/cloth-splatting):
  GaussianRasterizationSettings (:157-169), GaussianRasterizer (:171-220),
  rasterize_gaussians (:21-42), _RasterizeGaussians (:44-155) and a `_C` namespace with
  rasterize_gaussians / rasterize_gaussians_backward / mark_visible (RAST/ext.cpp:15-19).
Everything below the Python layer is libcloth_splatting_b200.so (hand-written sm_100a CUDA behind
the C-ABI of include/cloth_splatting_b200.h); there is no CPU fallback.

Extension (not in the reference): `rasterize_views` renders up to 16 views of the same Gaussians
with one launch set (train_utils.py:259 loops over the views sequentially).
"""
import ctypes
from typing import NamedTuple

import torch
import torch.nn as nn

import csb_cabi as _abi


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


# --------------------------------------------------------------------------------------
# `_C` level: same signatures and return tuples as the reference's pybind module
# --------------------------------------------------------------------------------------
_capacity_hint = {}  # (device, V, P, W, H) -> instances seen last time
_MIN_CAPACITY = 1 << 16  # first guess of the instance capacity: max(this, _PER_GAUSSIAN * V * P)
_PER_GAUSSIAN = 8
_pinned_counter = {}


_RING = 64
_COUNT_REFUSED = 0xFFFFFFFD  # per-tile binning refused: a tile list exceeds the per-tile sort's shared memory
_small_bins = {}
_ring_next = {}
_ring_owner = {}         # (device, slot) -> weakref of the PendingCount that still has to read the slot's word
_ring_slot_of_word = {}  # data_ptr of a ring word -> (device, slot)


_reserved_words = []  # pinned words handed out next (set by reserve_count_words for a CUDA-graph capture)
_captured_forwards = []  # PendingCount of every forward recorded into a CUDA graph since the list was last cleared


def reserve_count_words(n):
    """Pinned count words for the next `n` forwards, allocated NOW: a forward that is being captured in a CUDA graph must
    not allocate pinned memory, and its word must not be recycled by the ring while the graph lives.  Returns the
    tensor [n] (keep it alive with the graph; after a replay word i holds the instance count of the i-th forward)."""
    t = torch.full((n,), -1, dtype=torch.int32).pin_memory()
    _reserved_words[:] = [t[i:i + 1] for i in range(n)]
    return t


def _pinned_word(device):
    """One int32 word of a pinned (device-accessible) ring per device: calls in flight never share a word."""
    if _reserved_words:
        return _reserved_words.pop(0)
    t = _pinned_counter.get(device)
    if t is None:
        t = torch.zeros(_RING, dtype=torch.int32).pin_memory()
        _pinned_counter[device] = t
    i = _ring_next.get(device, 0)
    _ring_next[device] = (i + 1) % _RING
    owner = _ring_owner.get((device, i))
    owner = owner() if owner is not None else None
    if owner is not None and owner.value is None:
        # more than _RING deferred counts in flight: the slot's previous forward has not been read yet -- it keeps its word,
        # this call gets a fresh one (pinned allocation: slow, but correct)
        return torch.full((1,), -1, dtype=torch.int32).pin_memory()
    _ring_slot_of_word[t[i:i + 1].data_ptr()] = (device, i)
    return t[i:i + 1]


def _await_count(word, stream):
    """Spins on the pinned word the binning kernel writes (csb_rasterize_forward flags bit1).  The count arrives
    while the tile sort / render kernels are still running, so the caller goes on enqueueing (the backward pass,
    the loss) without ever draining the stream."""
    spins = 0
    while True:
        v = int(word.item())
        if v != -1:
            return v & 0xFFFFFFFF
        spins += 1
        if spins % 2048 == 0 and stream.query():
            v = int(word.item())  # the stream drained: the word is final
            if v == -1:
                raise RuntimeError("csb_rasterize_forward finished without reporting num_rendered")
            return v & 0xFFFFFFFF


class RenderAgain(RuntimeError):
    """A deferred instance count turned out larger than the binning workspace (or the per-tile binning path refused the
    scene): the images of that forward are incomplete.  The size hints have been updated; call the forward again."""


class PendingCount:
    """The instance count of a forward launched with `defer_count=True`: nothing has been read from the device yet.
    `wait()` (or `int()`) polls the pinned word -- normally long after the binning kernel wrote it, i.e. without
    waiting -- applies the overflow check the eager path does right after the launch, and returns the count."""

    def __init__(self, word, stream, capacity, key, small, n_tiles):
        self.word, self.stream, self.capacity, self.key, self.small, self.n_tiles = word, stream, capacity, key, small, n_tiles
        self.value = None
        slot = _ring_slot_of_word.get(word.data_ptr())
        if slot is not None:
            import weakref

            _ring_owner[slot] = weakref.ref(self)

    def wait(self):
        if self.value is None:
            v = _await_count(self.word, self.stream)
            if self.small and v == _COUNT_REFUSED:
                _small_bins[self.key] = False
                raise RenderAgain("the per-tile binning path refused this scene (tile list too long)")
            if v > self.capacity:
                _capacity_hint[self.key] = v
                raise RenderAgain(f"{v} instances do not fit the binning workspace sized for {self.capacity}")
            _capacity_hint[self.key] = v
            if self.small and v > 96 * self.n_tiles:
                _small_bins[self.key] = False
            self.value = v
        return self.value

    __int__ = wait
    __index__ = wait

    def check_replay(self):
        """After a replay of the graph this forward was captured into (and a synchronisation): did the frame fit the
        workspace the graph owns?  Updates the size hints like wait() does; False = the images of that replay are
        incomplete and the graph must be captured again."""
        v = int(self.word.item()) & 0xFFFFFFFF
        if v == 0xFFFFFFFF:
            return True  # not replayed yet
        if self.small and v == _COUNT_REFUSED:
            _small_bins[self.key] = False
            return False
        _capacity_hint[self.key] = max(v, _capacity_hint.get(self.key, 0))
        if self.small and v > 96 * self.n_tiles:
            _small_bins[self.key] = False
        return v <= self.capacity


def _f32c(t, name):
    if t is None or t.numel() == 0:
        return None
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    return t.contiguous()


def _views_array(views):
    arr = (_abi.CsbView * len(views))()
    keep = []
    for i, (viewmatrix, projmatrix, campos, bg, tanfovx, tanfovy) in enumerate(views):
        vm, pm, cp, b = _f32c(viewmatrix, "viewmatrix"), _f32c(projmatrix, "projmatrix"), _f32c(campos, "campos"), _f32c(bg, "bg")
        keep += [vm, pm, cp, b]
        arr[i].viewmatrix = vm.data_ptr()
        arr[i].projmatrix = pm.data_ptr()
        arr[i].campos = cp.data_ptr()
        arr[i].background = b.data_ptr()
        arr[i].tan_fovx = float(tanfovx)
        arr[i].tan_fovy = float(tanfovy)
    return arr, keep


def _on_device_of(t, fn, *args, **kwargs):
    """Runs `fn` with the CUDA device of tensor `t` current (the C-ABI launches on the current device's stream and sets
    per-device kernel attributes): a process that renders on cuda:1 while cuda:0 is current must not launch there."""
    if t.is_cuda and t.device.index is not None and torch.cuda.current_device() != t.device.index:
        with torch.cuda.device(t.device):
            return fn(*args, **kwargs)
    return fn(*args, **kwargs)


def _forward_views(views, means3D, *args, **kwargs):
    return _on_device_of(means3D, _forward_views_impl, views, means3D, *args, **kwargs)


def _backward_views(views, means3D, *args, **kwargs):
    return _on_device_of(means3D, _backward_views_impl, views, means3D, *args, **kwargs)


def _forward_views_impl(views, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        image_height, image_width, sh, degree, prefiltered, flags=0, defer_count=False):
    """Shared implementation: returns (rendered, color[V,3,H,W], depth[V,1,H,W], radii[V,P], geom, binning, img).
    `defer_count=True` returns a `PendingCount` instead of the int and does not touch the host word: the caller goes
    on enqueueing (the backward pass does not need the count) and calls `.wait()` when convenient; an overflow of the
    binning workspace then surfaces there as `RenderAgain` instead of being retried here."""
    lib = _abi.lib()
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise ValueError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:55-57
    V = len(views)
    P = means3D.size(0)
    H, W = int(image_height), int(image_width)
    dev = means3D.device
    if P == 0:
        # rasterize_points.cu:68-82: nothing is launched, the zero-filled outputs are returned as they are
        z = torch.zeros((0,), dtype=torch.uint8, device=dev)
        return (0, torch.zeros((V, 3, H, W), dtype=torch.float32, device=dev),
                torch.zeros((V, 1, H, W), dtype=torch.float32, device=dev),
                torch.zeros((V, 0), dtype=torch.int32, device=dev), z, z.clone(), z.clone())
    means3D = _f32c(means3D, "means3D")
    colors = _f32c(colors, "colors_precomp")
    opacity = _f32c(opacity, "opacities")
    scales = _f32c(scales, "scales")
    rotations = _f32c(rotations, "rotations")
    cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp")
    sh = _f32c(sh, "sh")
    M = sh.size(1) if sh is not None else 0

    out_color = torch.empty((V, 3, H, W), dtype=torch.float32, device=dev)
    out_depth = torch.empty((V, 1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((V, P), dtype=torch.int32, device=dev)
    geom = torch.empty(lib.csb_rasterize_geom_bytes(V, P), dtype=torch.uint8, device=dev)
    img = torch.empty(lib.csb_rasterize_image_bytes(V, W, H), dtype=torch.uint8, device=dev)
    counter = torch.empty(1, dtype=torch.int32, device=dev)
    host_word = _pinned_word(dev)
    arr, keep = _views_array(views)

    key = (dev, V, P, W, H)
    hint = _capacity_hint.get(key, 0)
    capacity = max(int(hint * 1.25) + 4096 if hint else 0, _PER_GAUSSIAN * V * P, _MIN_CAPACITY) if P else 0
    capturing = torch.cuda.is_current_stream_capturing()
    if capturing:
        # a captured forward is replayed on later frames with this workspace: no host word to poll, no retry, so the count
        # is deferred (the owner of the graph checks the pinned words after a replay) and the capacity gets head room
        defer_count = True
        capacity = max(capacity, 2 * hint + 4096)
    stream = _abi.current_stream_ptr()
    # short tile lists (cloth-sized scenes) are binned per tile (count / scan / fill / per-tile sort: flags bit2) instead
    # of going through the two global radix sorts; the library refuses when a tile list is too long and the global
    # path takes over for this problem size from then on
    n_tiles = V * ((W + 15) // 16) * ((H + 15) // 16)
    small = _small_bins.get(key, n_tiles <= (1 << 20) and V * P <= (1 << 19)) and P > 0
    while True:
        if not capturing:
            host_word.fill_(-1)
        nbytes = lib.csb_rasterize_binning_bytes(V, P, W, H, capacity)
        # canonical capacity of a buffer of that size: the backward pass recovers it from numel()
        capacity = lib.csb_rasterize_binning_capacity(V, P, W, H, nbytes)
        binning = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        rc = lib.csb_rasterize_forward(
            V, arr, P, int(degree), M, W, H, _abi.ptr(means3D), _abi.ptr(sh), _abi.ptr(colors), _abi.ptr(opacity),
            _abi.ptr(scales), float(scale_modifier), _abi.ptr(rotations), _abi.ptr(cov3D_precomp), int(bool(prefiltered)),
            geom.data_ptr(), geom.numel(), img.data_ptr(), img.numel(), _abi.ptr(binning), binning.numel(), capacity,
            out_color.data_ptr(), out_depth.data_ptr(), radii.data_ptr() if P else None, counter.data_ptr(),
            host_word.data_ptr(), flags | 2 | (4 if small else 0), stream)
        _abi.check(rc, "csb_rasterize_forward")
        if defer_count:
            del keep
            pending = PendingCount(host_word, torch.cuda.current_stream(), capacity, key, small, n_tiles)
            if capturing:
                _captured_forwards.append(pending)
            return (pending, out_color, out_depth, radii, geom, binning, img)
        # no stream synchronisation: the instance count is polled from pinned memory as soon as the binning
        # kernel knows it (the reference blocks on a cudaMemcpy in the middle, rasterizer_impl.cu:282)
        rendered = _await_count(host_word, torch.cuda.current_stream())
        if small and rendered == _COUNT_REFUSED:
            small = _small_bins[key] = False
            continue
        if rendered <= capacity:
            break
        capacity = rendered + 4096  # grow and render again
    _capacity_hint[key] = rendered
    if small and rendered > 96 * n_tiles:
        _small_bins[key] = False  # long lists: the global sorts are faster from here on
    del keep
    return rendered, out_color, out_depth, radii, geom, binning, img


def _backward_views_impl(views, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, image_height,
                         image_width, dL_dout_color, sh, degree, geom, binning, img, out=None, view_events=None):
    """`out`: optional dict of preallocated contiguous fp32 CUDA tensors ("means3D" [P,3], "scales" [P,3], "rotations"
    [P,4], "opacities" [P,1], "shs" [P,M,3]) the kernels write these gradients into -- e.g. views of the flat
    all-reduce bucket (csb_dist.FlatGradBucket), which saves the pack copies.
    `view_events`: optional list of V torch.cuda.Event (or None entries): view v's plane of dL_dout_color is complete
    once its event has fired (csb_rasterize_backward_staged)."""
    lib = _abi.lib()
    V = len(views)
    P = means3D.size(0)
    H, W = int(image_height), int(image_width)
    dev = means3D.device
    means3D = _f32c(means3D, "means3D") if P else means3D
    colors = _f32c(colors, "colors_precomp")
    scales = _f32c(scales, "scales")
    rotations = _f32c(rotations, "rotations")
    cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp")
    sh = _f32c(sh, "sh")
    M = sh.size(1) if sh is not None else 0
    dL_dpix = _f32c(dL_dout_color, "dL_dout_color")

    def _buf(name, shape):
        t = out.get(name) if out else None
        if t is None:
            return torch.empty(shape, dtype=torch.float32, device=dev)
        if tuple(t.shape) != tuple(shape) or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
            raise ValueError(f"out['{name}'] must be a contiguous float32 tensor of shape {tuple(shape)} on {dev}")
        if t.data_ptr() % 16:
            raise ValueError(f"out['{name}'] must be 16-byte aligned (the kernels use vector stores)")
        return t

    dL_dmeans2D = torch.empty((V, P, 3), dtype=torch.float32, device=dev)
    dL_dmeans3D = _buf("means3D", (P, 3))
    dL_dcolors = torch.empty((P, 3), dtype=torch.float32, device=dev)
    dL_dopacity = _buf("opacities", (P, 1))
    dL_dcov3D = torch.empty((P, 6), dtype=torch.float32, device=dev)
    dL_dsh = _buf("shs", (P, M, 3))
    dL_dscales = _buf("scales", (P, 3))
    dL_drotations = _buf("rotations", (P, 4))
    if P:
        arr, keep = _views_array(views)
        capacity = lib.csb_rasterize_binning_capacity(V, P, W, H, binning.numel())
        args = (V, arr, P, int(degree), M, W, H, _abi.ptr(means3D), _abi.ptr(sh), _abi.ptr(colors), _abi.ptr(scales),
                float(scale_modifier), _abi.ptr(rotations), _abi.ptr(cov3D_precomp), radii.data_ptr(), geom.data_ptr(),
                img.data_ptr(), _abi.ptr(binning), capacity, dL_dpix.data_ptr(), dL_dmeans2D.data_ptr(),
                dL_dcolors.data_ptr(), dL_dopacity.data_ptr(), dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(),
                _abi.ptr(dL_dsh), dL_dscales.data_ptr(), dL_drotations.data_ptr())
        if view_events is not None:
            if len(view_events) != V:
                raise ValueError("view_events needs one entry per view")
            import ctypes

            for ev in view_events:
                if ev is not None and not ev.cuda_event:  # torch creates the handle on the first record()
                    raise ValueError("view_events: every event must have been recorded before the backward is enqueued")
            evs = (ctypes.c_void_p * V)(*[None if ev is None else ev.cuda_event for ev in view_events])
            rc = lib.csb_rasterize_backward_staged(*args, evs, _abi.current_stream_ptr())
        else:
            rc = lib.csb_rasterize_backward(*args, _abi.current_stream_ptr())
        _abi.check(rc, "csb_rasterize_backward")
        del keep
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


class _CNamespace:
    """Mirror of the pybind module `diff_gaussian_rasterization._C` (RAST/ext.cpp:15-19)."""

    @staticmethod
    def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                            viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                            prefiltered, debug):
        views = [(viewmatrix, projmatrix, campos, background, tan_fovx, tan_fovy)]
        rendered, color, depth, radii, geom, binning, img = _forward_views(
            views, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, image_height,
            image_width, sh, degree, prefiltered)
        return rendered, color[0], depth[0], radii[0], geom, binning, img

    @staticmethod
    def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                     cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh,
                                     degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug):
        views = [(viewmatrix, projmatrix, campos, background, tan_fovx, tan_fovy)]
        H, W = dL_dout_color.shape[-2], dL_dout_color.shape[-1]
        g = _backward_views(views, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, H, W,
                            dL_dout_color, sh, degree, geomBuffer, binningBuffer, imageBuffer)
        dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations = g
        return dL_dmeans2D[0], dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations

    @staticmethod
    def mark_visible(means3D, viewmatrix, projmatrix):
        P = means3D.size(0)
        present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
        if P:
            m = _f32c(means3D, "means3D")
            vm, pm = _f32c(viewmatrix, "viewmatrix"), _f32c(projmatrix, "projmatrix")
            rc = _abi.lib().csb_mark_visible(P, m.data_ptr(), vm.data_ptr(), pm.data_ptr(), present.data_ptr(),
                                             _abi.current_stream_ptr())
            _abi.check(rc, "csb_mark_visible")
        return present


_C = _CNamespace()


# --------------------------------------------------------------------------------------
# Python operator API (identical to the reference's)
# --------------------------------------------------------------------------------------
def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        args = (raster_settings.bg, means3D, colors_precomp, opacities, scales, rotations,
                raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy,
                raster_settings.image_height, raster_settings.image_width, sh, raster_settings.sh_degree,
                raster_settings.campos, raster_settings.prefiltered, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians(*args)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth):
        # grad_depth is dropped: the reference's depth output is forward-only (RAST/README.md:5)
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (raster_settings.bg, means3D, radii, colors_precomp, scales, rotations, raster_settings.scale_modifier,
                cov3Ds_precomp, raster_settings.viewmatrix, raster_settings.projmatrix, raster_settings.tanfovx,
                raster_settings.tanfovy, grad_out_color, sh, raster_settings.sh_degree, raster_settings.campos,
                geomBuffer, num_rendered, binningBuffer, imgBuffer, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                grads = _C.rasterize_gaussians_backward(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            grads = _C.rasterize_gaussians_backward(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = grads
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = _C.mark_visible(positions, raster_settings.viewmatrix, raster_settings.projmatrix)
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, raster_settings)


# --------------------------------------------------------------------------------------
# Multi-view extension
# --------------------------------------------------------------------------------------
class _RasterizeViews(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, settings_list):
        s0 = settings_list[0]
        views = [(s.viewmatrix, s.projmatrix, s.campos, s.bg, s.tanfovx, s.tanfovy) for s in settings_list]
        rendered, color, depth, radii, geom, binning, img = _forward_views(
            views, means3D, colors_precomp, opacities, scales, rotations, s0.scale_modifier, cov3Ds_precomp,
            s0.image_height, s0.image_width, sh, s0.sh_degree, s0.prefiltered)
        ctx.settings_list = settings_list
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth):
        settings_list = ctx.settings_list
        s0 = settings_list[0]
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img = ctx.saved_tensors
        views = [(s.viewmatrix, s.projmatrix, s.campos, s.bg, s.tanfovx, s.tanfovy) for s in settings_list]
        staged = getattr(_staged, "grads", None)
        events = None
        if staged is not None and staged.dev.data_ptr() == grad_color.data_ptr() and len(staged.events) == len(views):
            events = staged.events  # the upstream gradient is the staged upload itself: wait view by view
        g = _backward_views(views, means3D, radii, colors_precomp, scales, rotations, s0.scale_modifier, cov3Ds_precomp,
                            s0.image_height, s0.image_width, grad_color, sh, s0.sh_degree, geom, binning, img,
                            view_events=events)
        dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations = g
        ctx.dL_dmeans2D = dL_dmeans2D
        return (dL_dmeans3D, dL_dsh, dL_dcolors, dL_dopacity, dL_dscales, dL_drotations, dL_dcov3D, None)


class _Staged:  # a plain global: autograd runs CUDA nodes on its own device thread, a thread-local would not be seen
    grads = None


_staged = _Staged()


class StagedPixelGrads:
    """dL/dpixel of a view batch on its way to the device, view by view: `upload(host)` enqueues one copy per view on
    `stream` (a side stream) and records an event behind each; `backward_staged(color, staged)` then runs the backward
    of `rasterize_views` so that view v is composited as soon as ITS plane has arrived, while the planes of the later
    views are still crossing PCIe."""

    def __init__(self, dev_buffer, stream):
        self.dev, self.stream = dev_buffer, stream
        self.events = [torch.cuda.Event() for _ in range(dev_buffer.shape[0])]

    def upload(self, host):
        with torch.cuda.stream(self.stream):
            for v, ev in enumerate(self.events):
                self.dev[v].copy_(host[v], non_blocking=True)
                ev.record(self.stream)
        return self


def backward_staged(color, staged):
    """color.backward(staged.dev) with the per-view events of `staged` honoured by the rasterizer's backward (the
    current stream never waits for the whole upload).  Only valid when `color` is the direct output of
    `rasterize_views` and staged.dev is passed as its upstream gradient."""
    _staged.grads = staged
    try:
        color.backward(staged.dev)
    finally:
        _staged.grads = None


def rasterize_views(settings_list, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                    cov3D_precomp=None):
    """Render len(settings_list) <= 16 views (same image size, sh_degree, scale_modifier) of the same
    Gaussians with one launch set.  Returns (color[V,3,H,W], radii[V,P], depth[V,1,H,W]); gradients of the
    Gaussian parameters are summed over the views, as autograd would over sequential renders."""
    if not 1 <= len(settings_list) <= _abi.MAX_VIEWS:
        raise ValueError(f"1..{_abi.MAX_VIEWS} views per call")
    s0 = settings_list[0]
    for s in settings_list:
        if (s.image_height, s.image_width, s.sh_degree, s.scale_modifier) != (
                s0.image_height, s0.image_width, s0.sh_degree, s0.scale_modifier):
            raise ValueError("all views of a batch must share image size, sh_degree and scale_modifier")
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or (
            (scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
    e = torch.Tensor([])
    return _RasterizeViews.apply(means3D, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
                                 opacities, e if scales is None else scales, e if rotations is None else rotations,
                                 e if cov3D_precomp is None else cov3D_precomp, tuple(settings_list))
