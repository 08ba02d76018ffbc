"""csb_image_loss_forward_static: the photometric-loss forward without a caller workspace and without any allocation
(partial sums in a file-scope `static __device__` array of CSB_IMAGE_LOSS_STATIC_BYTES).  No counterpart in the
reference; see include/cloth_splatting_b200.h.  Collected last on purpose.

Checked against the existing entry point csb_image_loss_forward on the same inputs, through the C-ABI, on two calls
in a row that share the array (different shapes, the second smaller, so stale partial sums of the first call sit behind
the second's).  Tolerance: the loss tests' 1e-5 absolute on loss values and SSIM map (tests/test_loss_gpu.py) -- and,
since both entry points run the same kernels and fold the same partial sums in the same order, bitwise equality."""
import ctypes

import pytest
import torch

import conftest  # noqa: F401  (puts the package on sys.path)

pytestmark = pytest.mark.gpu

ERR_STATIC_LIMIT = -4


def _bind():
    import csb_cabi as abi

    lib = abi.lib()
    c_int, c_void_p, c_float, c_size_t = ctypes.c_int, ctypes.c_void_p, ctypes.c_float, ctypes.c_size_t
    lib.csb_image_loss_workspace_bytes.restype = c_size_t
    lib.csb_image_loss_workspace_bytes.argtypes = [c_int] * 4
    lib.csb_image_loss_forward.restype = c_int
    lib.csb_image_loss_forward.argtypes = [c_int] * 4 + [c_void_p] * 3 + [c_int, c_float] + [c_void_p] * 4 + [c_size_t, c_void_p]
    lib.csb_image_loss_forward_static.restype = c_int
    lib.csb_image_loss_forward_static.argtypes = [c_int] * 4 + [c_void_p] * 3 + [c_int, c_float] + [c_void_p] * 3 + [c_void_p]
    lib.csb_image_loss_static_max_bytes.restype = c_size_t
    lib.csb_image_loss_static_max_bytes.argtypes = []
    lib.csb_launch_count.restype = ctypes.c_ulonglong
    lib.csb_launch_count.argtypes = []
    return abi, lib


def _inputs(n, C, H, W, masked, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    gt = torch.rand(n, C, H, W, device="cuda", generator=g)
    img = (gt + 0.1 * torch.randn(n, C, H, W, device="cuda", generator=g)).clamp(0, 1)
    mask = (torch.rand(n, 1, H, W, device="cuda", generator=g) > 0.3).float() if masked else None
    return img.contiguous(), gt.contiguous(), mask


def _outputs(x):
    return torch.full((3,), float("nan"), device="cuda"), torch.full_like(x, float("nan")), torch.full((3,) + tuple(x.shape), float("nan"), device="cuda")


def _both(abi, lib, shape, masked, seed, lam=0.2):
    n, C, H, W = shape
    x, y, m = _inputs(n, C, H, W, masked, seed)
    stream = abi.current_stream_ptr()
    # existing entry point, caller workspace
    l0, s0, d0 = _outputs(x)
    ws_bytes = lib.csb_image_loss_workspace_bytes(n, C, H, W)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    abi.check(lib.csb_image_loss_forward(n, C, H, W, abi.ptr(x), abi.ptr(y), abi.ptr(m), 1 if masked else 0, lam, abi.ptr(s0),
                                         abi.ptr(d0), abi.ptr(l0), abi.ptr(ws), ws_bytes, stream), "csb_image_loss_forward")
    # new entry point: outputs are allocated BEFORE the call, the call itself must not allocate
    l1, s1, d1 = _outputs(x)
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info()[0]
    abi.check(lib.csb_image_loss_forward_static(n, C, H, W, abi.ptr(x), abi.ptr(y), abi.ptr(m), 1 if masked else 0, lam, abi.ptr(s1),
                                                abi.ptr(d1), abi.ptr(l1), stream), "csb_image_loss_forward_static")
    torch.cuda.synchronize()
    after = torch.cuda.mem_get_info()[0]
    return (l0, s0, d0), (l1, s1, d1), before - after


def test_static_scratch_loss_matches_workspace_entry_point_twice_in_a_row():
    abi, lib = _bind()
    assert lib.csb_image_loss_static_max_bytes() == 1 << 20
    # free device memory is compared around the LAST call only (the first use of a kernel in a process makes the
    # driver load it; a cudaMalloc inside the call would show as >= 2 MiB less free memory)
    grew = []
    for k, (shape, masked) in enumerate([((4, 3, 800, 800), True), ((2, 3, 167, 211), False), ((4, 3, 800, 800), False)]):
        ref, got, used = _both(abi, lib, shape, masked, seed=100 + k)
        grew.append(used)
        for name, a, b in zip(("losses", "ssim_map", "dmaps"), ref, got):
            assert torch.isfinite(b).all(), name
            assert (a - b).abs().max().item() <= 1e-5, (k, name)
            assert torch.equal(a, b), (k, name)
    assert grew[-1] < (1 << 20), f"device memory in use grew across a no-allocation call: {grew}"


def test_static_scratch_rejects_sizes_beyond_its_array():
    abi, lib = _bind()
    # 16 planes x 91 x 91 tiles = 132 496 > 131 072: refused before anything is launched (the pointers are never read)
    n, C, H, W = 16, 1, 2900, 2900
    launches = lib.csb_launch_count
    count = launches()
    one = torch.zeros(1, device="cuda")
    rc = lib.csb_image_loss_forward_static(n, C, H, W, abi.ptr(one), abi.ptr(one), None, 0, 0.2, None, None, abi.ptr(one),
                                           abi.current_stream_ptr())
    assert rc == ERR_STATIC_LIMIT
    assert launches() == count
    assert b"no-allocation" in lib.csb_error_string(rc)
    with pytest.raises(RuntimeError):
        abi.check(rc, "csb_image_loss_forward_static")
    # a size just inside the limit is accepted: one plane of 362 x 362 tiles = 131 044 (11 584^2 px, 537 MB per image)
    H = W = 362 * 32
    x = torch.rand(1, 1, H, W, device="cuda")
    y = torch.rand(1, 1, H, W, device="cuda")
    l0 = torch.empty(3, device="cuda")
    l1 = torch.empty(3, device="cuda")
    ws_bytes = lib.csb_image_loss_workspace_bytes(1, 1, H, W)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    abi.check(lib.csb_image_loss_forward(1, 1, H, W, abi.ptr(x), abi.ptr(y), None, 0, 0.2, None, None, abi.ptr(l0), abi.ptr(ws), ws_bytes,
                                         abi.current_stream_ptr()), "csb_image_loss_forward")
    abi.check(lib.csb_image_loss_forward_static(1, 1, H, W, abi.ptr(x), abi.ptr(y), None, 0, 0.2, None, None, abi.ptr(l1),
                                                abi.current_stream_ptr()), "csb_image_loss_forward_static")
    torch.cuda.synchronize()
    assert torch.equal(l0, l1)
