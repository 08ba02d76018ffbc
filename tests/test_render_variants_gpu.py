"""The generations of compositing kernels (csb_debug_render_variant: low nibble = forward kernel, high nibble = backward
kernel; 0 = one pixel per lane, 1 = one vertical pixel pair per lane on packed fp32x2, 2 = two pairs per lane, forward 3 / 4 =
one / two pairs per lane without per-pixel branches (packed expf with libdevice's bits), backward 3..6
= register-capped / delayed-reduction builds, 7 = warp sums through shared memory; csrc/render_lanes.cuh) against each other and against the reference's own
CUDA (renderCUDA forward.cu:261-379, backward.cu:399-557).

Forward: every per-pixel output bit for bit across the variants (same operand order, same FMA contraction, expf kept).
Backward: the warp-reduced sums differ in association only -> 1e-4 relative, the north_star's gradient tolerance."""
import contextlib
import ctypes

import numpy as np
import pytest
import torch

import conftest  # noqa: F401
import csb_synth as S
import helpers as Hh

pytestmark = pytest.mark.gpu


@contextlib.contextmanager
def variant(v):
    import csb_cabi

    lib = csb_cabi.lib()
    old = lib.csb_debug_render_variant(v)
    try:
        yield
    finally:
        lib.csb_debug_render_variant(old)


VARIANTS = (0x00, 0x11, 0x22, 0x30, 0x43, 0x54, 0x60, 0x74)


def _scenes():
    g, cam = Hh.small_scene(seed=11)
    yield "small", g, [cam]
    g, cam = Hh.small_scene(P=2000, W=149, H=83, seed=31)  # ragged: partial tiles on both edges
    yield "ragged", g, [cam]
    g, cams = S.config_c2(3, 800)
    yield "c2", {k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")}, cams
    g, cams = S.config_c3(100_000, 1920, 1080)  # long tile lists: several staged batches per tile, partial last tile row
    yield "c3_100k", g, cams


def test_variants_agree():
    for name, g, cams in _scenes():
        gc = Hh.to_cuda(g)
        V, P = len(cams), g["means3D"].shape[0]
        H, W = cams[0].image_height, cams[0].image_width
        dl = torch.from_numpy(S.pixel_grad(V, H, W)).cuda()
        res = {}
        for v in VARIANTS:
            with variant(v):
                views, fwd = Hh.product_forward(gc, cams, 3)
                grads = Hh.product_backward(gc, cams, views, fwd, dl, 3)
                nc = Hh.product_state("n_contrib", torch.int32, (V, H, W), V, P, W, H, fwd)
                fT = Hh.product_state("final_T", torch.float32, (V, H, W), V, P, W, H, fwd)
                res[v] = (fwd[1].clone(), fwd[2].clone(), nc, fT, {k: t.clone() for k, t in grads.items()})
        for v in VARIANTS[1:]:
            assert torch.equal(res[v][0], res[0][0]), f"{name}: colour of variant {v}"
            assert torch.equal(res[v][1], res[0][1]), f"{name}: depth of variant {v}"
            assert torch.equal(res[v][2], res[0][2]), f"{name}: n_contrib of variant {v}"
            assert torch.equal(res[v][3], res[0][3]), f"{name}: final_T of variant {v}"
            for k in ("means2D", "means3D", "opacities", "shs", "scales", "rotations", "cov3D"):
                rel = Hh.rel_err(res[v][4][k].cpu().numpy(), res[0][4][k].cpu().numpy())
                # dL/dcov3D collects float atomics in a run-dependent order (two runs of the reference itself differ by ~1e-4
                # there); dL/dscales and dL/drotations are linear images of it and inherit that spread (measured between two
                # variants, run to run: 0.3 .. 1.13e-4)
                assert rel < (3e-4 if k in ("cov3D", "scales", "rotations") else 1e-4), f"{name}: grad {k} of variant {v}: rel {rel}"


@pytest.mark.parametrize("v", [0x00, 0x22, 0x43, 0x64])  # the default variant is what every other test runs
def test_variant_vs_reference_cuda(v):
    from oracle import ref_cuda

    if not ref_cuda.available():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    import test_rasterizer_gpu as T

    with variant(v):
        g, cam = Hh.small_scene(seed=13)
        T._compare_with_reference(g, cam)
        g, cams = S.config_c2(4, 800)
        T._compare_with_reference({k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")}, cams[2])


@contextlib.contextmanager
def cull_exact(on):
    import csb_cabi

    lib = csb_cabi.lib()
    lib.csb_debug_cull_exact.restype = ctypes.c_int
    lib.csb_debug_cull_exact.argtypes = [ctypes.c_int]
    old = lib.csb_debug_cull_exact(on)
    try:
        yield
    finally:
        lib.csb_debug_cull_exact(old)


def _needles():
    """Needle-shaped footprints (axis ratios up to 1 : 200, every orientation, opacities up to 1) in front of the C3 camera:
    the conics whose terms cancel most in q = a dx^2 + 2 b dx dy + c dy^2, i.e. the hardest case for a culling test that
    must never drop a quad in which a pixel blends."""
    g = S.random_gaussians(60_000, seed=77, log_scale=(0.001, 0.2))
    g["opacities"] = np.random.default_rng(78).uniform(0.004, 1.0, g["opacities"].shape).astype(np.float32)  # some below 1/255
    return g, [S.front_camera(640, 360, 1.2)]


def test_exact_quad_culling_changes_nothing():
    """csb_debug_cull_exact: the exact ellipse-vs-quad test only removes visits in which no pixel blends, so every
    per-pixel output is bit-identical with it on and off (and the gradients agree to the atomics' association)."""
    scenes = list(_scenes())
    g, cams = _needles()
    scenes.append(("needles", g, cams))
    for name, g, cams in scenes:
        gc = Hh.to_cuda(g)
        V, P = len(cams), g["means3D"].shape[0]
        H, W = cams[0].image_height, cams[0].image_width
        dl = torch.from_numpy(S.pixel_grad(V, H, W)).cuda()
        def run(on, v):
            with cull_exact(on), variant(v):
                views, fwd = Hh.product_forward(gc, cams, 3)
                grads = Hh.product_backward(gc, cams, views, fwd, dl, 3)
                nc = Hh.product_state("n_contrib", torch.int32, (V, H, W), V, P, W, H, fwd)
                fT = Hh.product_state("final_T", torch.float32, (V, H, W), V, P, W, H, fwd)
                return (fwd[1].clone(), fwd[2].clone(), nc, fT, {k: t.clone() for k, t in grads.items()})

        ref = run(0, 0x43)
        again = run(0, 0x43)   # the same launch twice: how far the float atomics of the backward move a gradient by themselves
        GRADS = ("means2D", "means3D", "opacities", "shs", "scales", "rotations", "cov3D")
        floor = {k: Hh.rel_err(again[4][k].cpu().numpy(), ref[4][k].cpu().numpy()) for k in GRADS}
        for key in ((0, 0x44), (1, 0x43), (1, 0x44)):       # one / two pixel pairs per lane: both quad layouts
            r = run(*key)
            for i, what in enumerate(("colour", "depth", "n_contrib", "final_T")):
                assert torch.equal(r[i], ref[i]), f"{name}: {what} with (cull, variant) = {key}"
            for k in GRADS:
                rel = Hh.rel_err(r[4][k].cpu().numpy(), ref[4][k].cpu().numpy())
                # the needles' dL/dscales cancels to ~1e-3 of its terms (two identical launches: 5.5e-4 apart on a B200): ten times
                # the usual bound there, and never less than four times the spread measured in this run
                bound = max((3e-4 if k in ("cov3D", "scales", "rotations") else 1e-4) * (10 if name == "needles" else 1), 4 * floor[k])
                assert rel < bound, f"{name}: grad {k} with {key}: rel {rel} (run-to-run {floor[k]})"
