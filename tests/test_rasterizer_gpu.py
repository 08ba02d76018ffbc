"""Parity of the CUDA rasterizer (through the C-ABI) with the CPU oracle and, where
oracle/_ref/libref_cuda.so travelled to the box, with the reference's own CUDA code.

Tolerances (BASELINE.json north_star): sort keys, tile ranges, radii bit-exact; forward colour / depth
within 1e-5 absolute; gradients within 1e-4 relative (atomic ordering differs)."""
import numpy as np
import pytest
import torch

import conftest  # noqa: F401
import csb_synth as S
import helpers as Hh

pytestmark = pytest.mark.gpu

COLOR_ATOL = 1e-5
GRAD_RTOL = 1e-4


def _grad_tol(name):
    # dL/dcov3D (an intermediate: cloth-splatting consumes dL/dscales, dL/drotations) sums ~1e3
    # sign-alternating per-pixel terms per Gaussian in fp32; two runs of the REFERENCE differ by ~1e-4
    # through atomic ordering alone, so it gets 3x the 1e-4 budget.
    return 3 * GRAD_RTOL if name == "cov3D" else GRAD_RTOL


def _ref_available():
    from oracle import ref_cuda

    return ref_cuda.available()


def _check_vs_oracle(g, cam, sh_degree=3, bg=(1.0, 1.0, 1.0), int_slack=2e-3):
    """The gcc-built oracle can round differently from nvcc for inputs sitting exactly on a boundary,
    so integer outputs are allowed `int_slack` mismatching Gaussians here; the exact check is the one
    against the reference CUDA below."""
    gc = Hh.to_cuda(g)
    P = g["means3D"].shape[0]
    H, W = cam.image_height, cam.image_width
    s = Hh.oracle_forward(g, cam, bg, sh_degree)
    views, fwd = Hh.product_forward(gc, cam, sh_degree, bg)
    R, color, depth, radii, geom, binning, img = fwd
    radii_np = radii[0].cpu().numpy()
    n_bad = int((radii_np != s.radii).sum())
    assert n_bad <= max(1, int(int_slack * P)), f"radii mismatches: {n_bad}"
    exact = n_bad == 0 and R == s.num_rendered
    if exact:
        pl = Hh.product_state("point_list", torch.int32, (R,), 1, P, W, H, fwd).cpu().numpy().astype(np.uint32)
        keys = Hh.product_state("keys", torch.int64, (R,), 1, P, W, H, fwd).cpu().numpy().view(np.uint64)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        rg = Hh.product_state("ranges", torch.int32, (tiles, 2), 1, P, W, H, fwd).cpu().numpy().astype(np.uint32)
        assert np.array_equal(keys, s.keys)
        assert np.array_equal(pl, s.point_list)
        assert np.array_equal(rg, s.ranges)
    # libm's expf (oracle) and CUDA's expf differ in the last ulp, which can flip an alpha >= 1/255 or a
    # T < 1e-4 test for a handful of (pixel, Gaussian) pairs; such a pixel is off by up to ~alpha*T.  The
    # zero-tolerance comparison is the one against the reference's own CUDA code below.
    err = np.abs(color[0].cpu().numpy() - s.color)
    assert (err > COLOR_ATOL).mean() <= (2e-4 if exact else 2e-3), f"colour max err {err.max()}"
    assert err.max() < 5e-3
    derr = np.abs(depth[0].cpu().numpy() - s.depth)
    assert (derr > COLOR_ATOL * max(1.0, float(s.depths.max()))).mean() <= (2e-4 if exact else 2e-3)
    dl = torch.from_numpy(S.pixel_grad(1, H, W)).cuda()
    got = Hh.product_backward(gc, cam, views, fwd, dl, sh_degree)
    want = __import__("oracle").rasterize_backward(s, dl[0].cpu().numpy())
    for k in ("means3D", "opacities", "shs", "scales", "rotations", "cov3D", "means2D"):
        if want.get(k) is None or k not in got or got[k].numel() == 0:
            continue
        a = got[k][0] if k == "means2D" else got[k]
        rel = Hh.rel_err(a.cpu().numpy(), want[k])
        assert rel < (_grad_tol(k) if exact else 5e-3), f"grad {k}: rel {rel}"
    return exact


@pytest.mark.parametrize("seed", [11, 12])
def test_small_cloud_vs_oracle(seed):
    g, cam = Hh.small_scene(seed=seed)
    _check_vs_oracle(g, cam)


@pytest.mark.parametrize("deg", [0, 1, 2])
def test_lower_sh_degrees_vs_oracle(deg):
    g, cam = Hh.small_scene(P=1500, seed=20 + deg)
    _check_vs_oracle(g, cam, sh_degree=deg, bg=(0.0, 0.3, 0.7))


def test_ragged_image_size_vs_oracle():
    g, cam = Hh.small_scene(P=2000, W=149, H=83, seed=31)  # neither dimension is a multiple of 16
    _check_vs_oracle(g, cam)


def test_mesh_scene_vs_oracle():
    g, cam = Hh.mesh_scene(res=208, view=1)
    _check_vs_oracle({k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")}, cam)


def _compare_with_reference(g, cam, sh_degree=3, bg=(1.0, 1.0, 1.0)):
    from oracle import ref_cuda

    gc = Hh.to_cuda(g)
    P = g["means3D"].shape[0]
    H, W = cam.image_height, cam.image_width
    t = Hh.cam_tensors(cam, bg)
    ref = ref_cuda.RefRasterizer()
    R_ref, c_ref, radii_ref, d_ref = ref.forward(
        gc["means3D"], gc["opacities"], t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"], cam.tanfovx,
        cam.tanfovy, H, W, shs=gc.get("shs"), colors_precomp=gc.get("colors_precomp"), scales=gc.get("scales"),
        rotations=gc.get("rotations"), cov3D_precomp=gc.get("cov3D_precomp"), sh_degree=sh_degree)
    views, fwd = Hh.product_forward(gc, cam, sh_degree, bg)
    R, color, depth, radii, geom, binning, img = fwd
    torch.cuda.synchronize()
    # bit-exact integers
    assert R == R_ref
    assert torch.equal(radii[0], radii_ref)
    if R:
        pl = Hh.product_state("point_list", torch.int32, (R,), 1, P, W, H, fwd)
        assert torch.equal(pl, ref.state("point_list", torch.int32, (R,)))
        keys = Hh.product_state("keys", torch.int64, (R,), 1, P, W, H, fwd)
        assert torch.equal(keys, ref.state("keys", torch.int64, (R,)))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    rg = Hh.product_state("ranges", torch.int32, (tiles, 2), 1, P, W, H, fwd)
    assert torch.equal(rg, ref.state("ranges", torch.int32, (tiles, 2)))
    # same keys, same lists, same expression tree, same expf: the last contributor of every pixel is the reference's
    nc = Hh.product_state("n_contrib", torch.int32, (H, W), 1, P, W, H, fwd)
    assert torch.equal(nc, ref.state("n_contrib", torch.int32, (H, W)))
    fT = Hh.product_state("final_T", torch.float32, (H, W), 1, P, W, H, fwd)
    assert torch.equal(fT, ref.state("accum_alpha", torch.float32, (H, W)))
    # colour / depth
    assert float((color[0] - c_ref).abs().max()) <= COLOR_ATOL
    zmax = max(1.0, float(gc["means3D"].abs().max()) * 4)
    assert float((depth[0] - d_ref).abs().max()) <= COLOR_ATOL * zmax
    # gradients
    dl = torch.from_numpy(S.pixel_grad(1, H, W)).cuda()
    g_ref = {k: (v.clone() if v is not None else None) for k, v in ref.backward(dl[0]).items()}
    # the reference accumulates with float atomics: its own run-to-run spread is part of the budget.  One extra run is a
    # noisy estimate of that spread (it made this test fail about once in 35 runs): the largest of three is used.
    g_refs = [{k: (v.clone() if v is not None else None) for k, v in ref.backward(dl[0]).items()} for _ in range(3)]
    got = Hh.product_backward(gc, cam, views, fwd, dl, sh_degree)
    for k in ("means2D", "opacities", "means3D", "cov3D", "shs", "scales", "rotations", "colors"):
        if g_ref.get(k) is None or got[k].numel() == 0:
            continue
        if k == "colors" and "colors_precomp" not in gc:
            continue
        if k in ("scales", "rotations") and "scales" not in gc:
            continue
        a = got[k][0] if k == "means2D" else got[k]
        rel = Hh.rel_err(a.cpu().numpy(), g_ref[k].cpu().numpy())
        spread = max(Hh.rel_err(r[k].cpu().numpy(), g_ref[k].cpu().numpy()) for r in g_refs)
        # two runs of the reference differ by up to ~5e-4 on the intermediate dL/dcov3D (float atomics over heavily
        # cancelling terms; measured 1.0e-4 .. 4.8e-4 over repeated runs): the spread is budget, capped so that a
        # broken reference build cannot hide a real error
        assert spread < 1e-3, f"reference run-to-run spread of grad {k}: {spread}"
        assert rel < _grad_tol(k) + 2 * spread, f"grad {k}: rel {rel} (reference vs itself: {spread})"


needs_ref = pytest.mark.skipif(not _ref_available(), reason="oracle/_ref/libref_cuda.so not built")


@needs_ref
@pytest.mark.parametrize("seed", [11, 13])
def test_small_cloud_vs_reference_cuda(seed):
    g, cam = Hh.small_scene(seed=seed)
    _compare_with_reference(g, cam)


@needs_ref
def test_c2_full_size_vs_reference_cuda():
    g, cams = S.config_c2(4, 800)
    gg = {k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    for v in (0, 3):
        _compare_with_reference(gg, cams[v])


@needs_ref
def test_c3_100k_vs_reference_cuda():
    g, cams = S.config_c3(100_000, 1920, 1080)
    _compare_with_reference(g, cams[0])


@needs_ref
def test_precomputed_colors_and_cov3d_vs_reference_cuda():
    g, cam = Hh.small_scene(P=2500, seed=41)
    s = Hh.oracle_forward(g, cam)
    gg = dict(means3D=g["means3D"], opacities=g["opacities"], colors_precomp=np.abs(g["shs"][:, 0, :]).copy(),
              cov3D_precomp=s.cov3D.copy())
    _compare_with_reference(gg, cam)


@needs_ref
def test_c4_launch_set_vs_reference_cuda():
    """One 16-view launch set of C4 (100 k Gaussians, 800x800; BASELINE.json configs[3]): V*P = 1.6 M entries and 40 000 tile
    ids go through the global two-sort path (depth sort of the entries, chained-scan emit, two onesweep passes over the
    16 tile bits).  Every view against the reference's own forward / backward of that view
    (rasterizer_impl.cu:198-339): radii, keys, point list, ranges, n_contrib bit for bit, colour / depth <= 1e-5, the
    gradients summed over the 16 views <= 1e-4 relative (+ the reference's own run-to-run spread)."""
    from oracle import ref_cuda

    g, cams = S.config_c4(100_000, 64, 800)
    cams = cams[16:32]
    g = {k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    gc = Hh.to_cuda(g)
    V, P, H, W = 16, 100_000, 800, 800
    tiles = 50 * 50
    views, fwd = Hh.product_forward(gc, cams, 3)
    R, color, depth, radii, geom, binning, img = fwd
    pl = Hh.product_state("point_list", torch.int32, (R,), V, P, W, H, fwd)
    keys = Hh.product_state("keys", torch.int64, (R,), V, P, W, H, fwd)
    rg = Hh.product_state("ranges", torch.int32, (V * tiles, 2), V, P, W, H, fwd).long()
    nc = Hh.product_state("n_contrib", torch.int32, (V, H, W), V, P, W, H, fwd)
    dl = torch.from_numpy(S.pixel_grad(V, H, W)).cuda()
    got = Hh.product_backward(gc, cams, views, fwd, dl, 3)
    ref = ref_cuda.RefRasterizer()
    tot, tot2, offset = None, None, 0
    for v, cam in enumerate(cams):
        t = Hh.cam_tensors(cam)
        R_ref, c_ref, radii_ref, d_ref = ref.forward(
            gc["means3D"], gc["opacities"], t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"], cam.tanfovx,
            cam.tanfovy, H, W, shs=gc["shs"], scales=gc["scales"], rotations=gc["rotations"], sh_degree=3)
        assert torch.equal(radii[v], radii_ref), f"view {v}: radii"
        assert torch.equal(pl[offset:offset + R_ref] - v * P, ref.state("point_list", torch.int32, (R_ref,))), v
        assert torch.equal(keys[offset:offset + R_ref], ref.state("keys", torch.int64, (R_ref,))), v
        rr = ref.state("ranges", torch.int32, (tiles, 2)).long()
        mine = rg[v * tiles:(v + 1) * tiles]
        mine = torch.where((mine[:, 1:] > mine[:, :1]).expand(-1, 2), mine - offset, mine)  # empty tiles stay (0, 0)
        assert torch.equal(mine, rr), f"view {v}: ranges"
        assert torch.equal(nc[v], ref.state("n_contrib", torch.int32, (H, W))), f"view {v}: n_contrib"
        assert float((color[v] - c_ref).abs().max()) <= COLOR_ATOL, v
        assert float((depth[v] - d_ref).abs().max()) <= COLOR_ATOL * 4, v
        g1 = ref.backward(dl[v])
        assert Hh.rel_err(got["means2D"][v].cpu().numpy(), g1["means2D"].cpu().numpy()) < 2 * GRAD_RTOL, v
        g2 = ref.backward(dl[v])
        if tot is None:
            tot = {k: g1[k].clone() for k in ("means3D", "opacities", "shs", "scales", "rotations")}
            tot2 = {k: g2[k].clone() for k in tot}
        else:
            for k in tot:
                tot[k] += g1[k]
                tot2[k] += g2[k]
        offset += R_ref
    assert offset == R
    for k in tot:
        spread = Hh.rel_err(tot2[k].cpu().numpy(), tot[k].cpu().numpy())
        rel = Hh.rel_err(got[k].cpu().numpy(), tot[k].cpu().numpy())
        assert spread < 1e-3 and rel < GRAD_RTOL + 2 * spread, f"grad {k}: rel {rel} (reference vs itself: {spread})"


@pytest.mark.parametrize("n_views", [4, 3, 9])  # 9: view lanes of the split per-Gaussian backward walk two views
def test_multi_view_batch_equals_sequential(n_views):
    g, cams = S.config_c2(n_views, 320)
    gg = Hh.to_cuda({k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")})
    H = W = 320
    dl = torch.from_numpy(S.pixel_grad(n_views, H, W)).cuda()
    views, fwd = Hh.product_forward(gg, cams, 3)
    gb = Hh.product_backward(gg, cams, views, fwd, dl, 3)
    tot = None
    for v, cam in enumerate(cams):
        v1, f1 = Hh.product_forward(gg, cam, 3)
        assert torch.equal(f1[1][0], fwd[1][v])  # same pixels, bit for bit
        assert torch.equal(f1[2][0], fwd[2][v])
        assert torch.equal(f1[3][0], fwd[3][v])
        g1 = Hh.product_backward(gg, cam, v1, f1, dl[v:v + 1], 3)
        assert Hh.rel_err(gb["means2D"][v].cpu().numpy(), g1["means2D"][0].cpu().numpy()) < GRAD_RTOL
        if tot is None:
            tot = {k: g1[k].clone() for k in g1 if k != "means2D"}
        else:
            for k in tot:
                tot[k] += g1[k]
    for k in ("means3D", "opacities", "shs", "scales", "rotations"):
        assert Hh.rel_err(gb[k].cpu().numpy(), tot[k].cpu().numpy()) < GRAD_RTOL, k


def test_full_size_c3_properties():
    """1M Gaussians at 1920x1080: size-independent properties (too big for the CPU oracle)."""
    g, cams = S.config_c3(1_000_000, 1920, 1080)
    gc = Hh.to_cuda(g)
    cam = cams[0]
    H, W, P = 1080, 1920, 1_000_000
    views, fwd = Hh.product_forward(gc, cam, 3)
    R, color, depth, radii, geom, binning, img = fwd
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    rg = Hh.product_state("ranges", torch.int32, (tiles, 2), 1, P, W, H, fwd).long()
    assert int((rg[:, 1] - rg[:, 0]).sum()) == R
    keys = Hh.product_state("keys", torch.int64, (R,), 1, P, W, H, fwd)
    assert bool((keys[1:] >= keys[:-1]).all())
    assert int(keys.max() >> 32) < tiles
    tt = Hh.product_state("tiles_touched", torch.int32, (P,), 1, P, W, H, fwd)
    assert int(tt.long().sum()) == R
    assert torch.equal(tt > 0, radii[0] > 0)
    pl = Hh.product_state("point_list", torch.int32, (R,), 1, P, W, H, fwd).long()
    rd = Hh.product_state("rgb_depth", torch.float32, (P, 4), 1, P, W, H, fwd)
    assert torch.equal(keys & 0xFFFFFFFF, rd[:, 3].contiguous().view(torch.int32)[pl].long() & 0xFFFFFFFF)
    same = keys[1:] == keys[:-1]
    assert bool((pl[1:][same] > pl[:-1][same]).all())  # stable
    assert torch.isfinite(color).all() and torch.isfinite(depth).all()
    fT = Hh.product_state("final_T", torch.float32, (H, W), 1, P, W, H, fwd)
    assert float(fT.min()) >= 0 and float(fT.max()) <= 1
    # linearity of the backward pass in dL/dpixel
    dl = torch.from_numpy(S.pixel_grad(1, H, W)).cuda()
    g1 = Hh.product_backward(gc, cam, views, fwd, dl, 3)
    g2 = Hh.product_backward(gc, cam, views, fwd, 2 * dl, 3)
    for k in ("means3D", "opacities", "shs", "scales", "rotations"):
        assert Hh.rel_err(g2[k].cpu().numpy(), 2 * g1[k].cpu().numpy()) < GRAD_RTOL, k


def test_empty_and_fully_culled_inputs():
    cam = S.front_camera(64, 48, 1.0)
    import diff_gaussian_rasterization as dgr

    settings = Hh.settings_for(cam, bg=(0.2, 0.4, 0.6))
    rast = dgr.GaussianRasterizer(settings)
    # P = 0 (rasterize_points.cu:82 short-circuit): nothing is launched, outputs stay zero (no background)
    z = lambda *s: torch.zeros(*s, device="cuda")
    color, radii, depth = rast(means3D=z(0, 3), means2D=z(0, 3), opacities=z(0, 1), shs=z(0, 16, 3), scales=z(0, 3),
                               rotations=z(0, 4))
    assert radii.numel() == 0 and color.shape == (3, 48, 64) and float(color.abs().max()) == 0
    # everything behind the camera
    g, _ = Hh.small_scene(P=500, seed=3)
    gc = Hh.to_cuda(g)
    gc["means3D"][:, 2] = -gc["means3D"][:, 2].abs() - 1
    color, radii, depth = rast(means3D=gc["means3D"], means2D=torch.zeros_like(gc["means3D"]),
                               opacities=gc["opacities"], shs=gc["shs"], scales=gc["scales"],
                               rotations=gc["rotations"])
    assert int(radii.abs().sum()) == 0
    assert torch.allclose(color[:, 5, 7], torch.tensor([0.2, 0.4, 0.6], device="cuda"))
    assert float(depth.abs().max()) == 0


def test_capacity_overflow_regrows():
    import diff_gaussian_rasterization as dgr

    g, cam = Hh.small_scene(P=3000, seed=17)
    gc = Hh.to_cuda(g)
    views, fwd_a = Hh.product_forward(gc, cam, 3)
    dgr._capacity_hint.clear()
    saved = dgr._MIN_CAPACITY, dgr._PER_GAUSSIAN
    try:
        dgr._MIN_CAPACITY, dgr._PER_GAUSSIAN = 16, 0  # force the first attempt to overflow
        views, fwd_b = Hh.product_forward(gc, cam, 3)
    finally:
        dgr._MIN_CAPACITY, dgr._PER_GAUSSIAN = saved
    assert fwd_a[0] == fwd_b[0]
    assert torch.equal(fwd_a[1], fwd_b[1])


def test_api_surface_and_autograd():
    import diff_gaussian_rasterization as dgr

    g, cam = Hh.small_scene(P=800, seed=19)
    gc = Hh.to_cuda(g)
    settings = Hh.settings_for(cam)
    assert settings._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier",
                                "viewmatrix", "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    rast = dgr.GaussianRasterizer(raster_settings=settings)
    m2 = torch.zeros_like(gc["means3D"], requires_grad=True)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        rast(means3D=gc["means3D"], means2D=m2, opacities=gc["opacities"], scales=gc["scales"],
             rotations=gc["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair"):
        rast(means3D=gc["means3D"], means2D=m2, opacities=gc["opacities"], shs=gc["shs"])
    leaves = {k: gc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    color, radii, depth = rast(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"],
                               shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
    assert color.shape == (3, cam.image_height, cam.image_width) and depth.shape == (1, cam.image_height, cam.image_width)
    assert radii.dtype == torch.int32 and radii.shape == (800,)
    assert not radii.requires_grad and not depth.requires_grad or depth.grad_fn is not None
    (color.sum() + 0 * depth.sum()).backward()  # grad_depth is dropped like the reference does
    for k, t in leaves.items():
        assert t.grad is not None and t.grad.shape == t.shape and torch.isfinite(t.grad).all(), k
    assert m2.grad is not None and float(m2.grad[:, 2].abs().max()) == 0
    vis = rast.markVisible(gc["means3D"])
    assert vis.dtype == torch.bool and torch.equal(vis, gc["means3D"][:, 2] > 0.2)


def test_deferred_count_matches_eager_and_reports_overflow():
    """defer_count=True: the forward returns without reading the instance count; .wait() gives the eager path's count,
    the images are the same, and a workspace that is too small surfaces as RenderAgain (then fits on the next call)."""
    import diff_gaussian_rasterization as dgr

    g, cam = Hh.small_scene(P=4000, W=176, H=128, seed=21)
    gc = Hh.to_cuda(g)
    e = torch.Tensor([])
    args = ([Hh.view_tuple(cam)], gc["means3D"], e, gc["opacities"], gc["scales"], gc["rotations"], 1.0, e,
            cam.image_height, cam.image_width, gc["shs"], 3, False)
    eager = dgr._forward_views(*args)
    key = (gc["means3D"].device, 1, 4000, cam.image_width, cam.image_height)
    deferred = dgr._forward_views(*args, defer_count=True)
    assert isinstance(deferred[0], dgr.PendingCount)
    assert deferred[0].wait() == eager[0] == int(deferred[0])
    for a, b in zip(eager[1:4], deferred[1:4]):
        assert torch.equal(a, b)
    # force a workspace that is too small: pretend the previous frame had almost no instances
    saved = (dgr._PER_GAUSSIAN, dgr._MIN_CAPACITY)
    try:
        dgr._PER_GAUSSIAN, dgr._MIN_CAPACITY = 0, 16
        dgr._capacity_hint.pop(key, None)
        small = dgr._forward_views(*args, defer_count=True)
        with pytest.raises(dgr.RenderAgain):
            small[0].wait()
        again = dgr._forward_views(*args, defer_count=True)  # the hint now holds the real count
        assert again[0].wait() == eager[0]
        assert torch.equal(again[1], eager[1])
    finally:
        dgr._PER_GAUSSIAN, dgr._MIN_CAPACITY = saved


def test_staged_pixel_grads_backward_matches_plain():
    """dL/dpixel uploaded view by view on a side stream (StagedPixelGrads) and consumed by backward_staged: the same
    gradients as color.backward(dl) with the whole tensor resident (the sums over pixels are atomic in both, hence the
    tolerance of the unordered float sums)."""
    import diff_gaussian_rasterization as dgr

    g = S.mesh_gaussians(8, 14, 1)
    center = g["verts"].mean(0)
    cams = S.ring_cameras(3, 208, 176, 0.6911, center)
    gc = Hh.to_cuda({k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")})
    settings = [Hh.settings_for(c) for c in cams]
    dl_host = torch.from_numpy(S.pixel_grad(3, 176, 208, seed=3)).pin_memory()

    def run(staged_mode):
        leaves = {k: gc[k].clone().requires_grad_(True) for k in gc}
        color, radii, depth = dgr.rasterize_views(settings, leaves["means3D"], leaves["opacities"], shs=leaves["shs"],
                                                  scales=leaves["scales"], rotations=leaves["rotations"])
        if staged_mode:
            dev = torch.empty_like(color)
            staged = dgr.StagedPixelGrads(dev, torch.cuda.Stream()).upload(dl_host)
            dgr.backward_staged(color, staged)
        else:
            color.backward(dl_host.cuda())
        torch.cuda.synchronize()
        return {k: t.grad.clone() for k, t in leaves.items()}

    plain, staged = run(False), run(True)
    assert dgr._staged.grads is None
    for k in plain:
        scale = float(plain[k].abs().max()) + 1e-20
        assert float((plain[k] - staged[k]).abs().max()) <= 2e-5 * scale, k
        assert float(plain[k].abs().max()) > 0
