"""CPU oracle: known-answer tests and invariants derived from the reference's code."""
import math

import numpy as np
import pytest

import conftest  # noqa: F401
import csb_synth as S
import oracle


def one_gaussian(px, py, W, H, z=3.0, scale=0.05, opacity=0.7, rgb=(0.2, 0.5, 0.9)):
    """A Gaussian whose centre projects exactly on pixel (px, py) of a camera at the origin."""
    cam = S.front_camera(W, H, 1.0)
    fx = W / (2 * cam.tanfovx)
    fy = H / (2 * cam.tanfovy)
    # pixel = ((ndc + 1) * S - 1) / 2  ->  ndc = (2 px + 1) / S - 1 ;  ndc = x / (z tan)
    x = ((2 * px + 1) / W - 1) * cam.tanfovx * z
    y = ((2 * py + 1) / H - 1) * cam.tanfovy * z
    g = dict(means3D=np.array([[x, y, z]], np.float32), scales=np.full((1, 3), scale, np.float32),
             rotations=np.array([[1, 0, 0, 0]], np.float32), opacities=np.array([[opacity]], np.float32),
             colors=np.array([rgb], np.float32))
    return g, cam, fx, fy


def run(g, cam, bg=(1, 1, 1), **kw):
    return oracle.rasterize_forward(g["means3D"], g["opacities"], cam.world_view_transform, cam.full_proj_transform,
                                    cam.camera_center, np.asarray(bg, np.float32), cam.tanfovx, cam.tanfovy,
                                    cam.image_height, cam.image_width, shs=g.get("shs"),
                                    colors_precomp=g.get("colors"), scales=g.get("scales"),
                                    rotations=g.get("rotations"), cov3D_precomp=g.get("cov3D"), **kw)


def test_single_gaussian_known_answer():
    W, H = 64, 48
    g, cam, fx, fy = one_gaussian(20, 17, W, H)
    s = run(g, cam, bg=(0.1, 0.2, 0.3))
    assert s.radii[0] > 0
    assert abs(s.means2D[0, 0] - 20) < 1e-3 and abs(s.means2D[0, 1] - 17) < 1e-3
    a = min(0.99, 0.7)  # exp(0) at the centre
    got = s.color[:, 17, 20]
    want = np.array([0.2, 0.5, 0.9]) * a + (1 - a) * np.array([0.1, 0.2, 0.3])
    assert np.allclose(got, want, atol=2e-4)
    assert abs(s.depth[0, 17, 20] - 3.0 * a) < 1e-3  # un-normalised alpha-weighted depth, no background
    assert s.n_contrib[17, 20] == 1
    assert abs(s.final_T[17, 20] - (1 - a)) < 2e-4
    # far from the Gaussian only the background is left
    assert np.allclose(s.color[:, 0, W - 1], [0.1, 0.2, 0.3], atol=1e-6)
    # radius: 3 sigma of cov2D = (fx * scale / z)^2 + 0.3
    lam = (fx * 0.05 / 3.0) ** 2 + 0.3
    assert s.radii[0] == math.ceil(3 * math.sqrt(lam))


def test_near_plane_cull_and_alpha_clamp():
    W, H = 32, 32
    g, cam, _, _ = one_gaussian(10, 10, W, H, z=0.2)  # view z <= 0.2 is culled (auxiliary.h:152-162)
    s = run(g, cam)
    assert s.radii[0] == 0 and s.num_rendered == 0
    assert np.allclose(s.color, 1.0)
    g, cam, _, _ = one_gaussian(10, 10, W, H, opacity=1.0)
    s = run(g, cam, bg=(0, 0, 0))
    assert abs(s.final_T[10, 10] - 0.01) < 1e-6  # alpha is clamped at 0.99


def test_binning_invariants():
    g = S.random_gaussians(4000, 3, box=((-1.2, 1.2), (-0.9, 0.9), (-0.5, 5.0)), log_scale=(0.01, 0.15))
    cam = S.front_camera(200, 136, 1.0)
    s = run(g, cam)
    R = s.num_rendered
    gx, gy = (200 + 15) // 16, (136 + 15) // 16
    assert R == int(s.tiles_touched.sum()) > 0
    assert int((s.ranges[:, 1] - s.ranges[:, 0]).sum()) == R
    assert np.all(np.diff(s.keys.astype(np.uint64)) >= 0) if R > 1 else True
    assert np.all(s.keys >> np.uint64(32) < gx * gy)
    assert np.array_equal(s.radii > 0, s.tiles_touched > 0)
    # the low word of a key is the depth bit pattern of its Gaussian
    assert np.array_equal((s.keys & np.uint64(0xFFFFFFFF)).astype(np.uint32), s.depths.view(np.uint32)[s.point_list])
    # stable: equal keys keep ascending Gaussian index
    same = s.keys[1:] == s.keys[:-1]
    assert np.all(s.point_list[1:][same] > s.point_list[:-1][same])
    # colour minus the background term is the blended sum (>= 0), T in [0, 1]
    assert s.final_T.min() >= 0 and s.final_T.max() <= 1
    assert oracle.higher_msb(2500) == 12 and oracle.higher_msb(8160) == 13 and oracle.higher_msb(1) == 1


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_backward_matches_finite_differences(deg):
    """Directional derivative of L = sum(color * dl) against the analytic backward, on a smooth scene:
    every Gaussian covers the whole image with alpha > 1/255, so neither tile membership nor the
    alpha threshold (both discontinuities of the forward pass) flips under the step."""
    rng = np.random.default_rng(5 + deg)
    P, W, H = 10, 32, 32
    g = dict(
        means3D=np.stack([rng.uniform(-0.8, 0.8, P), rng.uniform(-0.8, 0.8, P), rng.uniform(2.8, 3.5, P)], 1).astype(np.float32),
        scales=rng.uniform(2.2, 4.0, (P, 3)).astype(np.float32),
        rotations=(lambda q: (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32))(rng.standard_normal((P, 4))),
        opacities=rng.uniform(0.2, 0.45, (P, 1)).astype(np.float32),
        shs=(rng.standard_normal((P, 16, 3)) * 0.2 + 0.1).astype(np.float32))
    cam = S.front_camera(W, H, 1.0)
    dl = rng.standard_normal((3, H, W)).astype(np.float32)
    s = run(g, cam, sh_degree=deg)
    assert s.counts[2] == P * W * H  # every pair is blended: the scene is in the smooth regime
    grads = oracle.rasterize_backward(s, dl)

    def loss(gg):
        ss = run(gg, cam, sh_degree=deg)
        return float((ss.color.astype(np.float64) * dl).sum())

    for name, eps in (("means3D", 2e-3), ("scales", 2e-3), ("rotations", 2e-3), ("opacities", 2e-3), ("shs", 1e-2)):
        d = rng.standard_normal(g[name].shape).astype(np.float32)
        d /= np.linalg.norm(d)
        plus, minus = dict(g), dict(g)
        plus[name] = g[name] + eps * d
        minus[name] = g[name] - eps * d
        fd = (loss(plus) - loss(minus)) / (2 * eps)
        an = float((grads[name].reshape(d.shape).astype(np.float64) * d).sum())
        assert abs(fd - an) <= 0.03 * max(abs(an), abs(fd)) + 0.02, (name, fd, an)


def test_dist2_lattice_known_answer():
    h = 0.25
    ax = np.arange(6) * h
    pts = np.stack(np.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    d = oracle.dist2(pts)
    assert np.allclose(d, h * h, rtol=1e-6)  # every point has >= 3 neighbours at distance h
    # duplicates count as neighbours at distance 0
    pts2 = np.concatenate([pts, pts[:1]], 0)
    d2 = oracle.dist2(pts2)
    assert abs(d2[0] - (0 + h * h + h * h) / 3) < 1e-7 and abs(d2[-1] - d2[0]) < 1e-9


def test_dist2_tiny_clouds_keep_flt_max():
    # fewer than 4 points: the reference's best[] keeps FLT_MAX entries (simple_knn.cu:154,181)
    d = oracle.dist2(np.array([[0, 0, 0], [1, 0, 0]], np.float32))
    assert np.all(d > 1e37)
