"""Census of the compositing kernels' visit culling on the oracle's tile lists (CPU only):
    python tools/cull_census.py [c2|c3|c4|needles ...]
For every (Gaussian, tile) entry of a view, the four 8x8 pixel quads of the tile are tested the way
render_lanes.cuh::quad_mask does it (fp32 arithmetic restated in numpy): bounding box of the alpha >= 1/255 ellipse, then
the exact ellipse-vs-quad test.  Prints how many (Gaussian, quad) visits each test leaves, and checks on every entry that a
quad the exact test drops holds NO pixel that blends (alpha >= 1/255 with the reference's expression, forward.cu:330-345)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "cloth-splatting_b200"))
import csb_synth as S  # noqa: E402
import oracle  # noqa: E402

f32 = np.float32


def masks(xy, co, tx0, ty0):
    """(bounding-box mask, exact mask) per entry, bits 2*qy+qx -- quad_mask() in numpy fp32."""
    a, b, c, o = co[:, 0], co[:, 1], co[:, 2], co[:, 3]
    n = xy.shape[0]
    with np.errstate(all="ignore"):
        live = o * f32(255) >= f32(0.999)
        det = a * c - b * b
        odd = ~((det > 0) & (a > 0) & (c > 0))
        t = f32(2) * np.log(np.minimum(o, f32(1e30)) * f32(255)).astype(f32) * f32(1.001) + f32(1e-3)
        inv = t / det
        ex = np.sqrt(inv * c).astype(f32) * f32(1.001) + f32(0.02)
        ey = np.sqrt(inv * a).astype(f32) * f32(1.001) + f32(0.02)
        odd |= ~((ex < f32(1e30)) & (ey < f32(1e30)))
        x_lo, x_hi = xy[:, 0] - ex - tx0, xy[:, 0] + ex - tx0
        y_lo, y_hi = xy[:, 1] - ey - ty0, xy[:, 1] + ey - ty0
        cols = ((x_lo <= 7) & (x_hi >= 0)).astype(np.uint32) | (((x_lo <= 15) & (x_hi >= 8)).astype(np.uint32) << 1)
        m = np.where((y_lo <= 7) & (y_hi >= 0), cols, 0) | np.where((y_lo <= 15) & (y_hi >= 8), cols << 2, 0)
        m = np.where(odd, 0xF, m)
        m = np.where(live, m, 0).astype(np.uint32)
        rx, ry = tx0 - xy[:, 0], ty0 - xy[:, 1]
        nbc, nba = -b / c, -b / a
        X = np.maximum(np.abs(rx), np.abs(rx + f32(15)))
        Y = np.maximum(np.abs(ry), np.abs(ry + f32(15)))
        thr = t + f32(4e-6) * (a * X * X + c * Y * Y + f32(2) * np.abs(b) * X * Y)
        keep = np.zeros(n, np.uint32)
        for qy in (0, 1):
            for qx in (0, 1):
                bit = np.uint32(1 << (2 * qy + qx))
                x0 = rx + f32(8 * qx)
                x1 = x0 + f32(7)
                y0 = ry + f32(8 * qy)
                y1 = y0 + f32(7)
                inside = (x0 <= 0) & (x1 >= 0) & (y0 <= 0) & (y1 >= 0)
                qmin = np.full(n, f32(3e38))
                for xe, ye in ((x0, y0), (x1, y1)):
                    ys = np.minimum(np.maximum(nbc * xe, y0), y1)
                    qmin = np.fmin(qmin, a * xe * xe + (f32(2) * b * xe + c * ys) * ys)
                    xs = np.minimum(np.maximum(nba * ye, x0), x1)
                    qmin = np.fmin(qmin, c * ye * ye + (f32(2) * b * ye + a * xs) * xs)
                ok = inside | ~(qmin > thr)
                keep |= np.where(((m & bit) != 0) & ok, bit, 0).astype(np.uint32)
        keep = np.where(odd & live, 0xF, keep).astype(np.uint32)
    return m, keep


def blending_quads(xy, co, tx0, ty0):
    """bit q set iff some pixel of quad q has power <= 0 and alpha >= 1/255 (fp32, the reference's expression)."""
    n = xy.shape[0]
    out = np.zeros(n, np.uint32)
    with np.errstate(all="ignore"):
        for py in range(16):
            for px in range(16):
                dx = xy[:, 0] - (tx0 + f32(px))
                dy = xy[:, 1] - (ty0 + f32(py))
                power = f32(-0.5) * (co[:, 0] * dx * dx + co[:, 2] * dy * dy) - co[:, 1] * dx * dy
                alpha = np.minimum(f32(0.99), co[:, 3] * np.exp(power.astype(np.float64)).astype(f32))
                hit = ~(power > 0) & ~(alpha < f32(1.0 / 255.0))
                out |= np.where(hit, np.uint32(1 << (2 * (py // 8) + px // 8)), 0).astype(np.uint32)
    return out


def census(name, g, cam, chunk=2_000_000):
    H, W = cam.image_height, cam.image_width
    s = oracle.rasterize_forward(g["means3D"], g["opacities"], cam.world_view_transform, cam.full_proj_transform, cam.camera_center,
                                 np.zeros(3, f32), cam.tanfovx, cam.tanfovy, H, W, shs=g["shs"], scales=g["scales"], rotations=g["rotations"])
    gx = (W + 15) // 16
    tile = (s.keys >> np.uint64(32)).astype(np.int64)
    pid = s.point_list.astype(np.int64)
    n_box = n_exact = n_need = bad = 0
    for i in range(0, tile.shape[0], chunk):
        tl, pp = tile[i:i + chunk], pid[i:i + chunk]
        xy, co = s.means2D[pp], s.conic_opacity[pp]
        tx0, ty0 = ((tl % gx) * 16).astype(f32), ((tl // gx) * 16).astype(f32)
        m, keep = masks(xy, co, tx0, ty0)
        need = blending_quads(xy, co, tx0, ty0)
        pop = lambda v: int(np.unpackbits(v.astype(np.uint8)).sum())  # noqa: E731
        n_box += pop(m)
        n_exact += pop(keep)
        n_need += pop(need)
        bad += int(((need & ~keep) != 0).sum())
    print(f"{name}: {tile.shape[0]} entries; visits with the bounding box {n_box}, with the exact test {n_exact} ({n_exact / max(n_box, 1):.3f}), "
          f"quads with a blending pixel {n_need} ({n_need / max(n_box, 1):.3f}); entries with a blending quad dropped: {bad}")
    return bad


if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "c4", "c3"]
    bad = 0
    for w in which:
        if w == "c2":
            g, cams = S.config_c2()
            for v in (0, 2):
                bad += census(f"C2 view {v}", g, cams[v])
        elif w == "c4":
            g, cams = S.config_c4()
            bad += census("C4 view 0", g, cams[0])
        elif w == "needles":   # the adversarial scene of tests/test_render_variants_gpu.py (axis ratios up to 1 : 200)
            g = S.random_gaussians(60_000, seed=77, log_scale=(0.001, 0.2))
            g["opacities"] = np.random.default_rng(78).uniform(0.004, 1.0, g["opacities"].shape).astype(f32)
            bad += census("needles", g, S.front_camera(640, 360, 1.2))
        elif w == "c3":
            g, cams = S.config_c3(P=200_000)
            bad += census("C3 (200 k of the 1 M Gaussians)", g, cams[0])
    sys.exit(1 if bad else 0)
