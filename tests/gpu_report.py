"""Development aid (GPU box): compares product / reference CUDA / CPU oracle field by field and
times both GPU implementations.  Writes gpurun_out/report_<tag>.json.  Not a pytest file."""
import json
import os
import sys
import time
import traceback

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa: F401  (sets sys.path)
import csb_synth as S
import helpers as Hh
import oracle
from oracle import ref_cuda

OUT = os.path.join(conftest.ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def product_forward(g, cam, sh_degree=3, flags=1):
    import diff_gaussian_rasterization as dgr

    t = Hh.cam_tensors(cam)
    views = [(t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"], cam.tanfovx, cam.tanfovy)]
    e = torch.Tensor([])
    out = dgr._forward_views(views, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e,
                             cam.image_height, cam.image_width, g["shs"], sh_degree, False, flags=flags)
    return views, out


def product_state(name, dtype, shape, V, P, W, H, geom, binning, img):
    import csb_cabi

    lib = csb_cabi.lib()
    cap = lib.csb_rasterize_binning_capacity(V, P, W, H, binning.numel())
    ptr = lib.csb_rasterize_state_ptr(name.encode(), V, P, W, H, cap, geom.data_ptr(), img.data_ptr(),
                                      binning.data_ptr())
    n = int(np.prod(shape))
    out = torch.empty(shape, dtype=dtype, device="cuda")
    if n:
        import ctypes

        rt = ctypes.CDLL("libcudart.so.12")
        rc = rt.cudaMemcpy(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(ptr), ctypes.c_size_t(n * out.element_size()),
                           ctypes.c_int(3))
        assert rc == 0, rc
    return out


def compare_scene(tag, g_np, cam, with_oracle=True, sh_degree=3):
    import diff_gaussian_rasterization as dgr

    rep = {"tag": tag}
    g = Hh.to_cuda(g_np)
    P = g["means3D"].shape[0]
    H, W = cam.image_height, cam.image_width
    t = Hh.cam_tensors(cam)
    ref = ref_cuda.RefRasterizer()
    R_ref, c_ref, radii_ref, d_ref = ref.forward(g["means3D"], g["opacities"], t["viewmatrix"], t["projmatrix"],
                                                 t["campos"], t["bg"], cam.tanfovx, cam.tanfovy, H, W, shs=g["shs"],
                                                 scales=g["scales"], rotations=g["rotations"], sh_degree=sh_degree)
    torch.cuda.synchronize()
    views, (R, color, depth, radii, geom, binning, img) = product_forward(g, cam, sh_degree)
    torch.cuda.synchronize()
    rep.update(P=P, W=W, H=H, R_ref=int(R_ref), R=int(R))
    rep["radii_mismatch"] = int((radii[0] != radii_ref).sum())
    vis = radii_ref > 0
    m2 = product_state("means2D", torch.float32, (P, 2), 1, P, W, H, geom, binning, img)
    m2_ref = ref.state("means2D", torch.float32, (P, 2))
    rep["means2D_bits_mismatch"] = int((m2.view(torch.int32)[vis] != m2_ref.view(torch.int32)[vis]).any(-1).sum())
    co = product_state("conic_opacity", torch.float32, (P, 4), 1, P, W, H, geom, binning, img)
    co_ref = ref.state("conic_opacity", torch.float32, (P, 4))
    rep["conic_bits_mismatch"] = int((co.view(torch.int32)[vis] != co_ref.view(torch.int32)[vis]).any(-1).sum())
    rd = product_state("rgb_depth", torch.float32, (P, 4), 1, P, W, H, geom, binning, img)
    dep_ref = ref.state("depths", torch.float32, (P,))
    rgb_ref = ref.state("rgb", torch.float32, (P, 3))
    rep["depth_bits_mismatch"] = int((rd[:, 3].contiguous().view(torch.int32)[vis] != dep_ref.view(torch.int32)[vis]).sum())
    rep["rgb_bits_mismatch"] = int((rd[:, :3].contiguous().view(torch.int32)[vis] != rgb_ref.view(torch.int32)[vis]).any(-1).sum())
    rep["rgb_maxabs"] = float((rd[:, :3][vis] - rgb_ref[vis]).abs().max()) if vis.any() else 0.0
    c3 = product_state("cov3D", torch.float32, (P, 6), 1, P, W, H, geom, binning, img)
    c3_ref = ref.state("cov3D", torch.float32, (P, 6))
    rep["cov3D_bits_mismatch"] = int((c3.view(torch.int32)[vis] != c3_ref.view(torch.int32)[vis]).any(-1).sum())
    if R == R_ref and R > 0:
        pl = product_state("point_list", torch.int32, (R,), 1, P, W, H, geom, binning, img)
        pl_ref = ref.state("point_list", torch.int32, (R,))
        rep["point_list_mismatch"] = int((pl != pl_ref).sum())
        keys = product_state("keys", torch.int64, (R,), 1, P, W, H, geom, binning, img)
        keys_ref = ref.state("keys", torch.int64, (R,))
        rep["keys_mismatch"] = int((keys != keys_ref).sum())
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    rg = product_state("ranges", torch.int32, (tiles, 2), 1, P, W, H, geom, binning, img)
    rg_ref = ref.state("ranges", torch.int32, (tiles, 2))
    rep["ranges_mismatch"] = int((rg != rg_ref).any(-1).sum())
    nc = product_state("n_contrib", torch.int32, (H, W), 1, P, W, H, geom, binning, img)
    nc_ref = ref.state("n_contrib", torch.int32, (H, W))
    rep["n_contrib_mismatch"] = int((nc != nc_ref).sum())
    rep["color_maxabs"] = float((color[0] - c_ref).abs().max())
    rep["color_bits_mismatch"] = int((color[0].view(torch.int32) != c_ref.view(torch.int32)).sum())
    rep["depth_maxabs"] = float((depth[0] - d_ref).abs().max())

    # backward
    dl = torch.from_numpy(S.pixel_grad(1, H, W)).cuda()
    g_ref = ref.backward(dl[0])
    e = torch.Tensor([])
    gb = dgr._backward_views(views, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e, H, W, dl, g["shs"],
                             sh_degree, geom, binning, img)
    torch.cuda.synchronize()
    names = ["means2D", "colors", "opacities", "means3D", "cov3D", "shs", "scales", "rotations"]
    got = dict(zip(names, gb))
    got["means2D"] = got["means2D"][0]
    for k in names:
        if k == "colors":
            continue  # intermediate when SHs are used
        a, b = got[k].double().cpu().numpy(), g_ref[k].double().cpu().numpy()
        rep[f"grad_{k}_relnorm"] = Hh.rel_err(a, b)
        denom = np.abs(b).max() + 1e-30
        rep[f"grad_{k}_maxabs_over_max"] = float(np.abs(a.reshape(b.shape) - b).max() / denom)

    if with_oracle:
        s = oracle.rasterize_forward(g_np["means3D"], g_np["opacities"], cam.world_view_transform,
                                     cam.full_proj_transform, cam.camera_center, np.ones(3, np.float32), cam.tanfovx,
                                     cam.tanfovy, H, W, shs=g_np["shs"], scales=g_np["scales"],
                                     rotations=g_np["rotations"], sh_degree=sh_degree)
        rr = radii_ref.cpu().numpy()
        rep["oracle_R"] = s.num_rendered
        rep["oracle_radii_mismatch_vs_ref"] = int((s.radii != rr).sum())
        v = rr > 0
        rep["oracle_depth_bits_mismatch_vs_ref"] = int((s.depths.view(np.int32)[v] != dep_ref.cpu().numpy().view(np.int32)[v]).sum())
        rep["oracle_means2D_bits_mismatch_vs_ref"] = int((s.means2D.view(np.int32)[v] != m2_ref.cpu().numpy().view(np.int32)[v]).any(-1).sum())
        rep["oracle_conic_bits_mismatch_vs_ref"] = int((s.conic_opacity.view(np.int32)[v] != co_ref.cpu().numpy().view(np.int32)[v]).any(-1).sum())
        rep["oracle_cov3D_bits_mismatch_vs_ref"] = int((s.cov3D.view(np.int32)[v] != c3_ref.cpu().numpy().view(np.int32)[v]).any(-1).sum())
        rep["oracle_rgb_maxabs_vs_ref"] = float(np.abs(s.rgb[v] - rgb_ref.cpu().numpy()[v]).max())
        if s.num_rendered == R_ref:
            rep["oracle_point_list_mismatch_vs_ref"] = int((s.point_list.astype(np.int64) != pl_ref.cpu().numpy().astype(np.int64)).sum())
            rep["oracle_keys_mismatch_vs_ref"] = int((s.keys.view(np.int64) != keys_ref.cpu().numpy()).sum())
        rep["oracle_color_maxabs_vs_ref"] = float(np.abs(s.color - c_ref.cpu().numpy()).max())
        rep["oracle_color_gt1e-5"] = int((np.abs(s.color - c_ref.cpu().numpy()) > 1e-5).sum())
        rep["oracle_n_contrib_mismatch_vs_ref"] = int((s.n_contrib.astype(np.int64) != nc_ref.cpu().numpy().astype(np.int64)).sum())
        og = oracle.rasterize_backward(s, dl[0].cpu().numpy())
        for k in names:
            if k == "colors":
                continue
            rep[f"oracle_grad_{k}_relnorm_vs_ref"] = Hh.rel_err(og[k], g_ref[k].cpu().numpy().reshape(og[k].shape))
    return rep


def time_scene(tag, g_np, cams, iters=20, sh_degree=3):
    import diff_gaussian_rasterization as dgr

    g = Hh.to_cuda(g_np)
    H, W = cams[0].image_height, cams[0].image_width
    V = len(cams)
    dl = torch.from_numpy(S.pixel_grad(V, H, W)).cuda()
    rep = {"tag": tag, "V": V, "P": int(g["means3D"].shape[0]), "W": W, "H": H}
    e = torch.Tensor([])
    ts = [Hh.cam_tensors(c) for c in cams]

    def ev():
        return torch.cuda.Event(enable_timing=True)

    # reference: views sequentially
    ref = ref_cuda.RefRasterizer()

    def ref_step():
        for v, c in enumerate(cams):
            t = ts[v]
            ref.forward(g["means3D"], g["opacities"], t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"],
                        c.tanfovx, c.tanfovy, H, W, shs=g["shs"], scales=g["scales"], rotations=g["rotations"],
                        sh_degree=sh_degree)
            ref.backward(dl[v])

    def prod_step_seq():
        for v, c in enumerate(cams):
            t = ts[v]
            views = [(t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"], c.tanfovx, c.tanfovy)]
            R, color, depth, radii, geom, binning, img = dgr._forward_views(
                views, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e, H, W, g["shs"], sh_degree,
                False)
            dgr._backward_views(views, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e, H, W, dl[v:v + 1],
                                g["shs"], sh_degree, geom, binning, img)

    def prod_step_batched():
        for v0 in range(0, V, 16):
            vs = list(range(v0, min(V, v0 + 16)))
            views = [(ts[v]["viewmatrix"], ts[v]["projmatrix"], ts[v]["campos"], ts[v]["bg"], cams[v].tanfovx,
                      cams[v].tanfovy) for v in vs]
            R, color, depth, radii, geom, binning, img = dgr._forward_views(
                views, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e, H, W, g["shs"], sh_degree,
                False)
            dgr._backward_views(views, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e, H, W,
                                dl[vs[0]:vs[-1] + 1], g["shs"], sh_degree, geom, binning, img)

    for name, fn in (("ref", ref_step), ("prod_seq", prod_step_seq), ("prod_batched", prod_step_batched)):
        try:
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            a, b = ev(), ev()
            a.record()
            t0 = time.perf_counter()
            for _ in range(iters):
                fn()
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / iters
            wall = (time.perf_counter() - t0) / iters * 1e3
            rep[name + "_ms"] = ms
            rep[name + "_wall_ms"] = wall
            rep[name + "_mpix_s"] = V * H * W / (ms * 1e-3) / 1e6
        except Exception:
            rep[name + "_error"] = traceback.format_exc()[-600:]
    return rep


def main():
    which = sys.argv[1:] or ["small", "c2", "c3s", "time"]
    reports = []

    def run(fn, *a, **k):
        try:
            r = fn(*a, **k)
        except Exception:
            r = {"tag": str(a[0]), "error": traceback.format_exc()[-1500:]}
        print(json.dumps(r, indent=1), flush=True)
        reports.append(r)

    if "small" in which:
        g, cam = Hh.small_scene()
        run(compare_scene, "small", g, cam)
    if "c2" in which:
        g, cams = S.config_c2(4, 800)
        run(compare_scene, "c2_view0", g, cams[0])
        run(compare_scene, "c2_view2", g, cams[2])
    if "c3s" in which:
        g, cams = S.config_c3(100_000, 1920, 1080)
        run(compare_scene, "c3_100k", g, cams[0], with_oracle=False)
    if "c3" in which:
        g, cams = S.config_c3(1_000_000, 1920, 1080)
        run(compare_scene, "c3_1M", g, cams[0], with_oracle=False)
    if "time" in which:
        g, cams = S.config_c2(4, 800)
        run(time_scene, "time_c2", g, cams, 20)
        g, cams = S.config_c3(100_000, 1920, 1080)
        run(time_scene, "time_c3_100k", g, cams, 5)
    if "time3" in which:
        g, cams = S.config_c3(1_000_000, 1920, 1080)
        run(time_scene, "time_c3_1M", g, cams, 3)
    with open(os.path.join(OUT, "report_%s.json" % "_".join(which)), "w") as f:
        json.dump(reports, f, indent=1)


if __name__ == "__main__":
    main()
