"""Generates tests/golden/*.npz on the GPU box from the REFERENCE's own CUDA code
(oracle/_ref/libref_cuda.so = RAST/cuda_rasterizer/*.cu and KNN/simple_knn.cu of
KTH-RPL/cloth-splatting @ f7acc90, submodules f2d8fa99 / 44f76429, compiled unmodified for sm_100a).

    gpurun -- python tests/make_golden.py      # writes gpurun_out/golden/, copy into tests/golden/

The fixtures pin the CPU oracle (tests/test_golden.py, runs without a GPU) and the CUDA product
(tests/test_rasterizer_gpu.py).  Inputs are not stored: they are regenerated from the seeds below
(csb_synth, numpy Generator -> identical bits everywhere)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa: E402
import csb_synth as S  # noqa: E402
import helpers as Hh  # noqa: E402
from oracle import ref_cuda  # noqa: E402

REF_COMMIT = "KTH-RPL/cloth-splatting f7acc90; depth-diff-gaussian-rasterization f2d8fa99; simple-knn 44f76429"


def scenes():
    """name -> (gaussians, camera, sh_degree, bg); kept small: the CPU suite replays them."""
    out = {}
    g, cam = Hh.small_scene(P=3000, W=160, H=112, seed=11)
    out["cloud_3000_160x112_deg3"] = (g, cam, 3, (1.0, 1.0, 1.0))
    g, cam = Hh.small_scene(P=1500, W=149, H=83, seed=21)
    out["cloud_1500_149x83_deg1"] = (g, cam, 1, (0.0, 0.3, 0.7))
    g, cam = Hh.mesh_scene(res=208, view=1)
    out["cloth_10108_208x208_deg3"] = ({k: g[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")}, cam,
                                       3, (1.0, 1.0, 1.0))
    return out


def main():
    out_dir = os.path.join(conftest.ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, (g, cam, deg, bg) in scenes().items():
        gc = Hh.to_cuda(g)
        P = g["means3D"].shape[0]
        H, W = cam.image_height, cam.image_width
        t = Hh.cam_tensors(cam, bg)
        ref = ref_cuda.RefRasterizer()
        R, color, radii, depth = ref.forward(gc["means3D"], gc["opacities"], t["viewmatrix"], t["projmatrix"],
                                             t["campos"], t["bg"], cam.tanfovx, cam.tanfovy, H, W, shs=gc["shs"],
                                             scales=gc["scales"], rotations=gc["rotations"], sh_degree=deg)
        torch.cuda.synchronize()
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        vis = (radii > 0).cpu().numpy()

        def st(n, dt, shape):
            return ref.state(n, dt, shape).cpu().numpy()

        fix = dict(
            ref_commit=np.array(REF_COMMIT), num_rendered=np.int64(R), radii=radii.cpu().numpy(),
            color=color.cpu().numpy(), depth=depth.cpu().numpy(),
            means2D=st("means2D", torch.float32, (P, 2)) * vis[:, None],
            depths=st("depths", torch.float32, (P,)) * vis,
            conic_opacity=st("conic_opacity", torch.float32, (P, 4)) * vis[:, None],
            rgb=st("rgb", torch.float32, (P, 3)) * vis[:, None],
            cov3D=st("cov3D", torch.float32, (P, 6)),
            tiles_touched=st("tiles_touched", torch.int32, (P,)).astype(np.uint32),
            keys=st("keys", torch.int64, (R,)).view(np.uint64), point_list=st("point_list", torch.int32, (R,)).astype(np.uint32),
            ranges=st("ranges", torch.int32, (tiles, 2)).astype(np.uint32),
            n_contrib=st("n_contrib", torch.int32, (H, W)).astype(np.uint32),
            final_T=st("accum_alpha", torch.float32, (H, W)))
        dl = torch.from_numpy(S.pixel_grad(1, H, W)).cuda()
        gr = ref.backward(dl[0])
        for k in ("means2D", "opacities", "means3D", "cov3D", "shs", "scales", "rotations", "colors", "conic"):
            fix["grad_" + k] = gr[k].cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **fix)
        print(name, "R =", R, "visible =", int(vis.sum()))
    # distCUDA2
    g = S.mesh_gaussians(20, 14, 1)
    rng = np.random.default_rng(77)
    clouds = {"cloth_10108": g["means3D"], "normal_4096": rng.standard_normal((4096, 3)).astype(np.float32),
              "tiny_3": rng.standard_normal((3, 3)).astype(np.float32)}
    fix = {"ref_commit": np.array(REF_COMMIT)}
    for k, pts in clouds.items():
        fix[k] = ref_cuda.dist_cuda2(torch.from_numpy(pts).cuda()).cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, "dist2.npz"), **fix)
    print("golden written to", out_dir)


if __name__ == "__main__":
    main()
