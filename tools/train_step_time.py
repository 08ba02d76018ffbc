"""Time of one whole refinement iteration through the drop-in `train_step` (C2-sized scene: 20x20 cloth, 10108
mesh-anchored Gaussians, the 3 consecutive-time views of a training item at 800x800): simulator -> Gaussians -> render x3
-> loss + regularisers -> backward -> schedule -> both Adam steps, host work and the loss_dict `.item()` reads included.
Writes one JSON line.  Run on the GPU box: python tools/train_step_time.py > gpurun_out/train_step_time.json"""
import json
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "cloth-splatting_b200"), os.path.join(ROOT, "tests")]
import csb_synth as S  # noqa: E402
import fake_scene  # noqa: E402
from meshnet.meshnet_network import ResidualMeshSimulator  # noqa: E402
from scene_reconstruction.gaussian_mesh import MultiGaussianMesh  # noqa: E402
from scene_reconstruction.train_utils import train_step  # noqa: E402

W = H = 800
verts, faces, ei = S.cloth_mesh(20)
pos = torch.from_numpy(verts).cuda()
g = torch.Generator(device="cuda").manual_seed(0)
pc = MultiGaussianMesh.from_mesh(pos, torch.from_numpy(np.ascontiguousarray(faces.T)).cuda(), gaussian_init_factor=14, generator=g)
pc.set_edges(torch.from_numpy(ei).cuda())
pc.active_sh_degree = 3
with torch.no_grad():
    pc._opacity.add_(2.0)
    pc._features_dc.add_(torch.randn(pc._features_dc.shape, device="cuda", generator=g))
opt = fake_scene.options()
pc.training_setup(opt)
pc.init_densification_stats(0.01)
sim = ResidualMeshSimulator(torch.stack([pos, pos + 0.01, pos + 0.02]), device="cuda")
mesh_opt = torch.optim.Adam(sim.parameters(), lr=1e-4)
cams = S.ring_cameras(3, W, H, 0.6911, verts.mean(0))
for i, cam in enumerate(cams):
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, torch.from_numpy(getattr(cam, k)).cuda())
    cam.time, cam.mask = i / 2, None
    cam.original_image = torch.rand((3, H, W), generator=torch.Generator().manual_seed(i)).pin_memory()
bg = torch.ones(3, device="cuda")
pipe, args = types.SimpleNamespace(compute_cov3D_python=False), types.SimpleNamespace(no_shadow=True)

warm, steps = 10, 50
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
losses = []
for it in range(1, warm + steps + 1):
    if it > warm:
        ev[it - warm - 1][0].record()
    _, loss, _ = train_step(it, cams, pc, sim, mesh_opt, pipe, opt, 5.0, bg, user_args=args)
    if it > warm:
        ev[it - warm - 1][1].record()
    losses.append(loss.detach())
torch.cuda.synchronize()
ms = sorted(a.elapsed_time(b) for a, b in ev)
losses = [float(x) for x in losses]
print(json.dumps({"what": "train_step, 3 views 800x800, 10108 Gaussians, SH 3 (drop-in per-view loop)", "steps": steps, "warmup": warm,
                  "ms_per_iteration_median": ms[steps // 2], "ms_min": ms[0], "ms_max": ms[-1],
                  "mpix_per_s": 3 * W * H / ms[steps // 2] / 1e3, "gaussians": pc.num_gaussians,
                  "loss_first": losses[0], "loss_last": losses[-1]}))

if "--breakdown" in sys.argv:
    # where the iteration goes: the sections of train_step run one after another with a device synchronisation after each
    # (wall clock, so host work counts; the sum is larger than the free-running iteration above)
    import time as _t

    from gaussian_renderer import render
    from scene_reconstruction.train_utils import densification, image_losses, regularization
    from utils.image_utils import psnr

    acc = {}

    def section(name, t0):
        torch.cuda.synchronize()
        acc[name] = acc.get(name, 0.0) + (_t.perf_counter() - t0) * 1e3
        return _t.perf_counter()

    n = 30
    for it in range(100, 100 + n):
        torch.cuda.synchronize()
        t = _t.perf_counter()
        pk = [render(c, pc, sim, pipe, bg, no_shadow=True) for c in cams]
        t = section("render x3 (simulator, xyz, rotation, rasterizer forward, projections)", t)
        gt = torch.stack([c.original_image.to("cuda", non_blocking=True) for c in cams])
        t = section("ground-truth upload (3 x 7.7 MB pinned) + stack", t)
        img = torch.stack([p.render for p in pk])
        ps = psnr(img, gt).mean().double()
        loss, d = image_losses(img, gt, opt)
        t = section("stack + psnr + image_losses (2 x .item())", t)
        loss = loss + regularization(torch.stack([p.vertice_deform for p in pk]), pc, opt)
        t = section("regularization", t)
        loss.backward()
        t = section("backward (loss, rasterizer x3, rotation / xyz, simulator, regularisers)", t)
        with torch.no_grad():
            grad = sum(p.viewspace_points.grad for p in pk)
            densification(pc, it, torch.stack([p.visibility_filter for p in pk]).any(0),
                          torch.stack([p.radii for p in pk]).max(0).values, grad, opt, 5.0)
            t = section("densification statistics (boolean-mask indexing)", t)
            pc.optimizer.step()
            mesh_opt.step()
            pc.optimizer.zero_grad(set_to_none=True)
            mesh_opt.zero_grad()
            t = section("Adam steps (Gaussians: one launch; simulator: torch)", t)
    print(json.dumps({"breakdown_ms_per_iteration": {k: round(v / n, 3) for k, v in acc.items()},
                      "sum_ms": round(sum(acc.values()) / n, 3)}))
