"""Kernel-level timeline of the drop-in train_step (CUDA-graph replay) from CUPTI: python tools/dropin_profile.py"""
import collections
import json
import os
import sys
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "cloth-splatting_b200")]
sys.argv = ["bench.py"]
import bench  # noqa: E402
import csb_synth as S  # noqa: E402
from meshnet.meshnet_network import ResidualMeshSimulator  # noqa: E402
from scene_reconstruction.gaussian_mesh import MultiGaussianMesh  # noqa: E402
from scene_reconstruction.train_utils import train_step  # noqa: E402

dev = torch.device("cuda", 0)
W = H = 800
verts, faces, ei = S.cloth_mesh(20)
pos = torch.from_numpy(verts).to(dev)
g = torch.Generator(device=dev).manual_seed(0)
pc = MultiGaussianMesh.from_mesh(pos, torch.from_numpy(np.ascontiguousarray(faces.T)).to(dev), gaussian_init_factor=14, generator=g)
pc.set_edges(torch.from_numpy(ei).to(dev))
pc.active_sh_degree = 3
with torch.no_grad():
    pc._opacity.add_(2.0)
opt = bench._dropin_options()
pc.training_setup(opt)
pc.init_densification_stats(0.01)
sim = ResidualMeshSimulator(torch.stack([pos, pos + 0.01, pos + 0.02]), device=dev)
mesh_opt = torch.optim.Adam(sim.parameters(), lr=1e-4)
cams = S.ring_cameras(3, W, H, 0.6911, verts.mean(0))
for i, cam in enumerate(cams):
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, torch.from_numpy(getattr(cam, k)).to(dev))
    cam.time, cam.mask = i / 2, None
    cam.original_image = torch.rand((3, H, W), generator=torch.Generator().manual_seed(i)).pin_memory()
bg = torch.ones(3, device=dev)
pipe, uargs = types.SimpleNamespace(compute_cov3D_python=False), types.SimpleNamespace(no_shadow=True)
for it in range(1, 12):
    train_step(it, cams, pc, sim, mesh_opt, pipe, opt, 5.0, bg, user_args=uargs)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile

n = 5
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    t0 = time.perf_counter()
    for it in range(12, 12 + n):
        train_step(it, cams, pc, sim, mesh_opt, pipe, opt, 5.0, bg, user_args=uargs)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / n * 1e3
tot = collections.Counter()
cnt = collections.Counter()
t_first, t_last = None, None
for ev in prof.events():
    if "cuda" in str(getattr(ev, "device_type", "")).lower():
        name = ev.name[:110]
        tot[name] += ev.device_time / n
        cnt[name] += 1
rows = sorted(tot.items(), key=lambda kv: -kv[1])
print(json.dumps({"wall_ms_per_iteration_under_profiler": wall, "device_us_per_iteration": sum(tot.values()),
                  "kernels_per_iteration": sum(cnt.values()) / n,
                  "top": [{"name": k, "us": round(v, 1), "per_iter": cnt[k] / n} for k, v in rows[:45]]}, indent=1))
