"""Short image-loss + Adam workload for ncu / per-kernel timing: `python tests/loss_profile.py [views] [H] [W] [iters]`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa
import numpy as np
import torch

import csb_cabi
from csb_optim import Adam
from utils.loss_utils import image_loss

V = int(sys.argv[1]) if len(sys.argv) > 1 else 4
H = int(sys.argv[2]) if len(sys.argv) > 2 else 800
W = int(sys.argv[3]) if len(sys.argv) > 3 else 800
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
rng = np.random.default_rng(5)
gt = rng.random((V, 3, H, W), dtype=np.float32)
img = np.clip(gt + 0.05 * rng.standard_normal(gt.shape).astype(np.float32), 0, 1)
x = torch.from_numpy(img).cuda().requires_grad_(True)
y = torch.from_numpy(gt).cuda()
p = torch.nn.Parameter(torch.randn(10108 * 59, device="cuda"))
p.grad = torch.randn_like(p)
opt = Adam([p], lr=1e-3, eps=1e-15)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
lib = csb_cabi.lib()


def step():
    x.grad = None
    image_loss(x, y, 0.1)[0].backward()
    opt.step()


step()
torch.cuda.synchronize()
lib.csb_profile_enable(1)
for _ in range(iters):
    flush.fill_(1)
    step()
torch.cuda.synchronize()
prof = csb_cabi.profile_read()
lib.csb_profile_enable(0)
n = V * 3 * H * W
alg = {"image_loss_forward": 20.0 * n, "image_loss_backward": 24.0 * n, "adam_step": 28.0 * p.numel()}
for k, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    extra = f"  {alg[k] / (ms / cnt * 1e-3) / 1e9:8.0f} GB/s algorithmic" if k in alg else ""
    print(f"  {k:22s} {cnt / iters:4.1f} launches/step  {ms / cnt * 1e3:8.1f} us/launch{extra}")
