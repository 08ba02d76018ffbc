"""Small end-to-end workload for compute-sanitizer (memcheck / racecheck / synccheck): every hand-written kernel family once.
  rasterizer forward + backward on the per-tile binning path and on the global sort path (onesweep, chained-scan emit),
  every compositing kernel generation, kNN, image loss forward + backward, Adam, mesh transform, one GNN forward + backward."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import conftest  # noqa: F401,E402
import numpy as np  # noqa: E402
import torch  # noqa: E402

import csb_cabi  # noqa: E402
import csb_synth as S  # noqa: E402
import helpers as H  # noqa: E402

lib = csb_cabi.lib()
dev = torch.device("cuda", 0)


def raster(g, cams, flags):
    gc = H.to_cuda(g)
    V = len(cams)
    Hh, W = cams[0].image_height, cams[0].image_width
    dl = torch.from_numpy(S.pixel_grad(V, Hh, W)).cuda()
    views, fwd = H.product_forward(gc, cams, 3, flags=flags)
    grads = H.product_backward(gc, cams, views, fwd, dl, 3)
    torch.cuda.synchronize()
    return float(fwd[1].sum()), float(grads["means3D"].abs().sum())


g, cam = H.small_scene(P=1500, W=96, H=64, seed=3)
for variant in (0x00, 0x11, 0x22, 0x43, 0x64, 0x4F):
    lib.csb_debug_render_variant(variant)
    for flags in (1, 1 | 4):  # global sort path, per-tile binning path
        print("raster variant", hex(variant), "flags", flags, raster(g, [cam, cam], flags))
lib.csb_debug_render_variant(0x4F)

from simple_knn._C import distCUDA2  # noqa: E402

pts = torch.from_numpy(np.random.default_rng(0).standard_normal((3000, 3)).astype(np.float32)).cuda()
print("knn", float(distCUDA2(pts).sum()))

from scene_reconstruction.train_utils import image_losses  # noqa: E402
import types  # noqa: E402

img = torch.rand((2, 3, 70, 90), device=dev, requires_grad=True)
gt = torch.rand((2, 3, 70, 90), device=dev)
loss, _ = image_losses(img, gt, types.SimpleNamespace(lambda_dssim=0.2), None)
loss.backward()
print("loss", float(loss), float(img.grad.abs().sum()))

import csb_optim  # noqa: E402

ps = [torch.nn.Parameter(torch.randn(n, device=dev)) for n in (5, 1000, 4097)]
opt = csb_optim.Adam([{"params": [p], "lr": 1e-2, "name": str(i)} for i, p in enumerate(ps)], lr=0.0, eps=1e-15)
for p in ps:
    p.grad = torch.randn_like(p)
opt.step()
print("adam", float(ps[2].sum()))

from meshnet import EncodeProcessDecode  # noqa: E402

x, ei, ef, _, _ = S.gnn_inputs(9, 4)
torch.manual_seed(0)
m = EncodeProcessDecode(8, 3, 4, 128, 2, 2, 128).cuda()
y = m(torch.from_numpy(x).cuda().requires_grad_(True), torch.from_numpy(ei).cuda(), torch.from_numpy(ef).cuda())
y.square().sum().backward()
torch.cuda.synchronize()
print("gnn", float(y.abs().sum()))

from scene_reconstruction.gaussian_mesh import MultiGaussianMesh  # noqa: E402

verts, faces, _ = S.cloth_mesh(6)
pos = torch.from_numpy(verts).cuda()
face = torch.from_numpy(np.ascontiguousarray(faces.T)).cuda()
pc = MultiGaussianMesh.from_mesh(pos, face, gaussian_init_factor=2, generator=torch.Generator(device="cuda").manual_seed(0))
dv = (pos + 0.01 * torch.randn_like(pos)).requires_grad_(True)
(pc.get_xyz(dv).sum() + pc.get_rotation(dv).sum()).backward()
torch.cuda.synchronize()
print("mesh transform", float(dv.grad.abs().sum()))
print("sanitize target done")
