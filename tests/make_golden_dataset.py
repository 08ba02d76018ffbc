"""Goldens of the reference's `Camera` (scene_reconstruction/cameras.py:17-68) and `MDNerfDataset`
(scene_reconstruction/dataset.py:46-122): both classes lifted from the source with ast (the modules do not import here),
`utils.graphics_utils` imported from the reference as it is.  Run: python tests/make_golden_dataset.py ->
tests/golden/dataset_cameras.npz"""
import ast
import json
import os
import sys
import types

import numpy as np
import torch
from torch import nn

REF = "/root/reference"
sys.path.insert(0, REF)
from utils.graphics_utils import getProjectionMatrix, getWorld2View2  # noqa: E402

from fake_scene import camera_infos as infos  # noqa: E402


def lift_class(path, name, ns):
    tree = ast.parse(open(os.path.join(REF, path)).read())
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == name)
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    return ns[name]


Camera = lift_class("scene_reconstruction/cameras.py", "Camera",
                    {"torch": torch, "nn": nn, "np": np, "getWorld2View2": getWorld2View2, "getProjectionMatrix": getProjectionMatrix})
MDNerfDataset = lift_class("scene_reconstruction/dataset.py", "MDNerfDataset",
                           {"torch": torch, "np": np, "Dataset": object, "Camera": Camera})


if __name__ == "__main__":
    out = {}
    src = infos(3, 5)
    ds = MDNerfDataset(src, None)
    out["len"] = len(ds)
    picks = []
    np.random.seed(17)
    for k in range(12):
        item = ds[k % 3]
        picks.append([[c.view_id, c.time_id] for c in item])
        if k == 0:
            c = item[1]
            for a in ("world_view_transform", "projection_matrix", "full_proj_transform", "camera_center", "original_image", "mask"):
                out[f"cam/{a}"] = getattr(c, a).numpy()
            out["cam/scalars"] = np.array([c.FoVx, c.FoVy, c.time, c.image_width, c.image_height, c.znear, c.zfar, c.uid, c.colmap_id])
    out["picks"] = np.array(picks)
    short = MDNerfDataset(infos(2, 2, seed=5), None)
    out["short_picks"] = np.array([[[c.view_id, c.time_id] for c in short[i]] for i in range(2)])
    a = torch.rand((1, 6, 8), generator=torch.Generator().manual_seed(1))
    s0 = src[0]
    c = Camera(colmap_id=1, R=s0.R, T=s0.T, FoVx=0.7, FoVy=0.4, image=s0.image.clone(), gt_alpha_mask=a, image_name="x", uid=1,
               trans=np.array([0.1, -0.2, 0.3]), scale=1.5, data_device="cpu", time=0.25)
    for k in ("world_view_transform", "full_proj_transform", "camera_center", "original_image"):
        out[f"alpha/{k}"] = getattr(c, k).numpy()
    out["alpha/mask"] = a.numpy()
    commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dataset_cameras.npz")
    np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **out)
    print("wrote", path, out["picks"][:3].tolist())
