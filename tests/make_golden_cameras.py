"""Golden camera matrices from the REFERENCE's own utils/graphics_utils.py (imported from /root/reference in the build
container) for a small synthetic transforms file; the three numpy lines that turn `transform_matrix` into (R, T) are
quoted from scene_reconstruction/dataset_readers.py:352-359 (that module itself needs PIL / plyfile / h5py and cannot be
imported here).  Run: python tests/make_golden_cameras.py  ->  tests/golden/cameras_transforms.npz"""
import ast
import json
import math
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
from utils.graphics_utils import getProjectionMatrix, getWorld2View2  # noqa: E402

rng = np.random.default_rng(7)


def c2w_look_at(eye, target):  # OpenGL camera: looks along -z, y up
    f = target - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, np.array([0.0, 0.0, 1.0]))
    r /= np.linalg.norm(r)
    u = np.cross(r, f)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = r, u, -f, eye
    return m


frames = []
for v in range(3):
    ang = 2 * math.pi * v / 3 + 0.3
    eye = np.array([1.3 * math.cos(ang), 1.3 * math.sin(ang), 0.8 + 0.1 * v])
    m = c2w_look_at(eye, np.array([0.05, -0.02, 0.1]))
    for t in range(2):
        frames.append({"file_path": f"./train/r_{v}_{t}", "time": t / 1.0 * 0.5, "transform_matrix": m.tolist()})
contents = {"camera_angle_x": 0.6911, "camera_angle_y": 0.55, "frames": frames}

exp = {"wvt": [], "full": [], "center": []}
rts = []
for fr in frames:
    c2w = np.array(fr["transform_matrix"])
    c2w[:3, 1:3] *= -1                       # dataset_readers.py:352-354
    w2c = np.linalg.inv(c2w)                 # :357
    R = np.transpose(w2c[:3, :3])            # :358
    T = w2c[:3, 3]                           # :359
    # scene_reconstruction/cameras.py:57-68
    wvt = torch.tensor(getWorld2View2(R, T, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
    proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=contents["camera_angle_x"], fovY=contents["camera_angle_y"]).transpose(0, 1)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    center = wvt.inverse()[3, :3]
    exp["wvt"].append(wvt.numpy()); exp["full"].append(full.numpy()); exp["center"].append(center.numpy())
    rts.append(types.SimpleNamespace(R=R, T=T))

# getNerfppNorm (scene_reconstruction/dataset_readers.py:58-79), lifted from the source: -> Scene.cameras_extent (scene.py:120)
tree = ast.parse(open(os.path.join(REF, "scene_reconstruction/dataset_readers.py")).read())
node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "getNerfppNorm")
ns = {"np": np, "getWorld2View2": getWorld2View2}
exec(compile(ast.Module(body=[node], type_ignores=[]), "dataset_readers.py", "exec"), ns)
norm = ns["getNerfppNorm"](rts)
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]  # the checkout carries no .git
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cameras_transforms.npz")
np.savez(out, transforms=json.dumps(contents), ref_commit=f"KTH-RPL/cloth-splatting {commit}", nerf_translate=norm["translate"],
         nerf_radius=np.float64(norm["radius"]), **{k: np.stack(v) for k, v in exp.items()})
print("wrote", out)
