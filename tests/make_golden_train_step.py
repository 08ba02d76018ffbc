"""Golden results of the reference's `train_step` (scene_reconstruction/train_utils.py:240-321) on the CPU stand-in
scene of tests/fake_scene.py.  `train_step`, `densification`, `regularization` and `psnr` are lifted from the reference's
source with ast (the modules do not import here) and run unchanged; `render` and `image_losses` are the stand-ins, the
Gaussian object is the mirror class (its methods are pinned one by one in tests/test_densify.py).
Run: python tests/make_golden_train_step.py -> tests/golden/train_step_cases.npz"""
import ast
import json
import os

import numpy as np
import torch

import fake_scene

REF = "/root/reference"


def lift_function(path, name, ns):
    tree = ast.parse(open(os.path.join(REF, path)).read().replace('device="cuda"', 'device="cpu"'))
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            node.decorator_list = []
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
            return ns[name]
    raise KeyError(name)


ns = {"torch": torch, "np": np, "render": fake_scene.render, "image_losses": fake_scene.image_losses}
for a in ("Camera", "MultiGaussianMesh", "MeshSimulator", "PipelineParams", "OptimizationParams"):
    ns[a] = object
lift_function("utils/image_utils.py", "psnr", ns)
lift_function("scene_reconstruction/train_utils.py", "regularization", ns)
lift_function("scene_reconstruction/train_utils.py", "densification", ns)
train_step = lift_function("scene_reconstruction/train_utils.py", "train_step", ns)

out = {}
for name in fake_scene.CASES:
    for k, v in fake_scene.run_case(train_step, name).items():
        out[f"{name}/{k}"] = v
    print(name, "P =", out[f"{name}/face_ids"].shape[0], "loss", out[f"{name}/1/loss"])
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "train_step_cases.npz")
np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **out)
print("wrote", path, os.path.getsize(path), "bytes")
