"""Slot order of the reference's checkpoint tuple: `GaussianModel.capture` / `restore` (scene_reconstruction/gaussian_model.py:
62-94) lifted from the source and run on an object whose attributes are their own names.
Run: python tests/make_golden_checkpoint.py -> tests/golden/checkpoint_slots.npz"""
import ast
import json
import os

import numpy as np

REF = "/root/reference"
tree = ast.parse(open(os.path.join(REF, "scene_reconstruction/gaussian_model.py")).read())
cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussianModel")
ns = {}
for f in cls.body:
    if isinstance(f, ast.FunctionDef) and f.name in ("capture", "restore"):
        exec(compile(ast.Module(body=[f], type_ignores=[]), "gaussian_model.py", "exec"), ns)

from fake_scene import SLOTS, Named  # noqa: E402


if __name__ == "__main__":
    captured = ns["capture"](Named())
    obj = Named()
    ns["restore"](obj, tuple(f"slot{i}" for i in range(12)), "opt")
    restored = [f"{a}={getattr(obj, a)}" for a in SLOTS] + obj.log
    commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "checkpoint_slots.npz")
    np.savez(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", captured=np.array(captured), restored=np.array(restored))
    print(captured, restored, sep="\n")
