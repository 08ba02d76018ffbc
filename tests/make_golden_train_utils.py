"""Goldens of `regularization` (scene_reconstruction/train_utils.py:77-102,238) and `get_expon_lr_func`
(utils/general_utils.py:32-65), lifted from the reference source with ast (the modules need wandb / lpips / PIL ... to
import) and run on the CPU.  Run: python tests/make_golden_train_utils.py -> tests/golden/train_utils_seed4.npz"""
import ast
import json
import os
import types

import numpy as np
import torch

REF = "/root/reference"


def lift_function(path, name):
    tree = ast.parse(open(os.path.join(REF, path)).read().replace('device="cuda"', 'device="cpu"'))
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            ns = {"torch": torch, "np": np, "OptimizationParams": object}
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
            return ns[name]
    raise KeyError(name)


regularization = lift_function("scene_reconstruction/train_utils.py", "regularization")
get_expon_lr_func = lift_function("utils/general_utils.py", "get_expon_lr_func")

g = torch.Generator().manual_seed(4)
N, E = 16, 40
deform = torch.randn((3, N, 3), generator=g) * 0.05
edge_index = torch.randint(0, N, (2, E), generator=g)
edge_norm = torch.rand((E, 1), generator=g) * 0.1
out = {"deform": deform.numpy(), "edge_index": edge_index.numpy(), "edge_norm": edge_norm.numpy()}
gs = types.SimpleNamespace(mesh=types.SimpleNamespace(edge_index=edge_index), edge_norm=edge_norm)
for tag, lam in (("all", (0.3, 0.7, 0.2)), ("rigid_only", (0.0, 1.5, 0.0))):
    opt = types.SimpleNamespace(lambda_deform_mag=lam[0], lambda_rigid=lam[1], lambda_momentum=lam[2])
    for n_cams in (3, 2):
        d = deform.clone().requires_grad_(True)
        loss = regularization(d[:n_cams], gs, opt, static=False)
        loss.backward()
        out[f"{tag}/{n_cams}/loss"], out[f"{tag}/{n_cams}/grad"] = loss.detach().numpy(), d.grad.numpy()
steps = np.array([0, 1, 50, 99, 100, 1000, 1999, 2000, 30000, 40000])
f = get_expon_lr_func(1.6e-4, 1.6e-6, lr_delay_mult=0.01, max_steps=30000)
h = get_expon_lr_func(1e-3, 1e-5, lr_delay_steps=100, lr_delay_mult=0.1, max_steps=2000)
out["lr_steps"], out["lr_plain"], out["lr_delayed"] = steps, np.array([f(int(s)) for s in steps]), np.array([h(int(s)) for s in steps])
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "train_utils_seed4.npz")
np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **out)
print("wrote", path)
