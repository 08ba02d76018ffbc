"""Golden results of the reference's densification / pruning bookkeeping (second half).

The reference modules cannot be imported here (roma, h5py, plyfile, simple_knn ... are missing), so the METHODS are lifted
from the reference's source files at run time (ast) and bound to a stub object: `replace_tensor_to_optimizer`,
`_prune_optimizer`, `cat_tensors_to_optimizer`, `reset_opacity`, `add_densification_stats`, `prune` of
scene_reconstruction/gaussian_model.py and `prune_points`, `densification_postfix`, `densify_and_clone`, `densify_and_split` (+ `get_xyz`, `build_rotation`,
`compute_barycentric_coordinates`) of
MultiGaussianMesh (scene_reconstruction/gaussian_mesh.py:336-431).  They run on the CPU with torch.optim.Adam exactly as
written (only `device="cuda"` in the two `torch.zeros` calls of densification_postfix is rewritten to "cpu").
Run: python tests/make_golden_densify.py -> tests/golden/densify_seed9.npz"""
import ast
import json
import os
import types

import numpy as np
import torch
from torch import nn

REF = "/root/reference"


def lift(path, cls, names):
    src = open(os.path.join(REF, path)).read().replace('device="cuda"', 'device="cpu"')
    tree = ast.parse(src)
    out = {}
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == cls:
            for f in node.body:
                if isinstance(f, ast.FunctionDef) and f.name in names:
                    f.decorator_list = []
                    mod = ast.Module(body=[f], type_ignores=[])
                    ns = {"torch": torch, "nn": nn, "np": np, "Optional": __import__("typing").Optional,
                          "inverse_sigmoid": lambda x: torch.log(x / (1 - x))}  # utils/general_utils.py:20-21
                    exec(compile(mod, path, "exec"), ns)
                    out[f.name] = ns[f.name]
    assert set(out) == set(names), (cls, set(names) - set(out))
    return out


base = lift("scene_reconstruction/gaussian_model.py", "GaussianModel",
            ["replace_tensor_to_optimizer", "_prune_optimizer", "cat_tensors_to_optimizer", "reset_opacity",
             "add_densification_stats", "prune"])
mesh = lift("scene_reconstruction/gaussian_mesh.py", "MultiGaussianMesh",
            ["prune_points", "densification_postfix", "densify_and_clone", "densify_and_split", "get_xyz",
             "cleanup_barycentric_coordinates"])


def lift_function(path, name):
    tree = ast.parse(open(os.path.join(REF, path)).read().replace("device='cuda'", "device='cpu'"))
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            ns = {"torch": torch, "np": np, "OptimizationParams": object}
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
            return ns[name]
    raise KeyError(name)


# helpers densify_and_split calls: utils/general_utils.py:81-102, meshnet/data_utils.py:494-530
for f in mesh.values():
    f.__globals__["build_rotation"] = lift_function("utils/general_utils.py", "build_rotation")
    f.__globals__["compute_barycentric_coordinates"] = lift_function("meshnet/data_utils.py", "compute_barycentric_coordinates")
    f.__globals__["Optional"] = __import__("typing").Optional


class Stub:
    pass


for k, f in {**base, **mesh}.items():
    setattr(Stub, k, f)
Stub.get_opacity = property(lambda self: torch.sigmoid(self._opacity))
Stub.get_scaling = property(lambda self: torch.exp(self._scaling))

g = torch.Generator().manual_seed(9)
P = 23
init = {"face_bary": torch.rand((P, 3), generator=g), "face_offset": torch.randn((P, 1), generator=g),
        "f_dc": torch.randn((P, 1, 3), generator=g), "f_rest": torch.randn((P, 15, 3), generator=g),
        "opacity": torch.randn((P, 1), generator=g), "scaling": torch.randn((P, 3), generator=g) - 3.0,
        "rotation": torch.randn((P, 4), generator=g)}
face_ids = torch.randint(0, 7, (P,), generator=g)
m1 = {k: torch.randn(v.shape, generator=g) for k, v in init.items()}       # Adam state as after some steps
m2 = {k: torch.rand(v.shape, generator=g) for k, v in init.items()}
mesh_pos = torch.rand((9, 3), generator=g)
mesh_face = torch.tensor([[0, 1, 3, 4, 1, 2, 4], [1, 2, 4, 5, 4, 5, 7], [3, 4, 6, 7, 3, 4, 6]])
view_grad = torch.randn((P, 3), generator=g)
update_filter = torch.rand(P, generator=g) > 0.3
radii = torch.rand(P, generator=g) * 40


def fresh():
    s = Stub()
    s.face_bary, s.face_offset = nn.Parameter(init["face_bary"].clone()), nn.Parameter(init["face_offset"].clone())
    s._features_dc, s._features_rest = nn.Parameter(init["f_dc"].clone()), nn.Parameter(init["f_rest"].clone())
    s._opacity, s._scaling, s._rotation = (nn.Parameter(init[k].clone()) for k in ("opacity", "scaling", "rotation"))
    s.face_ids = face_ids.clone()
    s.mesh = types.SimpleNamespace(pos=mesh_pos.clone(), face=mesh_face.clone())
    s.scaling_inverse_activation = torch.log
    s.percent_dense = 0.01
    s.pos_gradient_accum, s.denom, s.max_radii2D = torch.zeros((P, 1)), torch.zeros((P, 1)), radii.clone()
    names = ["face_bary", "face_offset", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
    params = [s.face_bary, s.face_offset, s._features_dc, s._features_rest, s._opacity, s._scaling, s._rotation]
    s.optimizer = torch.optim.Adam([{"params": [p], "lr": 1e-3, "name": n} for p, n in zip(params, names)], lr=0.0, eps=1e-15)
    for p, n in zip(params, names):
        s.optimizer.state[p] = {"step": torch.tensor(5.0), "exp_avg": m1[n].clone(), "exp_avg_sq": m2[n].clone()}
    return s


def snapshot(s, tag, out):
    groups = {gr["name"]: gr["params"][0] for gr in s.optimizer.param_groups}
    for n, p in groups.items():
        st = s.optimizer.state[p]
        out[f"{tag}/{n}"] = p.detach().numpy().copy()
        out[f"{tag}/{n}/exp_avg"] = st["exp_avg"].numpy().copy()
        out[f"{tag}/{n}/exp_avg_sq"] = st["exp_avg_sq"].numpy().copy()
    for a in ("face_ids", "pos_gradient_accum", "denom", "max_radii2D"):
        out[f"{tag}/{a}"] = getattr(s, a).numpy().copy()
    # the attributes must BE the optimiser's parameters
    assert s._opacity is groups["opacity"] and s.face_bary is groups["face_bary"] and s._features_rest is groups["f_rest"]


out = {}
s = fresh()
s.add_densification_stats(view_grad, update_filter)
snapshot(s, "stats", out)
grads = s.pos_gradient_accum / s.denom
grads[grads.isnan()] = 0.0
s.densify_and_clone(grads, 1.2, 20.0)
snapshot(s, "clone", out)
s = fresh()
s.add_densification_stats(view_grad, update_filter)
grads = s.pos_gradient_accum / s.denom
grads[grads.isnan()] = 0.0
torch.manual_seed(1234)  # torch.normal inside draws from the global CPU generator
s.densify_and_split(grads, 1.2, 0.5)
snapshot(s, "split", out)
s = fresh()
s.reset_opacity()
snapshot(s, "reset_opacity", out)
s = fresh()
torch.cuda.empty_cache = lambda: None
s.prune(0.0, 0.45, 1.0, 25.0)
snapshot(s, "prune", out)
s = fresh()
s.prune(0.0, 0.45, 1.0, None)
snapshot(s, "prune_opacity_only", out)

# cleanup_barycentric_coordinates (gaussian_mesh.py:267-322): Gaussians that left their face are re-anchored
s = fresh()
bary = init["face_bary"].clone()
bary[[2, 5, 5, 11, 17], [0, 1, 2, 2, 1]] = torch.tensor([-0.2, -0.05, -0.3, -0.4, -0.01])
s.face_bary = bary                      # a plain tensor: the method writes into it in place
s.cleanup_barycentric_coordinates()
out["cleanup/face_bary"], out["cleanup/face_ids"] = s.face_bary.numpy().copy(), s.face_ids.numpy().copy()
# densification() of the training loop (scene_reconstruction/train_utils.py:324-345) over the methods above
densification = lift_function("scene_reconstruction/train_utils.py", "densification")
Stub.densify = lift("scene_reconstruction/gaussian_model.py", "GaussianModel", ["densify"])["densify"]
OPT = dict(opacity_threshold_fine_init=0.5, opacity_threshold_fine_after=0.3, densify_until_iter=1000,
           densify_grad_threshold_fine_init=1.4, densify_grad_threshold_after=0.8, densify_from_iter=100,
           densification_interval=100, opacity_reset_interval=300, pruning_from_iter=100, pruning_interval=200)
vis = torch.rand(P, generator=g) > 0.2
new_radii = (torch.rand(P, generator=g) * 60)
for it in (50, 300, 400):   # statistics only / densify (no size threshold) / densify + prune with size threshold
    s = fresh()
    s.pos_gradient_accum += 3.0 * torch.rand((P, 1), generator=torch.Generator().manual_seed(it))
    s.denom += 2
    torch.manual_seed(77 + it)
    densification(s, it, vis, new_radii, view_grad, types.SimpleNamespace(**OPT), 0.5)
    snapshot(s, f"loop{it}", out)
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
inputs = {f"in/{k}": v.numpy() for k, v in init.items()}
inputs.update({f"in/m1/{k}": v.numpy() for k, v in m1.items()})
inputs.update({f"in/m2/{k}": v.numpy() for k, v in m2.items()})
inputs.update({"in/mesh_pos": mesh_pos.numpy(), "in/mesh_face": mesh_face.numpy(), "in/face_ids": face_ids.numpy(), "in/view_grad": view_grad.numpy(), "in/update_filter": update_filter.numpy(),
               "in/radii": radii.numpy(), "in/vis": vis.numpy(), "in/new_radii": new_radii.numpy(),
               "in/opt": json.dumps(OPT)})
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "densify_seed9.npz")
np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **inputs, **out)
print("wrote", path, len(out), "arrays; clone ->", out["clone/opacity"].shape[0], "prune ->", out["prune/opacity"].shape[0],
      out["prune_opacity_only/opacity"].shape[0])
