"""Golden results of the reference's OWN `render()` (gaussian_renderer/__init__.py:39-206), `SinusoidalEncoder` and
`ResidualMeshSimulator` (meshnet/meshnet_network.py:256-380) on a CPU stand-in scene.

Neither module imports here: gaussian_renderer needs the CUDA rasterizer build, meshnet_network.py carries merge-conflict
markers.  So the `class` / `def` blocks are lifted from the source text and run unchanged except for
  * `device="cuda"` -> `device="cpu"` and `.cuda()` -> `.cpu()` (no GPU in the build container),
  * the conflict at meshnet_network.py:340-349 resolved to its newer side (commit 9b63d7a: `time_delta = 1.0` when
    `n_times == 1`; for n_times > 1 both sides are the same expression).
`GaussianRasterizer` is tests/fake_scene.StandInRasterizer (a smooth function of every argument and settings field), the
Gaussian object is the mirror class with the reference's torch expressions (get_xyz) and the rotation oracle.
Run: python tests/make_golden_render.py -> tests/golden/render_glue.npz"""
import ast
import math
import os
import re
import json
from typing import NamedTuple, Optional

import numpy as np
import torch
import torch.nn as nn

import fake_scene

REF = "/root/reference"


def lifted(path, names, ns, resolve_conflicts=False):
    src = open(os.path.join(REF, path)).read()
    if resolve_conflicts:
        src = re.sub(r"<<<<<<< [^\n]*\n.*?=======\n(.*?)>>>>>>> [^\n]*\n", r"\1", src, flags=re.S)
    src = src.replace('device="cuda"', 'device="cpu"').replace(".cuda()", ".cpu()")
    tree = ast.parse(src)
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


def reference_functions():
    from diff_gaussian_rasterization import GaussianRasterizationSettings  # the 12-field NamedTuple (pinned: test_api_surface)

    ns_sim = {"torch": torch, "nn": nn, "Optional": Optional}
    Enc, Sim = lifted("meshnet/meshnet_network.py", ["SinusoidalEncoder", "ResidualMeshSimulator"], ns_sim, resolve_conflicts=True)
    ns = {"torch": torch, "math": math, "np": np, "NamedTuple": NamedTuple, "GaussianRasterizationSettings": GaussianRasterizationSettings,
          "GaussianRasterizer": fake_scene.StandInRasterizer, "MultiGaussianMesh": object, "ResidualMeshSimulator": object, "eval_sh": None}
    Results, render = lifted("gaussian_renderer/__init__.py", ["RenderResults", "render"], ns)
    return Enc, Sim, Results, render


CASES = {  # camera index, kwargs
    "t0": (0, {}), "t026_rounds_up": (1, {}), "t05_vertices": (2, {"project_vertices": True}), "t074": (3, {}),
    "t1_scaled": (4, {"scaling_modifier": 0.7}), "static": (2, {"render_static": True, "project_vertices": True}),
    "override": (3, {"override_color": "colours"}), "no_shadow": (1, {"no_shadow": True})}


def run_case(render, Sim, name, reference):
    """One render + a backward of a fixed scalar -> every RenderResults field, gradients of every parameter it reaches."""
    cam_i, kw = CASES[name]
    m, sim, cams, bg = fake_scene.render_scene()
    if reference:  # the reference's class with the mirror's parameters
        ref_sim = Sim(sim.mesh_predictions, device="cpu")
        ref_sim.load_state_dict(sim.state_dict())
        sim = ref_sim
    kw = dict(kw)
    if kw.get("override_color") == "colours":
        kw["override_color"] = torch.rand((m.face_ids.shape[0], 3), generator=torch.Generator().manual_seed(9))
    pipe = type("Pipe", (), {"compute_cov3D_python": False, "convert_SHs_python": False, "debug": False})()
    res = render(cams[cam_i], m, sim, pipe, bg, **kw)
    out = {}
    for f in res._fields:
        v = getattr(res, f)
        out[f"is_none/{f}"] = np.array(v is None)
        if v is not None:
            out[f] = v.detach().numpy()
    g = torch.Generator().manual_seed(11)
    scalar = (res.render * torch.rand(res.render.shape, generator=g)).sum() + 0.1 * res.depth.sum() + 1e-3 * res.projections.sum()
    if res.vertice_projections is not None:
        scalar = scalar + 1e-3 * res.vertice_projections.sum()
    scalar.backward()
    out["grad/viewspace_points"] = res.viewspace_points.grad.numpy()
    for n, p in list(sim.named_parameters()) + [(a, getattr(m, a)) for a in ("face_bary", "_rotation", "_opacity", "_scaling", "_features_dc", "_features_rest")]:
        g_ = torch.zeros(p.shape) if p.grad is None else p.grad
        if g_.numel() > 4096:  # the big weight gradients as their two marginal sums
            out[f"grad/{n}/sum0"], out[f"grad/{n}/sum1"] = g_.sum(0).numpy(), g_.sum(1).numpy()
        else:
            out[f"grad/{n}"] = g_.numpy()
    return out


def encoder_and_simulator_cases(Enc, Sim):
    out = {}
    g = torch.Generator().manual_seed(21)
    for nf, ident, scale in ((6, True, 1.0), (4, False, 0.5), (0, True, 1.0)):
        e = Enc(input_dim=1, num_freqs=nf, scale=scale, use_identity=ident)
        x = torch.rand((5, 1), generator=g)
        out[f"enc/{nf}_{int(ident)}/x"], out[f"enc/{nf}_{int(ident)}/y"] = x.numpy(), e(x).numpy()
        out[f"enc/{nf}_{int(ident)}/output_dim"] = np.array(e.output_dim)
    for n_times in (1, 2, 5):
        preds = torch.rand((n_times, 7, 3), generator=g)
        s = Sim(preds, device="cpu")
        torch.manual_seed(n_times)
        with torch.no_grad():
            for p in s.parameters():
                p.copy_(torch.randn(p.shape) * 0.1)
        out[f"sim/{n_times}/preds"] = preds.numpy()
        out[f"sim/{n_times}/time_delta"] = np.array(s.time_delta)
        out[f"sim/{n_times}/state_sum"] = np.array([float(v.double().sum()) for v in s.state_dict().values()])
        times = [0.0, 0.124, 0.125, 0.126, 0.375, 0.5, 0.62, 0.875, 1.0, 1.2, 2.0]
        for t in times:
            try:
                out[f"sim/{n_times}/t{t}"] = s(torch.tensor(t).repeat(7, 1)).detach().numpy()
            except ValueError:
                out[f"sim/{n_times}/t{t}"] = np.array("ValueError")
            except IndexError:
                out[f"sim/{n_times}/t{t}"] = np.array("IndexError")
        out[f"sim/{n_times}/times"] = np.array(times)
    return out


if __name__ == "__main__":
    Enc, Sim, Results, render = reference_functions()
    out = {"fields": np.array(Results._fields),
           "ref_commit": np.array("KTH-RPL/cloth-splatting " + json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7])}
    out.update(encoder_and_simulator_cases(Enc, Sim))
    for name in CASES:
        for k, v in run_case(render, Sim, name, True).items():
            out[f"{name}/{k}"] = v
    # the reference raises for a time past the last mesh prediction (meshnet_network.py:367-369)
    m, sim, cams, bg = fake_scene.render_scene()
    rs = Sim(sim.mesh_predictions, device="cpu")
    pipe = type("Pipe", (), {"compute_cov3D_python": False})()
    try:
        render(cams[5], m, rs, pipe, bg)
        out["past_end"] = np.array("no error")
    except ValueError:
        out["past_end"] = np.array("ValueError")
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "render_glue.npz")
    np.savez_compressed(path, **out)
    print(path, len(out), "arrays", os.path.getsize(path), "bytes")
