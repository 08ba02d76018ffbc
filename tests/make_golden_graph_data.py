"""Goldens of `_data_to_graph` / `_graph_to_data` (meshnet/dataloader_sim.py:352-425): the two methods lifted from the
reference's source (the module has a SyntaxError elsewhere and needs torch_geometric) with `Data` bound to the mirror's
attribute bag.  Run: python tests/make_golden_graph_data.py -> tests/golden/graph_data_seed6.npz"""
import ast
import json
import os
import sys
import types

import numpy as np
import torch

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
from fake_scene import graph_case  # noqa: E402
from meshnet.graph_data import Data  # noqa: E402

REF = "/root/reference"
src = open(os.path.join(REF, "meshnet/dataloader_sim.py")).read()
ns = {"torch": torch, "np": np, "Data": Data}
for name in ("_data_to_graph", "_graph_to_data"):
    i = src.index(f"    def {name}(")
    j = min(k for k in (src.find("\n    def ", i + 1), src.find("\nclass ", i + 1)) if k > 0)
    body = "\n".join(ln[4:] if ln.startswith("    ") else ln for ln in src[i:j].split("\n"))
    exec(compile(ast.parse(body), "dataloader_sim.py", "exec"), ns)

out = {}
for tag, as_numpy in (("tensor", False), ("numpy", True)):
    args = graph_case(as_numpy)
    g = ns["_data_to_graph"](None, *args)
    for k in ("x", "face", "y", "pos", "pos_target", "vel", "particle_actions"):
        out[f"{tag}/{k}"] = getattr(g, k).numpy()
    g.edge_index, g.edge_attr = torch.zeros((2, 1), dtype=torch.long), torch.zeros((1, 4))
    parts = ns["_graph_to_data"](types.SimpleNamespace(_input_length_sequence=2), g)
    for n, p in enumerate(parts):
        out[f"{tag}/part{n}"] = p.numpy()
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "graph_data_seed6.npz")
np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **out)
print("wrote", path, {k: v.shape for k, v in out.items() if k.startswith("tensor/")})
