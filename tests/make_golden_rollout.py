"""Golden of the reference's open-loop `rollout` (train_meshnet_sim.py:92-280; the file has merge-conflict markers, so the
function's text is cut out, the marker lines dropped, and compiled on its own) with `_data_to_graph` / `_graph_to_data`
lifted from meshnet/dataloader_sim.py, the torch_geometric transform replaced by the mirror's restated one
(meshnet/mesh_data.py) and a deterministic toy simulator.  Run: python tests/make_golden_rollout.py ->
tests/golden/rollout_seed12.npz"""
import json
import os
import sys
import time
import types

import numpy as np
import torch
from torch import optim

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
import fake_scene  # noqa: E402
from meshnet.graph_data import Data  # noqa: E402
from meshnet.mesh_data import cartesian_distance, face_to_edge  # noqa: E402
from meshnet.model_utils import NodeType  # noqa: E402

REF = "/root/reference"


def cut(path, start, ends, indent):
    src = open(os.path.join(REF, path)).read()
    i = src.index(start)
    j = min(k for k in (src.find(e, i + 1) for e in ends) if k > 0)
    lines = [ln for ln in src[i:j].split("\n") if not ln.startswith(("<<<<<<<", "=======", ">>>>>>>"))]
    return "\n".join(ln[indent:] if ln.startswith(" " * indent) else ln for ln in lines)


def transformer(graph):  # T.Compose([T.FaceToEdge(), T.Cartesian(norm=False), T.Distance(norm=False)])
    graph.edge_index = face_to_edge(graph.face, graph.pos.shape[0])
    graph.edge_attr = cartesian_distance(graph.pos, graph.edge_index)
    return graph


ns = {"torch": torch, "np": np, "Data": Data, "time": time, "optim": optim, "NodeType": NodeType, "transformer": transformer,
      "print": lambda *a, **k: None}
for name in ("_data_to_graph", "_graph_to_data"):
    exec(compile(cut("meshnet/dataloader_sim.py", f"    def {name}(", ("\n    def ", "\nclass "), 4), "dataloader_sim.py", "exec"), ns)
exec(compile(cut("train_meshnet_sim.py", "def rollout(", ("\ndef ",), 0), "train_meshnet_sim.py", "exec"), ns)
dataset = types.SimpleNamespace(_data_to_graph=lambda *a: ns["_data_to_graph"](None, *a),
                                _graph_to_data=lambda g, n: ns["_graph_to_data"](None, g, n))
ds = types.SimpleNamespace(dataset=dataset)

out = {}
for tag, rw in (("sim", False), ("real_world", True)):
    f = fake_scene.rollout_features()
    r = ns["rollout"](fake_scene.ToySimulator(), ds, f, 4, torch.device("cpu"), 2, 0.01, real_world=rw)
    for k in ("predicted_rollout", "ground_truth_rollout", "mean_loss"):
        out[f"{tag}/{k}"] = np.asarray(r[k])
    print(tag, r["mean_loss"], np.abs(r["predicted_rollout"]).max())
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rollout_seed12.npz")
np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **out)
print("wrote", path)
