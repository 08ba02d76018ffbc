"""Golden of the simulator dataset's sample indexing: `TrajectoriesClothSimDataset.__getitem__` and `_data_to_graph`
(meshnet/dataloader_sim.py:150-170, 352-399) and `_compute_cumulative_lengths` (meshnet/dataloader.py:40-48), cut out of
the reference's sources (both files carry merge-conflict markers elsewhere) and bound to a stub.
Run: python tests/make_golden_trajectory_dataset.py -> tests/golden/trajectory_samples.npz"""
import json
import os
import sys

import numpy as np
import torch

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
import fake_scene  # noqa: E402
from meshnet.graph_data import Data  # noqa: E402

REF = "/root/reference"


def cut(path, start):
    src = open(os.path.join(REF, path)).read()
    i = src.index(start)
    j = min(k for k in (src.find("\n    def ", i + 1), src.find("\nclass ", i + 1)) if k > 0)
    return "\n".join(ln[4:] if ln.startswith("    ") else ln for ln in src[i:j].split("\n"))


ns = {"torch": torch, "np": np, "Data": Data}
exec(compile(cut("meshnet/dataloader_sim.py", "    def __getitem__(self, idx):"), "dataloader_sim.py", "exec"), ns)
exec(compile(cut("meshnet/dataloader_sim.py", "    def _data_to_graph("), "dataloader_sim.py", "exec"), ns)
exec(compile(cut("meshnet/dataloader.py", "    def _compute_cumulative_lengths(self):"), "dataloader.py", "exec"), ns)


class Stub:
    __getitem__ = ns["__getitem__"]
    _data_to_graph = ns["_data_to_graph"]
    _compute_cumulative_lengths = ns["_compute_cumulative_lengths"]


out = {}
for L, F in ((1, 1), (2, 3)):
    s = Stub()
    s._data, s._input_length_sequence, s._future_sequence_length = fake_scene.trajectories(), L, F
    s._compute_cumulative_lengths()
    out[f"L{L}F{F}/len"] = s._length
    for idx in range(s._length):
        g = s[idx]
        for k in ("x", "face", "y", "pos", "pos_target", "vel", "particle_actions"):
            out[f"L{L}F{F}/{idx}/{k}"] = getattr(g, k).numpy()
        out[f"L{L}F{F}/{idx}/grasped"] = g.grasped_particle
    print(L, F, s._length, s._precompute_cumlengths)
commit = json.load(open(os.path.join(REF, ".SUBMODULES.json")))["commit"][:7]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "trajectory_samples.npz")
np.savez_compressed(path, ref_commit=f"KTH-RPL/cloth-splatting {commit}", **out)
print("wrote", path, os.path.getsize(path))
