"""Host-side pieces of the refinement step next to the kernels: `regularization` and the learning-rate
schedule, against the reference's own functions.  The golden values come from tests/make_golden_train_utils.py (functions
lifted from the reference source, `device="cuda"` rewritten to "cpu")."""
import os
import sys
import types

import numpy as np
from helpers import same_as_golden
import torch

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "train_utils_seed4.npz"))


def _inputs():
    d = torch.from_numpy(G["deform"].copy()).requires_grad_(True)
    gs = types.SimpleNamespace(mesh=types.SimpleNamespace(edge_index=torch.from_numpy(G["edge_index"])),
                               edge_norm=torch.from_numpy(G["edge_norm"]))
    return d, gs


def test_regularization_value_and_gradient():
    from scene_reconstruction.train_utils import regularization

    for tag, lam in (("all", (0.3, 0.7, 0.2)), ("rigid_only", (0.0, 1.5, 0.0))):
        opt = types.SimpleNamespace(lambda_deform_mag=lam[0], lambda_rigid=lam[1], lambda_momentum=lam[2])
        for n_cams in (3, 2):
            d, gs = _inputs()
            loss = regularization(d[:n_cams], gs, opt, static=False)
            loss.backward()
            assert same_as_golden(loss.detach().numpy(), G[f"{tag}/{n_cams}/loss"]), (tag, n_cams)
            assert same_as_golden(d.grad.numpy(), G[f"{tag}/{n_cams}/grad"]), (tag, n_cams)
    d, gs = _inputs()
    assert float(regularization(d, gs, opt, static=True)) == 0.0


def test_learning_rate_schedule():
    from scene_reconstruction.train_utils import get_expon_lr_func

    f = get_expon_lr_func(1.6e-4, 1.6e-6, lr_delay_mult=0.01, max_steps=30000)
    h = get_expon_lr_func(1e-3, 1e-5, lr_delay_steps=100, lr_delay_mult=0.1, max_steps=2000)
    steps = G["lr_steps"]
    assert np.array_equal(np.array([f(int(s)) for s in steps]), G["lr_plain"])
    assert np.array_equal(np.array([h(int(s)) for s in steps]), G["lr_delayed"])
    assert get_expon_lr_func(0.0, 0.0)(5) == 0.0 and f(-1) == 0.0


def test_update_learning_rate_keeps_the_reference_behaviour():
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    m = MultiGaussianMesh(torch.zeros((3, 3)), torch.zeros((3, 1), dtype=torch.long), torch.zeros(2, dtype=torch.long),
                          torch.full((2, 3), 1 / 3))
    ps = [torch.nn.Parameter(torch.zeros(2, 3)) for _ in range(2)]
    m.optimizer = torch.optim.Adam([{"params": [ps[0]], "lr": 1e-3, "name": "face_bary"},
                                    {"params": [ps[1]], "lr": 2e-3, "name": "f_dc"}], lr=0.0)
    assert m.update_learning_rate(100) is None                      # no group is called "xyz" in the mesh model
    assert [g["lr"] for g in m.optimizer.param_groups] == [1e-3, 2e-3]
    m.optimizer.param_groups[0]["name"] = "xyz"
    m.xyz_scheduler_args = lambda it: 0.5 / it
    assert m.update_learning_rate(100) == 0.005 and m.optimizer.param_groups[0]["lr"] == 0.005
