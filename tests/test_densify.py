"""Densification / pruning bookkeeping of the mesh-anchored Gaussians: parameters, Adam moments and
the per-Gaussian statistics after `add_densification_stats`, `densify_and_clone`, `reset_opacity`, `prune`, against what
the reference's own methods produced for the same seeded state (tests/make_golden_densify.py lifts them from the reference
source; tests/golden/densify_seed9.npz).  Host-side tensor surgery: runs on the CPU with torch.optim.Adam standing in for
csb_optim.Adam (same state keys)."""
import os
import sys

import numpy as np
from helpers import same_as_golden
import torch

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "densify_seed9.npz"))
NAMES = ["face_bary", "face_offset", "f_dc", "f_rest", "opacity", "scaling", "rotation"]


def fresh():
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    T = lambda k: torch.from_numpy(G[k].copy())  # noqa: E731
    class CpuMesh(MultiGaussianMesh):  # get_xyz of the reference in torch (gaussian_mesh.py:151-169): no GPU here
        def get_xyz(self, deformed_vertices=None):
            vid = self.mesh.face[:, self.face_ids].transpose(0, 1)
            nb = self.face_bary / self.face_bary.sum(dim=1, keepdim=True)
            return (nb.unsqueeze(1) @ self.mesh.pos[vid, :]).squeeze(1)

    m = CpuMesh(T("in/mesh_pos"), T("in/mesh_face"), T("in/face_ids"), T("in/face_bary"))
    m.set_gaussians(T("in/f_dc"), T("in/f_rest"), T("in/scaling"), T("in/rotation"), T("in/opacity"))
    m.face_offset = torch.nn.Parameter(T("in/face_offset"))
    m.init_densification_stats(0.01)
    m.max_radii2D = T("in/radii")
    params = [getattr(m, m._GROUP_ATTR[n]) for n in NAMES]
    m.optimizer = torch.optim.Adam([{"params": [p], "lr": 1e-3, "name": n} for p, n in zip(params, NAMES)], lr=0.0, eps=1e-15)
    for p, n in zip(params, NAMES):
        m.optimizer.state[p] = {"step": torch.tensor(5.0), "exp_avg": T(f"in/m1/{n}"), "exp_avg_sq": T(f"in/m2/{n}")}
    return m


def check(m, tag):
    groups = {g["name"]: g["params"][0] for g in m.optimizer.param_groups}
    for n in NAMES:
        p = groups[n]
        assert p is getattr(m, m._GROUP_ATTR[n]), n          # the attribute IS the optimiser's parameter
        assert p.requires_grad and isinstance(p, torch.nn.Parameter)
        st = m.optimizer.state[p]
        assert same_as_golden(p.detach().numpy(), G[f"{tag}/{n}"]), (tag, n)
        assert same_as_golden(st["exp_avg"].numpy(), G[f"{tag}/{n}/exp_avg"]), (tag, n)
        assert same_as_golden(st["exp_avg_sq"].numpy(), G[f"{tag}/{n}/exp_avg_sq"], atol=1e-12), (tag, n)
        assert float(st["step"]) == 5.0
    assert len(m.optimizer.state) == len(NAMES)              # no stale entries of replaced parameters
    for a in ("face_ids", "pos_gradient_accum", "denom", "max_radii2D"):
        assert same_as_golden(getattr(m, a).numpy(), G[f"{tag}/{a}"]), (tag, a)


def test_stats_then_clone():
    m = fresh()
    m.add_densification_stats(torch.from_numpy(G["in/view_grad"]), torch.from_numpy(G["in/update_filter"]))
    check(m, "stats")
    grads = m.pos_gradient_accum / m.denom
    grads[grads.isnan()] = 0.0
    m.densify_and_clone(grads, 1.2, 20.0)
    assert m.num_gaussians == G["clone/opacity"].shape[0] > G["in/opacity"].shape[0]
    check(m, "clone")


def test_reset_opacity():
    m = fresh()
    m.reset_opacity()
    check(m, "reset_opacity")
    assert float(m.get_opacity.detach().max()) <= 0.01 + 1e-7


def test_prune_with_and_without_screen_size():
    m = fresh()
    m.prune(0.0, 0.45, 1.0, 25.0)
    assert m.num_gaussians == G["prune/opacity"].shape[0]
    check(m, "prune")
    m = fresh()
    m.prune(0.0, 0.45, 1.0, None)
    check(m, "prune_opacity_only")


def test_split_replaces_large_gaussians_by_samples():
    m = fresh()
    m.add_densification_stats(torch.from_numpy(G["in/view_grad"]), torch.from_numpy(G["in/update_filter"]))
    grads = m.pos_gradient_accum / m.denom
    grads[grads.isnan()] = 0.0
    torch.manual_seed(1234)  # the same global-generator state the golden run drew its samples from
    m.densify_and_split(grads, 1.2, 0.5)
    assert m.num_gaussians == G["split/opacity"].shape[0] != G["in/opacity"].shape[0]
    check(m, "split")


def test_cleanup_reanchors_gaussians_that_left_their_face():
    m = fresh()
    with torch.no_grad():
        m.face_bary[[2, 5, 5, 11, 17], [0, 1, 2, 2, 1]] = torch.tensor([-0.2, -0.05, -0.3, -0.4, -0.01])
    before = m.face_ids.clone()
    m.cleanup_barycentric_coordinates()
    assert same_as_golden(m.face_bary.detach().numpy(), G["cleanup/face_bary"])
    assert np.array_equal(m.face_ids.numpy(), G["cleanup/face_ids"])
    assert int((m.face_ids != before).sum()) == 3 and float(m.face_bary.detach().min()) > 0   # both branches were taken


def test_densification_schedule_of_the_training_loop():
    """train_utils.densification at three iterations: statistics only (50), clone + split without the screen-size
    criterion (300), clone + split + prune with it (400) -- the same final state as the reference's function."""
    import json
    import types

    from scene_reconstruction.train_utils import densification

    opt = types.SimpleNamespace(**json.loads(str(G["in/opt"])))
    for it in (50, 300, 400):
        m = fresh()
        m.pos_gradient_accum += 3.0 * torch.rand((m.num_gaussians, 1), generator=torch.Generator().manual_seed(it))
        m.denom += 2
        torch.manual_seed(77 + it)
        with torch.no_grad():
            densification(m, it, torch.from_numpy(G["in/vis"]), torch.from_numpy(G["in/new_radii"]),
                          torch.from_numpy(G["in/view_grad"]), opt, 0.5)
        assert m.num_gaussians == G[f"loop{it}/opacity"].shape[0]
        check(m, f"loop{it}")
