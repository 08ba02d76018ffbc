"""Open-loop `rollout` (meshnet/rollout.py) against the reference's own function (tests/make_golden_rollout.py cuts it out
of train_meshnet_sim.py) with the same toy simulator, and through to the mesh prediction files."""
import os

import numpy as np
from helpers import same_as_golden
import pytest
import torch

import fake_scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rollout_seed12.npz"))


@pytest.mark.parametrize("tag,real_world", [("sim", False), ("real_world", True)])
def test_rollout_matches_the_reference_function(tag, real_world):
    from meshnet.rollout import rollout

    f = fake_scene.rollout_features()
    keep = {k: v.clone() for k, v in f.items() if isinstance(v, torch.Tensor)}
    r = rollout(fake_scene.ToySimulator(), f, 4, torch.device("cpu"), 2, 0.01, real_world=real_world)
    for k in ("predicted_rollout", "ground_truth_rollout", "mean_loss"):
        assert same_as_golden(np.asarray(r[k]), G[f"{tag}/{k}"]), k
    assert all(torch.equal(f[k], v) for k, v in keep.items())               # the trajectory is left as it was
    assert np.array_equal(r["initial_pos"], keep["pos"][0].numpy()) and r["predicted_rollout"].shape == (4, 16, 3)
    assert np.array_equal(r["predicted_rollout"][:, 5], f["actions"][:4].numpy())   # the grasped particle follows the actions


def test_rollout_to_prediction_files(tmp_path):
    from meshnet.mesh_data import load_mesh, write_mesh_predictions
    from meshnet.rollout import rollout

    f = fake_scene.rollout_features()
    r = rollout(fake_scene.ToySimulator(), f, 4, torch.device("cpu"), 2, 0.01)
    paths = write_mesh_predictions(str(tmp_path), r["initial_pos"], r["predicted_rollout"], r["faces"])
    last = load_mesh(paths[-1], device="cpu")
    assert np.allclose(last.pos.numpy(), r["initial_pos"] + r["predicted_rollout"].sum(0), atol=1e-6)
    assert np.array_equal(last.edge_index.numpy(), r["edge_index"])


def test_position_losses_follow_the_reference_loop():
    from meshnet.mesh_data import position_losses
    from meshnet.rollout import rollout

    r = rollout(fake_scene.ToySimulator(), fake_scene.rollout_features(), 4, torch.device("cpu"), 2, 0.01)
    losses = position_losses(r)
    want = [0]
    for i in range(len(r["predicted_rollout"])):                        # train_meshnet_sim.py:296-301, as written there
        pred_points = r["initial_pos"] + r["predicted_rollout"][:i + 1].sum(axis=0)
        want.append(((r["node_coords"][i + 1] - pred_points) ** 2).mean().item())
    assert len(losses) == 5 and np.allclose(losses, want, rtol=1e-5, atol=0) and losses[0] == 0 and min(losses[1:]) > 0
