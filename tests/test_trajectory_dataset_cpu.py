"""Sample indexing of the simulator's training set (meshnet/trajectory_dataset.py) against the reference's
`TrajectoriesClothSimDataset.__getitem__` for every sample of three trajectories, at two (history, future) settings
(golden: tests/make_golden_trajectory_dataset.py)."""
import os

import numpy as np
import pytest

import fake_scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "trajectory_samples.npz"))


@pytest.mark.parametrize("L,F", [(1, 1), (2, 3)])
def test_every_sample_is_the_reference_graph(L, F):
    from meshnet.trajectory_dataset import TrajectorySamples

    ds = TrajectorySamples(fake_scene.trajectories(), L, F)
    assert len(ds) == int(G[f"L{L}F{F}/len"]) == sum(T - L - F + 1 for T in (9, 7, 12))
    for idx in range(len(ds)):
        g = ds[idx]
        for k in ("x", "face", "y", "pos", "pos_target", "vel", "particle_actions"):
            assert np.array_equal(getattr(g, k).numpy(), G[f"L{L}F{F}/{idx}/{k}"]), (idx, k)
        assert g.grasped_particle == int(G[f"L{L}F{F}/{idx}/grasped"])
        assert g.x.shape == (6, 3 * L + 1) and g.y.shape == (6, F, 3)
    # windows: consecutive inside a trajectory, restarting at L in the next one
    w = [ds.window(i) for i in range(len(ds))]
    assert w[0] == (0, L) and w[-1] == (2, 12 - F) and all(b == (a[0], a[1] + 1) or b == (a[0] + 1, L) for a, b in zip(w, w[1:]))
