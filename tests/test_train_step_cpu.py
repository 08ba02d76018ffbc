"""The glue of one refinement iteration (`scene_reconstruction.train_utils.train_step`) against the
reference's own `train_step` run on the same CPU stand-in scene (tests/fake_scene.py; golden from
tests/make_golden_train_step.py): returned psnr / loss / loss_dict, every Gaussian parameter and Adam moment, the
densification statistics, the SH degree and the simulator's parameters after two iterations, for nine iterations that
take different branches of the schedule.  The kernels behind render / image_losses are replaced by torch stand-ins here
(their parity is tests/test_*_gpu.py); tests/test_train_step_gpu.py runs the same function on the real ones."""
import os

import numpy as np
from helpers import same_as_golden
import pytest

import fake_scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "train_step_cases.npz"))


@pytest.fixture()
def train_step(monkeypatch):
    import gaussian_renderer
    from scene_reconstruction import train_utils

    monkeypatch.setattr(gaussian_renderer, "render", fake_scene.render)
    monkeypatch.setattr(train_utils, "image_losses", fake_scene.image_losses)
    return train_utils.train_step


@pytest.mark.parametrize("name", list(fake_scene.CASES))
def test_train_step_matches_the_reference_function(train_step, name):
    out = fake_scene.run_case(train_step, name)
    keys = {k[len(name) + 1:] for k in G.files if k.startswith(name + "/")}
    assert keys == set(out), keys ^ set(out)
    for k, v in out.items():
        assert same_as_golden(np.asarray(v), G[f"{name}/{k}"], rtol=2e-5, atol=2e-7), (name, k)


def test_cases_take_different_branches():
    P0 = G["plain/face_ids"].shape[0]
    assert G["densify/face_ids"].shape[0] > P0 and G["sh_up_reset/sh"] == 1 and G["plain/sh"] == 0
    assert G["sh_up_reset/p/opacity"].max() <= np.log(0.01 / 0.99) + 0.1     # reset, then one Adam step
    # static: the simulator is not stepped (its parameters stay what fake_scene.build() made them)
    assert not np.array_equal(G["plain/sim/output.bias"], G["static/sim/output.bias"])
    assert not np.array_equal(G["plain/p/f_dc"], G["static/p/f_dc"])
