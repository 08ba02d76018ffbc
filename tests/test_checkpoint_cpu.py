"""`MultiGaussianMesh.capture / restore`: the reference's 12-slot checkpoint tuple (gaussian_model.py:62-94; golden from
tests/make_golden_checkpoint.py) and a round trip through a second model."""
import os

import numpy as np
import torch

import fake_scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "checkpoint_slots.npz"))


def test_slot_order_and_restore_sequence_are_the_reference_ones():
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    assert list(MultiGaussianMesh.capture(fake_scene.Named())) == G["captured"].tolist()
    obj = fake_scene.Named()
    MultiGaussianMesh.restore(obj, tuple(f"slot{i}" for i in range(12)), "opt")
    assert [f"{a}={getattr(obj, a)}" for a in fake_scene.SLOTS] + obj.log == G["restored"].tolist()


def test_round_trip_carries_parameters_statistics_and_adam_moments():
    a, *_ = fake_scene.build(seed=3)
    for p in (a._opacity, a._scaling, a._features_dc):                  # one Adam step so that there are moments to carry
        p.grad = torch.ones_like(p)
    a.optimizer.step()
    a.pos_gradient_accum += 1.5
    a.active_sh_degree = 2
    a.spatial_lr_scale = 1.0
    saved = a.capture()
    assert len(saved) == 12 and saved[1].numel() == 0                    # the `_xyz` slot of a mesh-anchored model is empty

    b, *_ = fake_scene.build(seed=4)
    b.restore(saved, fake_scene.options())
    assert b.active_sh_degree == 2 and b._opacity is a._opacity and torch.equal(b.pos_gradient_accum, a.pos_gradient_accum)
    sa, sb = a.optimizer.state_dict()["state"], b.optimizer.state_dict()["state"]
    assert sa.keys() == sb.keys() and len(sb) == 3
    for k in sa:
        assert torch.equal(sa[k]["exp_avg"], sb[k]["exp_avg"]) and torch.equal(sa[k]["exp_avg_sq"], sb[k]["exp_avg_sq"])
    assert [g["name"] for g in b.optimizer.param_groups] == ["face_bary", "face_offset", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
