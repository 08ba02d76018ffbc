"""`Camera` and `MDNerfDataset` (the view x time grid whose items are the `viewpoint_cams` of one `train_step`) against the
reference's own classes (tests/make_golden_dataset.py lifts them from the source; tests/golden/dataset_cameras.npz)."""
import os
import types

import numpy as np
import pytest
import torch

import fake_scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dataset_cameras.npz"))


def _same(a, b, name):
    """Bit-identical, except what passes through a matrix inverse (`camera_center`, `full_proj_transform` with a translated
    camera): LAPACK's last bit depends on the host's vector units, so those are held to 4 ulp."""
    if name in ("camera_center",):
        return np.allclose(a, b, rtol=5e-7, atol=1e-7)
    return np.array_equal(a, b)


def test_items_are_three_consecutive_time_steps_of_one_view_drawn_like_the_reference():
    from scene_reconstruction.dataset import MDNerfDataset

    ds = MDNerfDataset(fake_scene.camera_infos(3, 5))
    assert len(ds) == int(G["len"]) == 3
    np.random.seed(17)
    for k in range(12):
        item = ds[k % 3]
        assert [[c.view_id, c.time_id] for c in item] == G["picks"][k].tolist()
        assert item[0].time < item[1].time < item[2].time and len({c.view_id for c in item}) == 1
        if k == 0:
            c = item[1]
            for a in ("world_view_transform", "projection_matrix", "full_proj_transform", "camera_center", "original_image", "mask"):
                assert _same(getattr(c, a).numpy(), G[f"cam/{a}"], a), a
            assert np.array_equal(np.array([c.FoVx, c.FoVy, c.time, c.image_width, c.image_height, c.znear, c.zfar, c.uid,
                                            c.colmap_id]), G["cam/scalars"])
            assert float(c.original_image.min()) >= 0 and float(c.original_image.max()) <= 1
    short = MDNerfDataset(fake_scene.camera_infos(2, 2, seed=5))
    assert [[[c.view_id, c.time_id] for c in short[i]] for i in range(2)] == G["short_picks"].tolist()


def test_camera_with_alpha_mask_translation_and_scale():
    from scene_reconstruction.cameras import Camera

    s0 = fake_scene.camera_infos(3, 5)[0]
    c = Camera(colmap_id=1, R=s0.R, T=s0.T, FoVx=0.7, FoVy=0.4, image=s0.image.clone(), gt_alpha_mask=torch.from_numpy(G["alpha/mask"]),
               image_name="x", uid=1, trans=np.array([0.1, -0.2, 0.3]), scale=1.5, data_device="cpu", time=0.25)
    for k in ("world_view_transform", "full_proj_transform", "camera_center", "original_image"):
        assert _same(getattr(c, k).numpy(), G[f"alpha/{k}"], k), k


def test_incomplete_grid_raises_like_the_reference():
    from scene_reconstruction.dataset import MDNerfDataset

    src = [s for s in fake_scene.camera_infos(3, 5) if not (s.view_id == 4 and s.time_id == 4)]
    ds = MDNerfDataset(src)
    np.random.seed(0)
    with pytest.raises(ValueError):
        for _ in range(20):
            ds[1]
