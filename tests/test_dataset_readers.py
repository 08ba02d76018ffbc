"""transforms_{train,test}.json -> per-view matrices, pinned on matrices the reference's own
utils/graphics_utils.py produced for the same file (tests/make_golden_cameras.py -> tests/golden/cameras_transforms.npz)."""
import json
import os
import sys

import numpy as np

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cameras_transforms.npz")


def _scene(tmp_path):
    g = np.load(GOLD)
    contents = json.loads(str(g["transforms"]))
    (tmp_path / "transforms_train.json").write_text(json.dumps(contents))
    test = dict(contents, frames=[dict(contents["frames"][0], time=0.75)])
    (tmp_path / "transforms_test.json").write_text(json.dumps(test))
    return g, contents


def test_matrices_match_reference_graphics_utils(tmp_path):
    from scene_reconstruction import dataset_readers as dr

    g, contents = _scene(tmp_path)
    cams = dr.read_cameras_from_transforms(str(tmp_path), "transforms_train.json")
    assert len(cams) == len(contents["frames"]) == 6
    for i, cam in enumerate(cams):
        m = dr.view_matrices(cam)
        assert m.world_view_transform.dtype == np.float32 and m.world_view_transform.flags["C_CONTIGUOUS"]
        assert np.array_equal(m.world_view_transform, g["wvt"][i])          # same numpy ops, same fp32 rounding
        assert np.allclose(m.full_proj_transform, g["full"][i], rtol=0, atol=2e-6)
        assert np.allclose(m.camera_center, g["center"][i], rtol=0, atol=2e-6)
        assert abs(m.tanfovx - np.tan(0.6911 / 2)) < 1e-12 and abs(m.tanfovy - np.tan(0.55 / 2)) < 1e-12
        # the rasterizer reads them column-major: last row of W2C^T is the translation
        assert abs(float(m.world_view_transform[3, 3]) - 1.0) < 1e-6 and np.allclose(m.world_view_transform[:3, 3], 0)


def test_ids_skips_and_timeline(tmp_path):
    from scene_reconstruction import dataset_readers as dr

    _scene(tmp_path)
    cams = dr.read_cameras_from_transforms(str(tmp_path), "transforms_train.json")
    assert [(c.view_id, c.time_id) for c in cams] == [(0, 0), (0, 1), (1, 0), (1, 1), (2, 0), (2, 1)]
    assert cams[0].image_path.endswith("train/r_0_0.png") and cams[0].image_name == "r_0_0"
    assert [c.uid for c in dr.read_cameras_from_transforms(str(tmp_path), "transforms_train.json", view_skip=2)] == [0, 1, 4, 5]
    assert [c.time for c in dr.read_cameras_from_transforms(str(tmp_path), "transforms_train.json", time_skip=2)] == [0.0] * 3
    mapper, mx = dr.read_timeline(str(tmp_path))
    assert mx == 0.75 and mapper == {0.0: 0.0, 0.5: 0.5 / 0.75, 0.75: 1.0}


def test_ids_from_matrices_when_names_carry_none(tmp_path):
    from scene_reconstruction import dataset_readers as dr

    g, contents = _scene(tmp_path)
    for i, fr in enumerate(contents["frames"]):
        fr["file_path"] = f"./train/img{i}.jpg"
    (tmp_path / "t.json").write_text(json.dumps(contents))
    cams = dr.read_cameras_from_transforms(str(tmp_path), "t.json")
    assert sorted({c.view_id for c in cams}) == [0, 1, 2] and [c.time_id for c in cams] == [0, 1] * 3
    assert cams[3].image_path.endswith("img3.jpg")
    # frames that share a transform share a view id
    assert cams[0].view_id == cams[1].view_id != cams[2].view_id


def test_raster_settings_agree_with_the_synthetic_cameras(tmp_path):
    """A csb_synth ring camera written out as a Blender-style transform and read back gives the settings the bench and the
    parity tests render with (same transposed matrices, same tan(fov/2), same camera centre)."""
    import csb_synth as S
    from scene_reconstruction import dataset_readers as dr

    cams = S.ring_cameras(3, 96, 64, 0.6911, np.array([0.1, 0.0, 0.05], np.float32), fovy=0.5)
    frames = []
    for v, c in enumerate(cams):
        w2c = c.world_view_transform.T.astype(np.float64)
        c2w = np.linalg.inv(w2c)
        c2w[:3, 1:3] *= -1  # COLMAP axes -> the OpenGL axes the file format uses
        frames.append({"file_path": f"./train/r_{v}_0", "time": 0.0, "transform_matrix": c2w.tolist()})
    (tmp_path / "t.json").write_text(json.dumps({"camera_angle_x": 0.6911, "camera_angle_y": 0.5, "frames": frames}))
    got = dr.read_cameras_from_transforms(str(tmp_path), "t.json")
    for c, g in zip(cams, got):
        s = dr.raster_settings(g, c.image_width, c.image_height, (1.0, 1.0, 1.0), 3, device="cpu")
        assert s._fields[:4] == ("image_height", "image_width", "tanfovx", "tanfovy")
        assert (s.image_height, s.image_width, s.sh_degree) == (64, 96, 3)
        assert abs(s.tanfovx - c.tanfovx) < 1e-12 and abs(s.tanfovy - c.tanfovy) < 1e-12
        assert np.allclose(s.viewmatrix.numpy(), c.world_view_transform, atol=2e-6)
        assert np.allclose(s.projmatrix.numpy(), c.full_proj_transform, atol=5e-6)
        assert np.allclose(s.campos.numpy(), c.camera_center, atol=2e-6)


def test_nerf_normalization_is_the_reference_cameras_extent(tmp_path):
    from scene_reconstruction import dataset_readers as dr

    g, _ = _scene(tmp_path)
    norm = dr.nerf_normalization(dr.read_cameras_from_transforms(str(tmp_path), "transforms_train.json"))
    assert np.array_equal(norm["translate"], g["nerf_translate"]) and norm["radius"] == float(g["nerf_radius"])
