"""On-disk format of the mesh-anchored Gaussians: `point_cloud.ply` as the reference writes it with
plyfile (scene_reconstruction/gaussian_mesh.py:433-465, gaussian_model.py:181-262).  CPU only: the byte layout, the
column order and the save -> load round trip of the mirror class (positions through a torch get_xyz: no GPU here)."""
import os
import sys

import numpy as np
import torch

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)


def test_header_and_payload_bytes(tmp_path):
    from scene_reconstruction import ply_io

    names = ply_io.gaussian_attribute_names(3, 45)
    assert names[:6] == ["x", "y", "z", "nx", "ny", "nz"] and names[6:9] == ["f_dc_0", "f_dc_1", "f_dc_2"]
    assert names[9] == "f_rest_0" and names[53] == "f_rest_44" and names[54] == "opacity"
    assert names[55:58] == ["scale_0", "scale_1", "scale_2"] and names[58:62] == ["rot_0", "rot_1", "rot_2", "rot_3"]
    assert names[62:] == ["b1", "b2", "b3", "o", "id"] and len(names) == 67
    rng = np.random.default_rng(0)
    table = rng.standard_normal((5, len(names))).astype(np.float32)
    path = str(tmp_path / "pc" / "point_cloud.ply")
    ply_io.write_vertex_ply(path, names, table)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    # exactly what plyfile emits for PlyData([PlyElement.describe(float32 struct array, 'vertex')])
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 5"]
    assert lines[3:] == [f"property float {n}" for n in names]
    assert body == table.astype("<f4").tobytes()
    col = ply_io.read_vertex_ply(path)
    assert list(col) == names
    for i, n in enumerate(names):
        assert np.array_equal(col[n], table[:, i])


def test_reads_ascii_and_other_scalar_types(tmp_path):
    from scene_reconstruction import ply_io

    p = tmp_path / "a.ply"
    p.write_text("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\nproperty float x\nproperty uchar red\n"
                 "element face 0\nproperty list uchar int vertex_indices\nend_header\n0.5 7\n-1.25 255\n")
    col = ply_io.read_vertex_ply(str(p))
    assert col["x"].tolist() == [0.5, -1.25] and col["red"].tolist() == [7, 255] and col["red"].dtype == np.uint8


def test_mesh_gaussians_round_trip(tmp_path):
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    class CpuMesh(MultiGaussianMesh):  # get_xyz of the reference in torch (gaussian_mesh.py:151-169)
        def get_xyz(self, deformed_vertices=None):
            vid = self.mesh.face[:, self.face_ids].T
            b = self.face_bary / self.face_bary.sum(dim=1, keepdim=True)
            return (b[:, None, :] @ self.mesh.pos[vid]).squeeze(1)

    g = torch.Generator().manual_seed(3)
    pos = torch.rand((9, 3), generator=g)
    face = torch.tensor([[0, 1, 3, 4], [1, 2, 4, 5], [3, 4, 6, 7]])
    P = 11
    m = CpuMesh(pos, face, torch.randint(0, 4, (P,), generator=g), torch.rand((P, 3), generator=g) + 0.1, sh_degree=3)
    m.set_gaussians(torch.randn((P, 1, 3), generator=g), torch.randn((P, 15, 3), generator=g),
                    torch.randn((P, 3), generator=g), torch.randn((P, 4), generator=g), torch.randn((P, 1), generator=g))
    m.face_offset = torch.nn.Parameter(torch.randn((P, 1), generator=g))
    m.save_ply(str(tmp_path))
    assert os.path.exists(tmp_path / "point_cloud.ply")
    from scene_reconstruction import ply_io

    col = ply_io.read_vertex_ply(str(tmp_path / "point_cloud.ply"))
    assert np.allclose(np.stack((col["x"], col["y"], col["z"]), 1), m.get_xyz().detach().numpy())
    assert float(np.abs(col["nx"]).max()) == 0.0
    # channel-major SH columns: f_rest_k = features_rest[:, k % 15, k // 15]
    assert np.array_equal(col["f_rest_16"], m._features_rest.detach().numpy()[:, 1, 1])
    assert np.array_equal(col["f_dc_2"], m._features_dc.detach().numpy()[:, 0, 2])
    r = CpuMesh.load_ply(str(tmp_path), sh_degree=3, device="cpu")
    assert torch.equal(r.face_ids, m.face_ids) and torch.equal(r.mesh.face, m.mesh.face)
    for a in ("face_bary", "face_offset", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity"):
        assert torch.equal(getattr(r, a).detach(), getattr(m, a).detach()), a
    assert torch.equal(r.mesh.pos, m.mesh.pos) and r.active_sh_degree == 3
    # the mesh file carries every dataset the reference's loader reads (meshnet/data_utils.py:450-457) ...
    from meshnet.mesh_data import read_datasets

    assert open(tmp_path / "mesh.hdf5", "rb").read(8) == b"\x89HDF\r\n\x1a\n"       # a real HDF5 container, h5py or not
    assert set(read_datasets(str(tmp_path / "mesh.hdf5"))) == {"pos", "norm", "face", "edge_index"}
    # ... and a loaded model can run the regularisers (they read mesh.edge_index and edge_norm, set by set_edges)
    import types

    from scene_reconstruction.train_utils import regularization

    opt = types.SimpleNamespace(lambda_deform_mag=0.1, lambda_rigid=0.5, lambda_momentum=0.1)
    deform = torch.stack([r.mesh.pos, r.mesh.pos * 1.01, r.mesh.pos * 1.03])
    assert r.mesh.edge_index.shape[0] == 2 and r.edge_norm.shape == (r.mesh.edge_index.shape[1], 1)
    assert float(regularization(deform, r, opt)) > 0
