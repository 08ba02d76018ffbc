"""Mesh files between the GNN rollout and the refinement (meshnet/mesh_data.py).  torch_geometric is not in the image, so
these are property tests of the restated transforms (parity unpinned), plus consistency with the synthetic cloth the
benches use and a write / read round trip."""
import numpy as np
import torch

import csb_synth as S


def test_face_to_edge_is_sorted_symmetric_unique_and_matches_the_synthetic_cloth():
    from meshnet.mesh_data import face_to_edge

    verts, faces, ei = S.cloth_mesh(9)
    e = face_to_edge(torch.from_numpy(np.ascontiguousarray(faces.T)), verts.shape[0])
    assert np.array_equal(e.numpy(), ei)                                  # the edge order the GNN kernels are fed with
    key = e[0] * verts.shape[0] + e[1]
    assert bool((key[1:] > key[:-1]).all())
    assert set(map(tuple, e.T.tolist())) == set(map(tuple, e.flip(0).T.tolist()))
    for a, b, c in faces[:20]:
        assert {(a, b), (b, c), (c, a)} <= set(map(tuple, e.T.tolist()))


def test_normals_and_edge_features():
    from meshnet.mesh_data import cartesian_distance, face_to_edge, mesh_normals

    verts, faces, _ = S.cloth_mesh(6)
    pos = torch.from_numpy(verts).float()
    face = torch.from_numpy(np.ascontiguousarray(faces.T))
    flat = pos.clone()
    flat[:, 2] = 0
    n = mesh_normals(flat, face)
    assert torch.allclose(n.abs(), torch.tensor([0.0, 0.0, 1.0]).expand_as(n)) and bool((n[:, 2] == n[0, 2]).all())
    n = mesh_normals(pos, face)
    assert torch.allclose(n.norm(dim=1), torch.ones(pos.shape[0]), atol=1e-6)
    q, _ = torch.linalg.qr(torch.randn((3, 3), generator=torch.Generator().manual_seed(0)))
    assert torch.allclose(mesh_normals(pos @ q.T * 2.0 + 1.0, face), n @ q.T * torch.sign(torch.det(q)), atol=1e-5)
    e = face_to_edge(face, pos.shape[0])
    f = cartesian_distance(pos, e)
    assert f.shape == (e.shape[1], 4) and torch.allclose(f[:, :3].norm(dim=1), f[:, 3])
    assert torch.equal(f[:, :3], pos[e[0]] - pos[e[1]])


def test_prediction_files_round_trip(tmp_path):
    from meshnet.mesh_data import load_mesh, predicted_positions, write_mesh_predictions

    verts, faces, ei = S.cloth_mesh(5)
    rng = np.random.default_rng(0)
    roll = rng.standard_normal((4, verts.shape[0], 3)).astype(np.float32) * 0.01
    pos = predicted_positions(verts.astype(np.float32), roll)
    assert pos.shape == (5, verts.shape[0], 3) and np.allclose(pos[3], verts + roll[:3].sum(0), atol=1e-6)
    paths = write_mesh_predictions(str(tmp_path), verts.astype(np.float32), roll, faces.T)
    assert len(paths) == 6 and paths[1].endswith("mesh_000.hdf5") and paths[0].endswith("init_mesh.hdf5") and "mesh_predictions" in paths[-1]
    first, start, last = (load_mesh(p, device="cpu") for p in (paths[0], paths[1], paths[-1]))
    assert torch.equal(first.pos, start.pos) and np.array_equal(last.pos.numpy(), pos[4])
    assert np.array_equal(last.edge_index.numpy(), ei) and np.array_equal(last.face.numpy(), faces.T)
    assert last.norm.shape == last.pos.shape and last.face.dtype == torch.int64
    # what the refinement side builds its simulator from
    from meshnet.meshnet_network import ResidualMeshSimulator

    sim = ResidualMeshSimulator(torch.stack([load_mesh(p, device="cpu").pos for p in paths[1:]]), device="cpu")
    out = sim(torch.full((verts.shape[0], 1), 0.5))
    assert out.shape == (verts.shape[0], 3) and torch.allclose(out, torch.from_numpy(pos[2]), atol=1e-2)
