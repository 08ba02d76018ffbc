"""GPU parity of the barycentric mesh -> Gaussian transform against the reference expression
(scene_reconstruction/gaussian_mesh.py:160-169) evaluated with plain torch in fp64/fp32 on the CPU."""
import numpy as np
import pytest
import torch

import csb_synth as S

pytestmark = pytest.mark.gpu


def _ref_xyz(verts, face, face_ids, bary):
    vid = face[:, face_ids].transpose(0, 1)
    nb = bary / bary.sum(dim=1, keepdim=True)
    return (nb.unsqueeze(1) @ verts[vid, :]).squeeze(1)


@pytest.mark.parametrize("n,per_face", [(20, 14), (5, 1), (40, 3)])
def test_get_xyz_forward_backward(n, per_face):
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    verts, faces, _ = S.cloth_mesh(n)
    rng = np.random.default_rng(n)
    F = faces.shape[0]
    face_ids = np.sort(np.tile(np.arange(F), per_face))
    bary = np.clip(rng.normal(1 / 3, 0.05, (face_ids.size, 3)), 0.01, 1).astype(np.float32)  # un-normalised on purpose
    deformed = (verts + rng.normal(0, 0.01, verts.shape)).astype(np.float32)
    gout = rng.standard_normal((face_ids.size, 3)).astype(np.float32)

    vt = torch.from_numpy(deformed).double().requires_grad_(True)
    bt = torch.from_numpy(bary).double().requires_grad_(True)
    ft, it = torch.from_numpy(np.ascontiguousarray(faces.T)), torch.from_numpy(face_ids)
    ref = _ref_xyz(vt, ft, it, bt)
    ref.backward(torch.from_numpy(gout).double())

    m = MultiGaussianMesh(torch.from_numpy(verts).cuda(), ft.cuda(), it.cuda(), torch.from_numpy(bary).cuda())
    dv = torch.from_numpy(deformed).cuda().requires_grad_(True)
    out = m.get_xyz(dv)
    out.backward(torch.from_numpy(gout).cuda())
    torch.cuda.synchronize()
    assert float((out.detach().cpu().double() - ref.detach()).abs().max()) < 1e-6
    gv, gb = dv.grad.cpu().double(), m.face_bary.grad.cpu().double()
    assert float((gv - vt.grad).norm() / vt.grad.norm()) < 1e-5
    assert float((gb - bt.grad).norm() / bt.grad.norm()) < 1e-5
    # undeformed path uses mesh.pos
    out0 = m.get_xyz()
    ref0 = _ref_xyz(torch.from_numpy(verts), ft, it, torch.from_numpy(bary))
    assert float((out0.detach().cpu() - ref0).abs().max()) < 1e-6


def test_get_xyz_empty():
    from scene_reconstruction.gaussian_mesh import barycentric_xyz

    verts, faces, _ = S.cloth_mesh(4)
    out = barycentric_xyz(torch.from_numpy(verts).cuda(), torch.from_numpy(np.ascontiguousarray(faces.T)).cuda(),
                          torch.zeros(0, dtype=torch.int64).cuda(), torch.zeros((0, 3)).cuda())
    assert out.shape == (0, 3)
