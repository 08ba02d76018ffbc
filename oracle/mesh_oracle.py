"""TEST INFRASTRUCTURE ONLY: torch (fp64-capable, autograd) restatement of the mesh -> Gaussian transform of
scene_reconstruction/gaussian_mesh.py:151-188 (KTH-RPL/cloth-splatting @ f7acc90).

get_xyz follows the reference's four torch lines (:160-169).  get_rotation (:171-188) calls `roma`
(requirements.txt:17, unpinned, not vendored, not installable here): the three functions it uses are restated
from roma's published algorithms (roma 1.x, `roma/utils.py`, `roma/mappings.py`):
  rigid_points_registration(x, y): centre both point sets, M = sum_i yhat_i xhat_i^T, R = special_procrustes(M)
  special_procrustes(M):           U S V^T = svd(M);  R = U diag(1, 1, det(U) det(V)) V^T
  rotmat_to_unitquat(R):           SciPy's branch on argmax(R00, R11, R22, trace), XYZW, normalised
  quat_composition([p, q]):        quat_product(p, q) = (p_w q_v + q_w p_v + p_v x q_v,  p_w q_w - <p_v, q_v>)
PARITY UNPINNED for get_rotation: no roma here, and the reference class does not import."""
import torch
import torch.nn.functional as F


def get_xyz(vertices, face, face_ids, face_bary):
    vid = face[:, face_ids].transpose(0, 1)
    face_pos = vertices[vid, :]
    nb = face_bary / face_bary.sum(dim=1, keepdim=True)
    return (nb.unsqueeze(1) @ face_pos).squeeze(1)


def special_procrustes(M):
    U, _, Vh = torch.linalg.svd(M)
    d = torch.det(U) * torch.det(Vh)
    D = torch.diag_embed(torch.stack([torch.ones_like(d), torch.ones_like(d), d], dim=-1))
    return U @ D @ Vh


def rigid_points_registration(x, y):
    xh = x - x.mean(dim=-2, keepdim=True)
    yh = y - y.mean(dim=-2, keepdim=True)
    M = torch.einsum("bik,bil->bkl", yh, xh)
    return special_procrustes(M)


def rotmat_to_unitquat(R):
    n = R.shape[0]
    dec = torch.cat([torch.diagonal(R, dim1=1, dim2=2), torch.diagonal(R, dim1=1, dim2=2).sum(1, keepdim=True)], 1)
    choice = dec.argmax(dim=1)
    cols = []
    ar = torch.arange(n, device=R.device)
    i = choice.clamp(max=2)
    j = (i + 1) % 3
    k = (j + 1) % 3
    qi = 1 - dec[:, 3] + 2 * R[ar, i, i]
    qj = R[ar, j, i] + R[ar, i, j]
    qk = R[ar, k, i] + R[ar, i, k]
    q3 = R[ar, k, j] - R[ar, j, k]
    quat_a = torch.zeros((n, 4), dtype=R.dtype, device=R.device)
    quat_a = quat_a.scatter(1, i[:, None], qi[:, None]) + torch.zeros_like(quat_a).scatter(1, j[:, None], qj[:, None]) \
        + torch.zeros_like(quat_a).scatter(1, k[:, None], qk[:, None])
    quat_a = torch.cat([quat_a[:, :3], q3[:, None]], 1)
    quat_b = torch.stack([R[:, 2, 1] - R[:, 1, 2], R[:, 0, 2] - R[:, 2, 0], R[:, 1, 0] - R[:, 0, 1], 1 + dec[:, 3]], 1)
    quat = torch.where((choice == 3)[:, None], quat_b, quat_a)
    del cols
    return quat / quat.norm(dim=1, keepdim=True)


def quat_product(p, q):
    v = p[:, 3:4] * q[:, :3] + q[:, 3:4] * p[:, :3] + torch.cross(p[:, :3], q[:, :3], dim=-1)
    w = p[:, 3] * q[:, 3] - (p[:, :3] * q[:, :3]).sum(-1)
    return torch.cat([v, w[:, None]], 1)


def get_rotation(rest_vertices, vertices, face, face_ids, rotation):
    """gaussian_mesh.py:171-188 with deformed_vertices given."""
    q = F.normalize(rotation)
    vid = face[:, face_ids].transpose(0, 1)
    R = rigid_points_registration(rest_vertices[vid, :], vertices[vid, :])
    return quat_product(q, rotmat_to_unitquat(R))
