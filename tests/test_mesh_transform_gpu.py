"""csb_mesh_transform_* (csrc/mesh_transform.cu): get_xyz + get_rotation of several deformed meshes in one launch, against
the two single-purpose kernel pairs (same arithmetic: forward bit for bit) and against the fp64 oracle of
scene_reconstruction/gaussian_mesh.py:151-188 (oracle/mesh_oracle.py, autograd)."""
import numpy as np
import pytest
import torch

import conftest  # noqa: F401
import csb_synth as S
from oracle import mesh_oracle

pytestmark = pytest.mark.gpu


def _scene(n=9, per_face=5, views=3, seed=0):
    verts, faces, _ = S.cloth_mesh(n)
    g = torch.Generator().manual_seed(seed)
    rest = torch.from_numpy(verts).float()
    face = torch.from_numpy(np.ascontiguousarray(faces.T)).long()
    F = face.shape[1]
    ids = torch.arange(F).repeat(per_face)[torch.randperm(F * per_face, generator=g)]   # unsorted, like after densification
    bary = torch.rand((ids.numel(), 3), generator=g) + 0.05
    rot = torch.randn((ids.numel(), 4), generator=g)
    ang = torch.linspace(0.1, 0.9, views)
    base = []
    for a in ang:   # a rigid motion + a bend per view
        c, s_ = torch.cos(a), torch.sin(a)
        R = torch.tensor([[c, -s_, 0.0], [s_, c, 0.0], [0.0, 0.0, 1.0]])
        v = rest @ R.T + 0.05 * a
        v[:, 2] += 0.1 * torch.sin(6 * rest[:, 0] + a)
        base.append(v)
    base = torch.stack(base)
    residual = 0.01 * torch.randn(base.shape, generator=g)
    return rest, face, ids, bary, rot, base, residual


@pytest.mark.parametrize("with_residual", [False, True])
def test_fused_transform_equals_the_separate_kernels(with_residual):
    from scene_reconstruction.gaussian_mesh import barycentric_xyz, mesh_rotation, mesh_transform

    rest, face, ids, bary, rot, base, residual = (t.cuda() for t in _scene())
    V = base.shape[0]
    gx = torch.randn((V, ids.numel(), 3), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    gr = torch.randn((V, ids.numel(), 4), device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))

    def leaves():
        return [t.clone().requires_grad_(True) for t in (base, residual, bary, rot)]

    b1, r1, ba1, q1 = leaves()
    xyz, rq = mesh_transform(rest, b1, face, ids, ba1, q1, residual=r1 if with_residual else None)
    ((xyz * gx).sum() + (rq * gr).sum()).backward()

    b2, r2, ba2, q2 = leaves()
    verts = b2 + r2 if with_residual else b2
    xs = torch.stack([barycentric_xyz(verts[v], face, ids, ba2) for v in range(V)])
    rs = torch.stack([mesh_rotation(rest, verts[v], face, ids, q2) for v in range(V)])
    ((xs * gx).sum() + (rs * gr).sum()).backward()

    assert torch.equal(xyz, xs) and torch.equal(rq, rs)          # same expression trees
    def close(a, b, tol=2e-5):
        return float((a - b).abs().max()) <= tol * max(float(b.abs().max()), 1e-6)
    assert close(ba1.grad, ba2.grad) and close(q1.grad, q2.grad) and close(b1.grad, b2.grad)
    if with_residual:
        assert close(r1.grad, r2.grad) and torch.equal(r1.grad, b1.grad)
    # a single mesh [Nv,3] gives [P,3] / [P,4]
    x0, q0 = mesh_transform(rest, verts[0].detach(), face, ids, bary, rot)
    assert x0.shape == (ids.numel(), 3) and torch.equal(x0, xs[0]) and torch.equal(q0, rs[0])


def test_fused_transform_against_the_fp64_oracle():
    from scene_reconstruction.gaussian_mesh import mesh_transform

    rest, face, ids, bary, rot, base, residual = _scene(n=7, per_face=3, views=2, seed=3)
    V = base.shape[0]
    g = torch.Generator().manual_seed(5)
    gx, gr = torch.randn((V, ids.numel(), 3), generator=g), torch.randn((V, ids.numel(), 4), generator=g)
    # oracle, fp64 autograd
    bo, bao, qo = base.double().requires_grad_(True), bary.double().requires_grad_(True), rot.double().requires_grad_(True)
    tot = 0
    ref_x, ref_q = [], []
    for v in range(V):
        x = mesh_oracle.get_xyz(bo[v], face, ids, bao)
        q = mesh_oracle.get_rotation(rest.double(), bo[v], face, ids, qo)
        ref_x.append(x.detach()); ref_q.append(q.detach())
        tot = tot + (x * gx[v].double()).sum() + (q * gr[v].double()).sum()
    tot.backward()
    bc, bac, qc = base.cuda().requires_grad_(True), bary.cuda().requires_grad_(True), rot.cuda().requires_grad_(True)
    xyz, rq = mesh_transform(rest.cuda(), bc, face.cuda(), ids.cuda(), bac, qc)
    ((xyz * gx.cuda()).sum() + (rq * gr.cuda()).sum()).backward()
    assert float((xyz.cpu().double() - torch.stack(ref_x)).abs().max()) < 1e-6
    # q and -q are the same rotation: the oracle's SVD may pick the other sign of the relative quaternion
    rq_c, rq_o = rq.cpu().double(), torch.stack(ref_q)
    sign = torch.sign((rq_c * rq_o).sum(-1, keepdim=True))
    assert float((rq_c - sign * rq_o).abs().max()) < 2e-5
    def rel(a, b):
        return float((a.cpu().double() - b).norm() / b.norm())
    assert rel(bac.grad, bao.grad) < 1e-4
    if float(sign.min()) > 0:   # gradients of the rotation output follow the sign convention
        assert rel(qc.grad, qo.grad) < 2e-3 and rel(bc.grad, bo.grad) < 2e-3


def test_empty_and_out_of_range_faces():
    from scene_reconstruction.gaussian_mesh import mesh_transform

    rest, face, ids, bary, rot, base, residual = (t.cuda() for t in _scene(n=5, per_face=2, views=2))
    x, q = mesh_transform(rest, base, face, ids[:0], bary[:0], rot[:0])
    assert x.shape == (2, 0, 3) and q.shape == (2, 0, 4)
    bad = ids.clone()
    bad[3] = face.shape[1] + 7
    b = base.clone().requires_grad_(True)
    qp = rot.clone().requires_grad_(True)
    x, q = mesh_transform(rest, b, face, bad, bary, qp)
    assert torch.equal(x[:, 3], torch.zeros_like(x[:, 3]))                                       # no face: position 0 ...
    assert torch.allclose(q[:, 3], torch.nn.functional.normalize(rot[3:4]).expand(2, 4), atol=1e-6)   # ... rotation = the parameter
    (x.sum() + q.sum()).backward()
    assert torch.isfinite(b.grad).all() and torch.isfinite(qp.grad).all()
