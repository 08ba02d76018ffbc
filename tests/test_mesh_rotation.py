"""get_rotation (scene_reconstruction/gaussian_mesh.py:171-188): the SVD-based oracle (oracle/mesh_oracle.py, roma's
published algorithms restated; PARITY UNPINNED, roma is not installable) checked on known answers on the CPU, and
the closed-form CUDA kernels (csb_mesh_rotation_*) checked against it on the GPU: forward <= 2e-5 absolute on unit
quaternions, gradients <= 2e-3 relative L2 (fp32 kernel vs fp64 autograd through the SVD)."""
import numpy as np
import pytest
import torch

import csb_synth as S
from oracle import mesh_oracle as MO


def _rand_rot(rng, n):
    q = rng.standard_normal((n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    x, y, z, w = q.T
    return np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                     np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                     np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], 1), q


def test_oracle_recovers_known_rigid_motion():
    rng = np.random.default_rng(0)
    n = 64
    R, q = _rand_rot(rng, n)
    x = rng.standard_normal((n, 3, 3))
    y = np.einsum("nab,nkb->nka", R, x) + rng.standard_normal((n, 1, 3))
    Rh = MO.rigid_points_registration(torch.from_numpy(x), torch.from_numpy(y))
    assert float((Rh - torch.from_numpy(R)).abs().max()) < 1e-9
    qh = MO.rotmat_to_unitquat(Rh).numpy()
    sign = np.sign((qh * q).sum(1, keepdims=True))
    assert np.abs(qh * sign - q).max() < 1e-9  # XYZW, up to the double cover
    ident = torch.tensor([[0.0, 0.0, 0.0, 1.0]]).double().repeat(n, 1)
    assert float((MO.quat_product(ident, torch.from_numpy(qh)) - torch.from_numpy(qh)).abs().max()) < 1e-12


def test_oracle_agrees_with_scipy():
    """roma is not installable here (no network, not in the wheelhouse), so the restated oracle is cross-checked against
    an independent third-party implementation that IS in the image: SciPy 1.18 `Rotation.align_vectors` (Kabsch with
    the det correction), `Rotation.from_matrix(...).as_quat()` (the algorithm roma's rotmat_to_unitquat is adapted
    from, XYZW) and rotation composition (= the quaternion product).  Noisy, non-rigid triangle pairs like the ones
    get_rotation sees."""
    from scipy.spatial.transform import Rotation

    rng = np.random.default_rng(1)
    n = 200
    R, _ = _rand_rot(rng, n)
    x = rng.standard_normal((n, 3, 3)) * 0.02          # rest triangle
    y = np.einsum("nab,nkb->nka", R, x) + rng.standard_normal((n, 3, 3)) * 0.004 + 0.3   # deformed, not rigid
    Rh = MO.rigid_points_registration(torch.from_numpy(x), torch.from_numpy(y)).numpy()
    qh = MO.rotmat_to_unitquat(torch.from_numpy(Rh)).numpy()
    p = rng.standard_normal((n, 4))
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    comp = MO.quat_product(torch.from_numpy(p), torch.from_numpy(qh)).numpy()
    for i in range(n):
        xc, yc = x[i] - x[i].mean(0), y[i] - y[i].mean(0)
        rs, _ = Rotation.align_vectors(yc, xc)          # y ~ R x
        assert np.abs(rs.as_matrix() - Rh[i]).max() < 1e-7, i
        qs = Rotation.from_matrix(Rh[i]).as_quat()
        assert min(np.abs(qs - qh[i]).max(), np.abs(qs + qh[i]).max()) < 1e-9, i
        cs = (Rotation.from_quat(p[i]) * Rotation.from_quat(qh[i])).as_quat()
        assert min(np.abs(cs - comp[i]).max(), np.abs(cs + comp[i]).max()) < 1e-9, i


def _case(n, per_face, seed, noise):
    verts, faces, _ = S.cloth_mesh(n)
    rng = np.random.default_rng(seed)
    F = faces.shape[0]
    face_ids = np.sort(np.tile(np.arange(F), per_face))
    R, _ = _rand_rot(rng, 1)
    deformed = (verts @ R[0].T + rng.normal(0, noise, verts.shape) + 0.1).astype(np.float32)
    rot = rng.standard_normal((face_ids.size, 4)).astype(np.float32)  # un-normalised on purpose
    gout = rng.standard_normal((face_ids.size, 4)).astype(np.float32)
    return verts, np.ascontiguousarray(faces.T), face_ids, deformed, rot, gout


@pytest.mark.gpu
@pytest.mark.parametrize("n,per_face,noise", [(20, 14, 0.01), (6, 1, 0.0), (40, 2, 0.03)])
def test_get_rotation_forward_backward_vs_oracle(n, per_face, noise):
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    verts, face, ids, deformed, rot, gout = _case(n, per_face, n, noise)
    ft, it = torch.from_numpy(face), torch.from_numpy(ids)
    vt = torch.from_numpy(deformed).double().requires_grad_(True)
    qt = torch.from_numpy(rot).double().requires_grad_(True)
    ref = MO.get_rotation(torch.from_numpy(verts).double(), vt, ft, it, qt)
    ref.backward(torch.from_numpy(gout).double())

    m = MultiGaussianMesh(torch.from_numpy(verts).cuda(), ft.cuda(), it.cuda(), torch.full((ids.size, 3), 1 / 3).cuda())
    z = torch.zeros((ids.size, 1, 3)).cuda()
    m.set_gaussians(z, z.repeat(1, 15, 1), torch.zeros((ids.size, 3)).cuda(), torch.from_numpy(rot).cuda(),
                    torch.zeros((ids.size, 1)).cuda())
    dv = torch.from_numpy(deformed).cuda().requires_grad_(True)
    out = m.get_rotation(dv)
    out.backward(torch.from_numpy(gout).cuda())
    torch.cuda.synchronize()
    assert float((out.detach().cpu().double() - ref.detach()).abs().max()) < 2e-5
    gv, gq = dv.grad.cpu().double(), m._rotation.grad.cpu().double()
    assert float((gq - qt.grad).norm() / qt.grad.norm()) < 2e-3
    assert float((gv - vt.grad).norm() / vt.grad.norm()) < 2e-3
    # without a deformed mesh: the normalised parameter
    assert float((m.get_rotation().detach().cpu() - torch.nn.functional.normalize(torch.from_numpy(rot))).abs().max()) < 1e-6


@pytest.mark.gpu
def test_render_glue_matches_plain_composition():
    """render() = simulator -> get_xyz / get_rotation -> rasterizer: same image as calling the pieces by hand, and
    the loss gradient reaches the residual MLP (the tensor the refinement loop optimises)."""
    from gaussian_renderer import render
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from meshnet.meshnet_network import ResidualMeshSimulator
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    verts, faces, _ = S.cloth_mesh(8)
    pos = torch.from_numpy(verts).cuda()
    face = torch.from_numpy(np.ascontiguousarray(faces.T)).cuda()
    g = torch.Generator(device="cuda").manual_seed(0)
    pc = MultiGaussianMesh.from_mesh(pos, face, gaussian_init_factor=3, generator=g)
    pc.active_sh_degree = 3
    with torch.no_grad():
        pc._features_rest.add_(0.1 * torch.randn(pc._features_rest.shape, device="cuda", generator=g))
        pc._opacity.add_(2.0)
    preds = torch.stack([pos, pos + 0.01, pos + 0.02])
    sim = ResidualMeshSimulator(preds, device="cuda")
    with torch.no_grad():
        sim.output.weight.normal_(0, 1e-3, generator=g)
    cam = S.ring_cameras(1, 160, 120, 0.69, verts.mean(0), fovy=0.55)[0]
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, torch.from_numpy(getattr(cam, k)))
    cam.time = 0.5

    class Pipe:
        compute_cov3D_python = False

    bg = torch.ones(3, device="cuda")
    res = render(cam, pc, sim, Pipe(), bg, project_vertices=True)
    assert res.render.shape == (3, 120, 160) and res.depth.shape == (1, 120, 160)
    assert res.vertice_projections.shape == (verts.shape[0], 2) and res.projections.shape == (pc.num_gaussians, 2)
    v = sim(torch.full((verts.shape[0], 1), 0.5, device="cuda"))
    assert torch.equal(v, res.vertice_deform)
    settings = GaussianRasterizationSettings(120, 160, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.world_view_transform.cuda(),
                                             cam.full_proj_transform.cuda(), 3, cam.camera_center.cuda(), False, False)
    img, radii, depth = GaussianRasterizer(settings)(
        means3D=pc.get_xyz(v), means2D=torch.zeros_like(pc.get_xyz()), opacities=pc.get_opacity, shs=pc.get_features,
        scales=pc.get_scaling, rotations=pc.get_rotation(v))
    assert torch.equal(img, res.render) and torch.equal(radii, res.radii)
    assert int((res.radii > 0).sum()) > pc.num_gaussians // 2 and torch.equal(res.visibility_filter, res.radii > 0)
    (res.render - 0.3).abs().mean().backward()
    for p in (sim.output.weight, sim.input.weight, pc.face_bary, pc._rotation, pc._scaling, pc._opacity):
        assert p.grad is not None and float(p.grad.abs().max()) > 0
    assert res.viewspace_points.grad is not None
