"""A CPU stand-in scene for the GLUE of `train_step` (tests/make_golden_train_step.py, tests/test_train_step_cpu.py): the
mirror's MultiGaussianMesh / ResidualMeshSimulator on the CPU, a differentiable toy `render` with the fields train_step
reads, and the image loss of oracle/loss_oracle.py.  The kernels are not involved -- their parity is tested elsewhere;
this pins the order of operations of the step (losses, schedule, optimiser calls) on the reference's own function."""
import os
import sys
import types

import numpy as np
import torch

import conftest  # noqa: F401

sys.path.insert(0, conftest.PKG)
H, W = 24, 32


class Img:
    """`cam.original_image`: the reference calls `.cuda()`, the mirror `.to(device)`."""
    def __init__(self, t):
        self.t = t

    def cuda(self):
        return self.t

    def to(self, *a, **k):
        return self.t


def options(**over):
    o = dict(lambda_dssim=0.2, lambda_deform_mag=0.1, lambda_rigid=0.5, lambda_momentum=0.05, densify_until_iter=15000,
             densify_from_iter=500, densification_interval=100, pruning_from_iter=500, pruning_interval=300,
             opacity_reset_interval=3000, bary_cleanup=700, iterations=30000, opacity_threshold_fine_init=0.5,
             opacity_threshold_fine_after=0.4, densify_grad_threshold_fine_init=1e-4, densify_grad_threshold_after=5e-5,
             position_lr_init=1.6e-4, feature_lr=2.5e-3, opacity_lr=0.05, scaling_lr=5e-3, rotation_lr=1e-3)
    o.update(over)
    return types.SimpleNamespace(**o)


def build(seed=3, masks=False):
    from meshnet.meshnet_network import ResidualMeshSimulator
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    class CpuMesh(MultiGaussianMesh):  # get_xyz of the reference in torch (gaussian_mesh.py:151-169): no GPU here
        get_xyz_rotation = None  # no fused CUDA transform on the CPU: render() then calls the two accessors
        def get_xyz(self, deformed_vertices=None):
            verts = self.mesh.pos if deformed_vertices is None else deformed_vertices
            vid = self.mesh.face[:, self.face_ids].transpose(0, 1)
            nb = self.face_bary / self.face_bary.sum(dim=1, keepdim=True)
            return (nb.unsqueeze(1) @ verts[vid, :]).squeeze(1)

    g = torch.Generator().manual_seed(seed)
    pos = torch.rand((9, 3), generator=g)
    face = torch.tensor([[0, 1, 3, 4, 1, 2, 4, 5], [1, 2, 4, 5, 4, 5, 7, 8], [3, 4, 6, 7, 3, 4, 6, 7]])
    P = 3 * face.shape[1]
    bary = torch.rand((P, 3), generator=g) + 0.05
    bary[5, 1] = -0.02                                        # one Gaussian has left its face (cleanup branch)
    m = CpuMesh(pos, face, torch.arange(face.shape[1]).repeat(3).sort().values, bary, sh_degree=3)
    m.set_gaussians(torch.randn((P, 1, 3), generator=g), torch.randn((P, 15, 3), generator=g) * 0.1,
                    torch.randn((P, 3), generator=g) * 0.5 - 3.0, torch.randn((P, 4), generator=g),
                    torch.randn((P, 1), generator=g))
    ei = torch.cat([face[[0, 1]], face[[1, 2]], face[[2, 0]]], dim=1)
    m.set_edges(torch.cat([ei, ei.flip(0)], dim=1))
    m.face_offset = torch.nn.Parameter(torch.zeros((P, 1)))
    m.init_densification_stats(0.01)
    names = ["face_bary", "face_offset", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
    m.optimizer = torch.optim.Adam([{"params": [getattr(m, m._GROUP_ATTR[n])], "lr": 1e-2, "name": n} for n in names],
                                   lr=0.0, eps=1e-15)
    sim = ResidualMeshSimulator(torch.stack([pos, pos + 0.02 * torch.randn(pos.shape, generator=g),
                                             pos + 0.04 * torch.randn(pos.shape, generator=g)]), device="cpu")
    with torch.no_grad():
        for p in sim.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (1e-3 if p.shape[0] == 27 else 0.2))
    mesh_opt = torch.optim.Adam(sim.parameters(), lr=1e-3)
    cams = []
    for i, t in enumerate((0.0, 0.5, 1.0)):
        cams.append(types.SimpleNamespace(
            time=t, idx=i, original_image=Img(torch.rand((3, H, W), generator=g)),
            mask=Img((torch.rand((1, H, W), generator=g) > 0.2).float()) if masks else None,
            pattern=torch.rand((3, H, W), generator=g), axis=torch.randn(3, generator=g)))
    return m, sim, mesh_opt, cams, torch.ones(3)


def render(cam, pc, simulator, pipe, bg, no_shadow=False, render_static=False):
    base = pc.get_xyz()
    screen = torch.zeros_like(base, requires_grad=True) + 0
    screen.retain_grad()
    if render_static:
        verts = pc.mesh.pos
        xyz = base
    else:
        verts = simulator(time_vector=torch.tensor(cam.time).repeat(pc.mesh.pos.shape[0], 1))
        xyz = pc.get_xyz(verts)
    rot = torch.nn.functional.normalize(pc._rotation)
    weight = pc.get_opacity * (1 + xyz @ cam.axis[:, None] + (screen * 0.3).sum(1, keepdim=True)
                               + 0.1 * pc.get_scaling.sum(1, keepdim=True) + 0.05 * rot[:, :1])
    colour = (weight * torch.sigmoid(pc.get_features[:, 0, :] + pc.get_features[:, 1:pc.active_sh_degree + 2, :].sum(1))).mean(0)
    image = torch.sigmoid(colour[:, None, None] * cam.pattern + 0.1 * bg[:, None, None])
    radii = ((xyz[:, 0].detach() * 37 + cam.idx * 3).abs() - 4).clamp_min(0).to(torch.int32)
    return types.SimpleNamespace(render=image, viewspace_points=screen, visibility_filter=radii > 0, radii=radii,
                                 vertice_deform=verts)


def image_losses(image, gt, opt, mask=None):
    from oracle import loss_oracle

    loss, l1, ssim_loss = loss_oracle.image_losses(image, gt, opt.lambda_dssim, mask)
    return loss, {"l1": l1.item(), "ssim_loss": ssim_loss.item()}


CASES = {  # iteration, static, white_background, masks -- each takes a different set of branches of the step
    "plain": (1, False, False, False), "masked": (7, False, False, True), "static": (3, True, False, False),
    "densify": (600, False, False, False), "prune_cleanup": (2100, False, False, False),
    "sh_up_reset": (3000, False, False, False), "white_bg_reset": (500, False, True, False),
    "past_densify": (16100, False, False, False), "past_end": (30000, False, False, False)}


def run_case(train_step, name):
    """Two iterations (the first fills the Adam state and the statistics at iteration-1, no schedule hit by construction
    of CASES) -> everything the step returns and leaves behind."""
    it, static, white, masks = CASES[name]
    m, sim, mesh_opt, cams, bg = build(masks=masks)
    opt = options()
    out = {}
    for k, i in enumerate((it - 1 if it > 1 else 1, it)):
        torch.manual_seed(11 + k)                              # densify_and_split draws from the global generator
        ps, loss, d = train_step(i, cams, m, sim, mesh_opt, types.SimpleNamespace(), opt, 5.0, bg, static=static,
                                 white_background=white, user_args=types.SimpleNamespace(no_shadow=True))
        out[f"{k}/psnr"], out[f"{k}/loss"] = ps.detach().numpy(), loss.detach().numpy()
        out[f"{k}/l1"], out[f"{k}/ssim_loss"] = d["l1"], d["ssim_loss"]
    for n, a in m._GROUP_ATTR.items():
        p = getattr(m, a)
        out[f"p/{n}"] = p.detach().numpy().copy()
        st = m.optimizer.state.get(p)
        if st:
            out[f"m1/{n}"], out[f"m2/{n}"] = st["exp_avg"].numpy().copy(), st["exp_avg_sq"].numpy().copy()
    for a in ("face_ids", "pos_gradient_accum", "denom", "max_radii2D"):
        out[a] = getattr(m, a).numpy().copy()
    out["sh"] = m.active_sh_degree
    for n, p in sim.named_parameters():                        # the large layers as float64 sums (golden file size)
        v = p.detach().double()
        out[f"sim/{n}"] = (p.detach().numpy().copy() if p.numel() <= 1024
                           else torch.stack([v.sum(), v.abs().sum(), (v * v).sum()]).numpy())
        out[f"sim_grad_none/{n}"] = p.grad is None
    return out


def camera_infos(n_views, n_times, seed=2, H=6, W=8):
    """Camera infos of a complete view x time grid, in shuffled order, with non-contiguous ids."""
    rng = np.random.default_rng(seed)
    out = []
    for v in range(n_views):
        q, _ = np.linalg.qr(rng.standard_normal((3, 3)))
        R, T = q * np.sign(np.linalg.det(q)), rng.standard_normal(3)
        for t in range(n_times):
            out.append(types.SimpleNamespace(
                R=R, T=T, FovX=0.6 + 0.05 * v, FovY=0.5 + 0.03 * v, time=t / max(n_times - 1, 1), view_id=3 * v + 1, time_id=2 * t,
                image=torch.from_numpy(rng.uniform(-0.2, 1.2, (3, H, W)).astype(np.float32)), flow=None,
                mask=torch.from_numpy((rng.uniform(size=(1, H, W)) > 0.3).astype(np.float32))))
    return [out[i] for i in rng.permutation(len(out))]


# ---- checkpoint tuple (tests/make_golden_checkpoint.py, tests/test_checkpoint_cpu.py) ----
SLOTS = ["active_sh_degree", "_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity", "max_radii2D",
         "pos_gradient_accum", "denom", "spatial_lr_scale"]


class Named:
    """every attribute holds its own name; the optimiser reports what was asked of it"""
    def __init__(self):
        self.log = []
        for a in SLOTS:
            setattr(self, a, a)
        self.optimizer = self

    def state_dict(self):
        return "optimizer.state_dict()"

    def load_state_dict(self, d):
        self.log.append(f"load_state_dict({d})")

    def training_setup(self, args):
        self.log.append(f"training_setup({args})")
        self.pos_gradient_accum, self.denom = "zeroed", "zeroed"


# ---- planner graphs (tests/make_golden_graph_data.py, tests/test_graph_data_cpu.py) ----
def graph_case(as_numpy, N=7, F=4, L=2, seed=6):
    """(action, grasped_particle, velocity, node_type, faces, velocity_target, position_target, positions) of one sample"""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(s, generator=g)  # noqa: E731
    action, velocity, vt, pt, pos = r(F, 3) * 0.1, r(N, 3 * L), r(F, N, 3), r(F, N, 3), r(N, 3)
    node_type = torch.zeros((N, 1))
    node_type[3] = 1.0
    faces = torch.randint(0, N, (3, 5), generator=g)
    c = (lambda t: t.numpy().copy()) if as_numpy else (lambda t: t.clone())
    return (c(action), 3, c(velocity), c(node_type), faces, c(vt), c(pt), c(pos))


# ---- open-loop rollout (tests/make_golden_rollout.py, tests/test_rollout_cpu.py) ----
class ToySimulator:
    """deterministic stand-in for ClothMeshSimulator.predict_velocity: sensitive to edge order, direction and features"""
    def predict_velocity(self, velocities, node_type, edge_index, edge_features):
        agg = torch.zeros((velocities.shape[0], 3)).index_add_(0, edge_index[1], edge_features[:, :3] * edge_features[:, 3:])
        w = torch.linspace(0.5, 1.5, velocities.shape[1])
        return (0.3 * velocities[:, -3:] + 0.05 * (velocities * w).sum(1, keepdim=True) + 0.2 * agg) * (1 - 0.5 * node_type)


def rollout_features(T=6, n=4, seed=12):
    import csb_synth as S

    verts, faces, ei = S.cloth_mesh(n)
    g = torch.Generator().manual_seed(seed)
    N = verts.shape[0]
    vel = torch.randn((T, N, 3), generator=g) * 0.01
    pos = torch.from_numpy(verts).float()[None] + torch.cumsum(vel, 0)
    node_type = torch.zeros((N, 1))
    node_type[5] = 1.0
    return {"pos": pos, "vel": vel, "actions": torch.randn((T, 3), generator=g) * 0.02, "node_type": node_type,
            "edge_index": torch.from_numpy(ei), "faces": torch.from_numpy(np.ascontiguousarray(faces.T)), "grasped_particle": 5}


# ---- simulator training samples (tests/make_golden_trajectory_dataset.py, tests/test_trajectory_dataset_cpu.py) ----
def trajectories(lengths=(9, 7, 12), N=6, seed=21):
    rng = np.random.default_rng(seed)
    out = []
    for k, T in enumerate(lengths):
        node_type = np.zeros((T, N, 1), np.float32)
        node_type[:, k + 1] = 1.0
        out.append({"pos": rng.standard_normal((T, N, 3)).astype(np.float32), "velocity": rng.standard_normal((T, N, 3)).astype(np.float32),
                    "node_type": node_type, "actions": rng.standard_normal((T, 3)).astype(np.float32) * 0.1,
                    "edge_faces": [torch.from_numpy(rng.integers(0, N, (3, 4))) for _ in range(T)], "grasped_particle": k + 1})
    return out


# ------------------------------------------------------------------------------------------------------------------
# stand-ins for pinning the `render()` glue (tests/make_golden_render.py, tests/test_render_glue_cpu.py)
# ------------------------------------------------------------------------------------------------------------------
class StandInRasterizer(torch.nn.Module):
    """A CPU stand-in with `GaussianRasterizer`'s call signature: a smooth, deterministic function of EVERY argument and of
    every field of the raster settings, so that a wrong argument, a swapped matrix or a dropped field of the glue changes
    the image.  It is not a rasterizer (the kernels' parity is tested on the GPU against the reference's CUDA)."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None):
        s = self.raster_settings
        H, W = s.image_height, s.image_width
        ph = torch.cat([means3D, torch.ones_like(means3D[:, :1])], 1) @ s.projmatrix
        view = torch.cat([means3D, torch.ones_like(means3D[:, :1])], 1) @ s.viewmatrix
        ndc = ph[:, :2] / (ph[:, 3:] + 1e-7)
        pix = ((ndc + 1.0) * torch.tensor([W, H], dtype=ndc.dtype) - 1.0) * 0.5 + means2D[:, :2]
        sigma = 1.5 + 40.0 * s.scale_modifier * scales.sum(1) + 0.2 * rotations[:, 0] + 0.1 * rotations[:, 1:].sum(1) * s.tanfovx / s.tanfovy
        ys, xs = torch.meshgrid(torch.arange(H, dtype=ndc.dtype), torch.arange(W, dtype=ndc.dtype), indexing="ij")
        d2 = (xs[None] - pix[:, 0, None, None]) ** 2 + (ys[None] - pix[:, 1, None, None]) ** 2
        w = opacities[:, :, None] * torch.exp(-0.5 * d2 / (sigma[:, None, None] ** 2 + 1.0))             # [P, H, W]
        if colors_precomp is not None:
            colour = colors_precomp
        else:
            n = (s.sh_degree + 1) ** 2
            direction = torch.nn.functional.normalize(means3D - s.campos[None], dim=1)
            colour = torch.sigmoid(shs[:, 0, :] + 0.3 * shs[:, 1:n, :].sum(1) * direction)
        total = w.sum(0)
        image = (w[:, None] * colour[:, :, None, None]).sum(0) / (1.0 + total) + s.bg[:, None, None] / (1.0 + total)
        depth = (w * view[:, 2, None, None]).sum(0, keepdim=True)
        radii = (torch.ceil(3.0 * sigma.detach()) * (view[:, 2].detach() > 0.2)).to(torch.int32)
        return image, radii, depth


def render_scene(seed=4, n_times=3, H=10, W=14):
    """A 3x3-vertex mesh with 24 CPU Gaussians, its residual simulator and cameras built by the mirror's Camera class."""
    from oracle import mesh_oracle
    from scene_reconstruction.cameras import Camera

    m, sim, _, _, bg = build(seed)

    def get_rotation(deformed_vertices=None):
        if deformed_vertices is None:
            return torch.nn.functional.normalize(m._rotation)
        return mesh_oracle.get_rotation(m.mesh.pos, deformed_vertices, m.mesh.face, m.face_ids, m._rotation)

    m.get_rotation = get_rotation
    m.active_sh_degree = 2
    g = torch.Generator().manual_seed(seed + 100)
    if n_times != 3:
        from meshnet.meshnet_network import ResidualMeshSimulator

        pos = m.mesh.pos
        sim = ResidualMeshSimulator(torch.stack([pos + 0.02 * k * torch.randn(pos.shape, generator=g) for k in range(n_times)]),
                                    device="cpu")
        with torch.no_grad():
            for p in sim.parameters():
                p.copy_(torch.randn(p.shape, generator=g) * (1e-3 if p.shape[0] == 27 else 0.2))
    cams = []
    for i, t in enumerate((0.0, 0.26, 0.5, 0.74, 1.0, 1.3)):
        ang = 0.4 * i
        R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
        cams.append(Camera(colmap_id=i, R=R, T=np.array([-0.4, -0.5, 2.0 + 0.1 * i]), FoVx=0.7, FoVy=0.55,
                           image=torch.rand((3, H, W), generator=g), gt_alpha_mask=None, image_name=str(i), uid=i,
                           data_device="cpu", time=t))
    return m, sim, cams, bg
