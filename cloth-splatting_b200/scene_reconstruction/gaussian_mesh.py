"""`scene_reconstruction.gaussian_mesh`: barycentric placement of mesh-anchored Gaussians.

Mirrors MultiGaussianMesh.get_xyz (scene_reconstruction/gaussian_mesh.py:151-169 of the reference):
    vertice_ids = mesh.face[:, face_ids].T ; face_pos = vertices[vertice_ids]
    pos = (face_bary / face_bary.sum(1, keepdim=True)).unsqueeze(1) @ face_pos
as one CUDA kernel forward and one backward (csb_mesh_xyz_* of libcloth_splatting_b200.so); gradients reach
`face_bary` and the (deformed) vertices, i.e. the mesh the GNN / residual MLP predicted.  No CPU fallback.
Only the arithmetic is rebuilt: densification, PLY / hdf5 IO and the optimiser set-up of the reference class
are out of scope."""
import ctypes

import torch

import csb_cabi as _abi


def _bind(lib):
    if getattr(lib, "_mesh_bound", False):
        return
    c_int, c_void_p = ctypes.c_int, ctypes.c_void_p
    lib.csb_mesh_xyz_forward.restype = c_int
    lib.csb_mesh_xyz_forward.argtypes = [c_int, c_int, c_int] + [c_void_p] * 6
    lib.csb_mesh_xyz_backward.restype = c_int
    lib.csb_mesh_xyz_backward.argtypes = [c_int, c_int, c_int] + [c_void_p] * 8
    lib.csb_mesh_rotation_forward.restype = c_int
    lib.csb_mesh_rotation_forward.argtypes = [c_int, c_int, c_int] + [c_void_p] * 7
    lib.csb_mesh_rotation_backward.restype = c_int
    lib.csb_mesh_rotation_backward.argtypes = [c_int, c_int, c_int] + [c_void_p] * 9
    lib.csb_mesh_transform_workspace_bytes.restype = ctypes.c_size_t
    lib.csb_mesh_transform_workspace_bytes.argtypes = [c_int, c_int]
    lib.csb_mesh_transform_forward.restype = c_int
    lib.csb_mesh_transform_forward.argtypes = [c_int] * 4 + [c_void_p] * 10
    lib.csb_mesh_transform_backward.restype = c_int
    lib.csb_mesh_transform_backward.argtypes = [c_int] * 4 + [c_void_p] * 13 + [ctypes.c_size_t, c_void_p]
    lib._mesh_bound = True


class _MeshXYZ(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vertices, face, face_ids, face_bary):
        lib = _abi.lib()
        _bind(lib)
        if not vertices.is_cuda:
            raise ValueError("get_xyz (B200) needs CUDA tensors; there is no CPU path")
        v = vertices.detach().float().contiguous()
        b = face_bary.detach().float().contiguous()
        f = face.to(torch.int64).contiguous()
        ids = face_ids.to(torch.int64).contiguous()
        P, Nv, F = ids.numel(), v.size(0), f.size(1)
        out = torch.empty((P, 3), dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            rc = lib.csb_mesh_xyz_forward(P, Nv, F, _abi.ptr(v), _abi.ptr(f), _abi.ptr(ids), _abi.ptr(b), _abi.ptr(out),
                                          _abi.current_stream_ptr())
        _abi.check(rc, "csb_mesh_xyz_forward")
        ctx.save_for_backward(v, f, ids, b)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        lib = _abi.lib()
        v, f, ids, b = ctx.saved_tensors
        g = grad_out.float().contiguous()
        P, Nv, F = ids.numel(), v.size(0), f.size(1)
        dv = torch.empty_like(v) if ctx.needs_input_grad[0] else None
        db = torch.empty_like(b) if ctx.needs_input_grad[3] else None
        with torch.cuda.device(v.device):
            rc = lib.csb_mesh_xyz_backward(P, Nv, F, _abi.ptr(v), _abi.ptr(f), _abi.ptr(ids), _abi.ptr(b), _abi.ptr(g),
                                           _abi.ptr(dv) if dv is not None else None,
                                           _abi.ptr(db) if db is not None else None, _abi.current_stream_ptr())
        _abi.check(rc, "csb_mesh_xyz_backward")
        return dv, None, None, db


class _MeshRotation(torch.autograd.Function):
    """get_rotation(deformed_vertices) (gaussian_mesh.py:171-188): closed-form Kabsch of the face triangle,
    rotmat -> XYZW quaternion, composed with the normalised parameter; csb_mesh_rotation_*."""

    @staticmethod
    def forward(ctx, rest_vertices, vertices, face, face_ids, rotation):
        lib = _abi.lib()
        _bind(lib)
        if not vertices.is_cuda:
            raise ValueError("get_rotation (B200) needs CUDA tensors; there is no CPU path")
        v0 = rest_vertices.detach().float().contiguous()
        v = vertices.detach().float().contiguous()
        q = rotation.detach().float().contiguous()
        f = face.to(torch.int64).contiguous()
        ids = face_ids.to(torch.int64).contiguous()
        P, Nv, F = ids.numel(), v.size(0), f.size(1)
        out = torch.empty((P, 4), dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            rc = lib.csb_mesh_rotation_forward(P, Nv, F, _abi.ptr(v0), _abi.ptr(v), _abi.ptr(f), _abi.ptr(ids),
                                               _abi.ptr(q), _abi.ptr(out), _abi.current_stream_ptr())
        _abi.check(rc, "csb_mesh_rotation_forward")
        ctx.save_for_backward(v0, v, f, ids, q)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        lib = _abi.lib()
        v0, v, f, ids, q = ctx.saved_tensors
        g = grad_out.float().contiguous()
        P, Nv, F = ids.numel(), v.size(0), f.size(1)
        dv = torch.empty_like(v) if ctx.needs_input_grad[1] else None
        dq = torch.empty_like(q) if ctx.needs_input_grad[4] else None
        with torch.cuda.device(v.device):
            rc = lib.csb_mesh_rotation_backward(P, Nv, F, _abi.ptr(v0), _abi.ptr(v), _abi.ptr(f), _abi.ptr(ids),
                                                _abi.ptr(q), _abi.ptr(g), _abi.ptr(dv) if dv is not None else None,
                                                _abi.ptr(dq) if dq is not None else None, _abi.current_stream_ptr())
        _abi.check(rc, "csb_mesh_rotation_backward")
        return None, dv, None, None, dq


class _MeshTransform(torch.autograd.Function):
    """get_xyz AND get_rotation (gaussian_mesh.py:151-188) of `V` deformed meshes in one launch: csb_mesh_transform_*.
    `vertices` [V,Nv,3]; `residual` None or [V,Nv,3] (the deformed mesh is their sum, formed inside the kernel)."""

    @staticmethod
    def forward(ctx, rest_vertices, vertices, residual, face, face_ids, face_bary, rotation):
        lib = _abi.lib()
        _bind(lib)
        if not vertices.is_cuda:
            raise ValueError("the mesh transform (B200) needs CUDA tensors; there is no CPU path")
        v0 = rest_vertices.detach().float().contiguous()
        v = vertices.detach().float().contiguous()
        r = None if residual is None else residual.detach().float().contiguous()
        b = face_bary.detach().float().contiguous()
        q = rotation.detach().float().contiguous()
        f = face.to(torch.int64).contiguous()
        ids = face_ids.to(torch.int64).contiguous()
        V, Nv, P, F = v.size(0), v.size(1), ids.numel(), f.size(1)
        xyz = torch.empty((V, P, 3), dtype=torch.float32, device=v.device)
        rot = torch.empty((V, P, 4), dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            rc = lib.csb_mesh_transform_forward(V, P, Nv, F, _abi.ptr(v0), _abi.ptr(v), _abi.ptr(r) if r is not None else None,
                                                _abi.ptr(f), _abi.ptr(ids), _abi.ptr(b), _abi.ptr(q), _abi.ptr(xyz),
                                                _abi.ptr(rot), _abi.current_stream_ptr())
        _abi.check(rc, "csb_mesh_transform_forward")
        ctx.save_for_backward(v0, v, f, ids, b, q, *([] if r is None else [r]))
        return xyz, rot

    @staticmethod
    def backward(ctx, g_xyz, g_rot):
        lib = _abi.lib()
        v0, v, f, ids, b, q, *rest = ctx.saved_tensors
        r = rest[0] if rest else None
        V, Nv, P, F = v.size(0), v.size(1), ids.numel(), f.size(1)
        gx = None if g_xyz is None else g_xyz.float().contiguous()
        gr = None if g_rot is None else g_rot.float().contiguous()
        need_v = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dv = torch.empty_like(v) if need_v else None
        db = torch.empty_like(b) if ctx.needs_input_grad[5] and gx is not None else None
        dq = torch.empty_like(q) if ctx.needs_input_grad[6] and gr is not None else None
        ws = torch.empty(lib.csb_mesh_transform_workspace_bytes(V, F), dtype=torch.uint8, device=v.device)
        with torch.cuda.device(v.device):
            rc = lib.csb_mesh_transform_backward(
                V, P, Nv, F, _abi.ptr(v0), _abi.ptr(v), _abi.ptr(r) if r is not None else None, _abi.ptr(f), _abi.ptr(ids),
                _abi.ptr(b), _abi.ptr(q), _abi.ptr(gx) if gx is not None else None, _abi.ptr(gr) if gr is not None else None,
                _abi.ptr(dv) if dv is not None else None, _abi.ptr(db) if db is not None else None,
                _abi.ptr(dq) if dq is not None else None, _abi.ptr(ws), ws.numel(), _abi.current_stream_ptr())
        _abi.check(rc, "csb_mesh_transform_backward")
        if db is None and ctx.needs_input_grad[5]:
            db = torch.zeros_like(b)
        if dq is None and ctx.needs_input_grad[6]:
            dq = torch.zeros_like(q)
        return (None, dv if ctx.needs_input_grad[1] else None, dv if ctx.needs_input_grad[2] else None, None, None, db, dq)


def mesh_transform(rest_vertices, vertices, face, face_ids, face_bary, rotation, residual=None):
    """(xyz [V,P,3], rotations [V,P,4]) of the Gaussians on `V` deformed meshes `vertices (+ residual)` [V,Nv,3] -- or
    ([P,3], [P,4]) for a single mesh [Nv,3] -- in one fused launch (csb_mesh_transform_forward)."""
    single = vertices.dim() == 2
    if single:
        vertices = vertices.unsqueeze(0)
        residual = None if residual is None else residual.unsqueeze(0)
    xyz, rot = _MeshTransform.apply(rest_vertices, vertices, residual, face, face_ids, face_bary, rotation)
    return (xyz[0], rot[0]) if single else (xyz, rot)


def mesh_rotation(rest_vertices, vertices, face, face_ids, rotation):
    """[P,4] XYZW quaternions: normalize(rotation) composed with the rest -> deformed rotation of each face."""
    return _MeshRotation.apply(rest_vertices, vertices, face, face_ids, rotation)


def barycentric_xyz(vertices, face, face_ids, face_bary):
    """[P,3] positions of Gaussians anchored at barycentric coordinates `face_bary` of faces `face_ids`."""
    return _MeshXYZ.apply(vertices, face, face_ids, face_bary)


def _build_rotation(r):
    """utils/general_utils.py:81-102 (quaternion r, x, y, z -> rotation matrix)."""
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r * z)
    R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y)
    R[:, 2, 1] = 2 * (y * z + r * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def _barycentric_coordinates(points, triangles):
    """meshnet/data_utils.py:494-530, including its component order (u for A, then the AC weight, then the AB weight)."""
    A, B, C = triangles[:, 0, :], triangles[:, 1, :], triangles[:, 2, :]
    AB, AC, AP = B - A, C - A, points - A
    dot00 = torch.sum(AC * AC, dim=1)
    dot01 = torch.sum(AC * AB, dim=1)
    dot02 = torch.sum(AC * AP, dim=1)
    dot11 = torch.sum(AB * AB, dim=1)
    dot12 = torch.sum(AB * AP, dim=1)
    denom = dot00 * dot11 - dot01 * dot01
    v = (dot11 * dot02 - dot01 * dot12) / denom
    w = (dot00 * dot12 - dot01 * dot02) / denom
    u = 1.0 - v - w
    return torch.stack([u, v, w], dim=1)


class _Mesh:
    def __init__(self, pos, face):
        self.pos, self.face = pos, face


def compute_edge_features(points, edge_index):
    """Receiver - sender displacement of every edge and its length [E,1] (meshnet/data_utils.py:443-448)."""
    displacement = points[edge_index[1]] - points[edge_index[0]]
    return displacement, torch.norm(displacement, dim=1, keepdim=True)


class MultiGaussianMesh(torch.nn.Module):
    """What `render()` reads from the reference class (scene_reconstruction/gaussian_mesh.py:105-262 and the
    GaussianModel accessors, gaussian_model.py:96-118): `mesh.pos`, `mesh.face`, `face_ids`, `face_bary`,
    `_features_dc/_features_rest/_scaling/_rotation/_opacity`, `get_xyz`, `get_rotation`, `get_scaling`,
    `get_opacity`, `get_features`, `active_sh_degree`, `training_setup`, `save_ply` / `load_ply` (the reference's PLY
    columns), and the densification / pruning bookkeeping `train_step` drives (host-side tensor
    surgery further down, pinned on the reference's own methods)."""

    def __init__(self, pos, face, face_ids, face_bary, sh_degree: int = 3):
        super().__init__()
        self.mesh = _Mesh(pos, face)
        self.face_ids = face_ids
        self.face_bary = torch.nn.Parameter(face_bary.clone()) if not isinstance(face_bary, torch.nn.Parameter) else face_bary
        self.max_sh_degree = sh_degree
        self.active_sh_degree = 0
        self._features_dc = self._features_rest = self._scaling = self._rotation = self._opacity = None

    @classmethod
    def from_mesh(cls, pos, face, gaussian_init_factor=2, sh_degree=3, generator=None):
        """Initialisation of `_setup_callback` (gaussian_mesh.py:206-262): `gaussian_init_factor` Gaussians per
        face at jittered barycentres, scale = sqrt of the mean squared distance to the 3 nearest Gaussians
        (distCUDA2), identity rotation parameter (1,0,0,0), opacity 0.1, DC colour ~ U(0,1)/255."""
        from simple_knn._C import distCUDA2

        dev = pos.device
        n_faces = face.shape[1]
        P = gaussian_init_factor * n_faces
        bary = torch.full((P, 3), 1.0 / 3.0, device=dev)
        if gaussian_init_factor > 1:
            bary = torch.clip(bary + 0.05 * torch.randn((P, 3), device=dev, generator=generator), 0.0, 1.0)
            bary = bary / bary.sum(dim=1, keepdim=True)
        face_ids = torch.arange(n_faces, dtype=torch.long, device=dev).repeat(gaussian_init_factor).sort().values
        self = cls(pos, face, face_ids, bary, sh_degree)
        rgb = torch.rand((P, 3), device=dev, generator=generator) / 255.0
        feats = torch.zeros((P, 3, (sh_degree + 1) ** 2), device=dev)
        feats[:, :3, 0] = (rgb - 0.5) / 0.28209479177387814
        with torch.no_grad():
            dist2 = torch.clamp_min(distCUDA2(self.get_xyz().detach()), 1e-7)
        rots = torch.zeros((P, 4), device=dev)
        rots[:, 0] = 1
        op = torch.full((P, 1), 0.1, device=dev)
        self.set_gaussians(feats[:, :, 0:1].transpose(1, 2).contiguous(), feats[:, :, 1:].transpose(1, 2).contiguous(),
                           torch.log(torch.sqrt(dist2))[..., None].repeat(1, 3), rots, torch.log(op / (1 - op)))
        return self

    def set_edges(self, edge_index):
        """The rest-pose edge features `_setup_callback` derives from the mesh graph (gaussian_mesh.py:214-218):
        `edge_displacement`, `edge_norm` (read by `train_utils.regularization`) and `mesh.edge_attr` [E,4]."""
        self.mesh.edge_index = edge_index
        self.edge_displacement, self.edge_norm = compute_edge_features(self.mesh.pos, edge_index)
        self.mesh.edge_attr = torch.hstack((self.edge_displacement.clone().detach().to(torch.float32).contiguous(),
                                            self.edge_norm.clone().detach().to(torch.float32).contiguous()))
        return self

    def set_gaussians(self, features_dc, features_rest, scaling, rotation, opacity):
        """Raw (pre-activation) parameters: [P,1,3], [P,(D+1)^2-1,3], log-scales [P,3], quaternion [P,4], logit [P,1]."""
        P = torch.nn.Parameter
        self._features_dc, self._features_rest = P(features_dc), P(features_rest)
        self._scaling, self._rotation, self._opacity = P(scaling), P(rotation), P(opacity)

    def training_setup(self, training_args):
        """The seven Adam groups of MultiGaussianMesh.training_setup (gaussian_mesh.py:121-135), same names, learning
        rates and `torch.optim.Adam(l, lr=0.0, eps=1e-15)` settings, stepped by ONE multi-tensor launch
        (csb_optim.Adam -> csb_adam_step).  Like the reference it also zeroes the densification statistics (when
        `training_args.percent_dense` is given) and builds the position schedule (when `position_lr_final` is)."""
        from csb_optim import Adam

        if hasattr(training_args, "percent_dense"):
            dev, P = self.face_bary.device, self.num_gaussians
            self.percent_dense = training_args.percent_dense
            self.pos_gradient_accum = torch.zeros((P, 1), device=dev)
            self.denom = torch.zeros((P, 1), device=dev)
            if not hasattr(self, "max_radii2D"):
                self.max_radii2D = torch.zeros((P,), device=dev)
        if not hasattr(self, "face_offset"):
            self.face_offset = torch.nn.Parameter(torch.zeros((self.num_gaussians, 1), device=self.face_bary.device))
        scale = getattr(self, "spatial_lr_scale", 1.0)
        groups = [
            {"params": [self.face_bary], "lr": training_args.position_lr_init * scale, "name": "face_bary"},
            {"params": [self.face_offset], "lr": training_args.position_lr_init * scale, "name": "face_offset"},
            {"params": [self._features_dc], "lr": training_args.feature_lr, "name": "f_dc"},
            {"params": [self._features_rest], "lr": training_args.feature_lr / 20.0, "name": "f_rest"},
            {"params": [self._opacity], "lr": training_args.opacity_lr, "name": "opacity"},
            {"params": [self._scaling], "lr": training_args.scaling_lr, "name": "scaling"},
            {"params": [self._rotation], "lr": training_args.rotation_lr, "name": "rotation"},
        ]
        self.optimizer = Adam(groups, lr=0.0, eps=1e-15)
        if hasattr(training_args, "position_lr_final"):
            from .train_utils import get_expon_lr_func

            self.xyz_scheduler_args = get_expon_lr_func(
                lr_init=training_args.position_lr_init * scale, lr_final=training_args.position_lr_final * scale,
                lr_delay_mult=training_args.position_lr_delay_mult, max_steps=training_args.position_lr_max_steps)
        return self.optimizer

    def update_learning_rate(self, iteration):
        """gaussian_model.py:173-179, kept as it is: the schedule is applied to a group named "xyz" -- which the
        mesh-anchored model does not have (its position groups are `face_bary` / `face_offset`), so this returns None
        and the learning rates of `training_setup` stay constant, exactly like in the reference."""
        for group in self.optimizer.param_groups:
            if group["name"] == "xyz":
                group["lr"] = self.xyz_scheduler_args(iteration)
                return group["lr"]

    @property
    def num_gaussians(self):
        return self.face_ids.shape[0]

    # The three activations below do not depend on the view.  The reference evaluates them once per render() call, i.e.
    # once per view of a step; inside `shared_activations()` (train_step's view loop) each is evaluated once per step and
    # the views share the tensor -- autograd then sums the views' gradients at that one node: same values, 2 x (3 forward
    # + 3 backward) small launches fewer per 3-view iteration.
    def _activation(self, name, fn):
        cache = self.__dict__.get("_act_cache")
        if cache is None:
            return fn()
        if name not in cache:
            cache[name] = fn()
        return cache[name]

    def shared_activations(self):
        """Context manager: within it get_scaling / get_opacity / get_features are evaluated once and shared."""
        import contextlib

        @contextlib.contextmanager
        def scope():
            outer = self.__dict__.get("_act_cache")
            self.__dict__["_act_cache"] = {} if outer is None else outer
            try:
                yield self
            finally:
                self.__dict__["_act_cache"] = outer

        return scope()

    @property
    def get_scaling(self):
        return self._activation("scaling", lambda: torch.exp(self._scaling))

    @property
    def get_opacity(self):
        return self._activation("opacity", lambda: torch.sigmoid(self._opacity))

    @property
    def get_features(self):
        return self._activation("features", lambda: torch.cat((self._features_dc, self._features_rest), dim=1))

    def oneupSHdegree(self):
        if self.active_sh_degree < self.max_sh_degree:
            self.active_sh_degree += 1

    def get_xyz(self, deformed_vertices=None):
        if deformed_vertices is not None:
            assert deformed_vertices.shape[0] == self.mesh.pos.shape[0]
        verts = self.mesh.pos if deformed_vertices is None else deformed_vertices
        return barycentric_xyz(verts, self.mesh.face, self.face_ids, self.face_bary)

    def get_xyz_rotation(self, deformed_vertices, residual=None):
        """get_xyz(v) and get_rotation(v) together (gaussian_mesh.py:151-188) through the fused transform kernel: one forward
        launch and two backward launches for any number of deformed meshes ([Nv,3] or [V,Nv,3]; `residual` is added to the
        vertices inside the kernel).  Same arithmetic as the two accessors."""
        return mesh_transform(self.mesh.pos, deformed_vertices, self.mesh.face, self.face_ids, self.face_bary,
                              self._rotation, residual)

    # ---- densification / pruning bookkeeping (gaussian_model.py:266-430, gaussian_mesh.py:336-431) ----
    # Host-side tensor surgery on the parameters and on the Adam state (csb_optim.Adam keeps torch's state keys, so
    # the moments move with the rows).  Pinned on the reference's own methods: tests/make_golden_densify.py.
    _GROUP_ATTR = {"face_bary": "face_bary", "face_offset": "face_offset", "f_dc": "_features_dc", "f_rest": "_features_rest",
                   "opacity": "_opacity", "scaling": "_scaling", "rotation": "_rotation"}

    def init_densification_stats(self, percent_dense=0.01):
        """The statistics `training_setup` zeroes (gaussian_mesh.py:122-124) + `max_radii2D`."""
        dev, P = self.face_bary.device, self.num_gaussians
        self.percent_dense = percent_dense
        self.pos_gradient_accum = torch.zeros((P, 1), device=dev)
        self.denom = torch.zeros((P, 1), device=dev)
        self.max_radii2D = torch.zeros((P,), device=dev)

    def _adopt(self, tensors):
        for name, t in tensors.items():
            setattr(self, self._GROUP_ATTR[name], t)

    def replace_tensor_to_optimizer(self, tensor, name):
        out = {}
        for group in self.optimizer.param_groups:
            if group["name"] == name:
                state = self.optimizer.state.get(group["params"][0], None)
                state["exp_avg"] = torch.zeros_like(tensor)
                state["exp_avg_sq"] = torch.zeros_like(tensor)
                del self.optimizer.state[group["params"][0]]
                group["params"][0] = torch.nn.Parameter(tensor.requires_grad_(True))
                self.optimizer.state[group["params"][0]] = state
                out[name] = group["params"][0]
        return out

    def _prune_optimizer(self, mask):
        out = {}
        for group in self.optimizer.param_groups:
            if len(group["params"]) > 1:
                continue
            old = group["params"][0]
            state = self.optimizer.state.get(old, None)
            if state is not None:
                state["exp_avg"] = state["exp_avg"][mask]
                state["exp_avg_sq"] = state["exp_avg_sq"][mask]
                del self.optimizer.state[old]
            group["params"][0] = torch.nn.Parameter(old[mask].requires_grad_(True))
            if state is not None:
                self.optimizer.state[group["params"][0]] = state
            out[group["name"]] = group["params"][0]
        return out

    def cat_tensors_to_optimizer(self, tensors_dict):
        out = {}
        for group in self.optimizer.param_groups:
            assert len(group["params"]) == 1
            ext = tensors_dict[group["name"]]
            old = group["params"][0]
            state = self.optimizer.state.get(old, None)
            if state is not None:
                state["exp_avg"] = torch.cat((state["exp_avg"], torch.zeros_like(ext)), dim=0)
                state["exp_avg_sq"] = torch.cat((state["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
                del self.optimizer.state[old]
            group["params"][0] = torch.nn.Parameter(torch.cat((old, ext), dim=0).requires_grad_(True))
            if state is not None:
                self.optimizer.state[group["params"][0]] = state
            out[group["name"]] = group["params"][0]
        return out

    def prune_points(self, mask):
        keep = ~mask
        self._adopt(self._prune_optimizer(keep))
        self.face_ids = self.face_ids[keep]
        self.denom = self.denom[keep]
        self.max_radii2D = self.max_radii2D[keep]
        self.pos_gradient_accum = self.pos_gradient_accum[keep]

    def densification_postfix(self, new_face_bary, new_face_offset, new_face_ids, new_features_dc, new_features_rest,
                              new_opacities, new_scaling, new_rotation):
        self._adopt(self.cat_tensors_to_optimizer({
            "face_bary": new_face_bary, "face_offset": new_face_offset, "f_dc": new_features_dc, "f_rest": new_features_rest,
            "opacity": new_opacities, "scaling": new_scaling, "rotation": new_rotation}))
        self.face_ids = torch.cat((self.face_ids, new_face_ids), dim=0)
        dev, P = self.face_bary.device, self.face_bary.shape[0]
        self.pos_gradient_accum = torch.zeros((P, 1), device=dev)
        self.denom = torch.zeros((P, 1), device=dev)
        self.max_radii2D = torch.zeros((P,), device=dev)

    def densify_and_clone(self, grads, grad_threshold, scene_extent):
        sel = torch.where(torch.norm(grads, dim=-1) >= grad_threshold, True, False)
        sel = torch.logical_and(sel, torch.max(self.get_scaling, dim=1).values <= self.percent_dense * scene_extent)
        self.densification_postfix(self.face_bary[sel], self.face_offset[sel], self.face_ids[sel], self._features_dc[sel],
                                   self._features_rest[sel], self._opacity[sel], self._scaling[sel], self._rotation[sel])

    def densify_and_split(self, grads, grad_threshold, scene_extent, N=2):
        """gaussian_mesh.py:379-412: large Gaussians with a large view-space gradient are replaced by N samples drawn from
        them (torch.normal on the global generator), re-anchored to their face by barycentric coordinates (in the
        reference's [A, C, B] component order, meshnet/data_utils.py:494-530) and shrunk by 0.8 N."""
        n_init = self.face_bary.shape[0]
        dev = self.face_bary.device
        padded = torch.zeros((n_init,), device=dev)
        padded[:grads.shape[0]] = grads.squeeze()
        sel = torch.where(padded >= grad_threshold, True, False)
        sel = torch.logical_and(sel, torch.max(self.get_scaling, dim=1).values > self.percent_dense * scene_extent)
        stds = self.get_scaling[sel].repeat(N, 1)
        samples = torch.normal(mean=torch.zeros((stds.size(0), 3), device=dev), std=stds)
        rots = _build_rotation(self._rotation[sel]).repeat(N, 1, 1)
        new_xyz = self.get_xyz()[sel].repeat(N, 1) + torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1)
        triangles = self.mesh.pos[self.mesh.face[:, self.face_ids[sel]]].transpose(0, 1).repeat(N, 1, 1)
        new_bary = _barycentric_coordinates(new_xyz, triangles)
        self.densification_postfix(
            new_bary, self.face_offset[sel].repeat(N, 1), self.face_ids[sel].repeat(N),
            self._features_dc[sel].repeat(N, 1, 1), self._features_rest[sel].repeat(N, 1, 1), self._opacity[sel].repeat(N, 1),
            torch.log(self.get_scaling[sel].repeat(N, 1) / (0.8 * N)), self._rotation[sel].repeat(N, 1))
        self.prune_points(torch.cat((sel, torch.zeros(N * int(sel.sum()), device=dev, dtype=bool))))

    def densify(self, max_grad, min_opacity, extent, max_screen_size):
        """gaussian_model.py:420-425."""
        grads = self.pos_gradient_accum / self.denom
        grads[grads.isnan()] = 0.0
        self.densify_and_clone(grads, max_grad, extent)
        self.densify_and_split(grads, max_grad, extent)

    @torch.no_grad()
    def cleanup_barycentric_coordinates(self):
        """gaussian_mesh.py:267-322 (called from the training loop, train_utils.py:307): a negative barycentric
        coordinate means the Gaussian has left its face through the edge opposite to that vertex; it is handed to the
        face on the other side of that edge, with the reference's re-anchoring weights (distances to the new face's
        vertices, normalised) -- or, on a mesh border, the coordinate is reset the way the reference does it (0.005
        divided by itself: 1.0).  Host-side loop over the few affected Gaussians, in the reference's order."""
        face = self.mesh.face
        faces_of_vertex = [torch.nonzero(torch.eq(face, i), as_tuple=False)[:, 1].tolist() for i in range(self.mesh.pos.shape[0])]
        rows, comps = torch.where(self.face_bary < 0)
        old_ids = self.face_ids[rows]
        old_faces = face[:, old_ids].transpose(0, 1)
        xyz = self.get_xyz()
        for row, comp, tri, fid in zip(rows, comps, old_faces, old_ids):
            keep = tri[tri != tri[comp]]  # the two vertices of the edge that was crossed
            across = set(faces_of_vertex[keep[0]]).intersection(set(faces_of_vertex[keep[1]]))
            across.discard(fid.item())
            if not across:
                self.face_bary[row, comp] = 0.005
                self.face_bary[row, comp] = self.face_bary[row, comp] / self.face_bary[row, comp].sum()
            else:
                new_face = list(across)[0]
                self.face_ids[row] = new_face
                d = torch.linalg.norm(xyz[row].unsqueeze(0) - self.mesh.pos[face[:, new_face]], dim=1)
                self.face_bary[row] = d / d.sum()

    def add_densification_stats(self, viewspace_point_tensor, update_filter):
        """gaussian_model.py:427-430, written without boolean-mask indexing (no host synchronisation; adding 0 to the rows
        that are not updated leaves them bit-identical)."""
        seen = update_filter.unsqueeze(1)
        norms = torch.norm(viewspace_point_tensor[:, :2], dim=-1, keepdim=True)
        self.pos_gradient_accum += torch.where(seen, norms, torch.zeros_like(norms))
        self.denom += seen.to(self.denom.dtype)

    def reset_opacity(self):
        op = self.get_opacity
        new = torch.min(op, torch.ones_like(op) * 0.01)
        self._adopt(self.replace_tensor_to_optimizer(torch.log(new / (1 - new)), "opacity"))

    def prune(self, max_grad, min_opacity, extent, max_screen_size):
        mask = (self.get_opacity < min_opacity).squeeze()
        if max_screen_size:
            big_vs = self.max_radii2D > max_screen_size
            big_ws = self.get_scaling.max(dim=1).values > 0.1 * extent
            mask = torch.logical_or(torch.logical_or(mask, big_vs), big_ws)
        self.prune_points(mask)

    # ---- checkpoint tuple of the training CLI (train.py:106,220) ------------------------------------------------
    def capture(self):
        """gaussian_model.py:62-76, same 12 slots in the same order.  Slot 1 is `_xyz`, which for the mesh-anchored model is
        the base class's empty tensor: the reference's checkpoint does not carry `face_bary` / `face_ids` / the mesh (those
        travel in `save_ply`); kept as it is so that the tuples are interchangeable."""
        return (self.active_sh_degree, getattr(self, "_xyz", torch.empty(0)), self._features_dc, self._features_rest,
                self._scaling, self._rotation, self._opacity, self.max_radii2D, self.pos_gradient_accum, self.denom,
                self.optimizer.state_dict(), getattr(self, "spatial_lr_scale", 1.0))

    def restore(self, model_args, training_args):
        """gaussian_model.py:78-94: adopt the tensors, rebuild the optimiser, then put the statistics and its state back."""
        (self.active_sh_degree, self._xyz, self._features_dc, self._features_rest, self._scaling, self._rotation,
         self._opacity, self.max_radii2D, pos_gradient_accum, denom, opt_dict, self.spatial_lr_scale) = model_args
        self.training_setup(training_args)
        self.pos_gradient_accum = pos_gradient_accum
        self.denom = denom
        self.optimizer.load_state_dict(opt_dict)

    # ---- on-disk format of the reference --------------------------------------------------
    def save_ply(self, path):
        """`<path>/point_cloud.ply` with the reference's columns (gaussian_mesh.py:438-460): positions, zero normals,
        f_dc / f_rest channel-major, opacity, log-scales, quaternion, barycentric coordinates, face offset, face id.
        The mesh goes to `<path>/mesh.hdf5` (:462-465; every dataset load_mesh_from_h5py reads) -- through h5py when it is
        importable, else through utils/hdf5_lite.py, the same container format."""
        import os

        import numpy as np

        from . import ply_io

        n = lambda t: t.detach().cpu().numpy()  # noqa: E731
        face_offset = getattr(self, "face_offset", None)
        face_offset = n(face_offset) if face_offset is not None else np.zeros((self.num_gaussians, 1), np.float32)
        f_dc = n(self._features_dc.detach().transpose(1, 2).flatten(start_dim=1).contiguous())
        f_rest = n(self._features_rest.detach().transpose(1, 2).flatten(start_dim=1).contiguous())
        xyz = n(self.get_xyz())
        table = np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, n(self._opacity), n(self._scaling), n(self._rotation),
                                n(self.face_bary), face_offset, n(self.face_ids).astype(np.float32)[:, None]), axis=1)
        names = ply_io.gaussian_attribute_names(f_dc.shape[1], f_rest.shape[1], self._scaling.shape[1], self._rotation.shape[1])
        ply_io.write_vertex_ply(os.path.join(path, "point_cloud.ply"), names, table)
        # every dataset the reference's loader reads (meshnet/data_utils.py:450-457): pos, norm, face, edge_index
        from meshnet.mesh_data import face_to_edge, mesh_normals

        edge_index = getattr(self.mesh, "edge_index", None)
        if edge_index is None:
            edge_index = face_to_edge(self.mesh.face, self.mesh.pos.shape[0])
        mesh = {"pos": n(self.mesh.pos), "norm": n(mesh_normals(self.mesh.pos, self.mesh.face)), "face": n(self.mesh.face),
                "edge_index": n(edge_index)}
        from meshnet.mesh_data import write_datasets

        write_datasets(os.path.join(path, "mesh.hdf5"), mesh)

    @classmethod
    def load_ply(cls, path, sh_degree=3, device="cuda"):
        """The inverse (gaussian_model.py:215-262 + gaussian_mesh.py:467-480): reads `point_cloud.ply` and the mesh
        next to it; SH columns are sorted by their index and reshaped channel-major like the reference does."""
        import os

        import numpy as np

        from . import ply_io

        col = ply_io.read_vertex_ply(os.path.join(path, "point_cloud.ply"))
        from meshnet.mesh_data import read_datasets

        mesh = read_datasets(os.path.join(path, "mesh.hdf5"))
        T = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=device)  # noqa: E731

        def stack(prefix):
            names = sorted((k for k in col if k.startswith(prefix)), key=lambda k: int(k.split("_")[-1]))
            return np.stack([col[k] for k in names], axis=1)

        rest = stack("f_rest_")
        if rest.shape[1] != 3 * (sh_degree + 1) ** 2 - 3:
            raise ValueError("f_rest columns do not match sh_degree")
        self = cls(T(mesh["pos"]), T(mesh["face"], torch.long), T(col["id"], torch.long),
                   T(np.stack((col["b1"], col["b2"], col["b3"]), axis=1)), sh_degree)
        self.set_gaussians(T(stack("f_dc_").reshape(-1, 3, 1)).transpose(1, 2).contiguous(),
                           T(rest.reshape(rest.shape[0], 3, -1)).transpose(1, 2).contiguous(),
                           T(stack("scale_")), T(stack("rot_")), T(col["opacity"][:, None]))
        self.face_offset = torch.nn.Parameter(T(col["o"][:, None]))
        self.active_sh_degree = sh_degree
        # the reference's load_ply rebuilds the rest-pose edge features regularization() reads (gaussian_mesh.py:467-480)
        if "edge_index" in mesh:
            self.set_edges(T(mesh["edge_index"], torch.long))
        else:
            from meshnet.mesh_data import face_to_edge

            self.set_edges(face_to_edge(self.mesh.face, self.mesh.pos.shape[0]))
        return self

    def get_rotation(self, deformed_vertices=None):
        """normalize(_rotation), composed with each face's rest -> deformed rotation when a deformed mesh is given
        (gaussian_mesh.py:171-188)."""
        if deformed_vertices is None:
            return torch.nn.functional.normalize(self._rotation)
        return mesh_rotation(self.mesh.pos, deformed_vertices, self.mesh.face, self.face_ids, self._rotation)
