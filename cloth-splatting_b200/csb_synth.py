"""Seeded synthetic workloads of BASELINE.json's configs.

Everything is generated on the CPU in fp32 (numpy / torch CPU) so that the CPU oracle and the GPU
see identical bits.  Matrix conventions follow the reference: world_view_transform = W2C^T,
full_proj_transform = W2C^T @ P^T, camera_center = inverse(W2C^T)[3, :3]
(scene_reconstruction/cameras.py:63-68, utils/graphics_utils.py:51-71).
"""
import math

import numpy as np


def projection_matrix(znear, zfar, fovx, fovy):
    tx, ty = math.tan(fovx / 2), math.tan(fovy / 2)
    top, right = ty * znear, tx * znear
    P = np.zeros((4, 4), dtype=np.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_camera(eye, target, up=(0.0, 0.0, 1.0)):
    """World-to-camera 4x4 (camera looks along +z, y down, COLMAP style like the reference's loaders)."""
    eye, target, up = np.asarray(eye, np.float64), np.asarray(target, np.float64), np.asarray(up, np.float64)
    f = target - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, up)
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    R = np.stack([r, d, f])  # rows = camera axes in world coordinates
    W2C = np.eye(4)
    W2C[:3, :3] = R
    W2C[:3, 3] = -R @ eye
    return W2C.astype(np.float32)


class Camera:
    def __init__(self, W2C, width, height, fovx, fovy, znear=0.01, zfar=100.0):
        self.image_width, self.image_height = int(width), int(height)
        self.FoVx, self.FoVy = float(fovx), float(fovy)
        self.tanfovx, self.tanfovy = math.tan(fovx * 0.5), math.tan(fovy * 0.5)
        self.world_view_transform = np.ascontiguousarray(W2C.T.astype(np.float32))
        proj = projection_matrix(znear, zfar, fovx, fovy).T
        self.full_proj_transform = np.ascontiguousarray((self.world_view_transform @ proj).astype(np.float32))
        self.camera_center = np.ascontiguousarray(np.linalg.inv(self.world_view_transform)[3, :3].astype(np.float32))


def ring_cameras(n, width, height, fovx, center, radius=1.2, elevation_deg=35.0, fovy=None):
    fovy = fovx if fovy is None else fovy
    cams = []
    el = math.radians(elevation_deg)
    for i in range(n):
        az = 2 * math.pi * i / n + 0.3
        eye = np.asarray(center, np.float64) + radius * np.array(
            [math.cos(el) * math.cos(az), math.cos(el) * math.sin(az), math.sin(el)])
        cams.append(Camera(look_at_camera(eye, center), width, height, fovx, fovy))
    return cams


# ---------------------------------------------------------------------------------------------
# cloth mesh (C1, C2, C5)
# ---------------------------------------------------------------------------------------------
def cloth_mesh(n=20, size=0.5):
    """n x n grid in the z=0 plane with a gentle wave; 2 triangles per cell; directed edges of
    {horizontal, vertical, one diagonal} in both directions sorted by (src, dst)."""
    xs = np.linspace(0.0, size, n, dtype=np.float64)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    Z = 0.02 * np.sin(2 * np.pi * X / size) * np.cos(2 * np.pi * Y / size)
    verts = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    idx = np.arange(n * n).reshape(n, n)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel()
    faces = np.concatenate([np.stack([a, b, d], 1), np.stack([a, d, c], 1)], 0).astype(np.int64)
    und = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]], 0)
    both = np.concatenate([und, und[:, ::-1]], 0)
    both = np.unique(both, axis=0)  # sorted by (src, dst)
    edge_index = np.ascontiguousarray(both.T.astype(np.int64))
    return verts, faces, edge_index


def gnn_inputs(n=20, seed=0):
    verts, faces, ei = cloth_mesh(n)
    rng = np.random.default_rng(seed)
    N = verts.shape[0]
    vel = (rng.standard_normal((N, 6)) * 0.01).astype(np.float32)
    node_type = np.zeros(N, dtype=np.int64)
    node_type[0] = 1
    onehot = np.eye(2, dtype=np.float32)[node_type]
    x = np.concatenate([vel, onehot], 1).astype(np.float32)
    rel = verts[ei[0]] - verts[ei[1]]
    edge_features = np.concatenate([rel, np.linalg.norm(rel, axis=1, keepdims=True)], 1).astype(np.float32)
    return x, ei, edge_features, verts, faces


# ---------------------------------------------------------------------------------------------
# Gaussian clouds
# ---------------------------------------------------------------------------------------------
def _unit_quats(rng, P):
    q = rng.standard_normal((P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.astype(np.float32)


def mesh_gaussians(n=20, per_face=14, seed=1, sh_coeffs=16, scale=0.004):
    """C2: Gaussians anchored on the faces of the n x n cloth by random barycentrics."""
    verts, faces, _ = cloth_mesh(n)
    rng = np.random.default_rng(seed)
    F = faces.shape[0]
    face_ids = np.repeat(np.arange(F), per_face)
    P = face_ids.shape[0]
    bary = np.clip(rng.normal(1 / 3, 0.05, (P, 3)), 0, 1)
    bary = (bary / bary.sum(1, keepdims=True)).astype(np.float32)
    tri = verts[faces[face_ids]]  # [P,3,3]
    means = np.einsum("pk,pkd->pd", bary, tri).astype(np.float32)
    g = dict(
        means3D=means,
        scales=np.full((P, 3), scale, dtype=np.float32) * rng.uniform(0.7, 1.3, (P, 3)).astype(np.float32),
        rotations=_unit_quats(rng, P),
        opacities=(1 / (1 + np.exp(-rng.standard_normal((P, 1))))).astype(np.float32),
        shs=(rng.standard_normal((P, sh_coeffs, 3)) * 0.3).astype(np.float32),
        face_ids=face_ids.astype(np.int64), face_bary=bary, verts=verts, faces=faces,
    )
    return g


def random_gaussians(P, seed=2, sh_coeffs=16, box=((-1, 1), (-1, 1), (2, 6)), log_scale=(0.005, 0.05)):
    """C3: random cloud in front of a camera at the origin looking along +z."""
    rng = np.random.default_rng(seed)
    lo = np.array([b[0] for b in box], np.float32)
    hi = np.array([b[1] for b in box], np.float32)
    means = (rng.random((P, 3), dtype=np.float32) * (hi - lo) + lo).astype(np.float32)
    ls = rng.uniform(math.log(log_scale[0]), math.log(log_scale[1]), (P, 3))
    return dict(
        means3D=means,
        scales=np.exp(ls).astype(np.float32),
        rotations=_unit_quats(rng, P),
        opacities=rng.uniform(0.05, 1.0, (P, 1)).astype(np.float32),
        shs=(rng.standard_normal((P, sh_coeffs, 3)) * 0.3).astype(np.float32),
    )


def front_camera(width, height, fovx):
    fovy = 2 * math.atan(math.tan(fovx / 2) * height / width)
    return Camera(np.eye(4, dtype=np.float32), width, height, fovx, fovy)


def config_c2(n_views=4, res=800, seed=1):
    g = mesh_gaussians(20, 14, seed)
    center = g["verts"].mean(0)
    cams = ring_cameras(n_views, res, res, 0.6911, center)
    return g, cams


def config_c3(P=1_000_000, width=1920, height=1080, seed=2):
    return random_gaussians(P, seed), [front_camera(width, height, 1.2)]


def config_c4(P=100_000, n_views=64, res=800, seed=3):
    g = random_gaussians(P, seed, box=((-0.5, 0.5), (-0.5, 0.5), (-0.5, 0.5)), log_scale=(0.003, 0.02))
    cams = ring_cameras(n_views, res, res, 0.6911, (0.0, 0.0, 0.0), radius=2.0)
    return g, cams


def pixel_grad(n_views, height, width, seed=5):
    rng = np.random.default_rng(seed)
    return rng.standard_normal((n_views, 3, height, width)).astype(np.float32)
