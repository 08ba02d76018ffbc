"""Camera side of the reference's scene files: `transforms_{train,test}.json` -> the per-view
matrices `GaussianRasterizationSettings` is built from.

Restates `readCamerasFromTransforms` / `read_timeline` (scene_reconstruction/dataset_readers.py:268-384, 385-400) without
the image decoding, and the matrix set-up of `Camera` (scene_reconstruction/cameras.py:57-68) over
`getWorld2View2` / `getProjectionMatrix` (utils/graphics_utils.py:38-71).  Conventions kept: `transform_matrix` is
camera-to-world in OpenGL axes (y up, z back) and is flipped to COLMAP axes (y down, z forward); `R` is stored
transposed ("due to glm"); the matrices handed to the rasterizer are the TRANSPOSES of world-to-camera and projection,
i.e. column-major when flattened; znear 0.01, zfar 100."""
import json
import math
import os
from typing import List, NamedTuple, Optional

import numpy as np

VIABLE_EXTENSIONS = (".png", ".jpg", ".jpeg")


class CameraInfo(NamedTuple):
    uid: int
    R: np.ndarray          # [3,3] transposed world-to-camera rotation
    T: np.ndarray          # [3]
    FovY: float
    FovX: float
    image_path: str
    image_name: str
    time: float
    view_id: int
    time_id: int


def read_timeline(path):
    """(timestamp -> timestamp / max, max) over the frames of both transforms files (dataset_readers.py:385-400)."""
    times = set()
    for name in ("transforms_train.json", "transforms_test.json"):
        with open(os.path.join(path, name)) as f:
            times.update(frame["time"] for frame in json.load(f)["frames"])
    line = sorted(times)
    mx = max(line)
    return {t: t / mx for t in line}, mx


def read_cameras_from_transforms(path, transformsfile, extension=".png", time_skip=None, view_skip=None) -> List[CameraInfo]:
    with open(os.path.join(path, transformsfile)) as f:
        contents = json.load(f)
    fovx, fovy = contents["camera_angle_x"], contents["camera_angle_y"]
    frames = contents["frames"]
    unique_times = np.unique([frame["time"] for frame in frames])
    unique_transforms = np.unique(np.stack([np.array(frame["transform_matrix"]) for frame in frames]), axis=0)
    kept_times = unique_times[::time_skip] if time_skip is not None else None
    out = []
    for idx, frame in enumerate(frames):
        time = frame["time"]
        if kept_times is not None and time not in kept_times:
            continue
        file_path = frame["file_path"]
        if not file_path.endswith(VIABLE_EXTENSIONS):
            file_path += extension
        file_name = file_path.split("/")[-1].split(".")[0]
        parts = file_name.split("_")
        if len(parts) > 2:  # r_<view>_<time>
            view_id, time_id = int(parts[-2]), int(parts[-1])
        else:
            m = np.array(frame["transform_matrix"])
            # the reference's expression, kept as it is (a transform counts as found when one COLUMN of it matches)
            view_id = int(np.where(np.all(unique_transforms == m, axis=1))[0][0])
            time_id = int(np.where(unique_times == time)[0][0])
        if view_skip is not None and view_id % view_skip != 0:
            continue
        c2w = np.array(frame["transform_matrix"], dtype=np.float64)
        c2w[:3, 1:3] *= -1  # OpenGL / Blender axes -> COLMAP axes
        w2c = np.linalg.inv(c2w)
        out.append(CameraInfo(uid=idx, R=np.transpose(w2c[:3, :3]), T=w2c[:3, 3], FovY=fovy, FovX=fovx,
                              image_path=os.path.join(path, file_path), image_name=file_name, time=time,
                              view_id=view_id, time_id=time_id))
    return out


def world_to_view(R, t, translate=(0.0, 0.0, 0.0), scale=1.0):
    """getWorld2View2 (utils/graphics_utils.py:38-49)."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = np.asarray(R).transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    C2W[:3, 3] = (C2W[:3, 3] + np.asarray(translate)) * scale
    return np.float32(np.linalg.inv(C2W))


def nerf_normalization(cam_infos):
    """getNerfppNorm (dataset_readers.py:58-79): `translate` = minus the mean camera centre, `radius` = 1.1 x the largest
    distance of a camera from it -- the `cameras_extent` the densification thresholds scale with (scene.py:120)."""
    centers = np.hstack([np.linalg.inv(world_to_view(c.R, c.T))[:3, 3:4] for c in cam_infos])
    center = np.mean(centers, axis=1, keepdims=True)
    diagonal = np.max(np.linalg.norm(centers - center, axis=0, keepdims=True))
    return {"translate": -center.flatten(), "radius": diagonal * 1.1}


def projection_matrix(znear, zfar, fovX, fovY):
    """getProjectionMatrix (utils/graphics_utils.py:51-71), fp32 like the torch original."""
    f = np.float32
    top = f(math.tan(fovY / 2) * znear)
    right = f(math.tan(fovX / 2) * znear)
    P = np.zeros((4, 4), dtype=np.float32)
    P[0, 0] = 2.0 * znear / (right - (-right))
    P[1, 1] = 2.0 * znear / (top - (-top))
    P[0, 2] = (right + (-right)) / (right - (-right))
    P[1, 2] = (top + (-top)) / (top - (-top))
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


class ViewMatrices(NamedTuple):
    world_view_transform: np.ndarray   # [4,4] fp32 = W2C^T
    full_proj_transform: np.ndarray    # [4,4] fp32 = W2C^T @ P^T
    camera_center: np.ndarray          # [3] fp32
    tanfovx: float
    tanfovy: float


def view_matrices(cam: CameraInfo, trans=(0.0, 0.0, 0.0), scale=1.0, znear=0.01, zfar=100.0) -> ViewMatrices:
    """What Camera.__init__ derives (cameras.py:57-68) and render() passes on (gaussian_renderer/__init__.py:60-76)."""
    wvt = world_to_view(cam.R, cam.T, trans, scale).transpose()
    proj = projection_matrix(znear, zfar, cam.FovX, cam.FovY).transpose()
    full = (wvt.astype(np.float32) @ proj.astype(np.float32)).astype(np.float32)
    center = np.linalg.inv(wvt.astype(np.float32))[3, :3].astype(np.float32)
    return ViewMatrices(np.ascontiguousarray(wvt, dtype=np.float32), np.ascontiguousarray(full), np.ascontiguousarray(center),
                        math.tan(cam.FovX * 0.5), math.tan(cam.FovY * 0.5))


def raster_settings(cam: CameraInfo, width: int, height: int, bg, sh_degree: int, device="cuda", scale_modifier=1.0,
                    trans=(0.0, 0.0, 0.0), scale=1.0):
    """GaussianRasterizationSettings of one view, as gaussian_renderer/__init__.py:60-76 builds it."""
    import torch

    from diff_gaussian_rasterization import GaussianRasterizationSettings

    m = view_matrices(cam, trans, scale)
    t = lambda a: torch.as_tensor(a, dtype=torch.float32, device=device)  # noqa: E731
    return GaussianRasterizationSettings(
        image_height=int(height), image_width=int(width), tanfovx=m.tanfovx, tanfovy=m.tanfovy, bg=t(bg),
        scale_modifier=scale_modifier, viewmatrix=t(m.world_view_transform), projmatrix=t(m.full_proj_transform),
        sh_degree=sh_degree, campos=t(m.camera_center), prefiltered=False, debug=False)
