"""`Camera` of the reference (scene_reconstruction/cameras.py:17-68): what `render()` and `train_step` read from a view --
`time`, `original_image` (clamped, times the alpha mask), `mask`, `image_width/height`, `FoVx/FoVy`, and the transposed
world-to-view / projection / full-projection matrices and the camera centre, built with the same operations (numpy float64
inverse -> fp32, projection entries in double -> fp32, torch bmm / inverse) so that they are bit-identical."""
import math

import numpy as np
import torch

from .dataset_readers import world_to_view


def get_projection_matrix(znear, zfar, fovX, fovY):
    """getProjectionMatrix (utils/graphics_utils.py:51-71): entries in Python doubles, stored into an fp32 tensor."""
    top = math.tan(fovY / 2) * znear
    right = math.tan(fovX / 2) * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = 1.0 * zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


class Camera(torch.nn.Module):
    def __init__(self, colmap_id, R, T, FoVx, FoVy, image, gt_alpha_mask, image_name, uid, trans=np.array([0.0, 0.0, 0.0]),
                 scale=1.0, data_device="cuda", time=0, view_id=None, time_id=None, flow=None, mask=None):
        super().__init__()
        self.uid, self.colmap_id, self.R, self.T, self.FoVx, self.FoVy = uid, colmap_id, R, T, FoVx, FoVy
        self.image_name, self.time, self.view_id, self.time_id, self.flow, self.mask = image_name, time, view_id, time_id, flow, mask
        self.data_device = torch.device(data_device)
        self.original_image = image.clamp(0.0, 1.0)                 # stays on the host; train_step uploads it
        self.image_width, self.image_height = self.original_image.shape[2], self.original_image.shape[1]
        if gt_alpha_mask is not None:
            self.original_image *= gt_alpha_mask
        else:
            self.original_image *= torch.ones((1, self.image_height, self.image_width))
        self.zfar, self.znear, self.trans, self.scale = 100.0, 0.01, trans, scale
        self.world_view_transform = torch.tensor(world_to_view(R, T, trans, scale)).transpose(0, 1)
        self.projection_matrix = get_projection_matrix(self.znear, self.zfar, self.FoVx, self.FoVy).transpose(0, 1)
        self.full_proj_transform = self.world_view_transform.unsqueeze(0).bmm(self.projection_matrix.unsqueeze(0)).squeeze(0)
        self.camera_center = self.world_view_transform.inverse()[3, :3]
