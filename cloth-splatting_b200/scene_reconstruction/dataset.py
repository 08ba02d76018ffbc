"""`MDNerfDataset` of the reference (scene_reconstruction/dataset.py:46-122): the camera infos of a scene ordered in a
view x time grid; item `idx` is view `idx` at three consecutive time steps around a random middle one (`np.random`, as in
the reference) -- the `viewpoint_cams` of one `train_step`, which is why its regularisers read views 0, 1, 2 as times
t-1, t, t+1.  With fewer than three time steps an item is all of them."""
import numpy as np
import torch
from torch.utils.data import Dataset

from .cameras import Camera


class MDNerfDataset(Dataset):
    def __init__(self, dataset, args=None):
        self.dataset, self.args = dataset, args
        self.viewpoint_ids = np.unique([d.view_id for d in dataset])
        self.time_ids = np.unique([d.time_id for d in dataset])
        self.times = np.unique([d.time for d in dataset])
        self.n_viewpoints, self.n_times = len(self.viewpoint_ids), len(self.time_ids)
        self.ordered_data = np.empty((self.n_viewpoints, self.n_times), dtype=object)
        for d in dataset:
            self.ordered_data[np.searchsorted(self.viewpoint_ids, d.view_id), np.searchsorted(self.time_ids, d.time_id)] = d

    def __len__(self):
        return self.n_viewpoints

    def __getitem__(self, idx):
        if self.n_times >= 3:
            mid = np.random.randint(1, self.n_times - 1)
            return [self.get_one_item(idx, t) for t in (mid - 1, mid, mid + 1)]
        return [self.get_one_item(idx, t) for t in range(self.n_times)]

    def get_one_item(self, view_id, time_id):
        info = self.ordered_data[view_id, time_id]
        if info is None:
            # the reference means to fall back to another view of that time step, but its selection
            # (`time_caminfos[time_caminfos is not None]`, dataset.py:101) makes np.random.choice raise: an incomplete
            # grid is an error there too
            raise ValueError(f"no camera for view {view_id} at time step {time_id}: the view x time grid must be complete")
        return Camera(colmap_id=info.view_id, R=info.R, T=info.T, FoVx=info.FovX, FoVy=info.FovY, image=info.image,
                      gt_alpha_mask=None, image_name=f"{info.view_id}", uid=info.view_id, view_id=info.view_id,
                      time_id=info.time_id, data_device=torch.device("cuda"), time=info.time,
                      flow=getattr(info, "flow", None), mask=getattr(info, "mask", None))
