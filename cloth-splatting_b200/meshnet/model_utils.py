"""`meshnet.model_utils`: the feature normalisers the simulator wraps around the GNN
(meshnet/model_utils.py:15-96 of KTH-RPL/cloth-splatting).  Same class names, constructor arguments,
attribute names (`_acc_sum`, `_acc_sum_squared`, `_acc_count`, `_num_accumulations`, ... - the reference's
checkpoints restore them with setattr, cloth_network.py:247-252) and `get_variable()` dictionaries.
A few [1, size] element-wise ops per call: plain torch on whatever device the data lives on."""
import enum

import torch
import torch.nn as nn


class NodeType(enum.IntEnum):
    CLOTH = 0


class IdentityNormalizer(nn.Module):
    """Pass-through with the Normalizer interface (model_utils.py:15-37)."""

    def __init__(self, size, name="IdentityNormalizer", device="cuda"):
        super().__init__()
        self.name = name
        self._size = size
        self._mean = torch.zeros((1, size), dtype=torch.float32, device=device)
        self._std = torch.ones((1, size), dtype=torch.float32, device=device)

    def forward(self, batched_data, accumulate=True):
        return batched_data

    def inverse(self, normalized_batch_data):
        return normalized_batch_data

    def get_variable(self):
        return {"_size": self._size, "_mean": self._mean, "_std": self._std, "name": self.name}


class Normalizer(nn.Module):
    """Running mean / std feature normaliser (model_utils.py:40-96): statistics are accumulated from every
    batch seen with accumulate=True (until max_accumulations batches), std is floored at std_epsilon."""

    def __init__(self, size, max_accumulations=10 ** 6, std_epsilon=1e-8, name="Normalizer", device="cuda"):
        super().__init__()
        self.name = name
        self._max_accumulations = max_accumulations
        f32 = dict(dtype=torch.float32, device=device)
        self._std_epsilon = torch.tensor(std_epsilon, **f32)
        self._acc_count = torch.tensor(0, **f32)
        self._num_accumulations = torch.tensor(0, **f32)
        self._acc_sum = torch.zeros((1, size), **f32)
        self._acc_sum_squared = torch.zeros((1, size), **f32)
        # Data-parallel training (graphs sharded over the ranks, csb_dist.sync_normalizers): every batch's sums are added
        # over the ranks before they are accumulated, so all replicas normalise with the statistics of the WHOLE batch
        # and a rank-0 checkpoint carries all of them.  Off by default: a single process behaves like the reference.
        self.sync_ranks = False

    def forward(self, batched_data, accumulate=True):
        if accumulate and self._num_accumulations < self._max_accumulations:
            self._accumulate(batched_data.detach())
        return (batched_data - self._mean()) / self._std_with_epsilon()

    def inverse(self, normalized_batch_data):
        return normalized_batch_data * self._std_with_epsilon() + self._mean()

    def _accumulate(self, batched_data):
        s1 = batched_data.sum(dim=0, keepdim=True)
        s2 = (batched_data ** 2).sum(dim=0, keepdim=True)
        count = batched_data.shape[0]
        if self.sync_ranks:
            import torch.distributed as dist

            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                packed = torch.cat([s1.flatten(), s2.flatten(), s1.new_tensor([float(count)])])
                dist.all_reduce(packed)  # 2 * size + 1 floats: plumbing, not a data-path collective
                n = s1.numel()
                s1, s2, count = packed[:n].view_as(s1), packed[n:2 * n].view_as(s2), packed[2 * n]
        self._acc_sum += s1
        self._acc_sum_squared += s2
        self._acc_count += count
        self._num_accumulations += 1

    def _safe_count(self):
        return torch.clamp(self._acc_count, min=1.0)

    def _mean(self):
        return self._acc_sum / self._safe_count()

    def _std_with_epsilon(self):
        var = self._acc_sum_squared / self._safe_count() - self._mean() ** 2
        return torch.maximum(torch.sqrt(var), self._std_epsilon)

    def get_variable(self):
        return {"_max_accumulations": self._max_accumulations, "_std_epsilon": self._std_epsilon,
                "_acc_count": self._acc_count, "_num_accumulations": self._num_accumulations,
                "_acc_sum": self._acc_sum, "_acc_sum_squared": self._acc_sum_squared, "name": self.name}


def get_velocity_noise(velocity_sequence, noise_std):
    """N(0, noise_std) noise of the shape of the velocity history (model_utils.py:98-105)."""
    return torch.randn_like(velocity_sequence) * noise_std
