"""Data-parallel plumbing of the hot path: views (rasterizer) and graphs (GNN) are independent units, so
they are sharded round-robin over one process per GPU and the only collective is ONE flat all-reduce of
the parameter gradients per step.  The reference is single-GPU
(`utils/general_utils.py:136`); its sequential view loop is `scene_reconstruction/train_utils.py:259`.
Backend: NCCL on the GPUs, gloo in the CPU tests."""
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist


def shard_units(n_units: int, rank: int, world: int) -> List[int]:
    """Indices of the views / graphs rank `rank` owns (round-robin, every unit exactly once)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return list(range(rank, n_units, world))


class FlatGradBucket:
    """One contiguous fp32 buffer for all parameter gradients of a step -> a single all-reduce."""

    def __init__(self, shapes: Dict[str, Sequence[int]], device):
        self.names = list(shapes)
        self.shapes = {k: tuple(shapes[k]) for k in self.names}
        self.offsets, n = {}, 0
        for k in self.names:
            self.offsets[k] = n
            m = 1
            for s in self.shapes[k]:
                m *= int(s)
            n += m
        self.numel = n
        self.flat = torch.zeros(n, dtype=torch.float32, device=device)

    def view(self, name):
        o = self.offsets[name]
        n = 1
        for s in self.shapes[name]:
            n *= int(s)
        return self.flat[o:o + n].view(self.shapes[name])

    def zero_(self):
        self.flat.zero_()

    def add_(self, grads: Dict[str, torch.Tensor]):
        for k, g in grads.items():
            self.view(k).add_(g.reshape(self.shapes[k]))

    def copy_(self, grads: Dict[str, torch.Tensor]):
        for k, g in grads.items():
            self.view(k).copy_(g.reshape(self.shapes[k]))

    def all_reduce(self):
        """Sum over ranks (no-op for a single process)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        return self.flat

    def as_dict(self):
        return {k: self.view(k) for k in self.names}


def attach_flat_grads(params) -> torch.Tensor:
    """Makes every parameter's .grad a view into ONE flat fp32 buffer (returned), so that the gradients autograd
    accumulates over a multi-step rollout are summed over the ranks with a single all-reduce
    (`all_reduce_flat`).  Use `flat.zero_()` (or optimizer.zero_grad(set_to_none=False)) between steps."""
    params = [p for p in params if p.requires_grad]
    n = sum(p.numel() for p in params)
    flat = torch.zeros(n, dtype=torch.float32, device=params[0].device)
    o = 0
    for p in params:
        p.grad = flat[o:o + p.numel()].view_as(p)
        o += p.numel()
    return flat


def all_reduce_flat(flat: torch.Tensor, average: bool = False) -> torch.Tensor:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        if average:
            flat.div_(dist.get_world_size())
    return flat
