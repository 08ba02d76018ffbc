"""Data-parallel plumbing of the hot path: views (rasterizer) and graphs (GNN) are independent units, so
they are sharded round-robin over one process per GPU and the only collective is ONE flat all-reduce of
the parameter gradients per step.  The reference is single-GPU
(`utils/general_utils.py:136`); its sequential view loop is `scene_reconstruction/train_utils.py:259`.
Process-group plumbing (rendezvous, handle exchange, barriers): torch.distributed -- NCCL on the GPUs, gloo in
the CPU tests.  The gradient sum itself runs over NVLink peer memory with the library's own kernel
(`PeerBuffer` -> csb_peer_all_reduce, csrc/peer_allreduce.cu) whenever the ranks are CUDA devices of one node."""
import ctypes
import os
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


class _CudaBlob:
    """Lets torch wrap device memory the library allocated (CUDA array interface)."""

    def __init__(self, ptr, numel):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                         "strides": None}


class PeerUnavailable(RuntimeError):
    """CUDA IPC between the ranks is not possible here (raised on every rank together)."""


def _peer_or_none(numel, device):
    """A PeerBuffer, or None (with one line on stderr) where the ranks cannot map each other's memory."""
    try:
        return PeerBuffer(numel, device)
    except PeerUnavailable as ex:
        if dist.get_rank() == 0:
            import sys
            print(f"csb_dist: {ex}; the gradient sum falls back to NCCL", file=sys.stderr)
        return None


class PeerBuffer:
    """A flat fp32 device buffer every rank of the node can load from and store to (cudaMalloc + CUDA IPC), with
    the flags of csb_peer_all_reduce in front of it.  `flat` is the payload as a torch tensor; `all_reduce()`
    enqueues the library's one-kernel NVLink all-reduce on the current stream (in place, sums added in rank order:
    bit-identical on every rank and from run to run)."""

    def __init__(self, numel: int, device):
        import csb_cabi

        self.lib = csb_cabi.lib()
        self.check = csb_cabi.check
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.device = torch.device(device)
        self.numel = int(numel)
        self.padded = (self.numel + 3) // 4 * 4
        self.header = int(self.lib.csb_peer_header_bytes())
        self.seq = 0
        # Two collective stages, each followed by an agreement: a rank that cannot allocate / export / map raises
        # PeerUnavailable on EVERY rank (nobody is left waiting in a collective), and the callers fall back to NCCL.
        self.base, self.opened, err = None, [], None
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            try:
                if os.environ.get("CSB_PEER_FAIL_RANK") == str(self.rank):  # test hook for the fallback path
                    raise RuntimeError("CSB_PEER_FAIL_RANK")
                base = ctypes.c_void_p()
                self.check(self.lib.csb_peer_alloc(self.header + self.padded * 4, ctypes.byref(base)), "csb_peer_alloc")
                self.base = base.value
                self.check(self.lib.csb_peer_export(self.base, handle), "csb_peer_export")
            except Exception as ex:  # noqa: BLE001
                err = ex
            every = [None] * self.world
            dist.all_gather_object(every, (err is None, handle.raw))
            if not all(ok for ok, _ in every):
                self._abandon()
                raise PeerUnavailable(f"peer buffer: allocation / export failed on a rank ({err})")
            self.ptrs = (ctypes.c_void_p * self.world)()
            try:
                for r in range(self.world):
                    if r == self.rank:
                        self.ptrs[r] = self.base
                        continue
                    p = ctypes.c_void_p()
                    self.check(self.lib.csb_peer_import(every[r][1], ctypes.byref(p)), "csb_peer_import")
                    self.ptrs[r] = p.value
                    self.opened.append(p.value)
            except Exception as ex:  # noqa: BLE001
                err = ex
            oks = [None] * self.world
            dist.all_gather_object(oks, err is None)  # also the barrier: every rank has mapped every buffer
            if not all(oks):
                self._abandon()
                raise PeerUnavailable(f"peer buffer: a rank could not map its peers ({err})")
            self._blob = _CudaBlob(self.base + self.header, self.padded)
            self.flat = torch.as_tensor(self._blob, device=self.device)[: self.numel]
            if os.environ.get("CSB_PEER_TIMEOUT_S"):  # seconds a rank waits for its peers inside the kernel (default 600)
                self.check(self.lib.csb_peer_set_timeout(float(os.environ["CSB_PEER_TIMEOUT_S"])), "csb_peer_set_timeout")

    def _abandon(self):
        for p in self.opened:
            self.lib.csb_peer_close(p)
        self.opened = []
        if self.base:
            self.lib.csb_peer_free(self.base)
        self.base = None

    def all_reduce(self):
        import csb_cabi

        self.seq += 1
        self.check(self.lib.csb_peer_all_reduce(self.world, self.rank, self.ptrs, self.padded, self.seq,
                                                csb_cabi.current_stream_ptr()), "csb_peer_all_reduce")
        return self.flat

    def check_timeout(self):
        """Raises if a call of all_reduce() gave up waiting for a peer (the kernel then left the bucket unsummed).
        Synchronises the device; stats() calls it, training loops may call it where they synchronise anyway."""
        seq = ctypes.c_uint(0)
        with torch.cuda.device(self.device):
            self.check(self.lib.csb_peer_timed_out(self.base, ctypes.byref(seq)), "csb_peer_timed_out")
        if seq.value:
            raise RuntimeError(f"peer all-reduce: call {seq.value} on rank {self.rank} timed out waiting for a peer "
                               "(CSB_PEER_TIMEOUT_S); the gradient bucket was not summed")

    def stats(self, reset=False):
        """Mean device-side ns per call of the three phases (waiting for the peers to arrive = rank skew, reduce +
        push, fence + departure barrier), as the kernel timed itself with %globaltimer."""
        torch.cuda.synchronize(self.device)
        self.check_timeout()
        words = torch.as_tensor(_CudaBlob(self.base, self.header // 4), device=self.device)
        q = words[64:76].view(torch.int64)
        t_arrive, t_reduce, t_depart, calls, t_fence, t_count = (int(v) for v in q.cpu())
        if reset:
            q.zero_()
        n = max(calls, 1)
        return {"calls": calls, "wait_peers_us": t_arrive / n / 1e3, "reduce_push_us": t_reduce / n / 1e3,
                "depart_us": t_depart / n / 1e3, "depart_fence_us": t_fence / n / 1e3,
                "depart_count_us": t_count / n / 1e3}

    def close(self):
        """Unmaps the peers' buffers and frees the own one (call on every rank; collective)."""
        if self.base is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier()
        for p in self.opened:
            self.lib.csb_peer_close(p)
        self.opened = []
        dist.barrier()
        self.flat = None
        self.lib.csb_peer_free(self.base)
        self.base = None


def sync_normalizers(module) -> int:
    """Graph-sharded training: makes every meshnet Normalizer under `module` accumulate the statistics of the whole
    (all-rank) batch instead of its own shard's (see Normalizer.sync_ranks).  Returns how many were switched."""
    n = 0
    for m in module.modules():
        if hasattr(m, "sync_ranks") and hasattr(m, "_acc_sum"):
            m.sync_ranks = True
            n += 1
    return n


def peer_capable(device) -> bool:
    """True when the gradient sum can run over peer memory: several ranks, all CUDA devices of this node."""
    return (_world() > 1 and torch.device(device).type == "cuda" and dist.get_backend() == "nccl"
            and _world() <= min(8, torch.cuda.device_count()))


def shard_units(n_units: int, rank: int, world: int) -> List[int]:
    """Indices of the views / graphs rank `rank` owns (round-robin, every unit exactly once)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return list(range(rank, n_units, world))


class FlatGradBucket:
    """One contiguous fp32 buffer for all parameter gradients of a step -> a single all-reduce."""

    def __init__(self, shapes: Dict[str, Sequence[int]], device, peer=None):
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
        # several CUDA ranks on one node: the bucket lives in peer memory and is summed by csb_peer_all_reduce
        self.peer = None
        if peer is None and peer_capable(device):
            self.peer = _peer_or_none(n, device)
        elif peer:
            self.peer = PeerBuffer(n, device)
        self.flat = self.peer.flat if self.peer else torch.zeros(n, dtype=torch.float32, device=device)

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
        if self.peer:
            return self.peer.all_reduce()
        if _world() > 1:  # gloo (CPU tests) or ranks that cannot map each other's memory
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        return self.flat

    def as_dict(self):
        return {k: self.view(k) for k in self.names}


_PEER_OF = {}  # data_ptr of a flat gradient buffer -> the PeerBuffer that owns it


def attach_flat_grads(params) -> torch.Tensor:
    """Makes every parameter's .grad a view into ONE flat fp32 buffer (returned), so that the gradients autograd
    accumulates over a multi-step rollout are summed over the ranks with a single all-reduce
    (`all_reduce_flat`).  Use `flat.zero_()` (or optimizer.zero_grad(set_to_none=False)) between steps."""
    params = [p for p in params if p.requires_grad]
    n = sum(p.numel() for p in params)
    pb = _peer_or_none(n, params[0].device) if peer_capable(params[0].device) else None
    if pb is not None:
        flat = pb.flat
        _PEER_OF[flat.data_ptr()] = pb
    else:
        flat = torch.zeros(n, dtype=torch.float32, device=params[0].device)
    o = 0
    for p in params:
        p.grad = flat[o:o + p.numel()].view_as(p)
        o += p.numel()
    return flat


def all_reduce_flat(flat: torch.Tensor, average: bool = False) -> torch.Tensor:
    if _world() > 1:
        pb = _PEER_OF.get(flat.data_ptr())
        if pb is not None:
            pb.all_reduce()
        else:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        if average:
            flat.div_(dist.get_world_size())
    return flat
