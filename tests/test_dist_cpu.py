"""world_size-2 gloo test of the data-parallel plumbing (view sharding + one flat gradient all-reduce)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import conftest  # noqa: F401


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_view_grads(view, shapes):
    g = torch.Generator().manual_seed(100 + view)
    return {k: torch.randn(shp, generator=g) for k, shp in shapes.items()}


def _worker(rank, world, port, n_views, out):
    sys.path.insert(0, conftest.PKG)
    import csb_dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shapes = {"means3D": (37, 3), "scales": (37, 3), "rotations": (37, 4), "opacities": (37, 1), "shs": (37, 16, 3)}
    assert not csb_dist.peer_capable("cpu")  # CPU ranks never take the peer-memory path
    bucket = csb_dist.FlatGradBucket(shapes, "cpu")
    assert bucket.peer is None
    mine = csb_dist.shard_units(n_views, rank, world)
    for v in mine:
        bucket.add_(_fake_view_grads(v, shapes))
    bucket.all_reduce()
    if rank == 0:
        torch.save({"flat": bucket.flat.clone(), "mine": mine, "numel": bucket.numel}, out)
    dist.barrier()
    dist.destroy_process_group()


def test_view_sharding_and_flat_allreduce(tmp_path):
    sys.path.insert(0, conftest.PKG)
    import csb_dist

    n_views, world = 7, 2
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(world, _free_port(), n_views, out), nprocs=world, join=True)
    got = torch.load(out)
    shapes = {"means3D": (37, 3), "scales": (37, 3), "rotations": (37, 4), "opacities": (37, 1), "shs": (37, 16, 3)}
    ref = csb_dist.FlatGradBucket(shapes, "cpu")
    for v in range(n_views):
        ref.add_(_fake_view_grads(v, shapes))
    assert got["numel"] == 37 * 59 and got["mine"] == [0, 2, 4, 6]
    assert torch.allclose(got["flat"], ref.flat, atol=1e-5)


def test_shards_partition_units():
    sys.path.insert(0, conftest.PKG)
    import csb_dist

    for n, w in ((64, 8), (5, 4), (3, 8), (0, 2)):
        seen = sorted(sum((csb_dist.shard_units(n, r, w) for r in range(w)), []))
        assert seen == list(range(n))


def _normalizer_worker(rank, world, port, out):
    sys.path.insert(0, conftest.PKG)
    import csb_dist
    from meshnet.model_utils import Normalizer

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    holder = torch.nn.Module()
    holder.a, holder.b = Normalizer(5, device="cpu"), Normalizer(3, device="cpu")
    assert csb_dist.sync_normalizers(holder) == 2
    g = torch.Generator().manual_seed(50 + rank)
    outs = []
    for it in range(3):  # shards of different sizes on the two ranks
        outs.append(holder.a(torch.randn((7 + 3 * rank + it, 5), generator=g) * (1 + rank) + rank))
        holder.b(torch.randn((4 + rank, 3), generator=g))
    state = {k: v.clone() for k, v in holder.a.get_variable().items() if isinstance(v, torch.Tensor)}
    every = [None] * world
    dist.all_gather_object(every, state)
    if rank == 0:
        torch.save(every, out)
    dist.barrier()
    dist.destroy_process_group()


def test_normalizer_statistics_are_those_of_all_ranks(tmp_path):
    """Graph-sharded GNN training: after csb_dist.sync_normalizers every rank holds the same accumulators, equal to one
    process that saw the ranks' shards concatenated."""
    sys.path.insert(0, conftest.PKG)
    from meshnet.model_utils import Normalizer

    out = str(tmp_path / "n.pt")
    mp.spawn(_normalizer_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r0, r1 = torch.load(out)
    assert all(torch.equal(r0[k], r1[k]) for k in r0)
    single = Normalizer(5, device="cpu")
    gens = [torch.Generator().manual_seed(50 + r) for r in range(2)]
    for it in range(3):
        parts = []
        for r in range(2):
            parts.append(torch.randn((7 + 3 * r + it, 5), generator=gens[r]) * (1 + r) + r)
            torch.randn((4 + r, 3), generator=gens[r])
        single(torch.cat(parts))
    want = single.get_variable()
    assert float(r0["_acc_count"]) == float(want["_acc_count"]) == 3 * (7 + 10) + 2 * 3
    assert torch.allclose(r0["_acc_sum"], want["_acc_sum"], rtol=1e-5, atol=1e-5)
    assert torch.allclose(r0["_acc_sum_squared"], want["_acc_sum_squared"], rtol=1e-5, atol=1e-5)
