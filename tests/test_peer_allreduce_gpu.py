"""csb_peer_all_reduce (csrc/peer_allreduce.cu): the flat gradient bucket summed over the ranks through CUDA-IPC peer
memory.  Two processes -- on two GPUs where the box has them, else sharing cuda:0 (IPC and the flag protocol are the
same; the kernels then time-slice) -- against the fixed-order sum of the gathered inputs, bit for bit."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import conftest  # noqa: F401

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, conftest.PKG)
    import csb_dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    ndev = torch.cuda.device_count()
    dev = torch.device("cuda", rank % ndev)
    torch.cuda.set_device(dev)
    # gloo carries the handle exchange and the barriers; the sum itself never touches it
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ok = True
    for P in (37, 10108):  # 37*59 floats is not a multiple of 4: the padded tail must stay out of the payload
        shapes = {"means3D": (P, 3), "scales": (P, 3), "rotations": (P, 4), "opacities": (P, 1), "shs": (P, 16, 3)}
        bucket = csb_dist.FlatGradBucket(shapes, dev, peer=True)
        assert bucket.peer is not None and bucket.flat.numel() == P * 59
        for it in range(4):  # several calls on the same buffer: the flag sequence must keep working
            g = torch.Generator().manual_seed(1000 * it + rank)
            mine = torch.randn(bucket.numel, generator=g)
            bucket.flat.copy_(mine.to(dev))
            bucket.all_reduce()
            got = bucket.flat.cpu()
            parts = [None] * world
            dist.all_gather_object(parts, mine)
            want = parts[0].clone()
            for r in range(1, world):
                want += parts[r]
            ok &= bool(torch.equal(got, want))
            # views into the bucket see the reduced values
            ok &= bool(torch.equal(bucket.view("shs").cpu().reshape(-1), want[bucket.offsets["shs"]:]))
        bucket.peer.close()
    res = [None] * world
    dist.all_gather_object(res, ok)
    if rank == 0:
        torch.save(res, out)
    dist.barrier()
    dist.destroy_process_group()


def test_peer_all_reduce_two_ranks(tmp_path):
    out = str(tmp_path / "ok.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert torch.load(out) == [True, True]


def _timeout_worker(rank, world, port, out):
    sys.path.insert(0, conftest.PKG)
    import csb_dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["CSB_PEER_TIMEOUT_S"] = "0.3"
    dev = torch.device("cuda", rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    buf = csb_dist.PeerBuffer(4096, dev)
    ok = True
    if rank == 0:  # rank 1 never enters the call: the kernel gives up after 0.3 s instead of trapping
        buf.flat.fill_(1.0)
        buf.all_reduce()
        try:
            buf.check_timeout()
            ok = False
        except RuntimeError as ex:
            ok &= "timed out" in str(ex)
        ok &= float(buf.flat.sum()) == 4096.0                      # the bucket was left alone
        ok &= float((torch.ones(8, device=dev) * 2).sum()) == 16.0  # and the context is alive
    res = [None] * world
    dist.all_gather_object(res, ok)
    if rank == 0:
        torch.save(res, out)
    dist.barrier()
    buf.close()
    dist.destroy_process_group()


def test_a_missing_peer_is_a_timeout_error_not_a_trap(tmp_path):
    out = str(tmp_path / "ok.pt")
    mp.spawn(_timeout_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert torch.load(out) == [True, True]
