"""CPU tests of the multi-GPU host logic on the gloo backend (world_size 2): shard arithmetic and the AR-2 gradient
bucket.  The kernels themselves are exercised by tests/test_parity_gpu.py; here only plumbing that does not need a GPU."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from coslam_b200.dist import GradBucket, equal_shard, shard_bounds


def test_shard_bounds_cover_in_order():
    for n in (0, 1, 7, 2048, 2148, 65536 + 3):
        for world in (1, 2, 4, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for a, b in zip(spans[:-1], spans[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
            eq = [equal_shard(n, world, r) for r in range(world)]
            assert all(hi - lo == n // world for lo, hi in eq)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        params = [torch.nn.Parameter(torch.zeros(1000)), torch.nn.Parameter(torch.zeros(32, 80)), torch.nn.Parameter(torch.zeros(3))]
        params[0].grad = torch.full((1000,), float(rank + 1))
        params[1].grad = torch.arange(32 * 80, dtype=torch.float32).reshape(32, 80) * (rank + 1)
        # params[2] has no gradient on any rank -> must come back as zeros
        b = GradBucket(params)
        b.pack()
        b.all_reduce()
        b.unpack()
        tot = sum(range(1, world + 1))
        ok = bool(torch.all(params[0].grad == tot)) and \
            bool(torch.equal(params[1].grad, torch.arange(32 * 80, dtype=torch.float32).reshape(32, 80) * tot)) and \
            bool(torch.all(params[2].grad == 0))
        # AR-1 shape: int64 counts and double sums reduce exactly
        ws = torch.zeros(8, dtype=torch.float64)
        ws[:4].view(torch.int64)[:3] = torch.tensor([10 + rank, 20 + rank, 30 + rank])
        ws[4:] = torch.tensor([0.5, 1.5, 2.5, 3.5]) * (rank + 1)
        dist.all_reduce(ws[:4].view(torch.int64))
        dist.all_reduce(ws[4:8])
        ok = ok and ws[:4].view(torch.int64)[:3].tolist() == [10 * world + sum(range(world)), 20 * world + sum(range(world)),
                                                             30 * world + sum(range(world))]
        ok = ok and torch.allclose(ws[4:], torch.tensor([0.5, 1.5, 2.5, 3.5], dtype=torch.float64) * tot)
        out[rank] = ok
    finally:
        dist.destroy_process_group()


def test_grad_bucket_allreduce_gloo_world2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert all(out[r] for r in range(world))
