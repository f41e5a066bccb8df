"""world_size-2 gloo test (CPU) of the data-parallel step's host logic: the packed
(gradient, SSE, count) all-reduce + rescale reproduces the single-process RMSE step on the
concatenated batch.  The model here is a tiny stand-in with the same duck-type (loss/optimizer)
because the real forward needs a GPU; the GPU twin is tests/test_gpu_module.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hcatgnet_b200 import parallel


class _Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(0)
        self.lin = torch.nn.Linear(6, 1)
        self.loss = torch.nn.MSELoss()
        self.optimizer = torch.optim.Adam(self.parameters(), lr=0.01, eps=1e-9)

    def forward(self, batch):
        return self.lin(batch.x)


class _B:
    def __init__(self, x, y):
        self.x, self.y = x, y


def _data():
    g = torch.Generator().manual_seed(3)
    return torch.randn(10, 6, generator=g), torch.randn(10, generator=g)


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x, y = _data()
    lo, hi = (0, 3) if rank == 0 else (3, 10)  # uneven shards on purpose
    m = _Toy()
    for _ in range(3):
        rmse = parallel.dp_train_step(m, _B(x[lo:hi], y[lo:hi]))
    ret[rank] = (rmse.item(), [p.detach().clone() for p in m.parameters()])
    dist.destroy_process_group()


def test_dp_step_equals_single_process_global_batch():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    x, y = _data()
    m = _Toy()
    for _ in range(3):
        loss = parallel.dp_train_step(m, _B(x, y))
    for r in (0, 1):
        rmse, ps = ret[r]
        assert abs(rmse - loss.item()) < 1e-6
        for p, q in zip(ps, m.parameters()):
            assert torch.allclose(p, q.detach(), atol=1e-6), r


def test_shard_range_covers_everything():
    for n, w in [(10, 3), (1_000_000, 8), (5, 8), (0, 2)]:
        got = [parallel.shard_range(n, r, w) for r in range(w)]
        assert got[0][0] == 0 and got[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
