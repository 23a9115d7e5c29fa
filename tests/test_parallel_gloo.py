"""World-size-2 gloo tests of the data-parallel host logic; they run on CPU.
The kernels themselves need a GPU: here only the exchange steps are exercised."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from explainn_b200 import parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    parallel.init_from_env("gloo")
    try:
        out[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, out), nprocs=world, join=True)
    return [out[r] for r in range(world)]


def _grad_bucket(rank, world):
    torch.manual_seed(0)
    ps = [torch.nn.Parameter(torch.randn(3, 5)), torch.nn.Parameter(torch.randn(7)), torch.nn.Parameter(torch.randn(2, 2))]
    ps[0].grad = torch.full((3, 5), float(rank + 1))
    ps[1].grad = torch.arange(7.0) * (rank + 1)
    # ps[2] has no gradient on any rank: it must come back as zeros, not stay None
    parallel.GradBucket(ps).allreduce_mean()
    return [p.grad.numpy().copy() for p in ps]


def test_grad_bucket_allreduce_mean():
    r0, r1 = _run(_grad_bucket)
    for a, b in zip(r0, r1):
        np.testing.assert_array_equal(a, b)
    np.testing.assert_allclose(r0[0], np.full((3, 5), 1.5))
    np.testing.assert_allclose(r0[1], np.arange(7.0) * 1.5)
    np.testing.assert_array_equal(r0[2], np.zeros((2, 2)))


def _broadcast(rank, world):
    torch.manual_seed(rank)
    m = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    parallel.broadcast_parameters(m)
    return [t.detach().numpy().copy() for t in list(m.parameters()) + list(m.buffers())]


def test_broadcast_parameters_makes_replicas_identical():
    r0, r1 = _run(_broadcast)
    for a, b in zip(r0, r1):
        np.testing.assert_array_equal(a, b)


def _gather(rank, world):
    n = 11
    lo, hi = parallel.shard_range(n, rank, world)
    full = torch.arange(n * 3, dtype=torch.float32).view(n, 3)
    return parallel.gather_rows(full[lo:hi].clone(), n, rank, world).numpy()


def test_shard_range_and_gather_rows():
    for n in (0, 1, 7, 8, 11):
        for w in (1, 2, 3, 8):
            r = [parallel.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1
    for got in _run(_gather):
        np.testing.assert_array_equal(got, np.arange(33, dtype=np.float32).reshape(11, 3))


# per-sequence site counts of a fake data set: (strand, sequence, unit)
_COUNTS = np.random.default_rng(5).integers(0, 9, size=(2, 13, 4))


def _carry(rank, world):
    lo, hi = parallel.shard_range(_COUNTS.shape[1], rank, world)
    local = torch.as_tensor(_COUNTS[:, lo:hi].sum(axis=1), dtype=torch.int64)       # (S,U)
    carry0, grand = parallel.exclusive_carry(local)
    mx = parallel.allreduce_max_(torch.as_tensor(_COUNTS[:, lo:hi].max(axis=(0, 1)), dtype=torch.int64))
    return carry0.numpy(), grand.numpy(), mx.numpy(), (lo, hi)


def test_exclusive_carry_reproduces_the_single_process_site_order():
    """The rank of every shard's first site must equal the serial loop's running count
    (interpret.py:388-417: strand, then ascending sequence)."""
    res = _run(_carry)
    serial = np.zeros((2, _COUNTS.shape[1], 4), dtype=np.int64)   # running count before (strand, sequence)
    run = np.zeros(4, dtype=np.int64)
    for s in range(2):
        for i in range(_COUNTS.shape[1]):
            serial[s, i] = run
            run = run + _COUNTS[s, i]
    for carry0, grand, mx, (lo, hi) in res:
        np.testing.assert_array_equal(carry0[0], serial[0, lo])
        np.testing.assert_array_equal(carry0[1], serial[1, lo])
        np.testing.assert_array_equal(grand, run)
        np.testing.assert_array_equal(mx, _COUNTS.max(axis=(0, 1)))
    # world_size 1 path: no process group
    c0, g = parallel.exclusive_carry(torch.as_tensor(_COUNTS.sum(axis=1), dtype=torch.int64))
    np.testing.assert_array_equal(c0[0].numpy(), 0)
    np.testing.assert_array_equal(c0[1].numpy(), _COUNTS[0].sum(axis=0))
    np.testing.assert_array_equal(g.numpy(), run)
