"""Problem sharding over ranks: partition properties and the one exchange of the
use case -- gathering packed solutions to every rank -- on the gloo backend, world_size 2 and 3,
CPU tensors.  The NCCL path is the same code on CUDA tensors (bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from iterative_learning_nmpc_b200 import sharding as S


@pytest.mark.parametrize("total,world", [(1024, 8), (10, 3), (7, 8), (0, 2), (16384, 8), (5, 1)])
@pytest.mark.parametrize("interleave", [False, True])
def test_partition_is_exact_cover(total, world, interleave):
    parts = [S.shard_indices(total, world, g, interleave) for g in range(world)]
    allidx = np.concatenate(parts) if parts else np.zeros(0, np.int64)
    assert sorted(allidx.tolist()) == list(range(total))
    sizes = [len(p) for p in parts]
    assert max(sizes) - min(sizes) <= 1
    if not interleave:
        for g in range(world):
            lo, hi = S.shard_bounds(total, world, g)
            assert parts[g].tolist() == list(range(lo, hi))


def test_bad_rank_raises():
    with pytest.raises(ValueError):
        S.shard_bounds(10, 2, 2)
    with pytest.raises(ValueError):
        S.shard_indices(10, 2, -1, True)


def test_shard_batch_rows():
    batch = dict(qv=np.arange(10 * 36.0).reshape(10, 36), i_node=np.arange(10, dtype=np.int32), tag="x")
    sub = S.shard_batch(batch, 3, 1)
    lo, hi = S.shard_bounds(10, 3, 1)
    assert np.array_equal(sub["qv"], batch["qv"][lo:hi]) and sub["tag"] == "x"
    sub = S.shard_batch(batch, 3, 2, interleave=True)
    assert np.array_equal(sub["i_node"], np.array([2, 5, 8], dtype=np.int32))


def test_pack_roundtrip():
    x = torch.randn(5, 26, 42, dtype=torch.float64)
    u = torch.randn(5, 25, 30, dtype=torch.float64)
    p = S.pack_solution(x, u)
    assert p.shape == (5, S.solution_width(25)) == (5, 1842)
    x2, u2 = S.unpack_solution(p, 25)
    assert torch.equal(x, x2) and torch.equal(u, u2)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, total: int, interleave: bool, n_nodes: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(1234)                  # same "global truth" on every rank
        X = torch.randn(total, n_nodes + 1, 42, dtype=torch.float64, generator=g)
        U = torch.randn(total, n_nodes, 30, dtype=torch.float64, generator=g)
        ST = torch.randint(0, 5, (total,), dtype=torch.int32, generator=g)
        idx = torch.from_numpy(S.shard_indices(total, world, rank, interleave))
        x, u, st = S.gather_solutions(X[idx], U[idx], ST[idx], total, interleave)
        ok = torch.equal(x, X) and torch.equal(u, U) and torch.equal(st, ST)
        act = torch.arange(total * 12, dtype=torch.float64).reshape(total, 12)
        ok = ok and torch.equal(S.gather_rows(act[idx], total, interleave), act)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,total,interleave", [(2, 9, False), (2, 8, True), (3, 10, True)])
def test_gather_solutions_gloo(world, total, interleave):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, interleave, 5, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(r, True) for r in range(world)]
