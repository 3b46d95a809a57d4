"""Problem-level sharding of a batch of OCPs over the GPUs of one node.

OCP b never reads OCP b': rank g owns the contiguous block [g*B/G, (g+1)*B/G) (or, with
`interleave=True`, the round-robin set {b : b % G == g}, which spreads gait types evenly when
full SQP makes the per-problem work uneven).  There is NO collective on the data path of a
solve.  The only exchange the use case has is handing the packed solutions / dataset tensors
(reference: `Behavior_Cloning/utils/Rollout_MPC.py:275,298` records state[44]/action[12] per
step on the single process that also solves) to the rank that writes the dataset: one
all_gather (NCCL over NVLink on device tensors; gloo on CPU tensors in the tests) per tick.

The packed solution record of one OCP is what `QuadrupedAcadosSolver.solve` hands back
(`mpc_controller/utils/solver.py:403-429`): x [N+1,42] | u [N,30] = 1842 fp64 at N = 25, plus
an int32 status.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np

NX, NU = 42, 30


def shard_bounds(total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block of rank `rank`: sizes differ by at most one, earlier ranks get the extra."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_indices(total: int, world: int, rank: int, interleave: bool = False) -> np.ndarray:
    """Global problem indices owned by `rank`."""
    if interleave:
        if not (0 <= rank < world):
            raise ValueError(f"rank {rank} outside world of {world}")
        return np.arange(rank, total, world, dtype=np.int64)
    lo, hi = shard_bounds(total, world, rank)
    return np.arange(lo, hi, dtype=np.int64)


def shard_batch(batch: Dict[str, np.ndarray], world: int, rank: int, interleave: bool = False
                ) -> Dict[str, np.ndarray]:
    """Rows of every per-problem array of `batch` (leading dimension = total) owned by `rank`."""
    total = len(next(iter(batch.values())))
    idx = shard_indices(total, world, rank, interleave)
    return {k: (np.ascontiguousarray(v[idx]) if isinstance(v, np.ndarray) and v.shape[:1] == (total,) else v)
            for k, v in batch.items()}


def solution_width(n_nodes: int) -> int:
    return (n_nodes + 1) * NX + n_nodes * NU


def pack_solution(x, u):
    """x [b,N+1,42], u [b,N,30] (torch tensors, same device) -> [b, 1842] contiguous."""
    import torch
    b = x.shape[0]
    return torch.cat([x.reshape(b, -1), u.reshape(b, -1)], dim=1).contiguous()


def unpack_solution(packed, n_nodes: int):
    b = packed.shape[0]
    nx = (n_nodes + 1) * NX
    return packed[:, :nx].reshape(b, n_nodes + 1, NX), packed[:, nx:].reshape(b, n_nodes, NU)


def gather_solutions(x, u, status, total: int, interleave: bool = False, group=None):
    """All ranks contribute their shard; every rank gets the full [total, ...] tensors in GLOBAL
    problem order.  x/u fp64, status int32 torch tensors on the backend's device (CUDA for nccl,
    CPU for gloo).  Uneven shards are padded to the largest one for the collective.
    Returns (x [total,N+1,42], u [total,N,30], status [total])."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_nodes = u.shape[1]
    mine = shard_indices(total, world, rank, interleave)
    if x.shape[0] != len(mine):
        raise ValueError(f"rank {rank} holds {x.shape[0]} problems, its shard has {len(mine)}")
    if world == 1:
        return x, u, status
    width = solution_width(n_nodes)
    cap = -(-total // world)
    send = torch.zeros(cap, width, dtype=torch.float64, device=x.device)
    send[:len(mine)] = pack_solution(x, u)
    send_st = torch.full((cap,), -1, dtype=torch.int32, device=x.device)
    send_st[:len(mine)] = status
    recv = torch.empty(world * cap, width, dtype=torch.float64, device=x.device)
    recv_st = torch.empty(world * cap, dtype=torch.int32, device=x.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    dist.all_gather_into_tensor(recv_st, send_st, group=group)
    out = torch.empty(total, width, dtype=torch.float64, device=x.device)
    out_st = torch.empty(total, dtype=torch.int32, device=x.device)
    for g in range(world):
        idx = torch.from_numpy(shard_indices(total, world, g, interleave)).to(x.device)
        out[idx] = recv[g * cap:g * cap + len(idx)]
        out_st[idx] = recv_st[g * cap:g * cap + len(idx)]
    xo, uo = unpack_solution(out, n_nodes)
    return xo, uo, out_st


def gather_rows(t, total: int, interleave: bool = False, group=None):
    """Same exchange for an arbitrary per-problem tensor [b, ...] (dataset tensors: state[44],
    action[12], contact vectors).  Returns [total, ...] in global order on every rank."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return t
    mine = shard_indices(total, world, rank, interleave)
    cap = -(-total // world)
    tail = tuple(t.shape[1:])
    send = torch.zeros((cap,) + tail, dtype=t.dtype, device=t.device)
    send[:len(mine)] = t
    recv = torch.empty((world * cap,) + tail, dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(recv, send.contiguous(), group=group)
    out = torch.empty((total,) + tail, dtype=t.dtype, device=t.device)
    for g in range(world):
        idx = torch.from_numpy(shard_indices(total, world, g, interleave)).to(t.device)
        out[idx] = recv[g * cap:g * cap + len(idx)]
    return out
