#!/usr/bin/env python
"""Multi-GPU check under torchrun (NCCL): every rank solves its shard of one global batch, the solutions
are gathered with `sharding.gather_solutions`, and rank 0 compares the gathered result with a single-GPU
solve of the whole batch (bit-identical: a problem's result does not depend on its position).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/gpu_multi_check.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from iterative_learning_nmpc_b200 import sharding as SH
from iterative_learning_nmpc_b200 import workloads as WL
from iterative_learning_nmpc_b200._lib import default_opts
from iterative_learning_nmpc_b200.core import NmpcSolver

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
TOTAL = 1001                                   # uneven shards on purpose
opts = dict(real_time_it=1, warm_start_qp=0)
batch = WL.make_batch(TOTAL, seed=5, gaits=("trot", "pace", "bound"))
for interleave in (False, True):
    mine = SH.shard_batch(batch, world, rank, interleave)
    b = len(mine["qv"])
    s = NmpcSolver(b, default_opts(**opts), device=local)
    _, _, hg = s.kinematics(mine["qv"])
    WL.load_into(s, mine, hg)
    st = s.solve()
    x, u, _, _ = s.get_solution(duals=False)
    xd, ud = torch.from_numpy(x).cuda(), torch.from_numpy(u).cuda()
    std = torch.from_numpy(np.asarray(st, dtype=np.int32)).cuda()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    xg, ug, stg = SH.gather_solutions(xd, ud, std, TOTAL, interleave)
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        ref = NmpcSolver(TOTAL, default_opts(**opts), device=local)
        _, _, hg = ref.kinematics(batch["qv"])
        WL.load_into(ref, batch, hg)
        st_r = ref.solve()
        xr, ur, _, _ = ref.get_solution(duals=False)
        ok = (np.array_equal(xg.cpu().numpy(), xr) and np.array_equal(ug.cpu().numpy(), ur)
              and np.array_equal(stg.cpu().numpy(), st_r))
        print(f"world={world} interleave={interleave}: gathered {TOTAL} solutions in {e0.elapsed_time(e1):.2f} ms, "
              f"bit-identical to the single-GPU solve: {ok}")
        assert ok
dist.destroy_process_group()
