#!/usr/bin/env python
"""Regenerates tests/golden/*.npz.

The reference holds no golden vectors for this path and cannot run here, so
these are REGRESSION vectors produced by the CPU oracle itself (oracle/nmpc_oracle.c), not
reference outputs: they pin the oracle and the CUDA path against silent drift, nothing more.
Inputs are regenerated from the seed by `workloads.make_batch`; the files store a digest of the
inputs next to the outputs so that a change of the workload generator is detected as such.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from iterative_learning_nmpc_b200 import workloads as WL  # noqa: E402
from oracle import oracle as O  # noqa: E402

CASES = {
    # name: (B, seed, gaits, solver options)
    "rti_trot": (4, 11, ("trot",), dict(real_time_it=1, max_iter=1, max_qp_iter=6, warm_start_qp=0)),
    # reference defaults (config_abstract.py:58-64): warm_start_qp = True -> HPIPM level 1, first QP of the solve cold
    "sqp_mixed": (3, 12, ("trot", "pace", "bound"), dict(real_time_it=0, max_iter=15, max_qp_iter=6, warm_start_qp=1)),
    # the same with HPIPM's hot start (level 2): the setting under which max_qp_iter = 6 converges
    "sqp_hot": (4, 14, ("trot", "pace", "bound"), dict(real_time_it=0, max_iter=15, max_qp_iter=6, warm_start_qp=2)),
    "qp_cap2": (2, 13, ("trot",), dict(real_time_it=1, max_iter=1, max_qp_iter=2, warm_start_qp=0)),
}


def digest(batch) -> str:
    h = hashlib.sha256()
    for k in sorted(batch):
        if isinstance(batch[k], np.ndarray):     # the expanded problem data; `compact` restates the same inputs
            h.update(np.ascontiguousarray(batch[k]).tobytes())
    return h.hexdigest()


def run_case(name):
    B, seed, gaits, opts = CASES[name]
    batch = WL.make_batch(B, seed=seed, gaits=gaits)
    feet, com, hg = O.kinematics(batch["qv"])
    s = O.OracleSolver(B, O.default_opts(**opts))
    WL.load_into(s, batch, hg)
    status = s.solve(1)
    x, u, pi, lam = s.get_solution()
    st = s.get_stats()
    return batch, dict(inputs_sha256=np.array(digest(batch)), feet=feet, com=com, hg=hg, status=status, x=x, u=u,
                       pi=pi, lam=lam, sqp_iter=st["sqp_iter"], qp_iter=st["qp_iter"], res=st["res"], qp_ncond=st["qp_ncond"])


if __name__ == "__main__":
    O.build()
    for name in CASES:
        _, out = run_case(name)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, {k: getattr(v, "shape", None) for k, v in out.items()}, "status", out["status"].tolist())
