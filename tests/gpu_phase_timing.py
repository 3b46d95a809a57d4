"""Per-phase cycle counts of the Riccati kernel (CTA 0), debug build only.
Run under gpurun:  NMPC_LIB=.../libnmpc_b200_timing.so python tests/gpu_phase_timing.py [B]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from iterative_learning_nmpc_b200 import workloads as WL
from iterative_learning_nmpc_b200._lib import default_opts, load_library
from iterative_learning_nmpc_b200.core import NmpcSolver

B = int(sys.argv[1]) if len(sys.argv) > 1 else 592
lib = load_library()
batch = WL.make_batch(B, seed=1)
s = NmpcSolver(B, default_opts(real_time_it=1, warm_start_qp=0))
_, _, hg = s.kinematics(batch["qv"])
names = ["0 vec loads", "1 value_fn", "2 wait G", "3 m+masks", "4 AL compute", "5 AL write", "6 Ht+syrk", "7 pre-potrf",
         "8 panel write+bar", "9 panel fact", "10 bar2", "11 trailing", "12 store+p", "13 forward", "14 diag 6x6 (part of 9)"]
out = (C.c_ulonglong * 16)()
for rep in range(3):
    WL.load_into(s, batch, hg)
    lib.nmpc_debug_phase_cycles(out, 1)
    s.solve()
    lib.nmpc_debug_phase_cycles(out, 0)
    v = np.array(list(out), dtype=float)
    tot = v.sum()
    print(f"rep {rep}: total {tot/1e6:.2f} Mcycles over 6 launches (CTA 0)")
    for n, c in zip(names, v):
        print(f"   {n:20s} {c/6/26:10.0f} cycles/stage  {100*c/tot:5.1f}%")
