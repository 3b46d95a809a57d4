"""Per-phase cycle counts of the fused QP kernel (CTA 0), debug build only (-DNMPC_TIMING).
Run under gpurun:  NMPC_LIB=.../libnmpc_b200_timing.so python tests/gpu_fused_timing.py [B]"""
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
names = ["0 update", "1 bwd: vec loads", "2 bwd: wait SR", "3 bwd: residual part", "4 fac: w,U,m,PGh", "5 fac: tile assembly",
         "6 fac: panels", "7 fac: P write+store", "8 decide", "9 fwd affine", "10 sigma", "11 corr backward (rest)", "12 corr forward", "13 cb: issue+vec loads", "14 cb: wait SR+MU", "15 cb: rmc,gamma,w", "16 cb: m", "17 cb: lu chain", "18 cb: p", "19 cb: wait P"]
out = (C.c_ulonglong * 24)()
for rep in range(2):
    WL.load_into(s, batch, hg)
    lib.nmpc_debug_phase_cycles(out, 1)
    s.solve()
    lib.nmpc_debug_phase_cycles(out, 0)
    v = np.array(list(out), dtype=float)[:20]
    tot = v.sum()
    print(f"rep {rep}: B={B} total {tot/1e6:.2f} Mcycles for one QP (CTA 0), {tot/6/1e3:.0f} kcycles per IPM iteration")
    for n, c in zip(names, v):
        print(f"   {n:24s} {c/6/26:10.0f} cycles/stage/iter  {100*c/tot:5.1f}%")
