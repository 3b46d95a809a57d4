"""Per-pass cycle counts of the fused QP kernel (CTA 0), debug build only (-DNMPC_TIMING).
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
names = ["0 iteration bookkeeping", "1 R: update+residuals", "2 F: wait stage", "3 F: pass end", "4 F: w,U,m", "5 F: tile assembly",
         "6 F: panels", "7 F: P, gains K W, store", "8 A1: affine chain", "9 A2: affine bulk", "10 C1: corrector bulk",
         "11 C2: corrector chain", "12 D1: corrector fwd chain", "13 D2: corrector fwd bulk", "14 A2: issue", "15 A2: wait", "16 A2: Cz", "17 A2: rows", "18 R: wait", "19 R: step 1 update", "20 R: step 2 Cz rb", "21 R: step 3 rows", "22 R: step 4 rg", "23 A1/D1: wait", "24 A1/D1: du chain (tail: store, barrier)",
         "25 A1/D1: rhs = lu + Lxu' dx", "26 A1/D1: blocked back substitution"]
out = (C.c_ulonglong * 32)()
for rep in range(2):
    WL.load_into(s, batch, hg)
    lib.nmpc_debug_phase_cycles(out, 1)
    s.solve()
    lib.nmpc_debug_phase_cycles(out, 0)
    v = np.array(list(out), dtype=float)[:len(names)]
    tot = v.sum()
    print(f"rep {rep}: B={B} total {tot/1e6:.2f} Mcycles for one QP (CTA 0), {tot/6/1e3:.0f} kcycles per IPM iteration")
    for n, c in zip(names, v):
        print(f"   {n:28s} {c/6/26:10.0f} cycles/stage/iter  {100*c/tot:5.1f}%")
