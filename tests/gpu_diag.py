"""Stage-by-stage GPU-vs-oracle diagnostic (run under gpurun; prints max abs/rel differences).
Not a pytest file: the parity tests proper live in tests/test_gpu_*.py."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
from iterative_learning_nmpc_b200 import workloads as WL
from iterative_learning_nmpc_b200._lib import default_opts
from iterative_learning_nmpc_b200.core import NmpcSolver

np.set_printoptions(precision=4, suppress=False, linewidth=220)


def cmp(name, a, b):
    a = np.asarray(a, float); b = np.asarray(b, float)
    d = np.abs(a - b)
    scale = max(np.abs(b).max(), 1e-300)
    idx = np.unravel_index(np.argmax(d), d.shape) if d.size else ()
    print(f"  {name:10s} max|diff|={d.max():.3e}  rel={d.max()/scale:.3e}  at {idx}  ref={b[idx] if d.size else 0:.6e} got={a[idx] if d.size else 0:.6e}", flush=True)
    return d.max() / scale


B = int(os.environ.get("DIAG_B", "8"))
batch = WL.make_batch(B, seed=1)
feet_o, com_o, hg_o = O.kinematics(batch["qv"])

for mode in ("rti6",):
    oo = O.default_opts(max_iter=1, real_time_it=1, warm_start_qp=0)
    go = default_opts(max_iter=1, real_time_it=1, warm_start_qp=0)
    so = O.OracleSolver(B, oo)
    sg = NmpcSolver(B, go)
    print("kinematics:")
    feet_g, com_g, hg_g = sg.kinematics(batch["qv"])
    cmp("feet", feet_g, feet_o); cmp("com", com_g, com_o); cmp("hg", hg_g, hg_o)
    WL.load_into(so, batch, hg_o); WL.load_into(sg, batch, hg_o)
    print("export_qp:")
    qo = so.export_qp(); qg = sg.export_qp()
    for k in ("BAbt", "RSQrq", "DCt", "lg", "ug", "mask"):
        cmp(k, qg[k], qo[k])
    print("solve_qp on the oracle's QP, by IPM iteration cap:")
    dx0 = np.zeros((B, 42))
    for cap in (0, 1, 2, 3, 6, 12):
        oo2 = O.default_opts(max_qp_iter=cap, warm_start_qp=0, qp_tol=1e-2)
        so.set_opts(oo2)
        sg.set_opts(max_qp_iter=cap, warm_start_qp=0, qp_tol=1e-2)
        ro = so.solve_qp(qo, dx0)
        t0 = time.time(); rg = sg.solve_qp(qo, dx0); t1 = time.time()
        print(f" cap={cap} iters oracle={ro['qp_iter']} gpu={rg['qp_iter']} status o={ro['qp_status']} g={rg['qp_status']} ({t1-t0:.3f}s)")
        for k in ("dz", "pi", "lam", "t"):
            cmp(k, rg[k], ro[k])
    print("full RTI solve:")
    so.set_opts(oo); sg.set_opts(go)
    st_o = so.solve(1); st_g = sg.solve()
    print(" status", st_o, st_g)
    xo, uo, pio, lamo = so.get_solution(); xg, ug, pig, lamg = sg.get_solution()
    cmp("x", xg, xo); cmp("u", ug, uo); cmp("pi", pig, pio); cmp("lam", lamg, lamo)
    print(" stats oracle", so.get_stats()); print(" stats gpu", sg.get_stats())
    print("full SQP (max_iter 5):")
    oo3 = O.default_opts(max_iter=5, real_time_it=0, warm_start_qp=1)
    so.set_opts(oo3); sg.set_opts(max_iter=5, real_time_it=0, warm_start_qp=1)
    WL.load_into(so, batch, hg_o); WL.load_into(sg, batch, hg_o)
    st_o = so.solve(1); st_g = sg.solve()
    print(" status", st_o, st_g)
    xo, uo, pio, lamo = so.get_solution(); xg, ug, pig, lamg = sg.get_solution()
    cmp("x", xg, xo); cmp("u", ug, uo); cmp("pi", pig, pio); cmp("lam", lamg, lamo)
    so_s, sg_s = so.get_stats(), sg.get_stats()
    for k in so_s:
        print(" ", k, "oracle", so_s[k].ravel()[:12], "gpu", sg_s[k].ravel()[:12])
    print("launches", sg.launch_count)
print("DIAG DONE")
