"""How a one-ulp difference in the planner outputs grows through the truncated SQP of the contact-restricted OCP:
BatchedLocomotionMPC(contact_planner="raibert") planned on the device and on the host, tick by tick.
Usage: python tools/dbg_raibert_gpu.py [n_sqp_first]"""
import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from iterative_learning_nmpc_b200 import problem as P, workloads as WL
import iterative_learning_nmpc_b200.mpc_controller.mpc as M
from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
from test_front_end import FEET
M.N_SQP_FIRST = int(sys.argv[1]) if len(sys.argv) > 1 else 15
B = 5
rng = np.random.default_rng(4)
q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
v0 = np.zeros((B, 18))
cmds = rng.uniform([0.05, -0.1, 0.0], [0.3, 0.1, 0.0], (B, 3))
bms = []
for on_device in (True, False):
    bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME, contact_planner="raibert")
    bm.plan_on_device = on_device
    bm.unpack_solution = True
    bm.reset(reset_solver=False, i_node0=np.arange(B) * 3)
    bm.set_command(cmds, 0.1)
    bms.append(bm)
q, v = q0.copy(), v0.copy()
for tick in range(3):
    sols = []
    for bm in bms:
        bm.current_opt_node = tick
        bm.set_convergence_on_first_iter()
        bm.optimize(q, v)
        bm.first_solve = False
        bm.sim_step += 40
        d = bm.solver.handle.get_problem_data()
        x, u, pi, lam = bm.solver.handle.get_solution()
        sols.append((d, x, u, bm.solver.handle.get_stats()))
    (da, xa, ua, sa), (db, xb, ub, sb) = sols
    print("n_first", M.N_SQP_FIRST, "tick", tick, "p %.1e x %.2e u %.2e" % (np.abs(da["p"] - db["p"]).max(), np.abs(xa - xb).max(), np.abs(ua - ub).max()),
          "res", np.asarray(sa["res"]).reshape(B, -1)[0], "status", np.array(bms[0].solver.status))
    q, v = xa[:, 1, :18].copy(), xa[:, 1, 18:36].copy()
