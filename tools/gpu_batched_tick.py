"""Host share of a BatchedLocomotionMPC replanning tick at B rollouts: wall time of `optimize`
(planning + init + solve + unpack) against the device time of the solve inside it, device planning vs host planning.
usage (under gpurun): python tools/gpu_batched_tick.py [B]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from iterative_learning_nmpc_b200 import problem as P
from iterative_learning_nmpc_b200 import workloads as WL
from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
FEET = ["FL", "FR", "RL", "RR"]
rng = np.random.default_rng(0)
q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
v0 = np.zeros((B, 18))
for on_device in (True, False):
    bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
    bm.plan_on_device = on_device
    bm.reset(reset_solver=False, i_node0=rng.integers(0, 12, B))
    bm.set_command(rng.uniform([0.0, -0.1, 0.0], [0.3, 0.1, 0.0], (B, 3)), 0.0)
    h = bm.solver.handle
    bm.first_solve = False                      # steady-state ticks: max_iter = 1
    bm.set_convergence_on_first_iter()
    walls, devs, adv = [], [], []
    q, v = q0, v0
    for tick in range(6):
        h.set_profiling(True)
        t0 = time.perf_counter()
        bm.optimize(q, v)
        h.synchronize()
        walls.append(time.perf_counter() - t0)
        devs.append(sum(ms for ms, n in h.kernel_times().values()) * 1e-3)
        h.set_profiling(False)
        t0 = time.perf_counter()
        bm._advance_reference(bm.replanning_steps)      # device planning: queued, applied by the next nmpc_plan_tick
        adv.append(time.perf_counter() - t0)
        bm.current_opt_node += 1
    w, d, a = np.median(walls[1:]), np.median(devs[1:]), np.median(adv[1:])
    print(f"B={B} planning on {'device' if on_device else 'host'}: tick wall {w*1e3:.2f} ms, solver kernels {d*1e3:.2f} ms, "
          f"host share {(w-d)/d*100:.1f} % of the solve; reference increment for {bm.replanning_steps} sim steps {a*1e3:.2f} ms")
