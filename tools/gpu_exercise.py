"""Small end-to-end exercise of the CUDA path (a few problems, every entry-point family): RTI solve, full SQP with time
optimisation, output streaming, the device planners (cyclic and Raibert), closed-loop simulation steps, the
allocation-free kinematics.  Meant to be run under a memory checker where one is available
(`compute-sanitizer --tool memcheck python tools/gpu_exercise.py`); on this pool compute-sanitizer is closed, so it
only serves as a quick smoke of all paths:  python tools/gpu_exercise.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch

from iterative_learning_nmpc_b200 import problem as P
from iterative_learning_nmpc_b200 import workloads as WL
from iterative_learning_nmpc_b200._lib import default_opts, load_library, ptr
from iterative_learning_nmpc_b200.core import NmpcSolver

FEET = ["FL_foot", "FR_foot", "RL_foot", "RR_foot"]
B = 3
batch = WL.make_batch(B, seed=7)
for o in (dict(real_time_it=1), dict(real_time_it=0, max_iter=2, enable_time_opt=1)):
    s = NmpcSolver(B, default_opts(**o))
    _, _, hg = s.kinematics(batch["qv"])
    WL.load_into(s, batch, hg)
    ox = torch.zeros(B, 26, 42, dtype=torch.float64).pin_memory()
    ou = torch.zeros(B, 25, 30, dtype=torch.float64).pin_memory()
    s.set_output(ox, ou)
    st = s.solve()
    x, u, _, _ = s.get_solution()
    assert np.array_equal(ox.numpy(), x) and np.array_equal(ou.numpy(), u)
    print("solve", o, st.tolist())
    s.set_output()
    del s

from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
for planner in ("", "raibert"):
    bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME, contact_planner=planner)
    bm.set_command(np.array([0.2, 0.0, 0.0]), 0.1)
    q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1))
    out = bm.open_loop(q0, np.zeros((B, 18)), 0.045)
    print("open_loop", planner or "cyclic", np.isfinite(out[0]).all())
bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
bm.set_command(np.array([0.2, 0.0, 0.0]), 0.0)
res = bm.closed_loop(np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)), np.zeros((B, 18)), 0.045)
print("closed_loop", np.isfinite(res["qv"]).all())

lib = load_library()
qv = np.ascontiguousarray(np.tile(P.nominal_state(0.3), (5, 1)))
com = np.zeros((5, 3))
assert lib.nmpc_kinematics_static(ptr(qv), 5, None, ptr(com), None, None) == 0
print("static kinematics", com[0])
