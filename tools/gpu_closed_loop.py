"""Closed-loop batched rollouts on the batched rigid-body simulation: diagnostics (under gpurun).
usage: python tools/gpu_closed_loop.py [B] [T]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from iterative_learning_nmpc_b200 import problem as P
from iterative_learning_nmpc_b200._lib import default_sim_params, load_library
from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
T = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
FEET = ["FL", "FR", "RL", "RR"]
rng = np.random.default_rng(5)
bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
print("nom_height", bm.config_gait.nom_height, "Kp Kd", bm.Kp, bm.Kd, "replanning steps", bm.replanning_steps)
q0 = np.tile(P.nominal_state(bm.config_gait.nom_height)[:18], (B, 1))
q0[:, 6:] += rng.uniform(-0.03, 0.03, (B, 12))
feet, _, _ = bm.solver.handle.kinematics(np.concatenate((q0, np.zeros((B, 18))), axis=1))
q0[:, 2] -= feet[:, :, 2].min(axis=1)
vx = np.linspace(0.0, 0.3, B)
bm.set_command(np.stack((vx, np.zeros(B), np.zeros(B)), axis=1), 0.0)
t0 = time.perf_counter()
res = bm.closed_loop(q0, np.zeros((B, 18)), T)
wall = time.perf_counter() - t0
qv = res["qv"]
print(f"B={B} T={T}s wall {wall:.2f} s ({B * T / wall:.1f} sim-seconds per second), diverged {int(res['diverged'].sum())}")
for t in range(0, qv.shape[0], 250):
    print(f"t={t*1e-3:.2f} z mean {qv[t,:,2].mean():.3f} min {qv[t,:,2].min():.3f} | pitch max {np.abs(qv[t,:,4]).max():.3f} roll max {np.abs(qv[t,:,5]).max():.3f} | x {np.round(qv[t,::5,0],3)}")
st = np.array(bm.status_log)
print("solver status counts per tick (first 5):", [np.bincount(s, minlength=5).tolist() for s in st[:5]], "overall", np.bincount(st.ravel(), minlength=5).tolist())
fz = res["fc"].reshape(-1, B, 4, 3)[500:, :, :, 2].sum(axis=2).mean(axis=0)
print("mean vertical contact force / weight:", np.round(fz / (load_library().nmpc_model_mass() * 9.81), 3))
print("x travelled:", np.round(res["q_end"][:, 0] - q0[:, 0], 3), "commanded", np.round(vx * T, 3))
print("tau max", np.abs(res["tau"]).max())
