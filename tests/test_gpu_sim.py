"""Batched rigid-body rollout on the GPU (nmpc_sim_steps, nmpc_rollout.cuh) against its CPU restatement
(oracle_sim_steps), and closed-loop MPC rollouts of BatchedLocomotionMPC on it."""
import numpy as np
import pytest

from iterative_learning_nmpc_b200 import problem as P
from iterative_learning_nmpc_b200 import workloads as WL

pytestmark = pytest.mark.gpu
FEET = ["FL", "FR", "RL", "RR"]


def test_sim_steps_parity(oracle_mod):
    """40 closed-loop sim steps (controller of compute_torques_dof + forward dynamics + contact + integration) of 24
    rollouts on the plan of an SQP_RTI solve: states, applied torques and contact forces against the oracle."""
    from iterative_learning_nmpc_b200._lib import default_opts, default_sim_params
    from iterative_learning_nmpc_b200.core import NmpcSolver
    O = oracle_mod
    B = 24
    batch = WL.make_batch(B, seed=31, perturb=0.3)
    _, _, hg = O.kinematics(batch["qv"])
    so = O.OracleSolver(B, O.default_opts(real_time_it=1))
    sg = NmpcSolver(B, default_opts(real_time_it=1))
    WL.load_into(so, batch, hg)
    WL.load_into(sg, batch, hg)
    assert (so.solve(0) == sg.solve()).all()
    sp = default_sim_params()
    qv0 = batch["qv"].copy()
    feet = O.kinematics(qv0)[0]
    qv0[:, 2] -= feet[:, :, 2].min(axis=1) + 0.001              # lowest foot 1 mm into the ground
    sg.sim_set_state(qv0)
    og = sg.sim_steps(sp, 1000, 0, 40)
    oo = so.sim_steps(sp, qv0, 1000, 0, 40)
    qv_end, div = sg.sim_get_state()
    assert not div.any() and np.isfinite(qv_end).all()
    assert (np.abs(oo["fc"]) > 1.0).any()                        # the feet do touch the ground
    for k, tol in (("qv", 1e-8), ("tau", 1e-6), ("fc", 1e-5)):
        err = np.abs(og[k] - oo[k]).max() / max(1.0, np.abs(oo[k]).max())
        print(f"sim parity {k}: {err:.2e}")
        assert err < tol, k
    assert np.abs(qv_end - oo["qv_end"]).max() < 1e-8


def test_closed_loop_trot_walks_forward():
    """BatchedLocomotionMPC.closed_loop: 16 rollouts, 1 s of trot at different commands on the batched simulation.
    The robots stay upright, move along the commanded direction at the commanded (0.1-rounded, mpc.py:206) speed, and
    the vertical contact force averages the weight.  (Over longer horizons the base height of this formulation sags
    the way it does in the sim-free open-loop rollouts, bench --config 1: a property of the plans, not of the
    simulation.)"""
    from iterative_learning_nmpc_b200._lib import load_library
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
    B, T = 16, 1.0
    rng = np.random.default_rng(5)
    bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
    q0 = np.tile(P.nominal_state(bm.config_gait.nom_height)[:18], (B, 1))
    q0[:, 6:] += rng.uniform(-0.03, 0.03, (B, 12))
    feet, _, _ = bm.solver.handle.kinematics(np.concatenate((q0, np.zeros((B, 18))), axis=1))
    q0[:, 2] -= feet[:, :, 2].min(axis=1)
    vx = np.linspace(0.0, 0.3, B)
    bm.set_command(np.stack((vx, np.zeros(B), np.zeros(B)), axis=1), 0.0)
    res = bm.closed_loop(q0, np.zeros((B, 18)), T)
    assert not res["diverged"].any()
    qe = res["q_end"]
    mass = load_library().nmpc_model_mass()
    print("closed loop: x travelled", np.round(qe[:, 0] - q0[:, 0], 3), "height", np.round(qe[:, 2], 3))
    assert (np.abs(qe[:, 2] - q0[:, 2]) < 0.06).all() and (np.abs(qe[:, 4:6]) < 0.2).all()
    fz = res["fc"].reshape(-1, B, 4, 3)[400:, :, :, 2].sum(axis=2).mean(axis=0)
    assert np.allclose(fz, mass * 9.81, rtol=0.05)
    moved = qe[:, 0] - q0[:, 0]
    assert np.abs(moved - np.round(vx, 1) * T).max() < 0.06 and moved[-1] > 0.2


def test_closed_loop_rollouts_written_in_the_reference_format(tmp_path):
    """Closed-loop rollouts of the batched simulation as the reference's per-rollout `.npz` (Rollout_MPC.py:183-216):
    keys, shapes, `state` / `action` construction, real contact flags from the contact forces."""
    from iterative_learning_nmpc_b200 import dataset as DS
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
    B, T = 4, 0.3
    bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
    q0 = np.tile(P.nominal_state(bm.config_gait.nom_height)[:18], (B, 1))
    feet, _, _ = bm.solver.handle.kinematics(np.concatenate((q0, np.zeros((B, 18))), axis=1))
    q0[:, 2] -= feet[:, :, 2].min(axis=1)
    v_des = np.array([[0.0, 0, 0], [0.1, 0, 0], [0.2, 0, 0], [0.3, 0, 0]])
    bm.set_command(v_des, 0.0)
    res = bm.closed_loop(q0, np.zeros((B, 18)), T)
    recs = DS.closed_loop_to_rollouts(bm.solver.handle, res, v_des, bm.sim_dt, bm.Kp, bm.Kd)
    n = res["qv"].shape[0]
    assert len(recs) == B and n == 301
    for b, r in enumerate(recs):
        assert set(r) == set(DS.KEYS)
        assert r["q"].shape == (n, 19) and r["v"].shape == (n, 18) and r["state"].shape == (n, 44)
        assert r["action"].shape == (n, 12) and r["ctrl"].shape == (n, 12) and r["contact_vec"].shape == (n, 4)
        # the PD target reproduces the applied torque: tau = kp (action - q_j) - kd v_j
        tau = bm.Kp * (r["action"] - r["q"][:, 7:]) - bm.Kd * r["v"][:, 6:]
        np.testing.assert_allclose(tau[:, np.r_[3:6, 0:3, 9:12, 6:9]], r["ctrl"], atol=1e-9)
        assert np.allclose(np.linalg.norm(r["q"][:, 3:7], axis=1), 1.0)
        assert 1.5 < r["contact_vec"][100:].sum(axis=1).mean() < 4.0       # a trot: about two feet on the ground
        path = DS.save_rollout(r, str(tmp_path), nominal=False, replanning_point=b, nth_traj=0)
        back = np.load(path)
        assert sorted(back.files) == sorted(DS.KEYS) and np.array_equal(back["state"], r["state"])
