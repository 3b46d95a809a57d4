"""Synthetic OCP batches of BASELINE.json's configurations.

Perturbation ranges follow the reference's data collection
(`Behavior_Cloning/utils/RolloutMPC_shift_phase_percentage.py:466-484`: base height, roll/pitch/
yaw, joints, velocities; `Behavior_Cloning/examples/cfgs/bc_config.yaml:11-16`: velocity goals).
The nominal state is the Go2 home pose standing on its feet (the synthetic model's feet are on
the ground at base height ~0.265 m).
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np

from . import problem as P
from .mpc_controller.config.quadruped.utils import get_quadruped_config
from .mpc_controller.utils.contact_planner import ContactPlanner

FEET = ["FL_foot", "FR_foot", "RL_foot", "RR_foot"]
STAND_HEIGHT = 0.2648058464833031     # base height at which the home pose has its feet at z = 0
MASS = 15.017


def make_batch(B: int, seed: int = 0, gaits: Sequence[str] = ("trot",), n_nodes: int = 25,
               dt_nodes: float = 0.04, perturb: float = 1.0, robot: str = "go2") -> Dict[str, np.ndarray]:
    """B independent OCPs: perturbed x0, random gait phase, random velocity goal.

    Returns qv [B,36], W [B,76], W_e [B,52], yref [B,N,76], yref_e [B,52], p [B,N+1,44],
    i_node [B] int32, x_guess [B,N+1,42] (q,v repeated, h filled by the caller or zero),
    u_guess [B,N,30] (zero accelerations, weight shared by the stance feet), gait_id [B].
    """
    rng = np.random.default_rng(seed)
    N = n_nodes
    qv = np.tile(P.nominal_state(STAND_HEIGHT), (B, 1))
    s = perturb
    qv[:, 2] += rng.uniform(-0.05, 0.05, B) * s
    qv[:, 3] += rng.uniform(-0.25, 0.25, B) * s            # yaw
    qv[:, 4] += rng.uniform(-0.1, 0.1, B) * s              # pitch
    qv[:, 5] += rng.uniform(-0.1, 0.1, B) * s              # roll
    qv[:, 6:18] += rng.uniform(-0.3, 0.3, (B, 12)) * s
    qv[:, 18:24] += rng.uniform(-0.5, 0.5, (B, 6)) * s     # base twist
    qv[:, 24:36] += rng.uniform(-2.0, 2.0, (B, 12)) * s    # joint velocities
    v_des = np.zeros((B, 3))
    v_des[:, 0] = rng.uniform(0.0, 0.3, B)
    v_des[:, 1] = rng.uniform(-0.1, 0.1, B)
    gait_id = rng.integers(0, len(gaits), B)
    i_node = rng.integers(0, 12, B).astype(np.int32)

    W = np.zeros((B, P.NY)); We = np.zeros((B, P.NYE))
    p = np.zeros((B, N + 1, P.NP))
    yref = np.zeros((B, N, P.NY)); yref_e = np.zeros((B, P.NYE))
    u_guess = np.zeros((B, N, P.NU))
    for g, name in enumerate(gaits):
        idx = np.nonzero(gait_id == g)[0]
        if idx.size == 0:
            continue
        gait, opt, cost = get_quadruped_config(name, robot, fallback_to_trot_cost=True, fix_swapped_heights=True)
        planner = ContactPlanner(FEET, dt_nodes, gait)
        Wg, Weg = P.pack_weights(cost)
        W[idx], We[idx] = Wg, Weg
        cnt = np.stack([planner.get_contacts(int(i), N + 1) for i in i_node[idx]])
        peak = np.stack([planner.get_peaks(int(i), N + 1) for i in i_node[idx]])
        p[idx] = P.pack_params(cnt, cost, 0.0, peak)
        tracking = np.zeros((idx.size, 12))
        tracking[:, :2] = qv[idx, :2]
        tracking[:, 3] = qv[idx, 3]
        ref, ref_e = P.base_ref_vel_tracking(qv[idx, :18], v_des[idx], 0.0, tracking, gait.nom_height)
        yr, yre = P.pack_reference(ref, ref_e, P.GO2_JOINT_HOME, gait.step_height, N)
        yref[idx], yref_e[idx] = yr, yre
        act = cnt[:, :, :N].transpose(0, 2, 1).astype(float)          # [b, k, foot]
        n_act = np.maximum(act.sum(-1, keepdims=True), 1.0)
        fz = act * (MASS * 9.81) / n_act
        u_guess[idx, :, 18 + 2::3] = fz
    x_guess = np.zeros((B, N + 1, P.NX))
    x_guess[:, :, :36] = qv[:, None, :]
    return dict(qv=qv, W=W, W_e=We, yref=yref, yref_e=yref_e, p=p, i_node=i_node, x_guess=x_guess,
                u_guess=u_guess, gait_id=gait_id.astype(np.int32), v_des=v_des)


def load_into(solver, batch: Dict[str, np.ndarray], hg: Optional[np.ndarray] = None) -> None:
    """Push a batch into a solver object (product NmpcSolver or the oracle's OracleSolver)."""
    solver.set_weights(batch["W"], batch["W_e"])
    solver.set_yref(batch["yref"], batch["yref_e"])
    solver.set_params(batch["p"])
    xg = batch["x_guess"].copy()
    if hg is not None:
        xg[:, :, 36:] = hg[:, None, :]
    solver.set_guess(xg, batch["u_guess"])
    solver.set_x0(batch["qv"])
    solver.pin_stance_feet(batch["i_node"])
    solver.reset_duals()
