"""Batched host-side packing of OCP data (weights, references, per-node contact parameters).

Vectorised restatement, with a leading batch dimension, of what `QuadrupedAcadosSolver.init`
does one problem at a time in the reference (`mpc_controller/utils/solver.py`):
  set_cost_weights :108-141, setup_reference :153-177, init_contacts_parameters :212-225,
  setup_cnt_status :227-252, setup_contact_loc :254-275,
and of `LocomotionMPC.compute_base_ref_vel_tracking` (`mpc_controller/mpc.py:210-272`).
Layouts are those of include/nmpc_b200.h.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .mpc_controller.config.config_abstract import MPCCostConfig
from .mpc_controller.utils.transform import rpy_to_matrix

NX, NU, NZ, NP, NY, NYE, NG = 42, 30, 72, 44, 76, 52, 42
NQ, NFEET, NP_FOOT = 18, 4, 11
# y layout (include/nmpc_b200.h)
Y_BASE, Y_JOINT, Y_ACC, Y_SWING, Y_FREG, Y_POS = 0, 12, 36, 48, 52, 64
YE_BASE, YE_JOINT, YE_SWING, YE_POS = 0, 12, 36, 40
# per-foot parameter layout
P_ACTIVE, P_POINT, P_NORMAL, P_GAIN, P_RESTRICT, P_RADIUS, P_PEAK = 0, 1, 4, 7, 8, 9, 10

# Go2 "home" joint configuration (robot_descriptions q0: hips 0, thighs 0.9, calves -1.8)
GO2_JOINT_HOME = np.tile([0.0, 0.9, -1.8], 4)


def pack_weights(cost: MPCCostConfig, restrict_cnt: bool = False) -> Tuple[np.ndarray, np.ndarray]:
    """Diagonal weights W [76], W_e [52]  (solver.py:108-141)."""
    W, We = np.zeros(NY), np.zeros(NYE)
    W[Y_BASE:Y_BASE + 12] = cost.W_base
    W[Y_JOINT:Y_JOINT + 24] = cost.W_joint
    W[Y_ACC:Y_ACC + 12] = cost.W_acc
    W[Y_SWING:Y_SWING + 4] = cost.W_swing
    W[Y_FREG:Y_FREG + 12] = np.asarray(cost.W_cnt_f_reg).reshape(-1)
    We[YE_BASE:YE_BASE + 12] = cost.W_e_base
    We[YE_JOINT:YE_JOINT + 24] = cost.W_e_joint
    We[YE_SWING:YE_SWING + 4] = cost.W_swing
    if restrict_cnt:   # only x, y of the foot target are penalised (solver.py:127-129)
        for f in range(NFEET):
            W[Y_POS + 3 * f:Y_POS + 3 * f + 2] = cost.W_foot_displacement[0]
            We[YE_POS + 3 * f:YE_POS + 3 * f + 2] = cost.W_foot_displacement[0]
    return W, We


def pack_reference(base_ref: np.ndarray, base_ref_e: Optional[np.ndarray], joint_ref: np.ndarray,
                   step_height, n_nodes: int, cnt_locations: Optional[np.ndarray] = None
                   ) -> Tuple[np.ndarray, np.ndarray]:
    """yref [B,N,76], yref_e [B,52] from base_ref [B,12], joint_ref [B,12] or [12], step_height
    [B] or scalar  (solver.py:153-177; contact targets solver.py:273-275)."""
    base_ref = np.atleast_2d(np.asarray(base_ref, dtype=float))
    B = base_ref.shape[0]
    base_ref_e = base_ref if base_ref_e is None else np.atleast_2d(np.asarray(base_ref_e, dtype=float))
    joint_ref = np.broadcast_to(np.asarray(joint_ref, dtype=float), (B, 12))
    step_height = np.broadcast_to(np.asarray(step_height, dtype=float), (B,))
    yref = np.zeros((B, n_nodes, NY))
    yref_e = np.zeros((B, NYE))
    yref[:, :, Y_BASE:Y_BASE + 12] = base_ref[:, None, :]
    yref[:, :, Y_JOINT:Y_JOINT + 12] = joint_ref[:, None, :]          # joint velocities ref = 0
    yref[:, :, Y_SWING:Y_SWING + 4] = step_height[:, None, None]
    yref_e[:, YE_BASE:YE_BASE + 12] = base_ref_e
    yref_e[:, YE_JOINT:YE_JOINT + 12] = joint_ref
    yref_e[:, YE_SWING:YE_SWING + 4] = step_height[:, None]
    if cnt_locations is not None:     # [B, 4, N+1, 3]
        loc = np.asarray(cnt_locations, dtype=float).reshape(B, NFEET, n_nodes + 1, 3)
        yref[:, :, Y_POS:Y_POS + 12] = loc[:, :, 1:, :].transpose(0, 2, 1, 3).reshape(B, n_nodes, 12)
        yref_e[:, YE_POS:YE_POS + 12] = loc[:, :, -1, :].reshape(B, 12)
    return yref, yref_e


def pack_params(cnt_sequence: np.ndarray, cost: MPCCostConfig, height_offset: float = 0.0,
                swing_peak: Optional[np.ndarray] = None, restrict_cnt: bool = False,
                cnt_locations: Optional[np.ndarray] = None, opt_peak: bool = True) -> np.ndarray:
    """Per-node parameters p [B,N+1,44] from cnt_sequence [B,4,N+1]
    (solver.py:212-252; contact locations :254-275)."""
    cnt = np.asarray(cnt_sequence)
    if cnt.ndim == 2:
        cnt = cnt[None]
    B, nf, n1 = cnt.shape
    assert nf == NFEET, "Invalid contact plan shape. Wrong number of end effectors."
    p = np.zeros((B, n1, NFEET, NP_FOOT))
    p[..., P_ACTIVE] = cnt.transpose(0, 2, 1)
    p[..., P_POINT + 2] = height_offset
    p[..., P_NORMAL + 2] = 1.0
    p[..., P_GAIN] = np.asarray(cost.W_foot_pos_constr_stab)[None, None, :]
    p[..., P_RADIUS] = cost.cnt_radius if restrict_cnt else 1.0e10
    if opt_peak and swing_peak is not None:
        pk = np.asarray(swing_peak)
        if pk.ndim == 2:
            pk = pk[None]
        p[..., P_PEAK] = pk.transpose(0, 2, 1)
    if restrict_cnt:
        d = np.diff(cnt.astype(np.int64), axis=-1, prepend=cnt[..., :1].astype(np.int64))
        d[d == -1] = 0
        p[..., P_RESTRICT] = d.transpose(0, 2, 1)
        if cnt_locations is not None:
            loc = np.asarray(cnt_locations, dtype=float).reshape(B, NFEET, n1, 3)
            p[..., P_POINT:P_POINT + 3] = loc.transpose(0, 2, 1, 3)
    return p.reshape(B, n1, NP)


def base_ref_vel_tracking(q: np.ndarray, v_des: np.ndarray, w_yaw, tracking: np.ndarray,
                          nom_height, height_offset: float = 0.0, t_horizon: float = 1.0
                          ) -> Tuple[np.ndarray, np.ndarray]:
    """Batched `LocomotionMPC.compute_base_ref_vel_tracking` (mpc.py:210-272).

    q [B,18] Euler state, v_des [B,3] in the base frame, w_yaw [B], tracking [B,12] =
    `base_ref_vel_tracking` (integrated reference pose), nom_height [B] or scalar.
    Returns base_ref, base_ref_e, each [B,12] = [pos, ypr, lin vel, ang vel(ypr order)].
    """
    q = np.atleast_2d(q)
    B = q.shape[0]
    v_des = np.broadcast_to(np.asarray(v_des, dtype=float), (B, 3))
    w_yaw = np.broadcast_to(np.asarray(w_yaw, dtype=float), (B,))
    tracking = np.broadcast_to(np.asarray(tracking, dtype=float), (B, 12))
    nom_height = np.broadcast_to(np.asarray(nom_height, dtype=float), (B,))
    # vectorised over the batch (no per-problem Python): the same operations in the same order as the reference
    r = np.zeros((B, 12))
    r[:, :2] = np.round(q[:, :2], 2)
    r[:, 2] = nom_height + height_offset
    r[:, 3] = py_round1(q[:, 3])
    R_WB = rpy_to_matrix_batch(tracking[:, 5], tracking[:, 4], tracking[:, 3])
    v_glob = np.round(matvec3(R_WB, v_des), 1)
    r[:, 6:9] = v_glob
    r[:, 9] = w_yaw                                  # w_des[::-1] with w_des = (0, 0, w_yaw)
    e = r.copy()
    zeros = np.zeros(B)
    e[:, 6:9] = matvec3(rpy_to_matrix_batch(zeros, zeros, w_yaw * t_horizon), r[:, 6:9])
    e[:, :2] = tracking[:, :2] + v_glob[:, :2] * t_horizon
    e[:, :2] = np.clip(e[:, :2], -r[:, :2] + v_glob[:, :2] * t_horizon * 1.2, r[:, :2] + v_glob[:, :2] * t_horizon * 1.2)
    yaw_ref = tracking[:, 3]
    e[:, 3] = yaw_ref + w_yaw * t_horizon
    e[:, 3] = np.clip(e[:, 3], -yaw_ref + w_yaw * t_horizon * 1.5, yaw_ref + w_yaw * t_horizon * 1.5)
    r[:, :2] += (e[:, :2] - r[:, :2]) * 0.75
    r[:, 3] += (e[:, 3] - r[:, 3]) * 0.75
    e[:, 8] = 0.0
    e[:, 4:6] = 0.0
    r[:, 4:6] = 0.0
    e[:, -2:] = 0.0
    return r, e


def py_round1(x: np.ndarray) -> np.ndarray:
    """Python's `round(x, 1)` (correctly rounded decimal, mpc.py:222) over an array: numpy's `round` rounds the
    rounded product x * 10 instead, which differs when that product lands on a tie.  x * 10 is exact in the x87
    extended format (53 + 4 bits)."""
    x = np.asarray(x, dtype=np.float64)
    if np.finfo(np.longdouble).nmant >= 57:
        return (np.rint(x.astype(np.longdouble) * np.longdouble(10)) / np.longdouble(10)).astype(np.float64)
    return np.array([round(float(v), 1) for v in x.ravel()]).reshape(x.shape)


def rpy_to_matrix_batch(roll, pitch, yaw) -> np.ndarray:
    """`transform.rpy_to_matrix` (pin.rpy.rpyToMatrix) for arrays: [B,3,3]."""
    roll, pitch, yaw = np.broadcast_arrays(np.asarray(roll, dtype=float), np.asarray(pitch, dtype=float),
                                           np.asarray(yaw, dtype=float))
    cr, sr, cp, sp, cy, sy = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    R = np.empty(np.shape(cr) + (3, 3))
    R[..., 0, 0], R[..., 0, 1], R[..., 0, 2] = cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr
    R[..., 1, 0], R[..., 1, 1], R[..., 1, 2] = sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr
    R[..., 2, 0], R[..., 2, 1], R[..., 2, 2] = -sp, cp * sr, cp * cr
    return R


def matvec3(R: np.ndarray, v: np.ndarray) -> np.ndarray:
    """Row-by-row R @ v with the summation order of a 3-term dot product, batched."""
    return (R[..., 0] * v[..., None, 0] + R[..., 1] * v[..., None, 1]) + R[..., 2] * v[..., None, 2]


def nominal_state(nom_height: float = 0.30, joint_ref: np.ndarray = GO2_JOINT_HOME) -> np.ndarray:
    """qv [36]: base at (0,0,nom_height), level, joints at `joint_ref`, zero velocity."""
    qv = np.zeros(36)
    qv[2] = nom_height
    qv[6:18] = joint_ref
    return qv
