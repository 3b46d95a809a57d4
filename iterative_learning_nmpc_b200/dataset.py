"""Dataset writer in the reference's formats, fed by batched device outputs.

Reference (one rollout, one MuJoCo step at a time):
  `Behavior_Cloning/utils/Rollout_MPC.py:183-194` keys of the per-rollout `.npz`
      time, q, v, ctrl, feet_pos_w, base_wrt_feet, state[44], action[12], vc_goals, cc_goals, contact_vec
  `:218-298` how each entry is formed:
      q, v             MuJoCo vectors (q: pos 3 | quat wxyz 4 | joints 12; v: 18), leg order FL FR RL RR
      ctrl             applied torque in MuJoCo actuator order FR FL RR RL
      feet_pos_w       12, world frame, FL FR RL RR
      base_wrt_feet    8 = (base - foot)[:2] per foot
      state            [phase(1) | v(18) | q[2:](17) | base_wrt_feet(8)] = 44
      action           (tau_FLFRRLRR + kd v[6:]) / kp + q[7:]            (PD target)
  `:205-207` file names `traj_nominal_<timestamp>.npz` / `traj_<replanning_point>_<nth>.npz`
  `Behavior_Cloning/utils/database.py:156-203` HDF5 with states, actions, vc_goals, cc_goals, traj_ids, traj_times
  (through `h5py` when it is importable, else through the built-in writer `hdf5_min.py`: h5py is not part of this image).

The rollout here is the sim-free one of `LocomotionMPC.open_loop` (mpc.py:416-462): the recorded
states are the plan's own samples (`nmpc_plan_torques`, evaluated on the plan), `contact_vec` is the
planned contact mask of the node a sample falls in, MuJoCo physics being unavailable on this box.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List, Optional, Sequence

import numpy as np

from .mpc_controller.utils.transform import euler_derivative_to_local_angular, matrix_to_quat, rpy_to_matrix

KEYS = ("time", "q", "v", "ctrl", "feet_pos_w", "base_wrt_feet", "state", "action", "vc_goals", "cc_goals",
        "contact_vec")
# FL FR RL RR (model / dataset order) -> FR FL RR RL (MuJoCo actuator order of `ctrl`)
_TO_ACTUATOR = np.r_[3:6, 0:3, 9:12, 6:9]


def euler_to_mujoco(q: np.ndarray, v: np.ndarray):
    """Batched `QuadrupedDynamics.convert_to_mujoco` (dynamics.py:86-99): q [n,18] Euler (pos | yaw pitch roll |
    joints), v [n,18] -> q_mj [n,19] (pos | quat wxyz | joints), v_mj [n,18] (body angular velocity)."""
    q, v = np.atleast_2d(np.asarray(q, float)), np.atleast_2d(np.asarray(v, float))
    n = q.shape[0]
    q_mj = np.zeros((n, 19))
    v_mj = v.copy()
    q_mj[:, :3] = q[:, :3]
    q_mj[:, 7:] = q[:, 6:]
    # vectorised over the samples (a closed-loop batch is 10^6 of them): Shepperd's branches of `matrix_to_quat`
    # on R = Rz(yaw) Ry(pitch) Rx(roll), the Euler-rate map of `euler_derivative_to_local_angular`
    from .problem import rpy_to_matrix_batch
    R = rpy_to_matrix_batch(q[:, 5], q[:, 4], q[:, 3])
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    quat = np.zeros((n, 4))                                   # x y z w
    pos = tr > 0
    with np.errstate(invalid="ignore", divide="ignore"):
        s4 = 2.0 * np.sqrt(np.where(pos, 1.0 + tr, 1.0))
        quat[pos] = np.stack(((R[:, 2, 1] - R[:, 1, 2]) / s4, (R[:, 0, 2] - R[:, 2, 0]) / s4,
                              (R[:, 1, 0] - R[:, 0, 1]) / s4, 0.25 * s4), axis=1)[pos]
    for idx in np.nonzero(~pos)[0]:                           # rotations beyond 120 degrees: rare, done one by one
        quat[idx] = matrix_to_quat(R[idx])
    quat[pos] /= np.linalg.norm(quat[pos], axis=1, keepdims=True)
    q_mj[:, 3], q_mj[:, 4:7] = quat[:, 3], quat[:, :3]
    cx, sx, cy, sy = np.cos(q[:, 5]), np.sin(q[:, 5]), np.cos(q[:, 4]), np.sin(q[:, 4])
    yd, pd, rd = v[:, 3], v[:, 4], v[:, 5]
    v_mj[:, 3] = (-sy * yd + 0.0 * pd) + 1.0 * rd
    v_mj[:, 4] = (cy * sx * yd + cx * pd) + 0.0 * rd
    v_mj[:, 5] = (cx * cy * yd + -sx * pd) + 0.0 * rd
    return q_mj, v_mj


def rollout_record(times: np.ndarray, q_euler: np.ndarray, v_euler: np.ndarray, tau: np.ndarray, feet_w: np.ndarray,
                   contact_vec: np.ndarray, vc_goal: Sequence[float], kp: float = 20.0, kd: float = 1.5,
                   phase: Optional[np.ndarray] = None, rng: Optional[np.random.Generator] = None
                   ) -> Dict[str, np.ndarray]:
    """One rollout in the reference's record layout.  times [n]; q_euler, v_euler [n,18]; tau [n,12] (FL FR RL RR);
    feet_w [n,4,3]; contact_vec [n,4]; vc_goal [3].  `phase` defaults to 0 (the reference's current
    `get_phase_percentage`, RolloutMPC_force_perturbation.py:259-269); `cc_goals` is the reference's N(0, 0.1)
    placeholder noise (Rollout_MPC.py:160,307)."""
    n = len(times)
    rng = rng or np.random.default_rng(0)
    q_mj, v_mj = euler_to_mujoco(q_euler, v_euler)
    feet_w = np.asarray(feet_w, float).reshape(n, 4, 3)
    base_wrt_feet = (q_mj[:, None, :3] - feet_w)[:, :, :2].reshape(n, 8)
    ph = np.zeros((n, 1)) if phase is None else np.round(np.asarray(phase, float).reshape(n, 1), 4)
    tau = np.asarray(tau, float).reshape(n, 12)
    return {
        "time": np.round(np.asarray(times, float), 4),
        "q": q_mj, "v": v_mj,
        "ctrl": tau[:, _TO_ACTUATOR],
        "feet_pos_w": feet_w.reshape(n, 12),
        "base_wrt_feet": base_wrt_feet,
        "state": np.concatenate([ph, v_mj, q_mj[:, 2:], base_wrt_feet], axis=1),
        "action": (tau + kd * v_mj[:, 6:]) / kp + q_mj[:, 7:],
        "vc_goals": np.tile(np.asarray(vc_goal, float), (n, 1)),
        "cc_goals": rng.normal(0.0, 0.1, (n, 8)),
        "contact_vec": np.asarray(contact_vec).reshape(n, 4).astype(np.int64),
    }


def save_rollout(record: Dict[str, np.ndarray], record_dir: str, nominal: bool = True, replanning_point: int = 0,
                 nth_traj: int = 0, timestamp: Optional[str] = None) -> str:
    """`DataRecorder.save` (Rollout_MPC.py:196-216): one `.npz` per rollout, reference file names."""
    missing = [k for k in KEYS if k not in record]
    if missing:
        raise KeyError(f"rollout record lacks {missing}")
    os.makedirs(record_dir, exist_ok=True)
    if nominal:
        name = f"traj_nominal_{timestamp or time.strftime('%Y%m%d_%H%M%S')}.npz"
    else:
        name = f"traj_{replanning_point}_{nth_traj}.npz"
    path = os.path.join(record_dir, name)
    np.savez(path, **{k: record[k] for k in KEYS})
    return path


def batch_to_rollouts(solver, batch: Dict[str, np.ndarray], sim_dt: float = 1.0e-3, n_steps: int = 20,
                      kp: float = 20.0, kd: float = 1.5, t0: float = 0.0) -> List[Dict[str, np.ndarray]]:
    """Open-loop expert labels for every problem of a solved batch: the first `n_steps` sim steps of each plan
    (one replanning interval), recorded in the reference's layout.  `solver` is a solved `NmpcSolver`; the
    interpolation, inverse dynamics and PD run on the device (`nmpc_plan_torques`), foot positions come from
    the device kinematics."""
    B, N = solver.B, solver.N
    n_interp = round(N * float(solver.opts.dt_nodes) / sim_dt)
    out = solver.plan_torques(n_interp, 0, n_steps, kp=kp, kd=kd)
    qp, vp = out["q_plan"], out["v_plan"]
    feet, _, _ = solver.kinematics(np.concatenate([qp, vp], axis=2).reshape(B * n_steps, 36))
    feet = np.asarray(feet).reshape(B, n_steps, 4, 3)
    times = t0 + sim_dt * np.arange(1, n_steps + 1)
    node = np.minimum((times - t0) / float(solver.opts.dt_nodes), N).astype(int)
    cnt = batch["compact"]["cnt_sequence"]                       # [B,4,N+1]
    recs = []
    for b in range(B):
        recs.append(rollout_record(times, qp[b], vp[b], out["tau"][b], feet[b], cnt[b][:, node].T,
                                   batch["v_des"][b], kp, kd, rng=np.random.default_rng(b)))
    return recs


def save_hdf5(path: str, records: Sequence[Dict[str, np.ndarray]]) -> str:
    """`Database` file of the reference's training code (database.py:156-203): states, actions, vc_goals,
    cc_goals, traj_ids, traj_times stacked over rollouts.  Written through h5py when it is installed, else by the
    built-in writer `hdf5_min.write_hdf5` (same dataset names, shapes and dtypes; contiguous, uncompressed)."""
    if not len(records):
        raise ValueError("save_hdf5 needs at least one rollout record")
    data = {
        "states": np.concatenate([r["state"] for r in records]).astype(np.float64),
        "actions": np.concatenate([r["action"] for r in records]).astype(np.float64),
        "vc_goals": np.concatenate([r["vc_goals"] for r in records]).astype(np.float64),
        "cc_goals": np.concatenate([r["cc_goals"] for r in records]).astype(np.float64),
        "traj_ids": np.concatenate([np.full(len(r["time"]), i, dtype=np.int64) for i, r in enumerate(records)]),
        "traj_times": np.concatenate([r["time"] for r in records]).astype(np.float64),
    }
    try:
        import h5py
    except ImportError:                                          # not in this image
        from .hdf5_min import write_hdf5
        return write_hdf5(path, data)
    with h5py.File(path, "w") as hf:
        for name, arr in data.items():
            hf.create_dataset(name, data=arr)
    return path


def load_hdf5(path: str) -> Dict[str, np.ndarray]:
    """Read a `Database` file back the way `Database.load_saved_database` does (database.py:172-198).  Needs h5py:
    the built-in writer has no reader in the product (tests/hdf5_reader.py is the checker)."""
    import h5py
    with h5py.File(path, "r") as hf:
        return {k: hf[k][:] for k in hf.keys()}


def closed_loop_to_rollouts(solver, res: Dict[str, np.ndarray], v_des: np.ndarray, sim_dt: float = 1.0e-3,
                            kp: float = 20.0, kd: float = 1.5, t0: float = 0.0, contact_threshold: float = 1.0
                            ) -> List[Dict[str, np.ndarray]]:
    """Expert rollouts of `BatchedLocomotionMPC.closed_loop` in the reference's record layout -- what
    `Rollout_MPC.run` records one MuJoCo rollout at a time (Rollout_MPC.py:218-298): the states are the simulated ones
    the controller saw, `ctrl` the torques it applied, `action` the PD target that reproduces them, `contact_vec`
    the feet whose contact force exceeds `contact_threshold` [N] (MuJoCo: the geoms in contact).
    `res` = dict(qv [n,B,36], tau [n,B,12], fc [n,B,12]); `solver` supplies the device kinematics."""
    qv, tau, fc = np.asarray(res["qv"]), np.asarray(res["tau"]), np.asarray(res["fc"])
    n, B = qv.shape[:2]
    feet, _, _ = solver.kinematics(qv.reshape(n * B, 36))
    feet = np.asarray(feet).reshape(n, B, 4, 3)
    times = t0 + sim_dt * np.arange(n)
    v_des = np.broadcast_to(np.asarray(v_des, float), (B, 3))
    contact = (fc.reshape(n, B, 4, 3)[..., 2] > contact_threshold)
    return [rollout_record(times, qv[:, b, :18], qv[:, b, 18:], tau[:, b], feet[:, b], contact[:, b], v_des[b], kp, kd,
                           rng=np.random.default_rng(b)) for b in range(B)]
