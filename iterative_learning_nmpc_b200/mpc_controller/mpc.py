"""`LocomotionMPC` -- the caller of the hot path, same public surface as the reference's
(mpc_controller/mpc.py:19-693): replanning cadence, base-reference generation, (a)synchronous solve,
interpolation of the plan to the simulation rate, feed-forward + PD torques.

Differences that follow from the environment, not from choice: there is no mj_pin / MuJoCo here, so
the `PinController` base is a two-attribute stand-in and `compute_torques_dof` accepts any object
with `.time/.qpos/.qvel`; plotting and the keyboard goal (matplotlib / mj_pin `Keyboard`) are left
out.  Everything numeric goes through `QuadrupedAcadosSolver`, i.e. the
CUDA library.
"""
from __future__ import annotations

import math
import time
import traceback
from collections import defaultdict
from concurrent.futures import Future, ThreadPoolExecutor
from typing import Any, List, Optional, Tuple

import numpy as np
from scipy.interpolate import CubicHermiteSpline

from .config.quadruped.utils import get_quadruped_config
from .utils.contact_planner import ContactPlanner, CustomContactPlanner, RaiberContactPlanner
from .utils.profiling import print_timings, time_fn
from .utils.solver import QuadrupedAcadosSolver
from .utils.transform import rpy_to_matrix


class PinController:
    """What the reference inherits from `mj_pin.abstract.PinController`: the torque vector the
    simulator reads back."""

    def __init__(self, pin_model):
        self.pin_model = pin_model
        self.torques_dof = np.zeros(pin_model.nv)

    def get_torque_map(self):
        return {i: float(t) for i, t in enumerate(self.torques_dof[6:])}


N_SQP_FIRST = 15            # mpc.py:465


class LocomotionMPC(PinController):

    def __init__(self, path_urdf: Optional[str], feet_frame_names: List[str], robot_name: str,
                 gait_name: str = "trot", joint_ref: np.ndarray = None, interactive_goal: bool = False,
                 sim_dt: float = 1.0e-3, height_offset: float = 0., contact_planner: str = "",
                 print_info: bool = True, compute_timings: bool = True, solve_async: bool = True) -> None:
        if interactive_goal:
            raise NotImplementedError("keyboard velocity goals need mj_pin's viewer (out of scope, SURVEY.md 2.1 #8)")
        self.gait_name, self.print_info, self.height_offset = gait_name, print_info, height_offset
        self.config_gait, self.config_opt, self.config_cost = get_quadruped_config(gait_name, robot_name)
        self.solver = QuadrupedAcadosSolver(path_urdf, feet_frame_names, self.config_opt, self.config_cost,
                                            height_offset, print_info, compute_timings)
        dyn = self.solver.dyn
        super().__init__(pin_model=dyn.pin_model)
        self.nu, self.nq, self.nv = dyn.pin_model.nv - 6, dyn.pin_model.nq, dyn.pin_model.nv
        if joint_ref is None:
            home = dyn.pin_model.referenceConfigurations.get("home")
            if home is not None:
                joint_ref = np.asarray(home)[-self.nu:]
            else:
                print("Joint reference not found in pinocchio model. Set to zero.")
                joint_ref = np.zeros(self.nu)
        self.joint_ref = np.asarray(joint_ref, dtype=float)[-self.nu:]

        # nominal pose: base at the origin, joints at the reference (mpc.py:69-72); Euler state
        q0, v0 = np.zeros(self.nv), np.zeros(self.nv)
        q0[-self.nu:] = self.joint_ref
        dyn.update_pin(q0, v0)

        self.n_foot = len(feet_frame_names)
        self._contact_planner_str = contact_planner
        kind = contact_planner.lower()
        if kind == "raibert":
            offset_hip_b = dyn.get_feet_position_w()
            offset_hip_b[:, -1] = 0.
            self.contact_planner = RaiberContactPlanner(feet_frame_names, self.solver.dt_nodes, self.config_gait,
                                                        offset_hip_b, y_offset=0.02, x_offset=0.04,
                                                        foot_size=0.0085, cache_cnt=False)
            self.restrict_cnt = True
        elif kind == "custom":
            self.contact_planner = CustomContactPlanner(feet_frame_names, self.solver.dt_nodes, self.config_gait)
            self.restrict_cnt = True
        else:
            self.contact_planner = ContactPlanner(feet_frame_names, self.solver.dt_nodes, self.config_gait)
            self.restrict_cnt = False
        self.solver.set_contact_restriction(self.restrict_cnt)

        self.Kp, self.Kd = self.config_opt.Kp, self.config_opt.Kd
        self.sim_dt = sim_dt
        self.dt_nodes: float = self.solver.dt_nodes
        self.replanning_freq: int = self.config_opt.replanning_freq
        self.replanning_steps: int = int(1 / (self.replanning_freq * sim_dt))
        self.solve_async, self.compute_timings, self.interactive_goal = solve_async, compute_timings, False
        self.reset(reset_solver=False)

    # -- state ----------------------------------------------------------------------------------
    def reset(self, reset_solver: bool = True) -> None:
        """mpc.py:121-169."""
        if reset_solver:
            self.solver.reset()
        self.first_solve, self.diverged = True, False
        self.t0, self.sim_step, self.plan_step, self.current_opt_node, self.delay = 0., 0, 0, 0, 0
        self.v_des, self.w_des = np.zeros(3), np.zeros(3)
        self.base_ref_vel_tracking = np.zeros(12)
        n = self.n_interp_plan = round(self.config_opt.time_horizon / self.sim_dt)
        self.id_repeat = np.int32(np.linspace(0, 1, n) * (self.config_opt.n_nodes - 1))
        self.q_plan, self.v_plan, self.a_plan = (np.zeros((n, self.nv)) for _ in range(3))
        self.f_plan = np.zeros((n, self.n_foot, 3))
        self.time_traj = np.zeros(n)
        self.qref_pd = np.zeros(self.nq)
        self.q_plan_full, self.q_full, self.v_plan_full, self.v_full = [], [], [], []
        self.a_full, self.f_full, self.tau_full, self.dt_full = [], [], [], []
        self.timings = defaultdict(list)
        self.executor = ThreadPoolExecutor(max_workers=1)       # one worker: the asynchronous solve
        self.optimize_future: Future = Future()
        self.plan_submitted = False
        self.velocity_goal = None

    def _replan(self) -> bool:
        replan = self.sim_step % self.replanning_steps == 0
        if self.solve_async:
            replan &= not self.plan_submitted
        return replan

    def _step(self) -> None:
        self.increment_base_ref_position()
        self.sim_step += 1
        self.plan_step += 1

    def _record_plan(self) -> None:
        sl = slice(self.delay, self.plan_step)
        self.q_full.append(self.q_plan[sl].copy())
        self.v_full.append(self.v_plan[sl].copy())
        self.a_full.append(self.a_plan[sl].copy())
        self.f_full.append(self.f_plan[sl].copy())

    def set_command(self, v_des: np.ndarray = np.zeros((3,)), w_yaw: float = 0.) -> None:
        self.v_des = np.asarray(v_des, dtype=float)
        self.w_des[2] = w_yaw

    # -- base references (mpc.py:204-315) -------------------------------------------------------------
    def _v_des_world(self) -> np.ndarray:
        R_WB = rpy_to_matrix(self.base_ref_vel_tracking[3:6][::-1])
        return np.round(R_WB @ self.v_des, 1)

    def increment_base_ref_position(self):
        v = self._v_des_world()
        self.base_ref_vel_tracking[:2] += v[:2] * self.sim_dt
        self.base_ref_vel_tracking[3] += self.w_des[-1] * self.sim_dt

    def compute_base_ref_vel_tracking(self, q: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """12-d running and terminal base references from the velocity command (mpc.py:210-272);
        the batched twin is `problem.base_ref_vel_tracking`."""
        from ..problem import base_ref_vel_tracking
        ref, ref_e = base_ref_vel_tracking(np.asarray(q, dtype=float)[None, :18], self.v_des, self.w_des[-1],
                                           self.base_ref_vel_tracking, self.config_gait.nom_height,
                                           self.height_offset, self.config_opt.time_horizon)
        return ref[0], ref_e[0]

    def compute_base_ref_cnt_restricted(self, q_mj: np.ndarray, contact_locations: np.ndarray):
        """References centred on the first / last complete set of planned contacts (mpc.py:274-315)."""
        loc = np.unique(contact_locations, axis=1)
        nz = np.argwhere(np.all(loc != np.zeros(3), axis=-1))
        count = np.bincount(nz[:, 1]) if len(nz) else np.zeros(0, dtype=int)
        if len(count) > 0:
            first = int(np.argmax(count))
            last = len(count) - int(np.argmax(count[::-1])) - 1
            c_first, c_last = loc[:, first, :].mean(axis=0), loc[:, last, :].mean(axis=0)
        else:
            c_first, c_last = contact_locations[:, 0, :].mean(axis=0), contact_locations[:, -1, :].mean(axis=0)
        base_ref, base_ref_e = np.zeros(12), np.zeros(12)
        alpha = 0.35
        base_ref[:2] = alpha * c_first[:2] + (1 - alpha) * c_last[:2]
        base_ref_e[:2] = c_last[:2]
        base_ref[2] = base_ref_e[2] = self.config_gait.nom_height + self.height_offset
        return base_ref, base_ref_e

    # -- one replanning tick (mpc.py:317-369) -----------------------------------------------------
    @time_fn("optimize")
    def optimize(self, q: np.ndarray, v: np.ndarray):
        """Contact plan + references for the current node, then `solver.init` and `solver.solve`."""
        N1 = self.config_opt.n_nodes + 1
        self.solver.dyn.update_pin(q, v)
        cnt_sequence = self.contact_planner.get_contacts(self.current_opt_node, N1)
        swing_peak = self.contact_planner.get_peaks(self.current_opt_node, N1) if self.config_opt.opt_peak else None
        cnt_locations = None
        if self.restrict_cnt:
            if self._contact_planner_str.lower() == "raibert":
                com_xyz = self.solver.dyn.pin_data.com
                self.contact_planner.set_state(q[:3], v[:3], q[3:6][::-1], com_xyz, self.v_des, self.w_des[-1])
            cnt_locations = self.contact_planner.get_locations(self.current_opt_node, N1)
            base_ref, base_ref_e = self.compute_base_ref_cnt_restricted(q, cnt_locations)
        else:
            base_ref, base_ref_e = self.compute_base_ref_vel_tracking(q)
        self.solver.init(self.current_opt_node, q, v, base_ref, base_ref_e, self.joint_ref,
                         self.config_gait.step_height, cnt_sequence, cnt_locations, swing_peak)
        return self.solver.solve()

    # -- plan interpolation (mpc.py:371-414) ----------------------------------------------------------
    def interpolate_state_trajectory(self, q_sol, v_sol, a_sol, dt_sol) -> Tuple[np.ndarray, np.ndarray]:
        """Node solution -> sim-rate samples; sample 0 (the current state) is dropped."""
        time_traj = np.concatenate(([0.], np.cumsum(dt_sol)))
        q_plan, v_plan = self.interpolate_trajectory_with_derivatives(time_traj, q_sol, v_sol, a_sol)
        return q_plan[1:], v_plan[1:]

    def interpolate_trajectory_with_derivatives(self, time_traj, positions, velocities, accelerations):
        """Cubic Hermite in (q, v) and in (v, a) with a_0 repeated for the last knot's slope
        convention of the reference (mpc.py:388-414)."""
        t = np.linspace(time_traj[0], time_traj[-1], self.n_interp_plan + 1)
        pos = CubicHermiteSpline(time_traj, positions, velocities)(t)
        acc = np.concatenate((accelerations[:1], accelerations))
        vel = CubicHermiteSpline(time_traj, velocities, acc)(t)
        return pos, vel

    # -- simulator-free MPC loop (mpc.py:416-462) ---------------------------------------------------
    def open_loop(self, q_mj: np.ndarray, v_mj: np.ndarray, trajectory_time: float) -> np.ndarray:
        """Roll the MPC forward on its own plan from (q_mj, v_mj); returns the MuJoCo-format
        configurations visited, one per sim step."""
        dyn = self.solver.dyn
        out = []
        sim_time = 0.
        while sim_time <= trajectory_time:
            if sim_time >= (self.current_opt_node + 1) * self.dt_nodes:
                self.current_opt_node += 1
            if self._replan():
                if self.sim_step > 0:
                    self._record_plan()
                self.set_convergence_on_first_iter()
                q, v = dyn.convert_from_mujoco(q_mj, v_mj)
                q_sol, v_sol, a_sol, _, dt_sol = self.optimize(q, v)
                self.q_plan[:], self.v_plan[:] = self.interpolate_state_trajectory(q_sol, v_sol, a_sol, dt_sol)
                self.plan_step = 0
                self.first_solve = False
            q_mj, v_mj = dyn.convert_to_mujoco(self.q_plan[self.plan_step], self.v_plan[self.plan_step])
            out.append(q_mj)
            self._step()
            sim_time += self.sim_dt
        return np.array(out)

    def set_convergence_on_first_iter(self):
        """First solve: 15 SQP iterations at a tenth of the tolerances; afterwards the configured
        values (mpc.py:464-473)."""
        o = self.solver.config_opt
        if self.first_solve:
            self.solver.set_max_iter(N_SQP_FIRST)
            self.solver.set_nlp_tol(o.nlp_tol / 10.)
            self.solver.set_qp_tol(o.qp_tol / 10.)
        elif self.sim_step <= self.replanning_steps:
            self.solver.set_max_iter(o.max_iter)
            self.solver.set_nlp_tol(o.nlp_tol)
            self.solver.set_qp_tol(o.qp_tol)

    # -- closed-loop torque interface (mpc.py:475-599) ------------------------------------------------
    def compute_torques_dof(self, mj_data: Any) -> None:
        t, q_mj, v_mj = mj_data.time, np.asarray(mj_data.qpos, dtype=float), np.asarray(mj_data.qvel, dtype=float)
        torques_ff = self._compute_torques_ff(t, q_mj, v_mj)
        torques_pd = self._compute_pd_torques(q_mj, v_mj, torques_ff)
        self.tau_full.append(torques_pd.copy())
        self.torques_dof[-self.nu:] = torques_pd

    def _adopt_plan(self, t: float):
        """Take the finished solve: interpolate states, hold inputs, compensate the solve delay."""
        q_sol, v_sol, a_sol, f_sol, dt_sol = self.optimize_future.result()
        if not self.first_solve:
            self._record_plan()
        self.q_plan[:], self.v_plan[:] = self.interpolate_state_trajectory(q_sol, v_sol, a_sol, dt_sol)
        self.a_plan[:] = a_sol[self.id_repeat]
        self.f_plan[:] = f_sol[self.id_repeat]
        if self.solve_async and not self.first_solve:
            self.delay = math.ceil((t - self.start_time) / self.sim_dt) - 1
        else:
            self.delay = 0
        self.plan_step = self.delay
        self.plan_submitted = False
        self.first_solve = False

    def _compute_torques_ff(self, sim_time: float, q_mj: np.ndarray, v_mj: np.ndarray) -> np.ndarray:
        t = round(sim_time - self.t0, 4)
        q, v = self.solver.dyn.convert_from_mujoco(q_mj, v_mj)
        if not self.first_solve and t >= (self.current_opt_node + 1) * self.dt_nodes:
            self.current_opt_node += 1
        if self._replan():
            self.set_convergence_on_first_iter()
            self.start_time = t
            self.optimize_future = self.executor.submit(self.optimize, q, v)
            self.plan_submitted = True
            if self.print_info:
                print("\n" + "#" * 10, "Replan", "#" * 10)
                print("Current node:", self.current_opt_node, "Sim time:", t, "Sim step:", self.sim_step, "\n")
            while not self.solve_async and not self.optimize_future.done():
                time.sleep(5.0e-4)
        if self.plan_submitted and self.optimize_future.done():
            try:
                self._adopt_plan(t)
            except Exception:
                print("Optimization error:\n")
                print(traceback.format_exc())
                self.optimize_future = Future()
                self.diverged = True
                self.plan_submitted = False
                self.executor.shutdown(wait=False, cancel_futures=True)
                time.sleep(0.1)
        if self.first_solve:                      # no plan yet: PD on the current pose
            self.t0 = t
            if np.all(self.q_plan[0, :] == 0.):
                self.q_plan[:] = q.reshape(1, -1)
            return np.zeros(self.nu)
        self.q_plan_full.append(q)
        self.v_plan_full.append(v)
        tau = self.solver.dyn.id_torques(q, v, self.a_plan[self.plan_step], self.f_plan[self.plan_step])
        self._step()
        return tau

    def _compute_pd_torques(self, q: np.ndarray, v: np.ndarray, torques_ff: np.ndarray) -> np.ndarray:
        """Feed-forward + joint PD on the plan; stiffer gains until the first plan exists
        (mpc.py:592-599; `q`, `v` are the MuJoCo vectors, whose last `nu` entries are the joints)."""
        Kp, Kd = (44, 5) if self.first_solve else (self.Kp, self.Kd)
        k = self.plan_step
        return (torques_ff + Kp * (self.q_plan[k, -self.nu:] - q[-self.nu:])
                + Kd * (self.v_plan[k, -self.nu:] - v[-self.nu:]))

    def print_timings(self):
        print()
        print_timings(self.timings)
        print_timings(self.solver.timings)

    def __del__(self):
        ex = getattr(self, "executor", None)
        if ex is not None:
            ex.shutdown(wait=False, cancel_futures=True)


class BatchedLocomotionMPC:
    """B sim-free MPC rollouts in lockstep: `LocomotionMPC.open_loop` (mpc.py:416-462) with a leading batch
    dimension, which is what the data-collection loops (`Behavior_Cloning/utils/Rollout_MPC.py`,
    `DAgger/utils/data_collection_locosafedagger.py`) need from the controller when thousands of rollouts are
    run at once.  Every rollout has its own start state, gait phase offset and velocity command; all share the
    gait, the options and the sim clock.  Per replanning tick: one batched `init` + `solve`
    (`BatchedQuadrupedSolver`), one batched plan -> torque call on the device (`nmpc_plan_torques`) for the
    sim steps until the next tick; between ticks the state follows the plan, exactly as in `open_loop`.

    Contact planners: the cyclic `ContactPlanner` (default) and `contact_planner="raibert"` (Raibert footholds,
    contact-patch restriction, references centred on the planned contacts), both planned on the device; states are Euler vectors `[B,18]` (use
    `dataset.euler_to_mujoco` / `QuadrupedDynamics.convert_from_mujoco` at the boundary).
    """

    def __init__(self, batch: int, path_urdf: Optional[str], feet_frame_names: List[str], robot_name: str,
                 gait_name: str = "trot", joint_ref: Optional[np.ndarray] = None, sim_dt: float = 1.0e-3,
                 height_offset: float = 0., print_info: bool = False, device: int = 0, contact_planner: str = ""):
        from .utils.solver import BatchedQuadrupedSolver
        self.B = int(batch)
        self.config_gait, self.config_opt, self.config_cost = get_quadruped_config(gait_name, robot_name)
        self.solver = BatchedQuadrupedSolver(self.B, path_urdf, feet_frame_names, self.config_opt, self.config_cost,
                                             height_offset, print_info, device=device)
        self.height_offset, self.sim_dt = height_offset, sim_dt
        self.dt_nodes = self.solver.dt_nodes
        self.replanning_steps = int(1 / (self.config_opt.replanning_freq * sim_dt))
        self.n_interp_plan = round(self.config_opt.time_horizon / sim_dt)
        nu = 12
        self.joint_ref = np.zeros(nu) if joint_ref is None else np.asarray(joint_ref, dtype=float)[-nu:]
        self._contact_planner_str = contact_planner.lower()
        if self._contact_planner_str == "raibert":           # mpc.py:76-90
            q0 = np.zeros((1, 36))
            q0[0, 6:18] = self.joint_ref
            offset_hip_b = np.array(self.solver.handle.kinematics(q0)[0], dtype=float).reshape(4, 3)
            offset_hip_b[:, -1] = 0.
            self.contact_planner = RaiberContactPlanner(feet_frame_names, self.dt_nodes, self.config_gait, offset_hip_b,
                                                        y_offset=0.02, x_offset=0.04, foot_size=0.0085, cache_cnt=False)
            self.restrict_cnt = True
        elif self._contact_planner_str in ("", "cyclic"):
            self.contact_planner = ContactPlanner(feet_frame_names, self.dt_nodes, self.config_gait)
            self.restrict_cnt = False
        else:
            raise NotImplementedError(f"batched contact planner '{contact_planner}' (cyclic and raibert are built)")
        self.solver.set_contact_restriction(self.restrict_cnt)
        self.Kp, self.Kd = self.config_opt.Kp, self.config_opt.Kd
        self.plan_on_device = True           # False: the host planners + the reference's `init` sequence (same results)
        self.unpack_solution = False         # device planning: keep the solution resident, do not copy 15 KB per rollout back
        self.reset(reset_solver=False)

    def reset(self, reset_solver: bool = True, i_node0=0) -> None:
        if reset_solver:
            self.solver.reset()
        self.first_solve, self.sim_step, self.current_opt_node = True, 0, 0
        self.i_node0 = np.broadcast_to(np.asarray(i_node0, dtype=np.int64), (self.B,)).copy()
        self.v_des, self.w_yaw = np.zeros((self.B, 3)), np.zeros(self.B)
        self.base_ref_vel_tracking = np.zeros((self.B, 12))
        self._tracking_on_device, self._pending_steps = False, 0
        self.status_log: List[np.ndarray] = []

    def set_command(self, v_des=np.zeros(3), w_yaw=0.) -> None:
        self.v_des = np.broadcast_to(np.asarray(v_des, dtype=float), (self.B, 3)).copy()
        self.w_yaw = np.broadcast_to(np.asarray(w_yaw, dtype=float), (self.B,)).copy()

    def _advance_reference(self, n_steps: int) -> None:
        """`increment_base_ref_position` (mpc.py:204-208), one sim step at a time, for every rollout."""
        if n_steps <= 0:
            return
        if self.plan_on_device and self._tracking_on_device:
            self._pending_steps += n_steps           # applied by nmpc_plan_tick at the next tick, on the device
            return
        tr = self.base_ref_vel_tracking
        # Vectorised over the rollouts AND over the sim steps: the yaw of step j is the running sum
        # yaw + w dt + ... + w dt (numpy's cumsum adds in that order, like the loop), the position the running sum
        # of the per-step increments -- the same additions in the same order as n_steps calls of the reference.
        yaw = np.cumsum(np.concatenate((tr[:, 3:4], np.repeat((self.w_yaw * self.sim_dt)[:, None], n_steps, axis=1)),
                                       axis=1), axis=1)                      # [B, n+1]
        # rows x, y of R_WB(roll, pitch, yaw_j) v_des with the reference's operation order (matvec3 / rpy_to_matrix)
        cy, sy = np.cos(yaw[:, :-1]), np.sin(yaw[:, :-1])
        cr, sr, cp, sp = (f(tr[:, i, None]) for i, f in ((5, np.cos), (5, np.sin), (4, np.cos), (4, np.sin)))
        v0, v1, v2 = (self.v_des[:, i, None] for i in range(3))
        vx = np.round(((cy * cp) * v0 + (cy * sp * sr - sy * cr) * v1) + (cy * sp * cr + sy * sr) * v2, 1)
        vy = np.round(((sy * cp) * v0 + (sy * sp * sr + cy * cr) * v1) + (sy * sp * cr - cy * sr) * v2, 1)
        inc = np.stack((vx, vy), axis=-1) * self.sim_dt                              # [B, n, 2]
        pos = np.cumsum(np.concatenate((tr[:, None, :2], inc), axis=1), axis=1)
        tr[:, :2] = pos[:, -1]
        tr[:, 3] = yaw[:, -1]

    def set_convergence_on_first_iter(self) -> None:
        o = self.config_opt
        if self.first_solve:
            self.solver.set_max_iter(N_SQP_FIRST)
            self.solver.set_nlp_tol(o.nlp_tol / 10.)
            self.solver.set_qp_tol(o.qp_tol / 10.)
        elif self.sim_step <= self.replanning_steps:
            self.solver.set_max_iter(o.max_iter)
            self.solver.set_nlp_tol(o.nlp_tol)
            self.solver.set_qp_tol(o.qp_tol)

    def optimize(self, q: np.ndarray, v: np.ndarray):
        """One replanning tick for all rollouts (mpc.py:317-369): contact plan and references at each rollout's
        own node, batched `init` and `solve`."""
        from ..problem import base_ref_vel_tracking
        N1 = self.config_opt.n_nodes + 1
        node = self.i_node0 + self.current_opt_node
        if self.plan_on_device:
            # contact / swing windows, base references and the expansion of `init` on the device (nmpc_plan_tick):
            # 0.3 KB per rollout cross PCIe and the host does nothing per problem
            # the tracked reference pose is resident too: uploaded once, advanced by the kernel for the elapsed sim steps
            tr = None if self._tracking_on_device else self.base_ref_vel_tracking
            self.solver.init_planned(node, q, v, self.v_des, self.w_yaw, tr, self.contact_planner,
                                     self.config_gait.nom_height, self.joint_ref, self.config_gait.step_height,
                                     self._pending_steps, self.sim_dt)
            self._tracking_on_device, self._pending_steps = True, 0
            # the solution stays on the device (plan -> torque reads it there); `self.solver.fetch_solution()` unpacks it
            sol = self.solver.solve(unpack=self.unpack_solution)
            self.status_log.append(np.array(self.solver.status).copy())
            return sol
        cnt = self.contact_planner.get_contacts_batch(node, N1)
        peak = self.contact_planner.get_peaks_batch(node, N1) if self.config_opt.opt_peak else None
        loc = None
        if self.restrict_cnt:
            # the host planners one rollout at a time, as LocomotionMPC.optimize runs them (the checker of the device path)
            _, com, _ = self.solver.handle.kinematics(np.concatenate((q, v), axis=1))
            loc, ref, ref_e = np.zeros((self.B, 4, N1, 3)), np.zeros((self.B, 12)), np.zeros((self.B, 12))
            for b in range(self.B):
                self.contact_planner.set_state(q[b, :3], v[b, :3], q[b, 3:6][::-1], com[b], self.v_des[b], self.w_yaw[b])
                loc[b] = self.contact_planner.get_locations(int(node[b]), N1)
                ref[b], ref_e[b] = LocomotionMPC.compute_base_ref_cnt_restricted(self, q[b], loc[b])
        else:
            ref, ref_e = base_ref_vel_tracking(q, self.v_des, self.w_yaw, self.base_ref_vel_tracking,
                                               self.config_gait.nom_height, self.height_offset, self.config_opt.time_horizon)
        self.solver.init(node, q, v, ref, ref_e, self.joint_ref, self.config_gait.step_height, cnt, loc, peak)
        sol = self.solver.solve()
        self.status_log.append(np.array(self.solver.status).copy())
        return sol

    def closed_loop(self, q: np.ndarray, v: np.ndarray, trajectory_time: float, sim_params=None, record: bool = True):
        """B closed-loop rollouts in lockstep on the batched rigid-body simulation (nmpc_rollout.cuh), the MuJoCo-free
        counterpart of `Rollout_MPC.run` (Behavior_Cloning/utils/Rollout_MPC.py:465) for thousands of rollouts at
        once: every replanning tick one batched `optimize` on the simulated states, between ticks the low-level
        controller of `compute_torques_dof` (feed-forward inverse dynamics on the measured state + PD on the plan,
        mpc.py:503-599) and the physics run on the device, `replanning_steps` sim steps per launch.  The solve is
        taken to be instantaneous (the reference's `delay` is zero, as with `solve_async=False`).
        Returns dict(qv [n_steps,B,36] states seen by the controller, tau [n_steps,B,12], fc [n_steps,B,12] contact
        forces, diverged [B])."""
        from .._lib import default_sim_params
        sp = sim_params or default_sim_params(sim_dt=self.sim_dt, kp=self.Kp, kd=self.Kd, ground_z=self.height_offset)
        h = self.solver.handle
        q, v = np.array(q, dtype=float).reshape(self.B, 18), np.array(v, dtype=float).reshape(self.B, 18)
        h.sim_set_state(np.concatenate((q, v), axis=1))
        R = self.replanning_steps
        n_total = int(np.floor(trajectory_time / self.sim_dt)) + 1
        logs = {k: [] for k in ("qv", "tau", "fc")}
        while self.sim_step < n_total:
            sim_time = self.sim_step * self.sim_dt
            while sim_time >= (self.current_opt_node + 1) * self.dt_nodes:
                self.current_opt_node += 1
            self.set_convergence_on_first_iter()
            self.optimize(q, v)
            self.first_solve = False
            n = min(R, n_total - self.sim_step)
            out = h.sim_steps(sp, self.n_interp_plan, 0, n, record=record)
            if record:
                for k in logs:
                    logs[k].append(np.moveaxis(out[k], 0, 1))
            qv, _ = h.sim_get_state()
            q, v = qv[:, :18].copy(), qv[:, 18:].copy()
            self._advance_reference(n)
            self.sim_step += n
        _, div = h.sim_get_state()
        res = {k: np.concatenate(a, axis=0) for k, a in logs.items()} if record else {}
        res["diverged"], res["q_end"], res["v_end"] = div, q, v
        return res

    def open_loop(self, q: np.ndarray, v: np.ndarray, trajectory_time: float, record_torques: bool = False):
        """Roll all B rollouts forward on their own plans from Euler states q, v [B,18].
        Returns q_traj, v_traj [n_steps,B,18] (and tau [n_steps,B,12] with `record_torques`)."""
        q, v = np.array(q, dtype=float).reshape(self.B, 18), np.array(v, dtype=float).reshape(self.B, 18)
        R = self.replanning_steps
        qs, vs, taus = [], [], []
        sim_time = 0.0
        n_total = 0
        pending = 0                      # sim steps whose reference increment has not been applied yet
        while sim_time <= trajectory_time:
            if sim_time >= (self.current_opt_node + 1) * self.dt_nodes:
                self.current_opt_node += 1
            if self.sim_step % R == 0:
                self._advance_reference(pending)      # only `optimize` reads the tracked pose
                pending = 0
                self.set_convergence_on_first_iter()
                self.optimize(q, v)
                self.first_solve = False
                # the samples up to the next tick, on the device: Hermite interpolation (+ inverse dynamics, PD)
                out = self.solver.handle.plan_torques(self.n_interp_plan, 0, R, self.Kp, self.Kd)
                plan_q, plan_v, plan_tau, k = out["q_plan"], out["v_plan"], out["tau"], 0
            q, v = plan_q[:, k].copy(), plan_v[:, k].copy()
            qs.append(q); vs.append(v)
            if record_torques:
                taus.append(plan_tau[:, k].copy())
            k += 1
            pending += 1
            self.sim_step += 1
            n_total += 1
            sim_time += self.sim_dt
        self._advance_reference(pending)
        res = (np.array(qs), np.array(vs))
        return res + (np.array(taus),) if record_torques else res
