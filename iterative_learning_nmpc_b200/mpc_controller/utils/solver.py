"""`QuadrupedAcadosSolver` -- the OCP solver front-end of the hot path, same public surface as the
reference's (mpc_controller/utils/solver.py:15-431), on top of the CUDA solver instead of acados.

What each method has to do is fixed by the reference (cited per method); how it is done here is
not: the name -> slice bookkeeping lives in `acados_helper.AcadosSolverHelper`, which forwards to
one batch handle of libnmpc_b200.so.  `batch=None` (default) gives the reference's unbatched
shapes on a one-problem handle; `batch=B` gives every array a leading batch dimension and solves
B independent OCPs per `solve()` -- that is the data-collection use case this repo exists for.
"""
from __future__ import annotations

from collections import defaultdict
from typing import List, Optional, Tuple

import numpy as np

from ..config.config_abstract import MPCCostConfig, MPCOptConfig
from .acados_helper import AcadosSolverHelper, ProblemFormulation
from .dynamics import QuadrupedDynamics
from .profiling import print_timings, time_fn


class QuadrupedAcadosSolver(AcadosSolverHelper):
    NAME = "quadruped_solver"

    def __init__(self, path_urdf: Optional[str], feet_frame_names: List[str], config_opt: MPCOptConfig,
                 config_cost: MPCCostConfig, height_offset: float = 0., print_info: bool = False,
                 compute_timings: bool = True, batch: Optional[int] = None, device: int = 0):
        self.feet_frame_names = feet_frame_names
        self.config_opt, self.config_cost = config_opt, config_cost
        self.height_offset = height_offset
        self.print_info = print_info
        self.restrict_cnt: bool = False
        self.dyn = QuadrupedDynamics(path_urdf, feet_frame_names, True, mu_contact=0.8)     # solver.py:34-39

        dt_min, dt_max = config_opt.get_dt_bounds()
        self.dt_nodes = config_opt.get_dt_nodes()
        self.enable_time_opt = config_opt.enable_time_opt
        problem = ProblemFormulation(self.dt_nodes, dt_min, dt_max, self.enable_time_opt)
        self.dyn.setup(problem)
        super().__init__(problem, config_opt.n_nodes, QuadrupedAcadosSolver.NAME, config_cost.reg_eps,
                         config_cost.reg_eps_e, batch=batch, device=device)
        self._opts.mu_contact = self.dyn.mu_contact
        self.reset()
        self.default_normal = np.array([0., 0., 1.])
        self.compute_timings = compute_timings
        self.timings = defaultdict(list)

    # ------------------------------------------------------------------------------------------
    def reset(self):
        """solver.py:66-93."""
        self.last_node = 0
        o = self.config_opt
        self.setup(o.recompile, o.use_cython, o.real_time_it, o.max_qp_iter, o.hpipm_mode)
        self.dyn.bind_device(self.handle.kinematics, self.handle.inverse_dynamics)
        self.timings = defaultdict(list)
        self.set_max_iter(o.max_iter)
        self.set_warm_start_inner_qp(o.warm_start_qp)
        self.set_warm_start_nlp(o.warm_start_nlp)
        self.set_qp_tol(o.qp_tol)
        self.set_nlp_tol(o.nlp_tol)
        self.data = self.get_data_template()
        self.set_cost_weights()
        self.update_cost_weights()
        N, ld = o.n_nodes, self._lead
        self.q_sol_euler = np.zeros(ld + (N + 1, 18))
        self.v_sol_euler = np.zeros(ld + (N + 1, 18))
        self.h_sol = np.zeros(ld + (N + 1, 6))
        self.a_sol = np.zeros(ld + (N, 18))
        self.f_sol = np.zeros(ld + (N, 4, 3))
        self.dt_node_sol = np.zeros(ld + (N,))
        self._has_solution = False
        self._device_warm_started = False

    cold_guess = "zero"      # what the iterate holds before the first solve; the reference pushes zeros

    def _fill_cold_guess(self, q, v):
        if self.cold_guess != "hold":
            return
        d = self.dyn
        self.states[d.q.name][:] = np.asarray(q, dtype=float)[..., None]
        self.states[d.v.name][:] = np.asarray(v, dtype=float)[..., None]
        self.states[d.h.name][:] = np.asarray(self.data["x"][d.h.name], dtype=float)[..., None]
        self.inputs[d.a.name][:] = 0.
        weight = 9.81 * self.handle.total_mass
        # the planner's mask, before setup_initial_feet_pos forces node 0 into stance (what nmpc_cold_guess sees)
        act = np.asarray(self._cnt_mask, dtype=float)[..., :-1]                                   # [..,4,N]
        share = weight / np.maximum(act.sum(axis=-2, keepdims=True), 1.0)
        for i, foot in enumerate(d.feet):
            self.inputs[foot.f.name][:] = 0.
            self.inputs[foot.f.name][..., 2, :] = act[..., i, :] * share[..., 0, :]

    def set_contact_restriction(self, restrict: bool = True):
        self.restrict_cnt = restrict
        self.set_cost_weights()
        self.update_cost_weights()

    def update_cost(self, config_cost: MPCCostConfig):
        self.config_cost = config_cost
        self.set_cost_weights()
        self.update_cost_weights()

    def set_cost_weights(self):
        """Diagonal NLS weights from the cost config (solver.py:108-141)."""
        c, d = self.config_cost, self.dyn
        W, We = self.data["W"], self.data["W_e"]
        W[d.base_cost.name] = np.array(c.W_base)
        W[d.joint_cost.name] = np.array(c.W_joint)
        W[d.acc_cost.name] = np.array(c.W_acc)
        W[d.swing_cost.name] = np.array(c.W_swing)
        We[d.base_cost.name] = np.array(c.W_e_base)
        We[d.joint_cost.name] = np.array(c.W_e_joint)
        We[d.swing_cost.name] = np.array(c.W_swing)
        for i, foot in enumerate(d.feet):
            W[foot.f_reg.name] = np.array(c.W_cnt_f_reg[i])
            w_pos = np.zeros(3)
            if self.restrict_cnt:            # only x, y of the foot target are penalised (:127-129)
                w_pos[:2] = c.W_foot_displacement[0]
            W[foot.pos_cost.name] = w_pos
            We[foot.pos_cost.name] = w_pos.copy()
        if self.enable_time_opt:                 # solver.py:124-125
            W["dt"] = np.array(c.time_opt)
        self.set_cost_weight_constant(W)
        self.set_cost_weight_terminal(We)

    @staticmethod
    def _repeat_last(arr: np.ndarray, n_repeat: int, axis=-1):
        tail = np.repeat(np.take(arr, [-1], axis=axis), n_repeat, axis=axis)
        return np.concatenate([arr, tail], axis=axis)

    # -- per-tick packing (solver.py:153-275) ---------------------------------------------------
    def _col(self, a, dim: int) -> np.ndarray:
        """[dim] or [B,dim] -> column(s) broadcastable over the node axis."""
        return np.asarray(a, dtype=float).reshape(self._lead + (dim, 1))

    def setup_reference(self, base_ref: np.ndarray, base_ref_e: Optional[np.ndarray], joint_ref: np.ndarray,
                        step_height):
        """solver.py:153-177."""
        d, ld = self.dyn, self._lead
        if base_ref_e is None:
            base_ref_e = np.array(base_ref, copy=True)
        joint_ref = np.broadcast_to(np.asarray(joint_ref, dtype=float), ld + (12,))
        joint_ref_vel = np.concatenate((joint_ref, np.zeros_like(joint_ref)), axis=-1)
        sh = np.broadcast_to(np.asarray(step_height, dtype=float), ld) if ld else float(step_height)
        self.cost_ref[d.base_cost.name][:] = self._col(base_ref, 12)
        self.cost_ref[d.swing_cost.name][:] = np.reshape(sh, ld + (1, 1))
        self.cost_ref_terminal[d.base_cost.name] = np.asarray(base_ref_e, dtype=float).reshape(ld + (12,))
        self.cost_ref_terminal[d.swing_cost.name][:] = np.reshape(sh, ld + (1,))
        self.cost_ref[d.joint_cost.name] = self._col(joint_ref_vel, 24)
        self.cost_ref_terminal[d.joint_cost.name] = joint_ref_vel.copy()
        if self.enable_time_opt:                 # solver.py:166-167 (the kernel's reference for dt is dt_nodes)
            self.cost_ref["dt"][:] = self.dt_nodes

    def setup_initial_state(self, q_euler: np.ndarray, v_global: np.ndarray, h: Optional[np.ndarray] = None):
        """solver.py:179-192; h defaults to the momentum of the last `dyn.update_pin` (unbatched) or
        is evaluated on the device for the whole batch."""
        d = self.dyn
        q_euler, v_global = np.asarray(q_euler, dtype=float), np.asarray(v_global, dtype=float)
        if h is None:
            if self._lead:
                _, _, h = self.handle.kinematics(np.concatenate((q_euler, v_global), axis=-1))
            else:
                h = d.pin_data.hg.np
        self.data["x"][d.q.name] = q_euler
        self.data["x"][d.v.name] = v_global
        self.data["x"][d.h.name] = np.asarray(h, dtype=float)
        if self.enable_time_opt and not self._has_solution:      # solver.py:189-190
            self.inputs["dt"][:] = self.dt_nodes
        self.set_initial_state(self.data["x"])

    def init_contacts_parameters(self):
        """solver.py:212-225."""
        for i, foot in enumerate(self.dyn.feet):
            self.params[foot.active.name][:] = 1.
            self.params[foot.plane_normal.name][:] = self.default_normal[:, None]
            self.params[foot.plane_point.name][:] = 0.
            self.params[foot.plane_point.name][..., -1, :] = self.height_offset
            self.params[foot.p_gain.name][:] = self.config_cost.W_foot_pos_constr_stab[i]
            self.params[foot.restrict.name][:] = 0.
            self.params[foot.range_radius.name][:] = self.config_cost.cnt_radius if self.restrict_cnt else 1.0e10

    def setup_cnt_status(self, cnt_sequence: np.ndarray, peak_plan: Optional[np.ndarray] = None):
        """cnt_sequence [4,N+1] (or [B,4,N+1]) of 0/1 (solver.py:227-252)."""
        cnt_sequence = np.asarray(cnt_sequence)
        n1 = self.config_opt.n_nodes + 1
        assert cnt_sequence.shape[-1] == n1, \
            f"Invalid contact plan shape. Wrong number of optimization nodes. ({cnt_sequence.shape[-1]} vs {n1})"
        assert cnt_sequence.shape[-2] == len(self.feet_frame_names), \
            "Invalid contact plan shape. Wrong number of end effectors."
        self._cnt_mask = cnt_sequence.copy()
        for i, foot in enumerate(self.dyn.feet):
            seq = cnt_sequence[..., i, :]
            self.params[foot.active.name][..., 0, :] = seq
            if self.config_opt.opt_peak and peak_plan is not None:
                self.params[foot.peak.name][..., 0, :] = np.asarray(peak_plan)[..., i, :]
            if self.restrict_cnt:
                made = np.diff(seq.astype(np.int64), axis=-1, prepend=seq[..., :1].astype(np.int64))
                made[made == -1] = 0
                self.params[foot.restrict.name][..., 0, :] = made

    @time_fn("setup_contact_plan")
    def setup_contact_loc(self, contact_loc_plan: np.ndarray):
        """contact_loc_plan [4,N+1,3] (or [B,4,N+1,3]) (solver.py:254-275)."""
        loc = np.asarray(contact_loc_plan, dtype=float)
        assert loc.shape[-2] == self.config_opt.n_nodes + 1, "Invalid contact plan shape. Wrong number of optimization nodes."
        assert loc.shape[-3] == len(self.feet_frame_names), "Invalid contact plan shape. Wrong number of end effectors."
        assert loc.shape[-1] == 3, "Invalid contact plan shape. 3D points required."
        for i, foot in enumerate(self.dyn.feet):
            pts = np.swapaxes(loc[..., i, :, :], -1, -2)            # [.., 3, N+1]
            self.params[foot.plane_point.name] = pts.copy()
            self.cost_ref[foot.pos_cost.name] = pts[..., 1:].copy()
            self.cost_ref_terminal[foot.pos_cost.name] = pts[..., -1].copy()

    def setup_initial_feet_pos(self, i_node=0):
        """Stance feet at node 0 keep their current (FK) position until they next swing
        (solver.py:194-210).  Unbatched: on the host dicts like the reference; batched: the device
        does it for every problem after the parameters are uploaded (nmpc_pin_stance_feet)."""
        if self._lead:
            self._pin_i_node = np.broadcast_to(np.asarray(i_node, dtype=np.int32), self._lead).copy()
            return
        feet_pos = self.dyn.get_feet_position_w()
        for foot, pos in zip(self.dyn.feet, feet_pos):
            act = self.params[foot.active.name]
            if i_node == 0:
                act[0, 0] = 1
            if act[0, 0]:
                next_swing = int(np.argmin(act[0, :]))
                self.params[foot.plane_point.name][:, :next_swing] = pos[:, None]
                if self.print_info:
                    print(f"Reset foot contact {foot.frame_name}, height {pos[2]}")

    def print_contact_constraints(self):
        print("\n" + "-" * 10, "Contacts", "-" * 10)
        for foot in self.dyn.feet:
            print("\n--- ", foot.frame_name, "contact")
            print(np.int8(self.params[foot.active.name]))
            print(np.unique(self.params[foot.plane_point.name], axis=-1).T)
            print(np.int8(self.params[foot.peak.name]))
            if self.restrict_cnt:
                print("Restriction")
                print(np.int8(self.params[foot.restrict.name]))

    # -- warm start (solver.py:290-342) -----------------------------------------------------------
    @time_fn("warm_start_solver")
    def warm_start_solver(self, i_node, repeat_last: bool = False):
        """Shift the previous solution by `start_node = i_node - last_node` nodes.

        Unbatched: the host arrays are shifted the way the reference does it and pushed by
        `update_solver`; the multipliers are shifted on the device.  Batched: `i_node` may differ
        per problem, so primal and dual are shifted together on the device (one kernel,
        nmpc_shift_warm_start) and the host copies of states / inputs are not pushed.
        """
        N, d = self.config_opt.n_nodes, self.dyn
        if self._lead:
            i_node = np.broadcast_to(np.asarray(i_node, dtype=np.int64), self._lead)
            start = (i_node - self.last_node).astype(np.int32)
            self.warm_start_multipliers(start, None, repeat_last=repeat_last)
            self._device_warm_started = True
            self.last_node = i_node.copy()
            return
        start_node = int(i_node - self.last_node)
        n_ws = N - start_node
        if self.print_info:
            print(f"Warm start size: {n_ws}, start_node {start_node}")
        S, U = self.states, self.inputs
        S[d.q.name][:, 1:n_ws + 1] = self.q_sol_euler[start_node + 1:].T
        S[d.v.name][:, 1:n_ws + 1] = self.v_sol_euler[start_node + 1:].T
        S[d.h.name][:, 1:n_ws + 1] = self.h_sol[start_node + 1:].T
        U[d.a.name][:, :n_ws] = self.a_sol[start_node:].T
        for i, foot in enumerate(d.feet):
            f = U[foot.f.name]
            f[:, :n_ws] = self.f_sol[start_node:, i, :].T
            f[:, n_ws:] = self.f_sol[-1, i][:, None] if repeat_last else 0.
        if self.enable_time_opt:                 # solver.py:324-336: shifted node durations, nominal on the tail
            U["dt"][:, :n_ws] = self.dt_node_sol[start_node:]
            U["dt"][:, n_ws:] = self.dt_nodes
        if repeat_last and n_ws < N:
            S[d.q.name][:, n_ws:] = self.q_sol_euler[-1][:, None]
            S[d.v.name][:, n_ws:] = self.v_sol_euler[-1][:, None]
            S[d.h.name][:, n_ws:] = self.h_sol[-1][:, None]
            U[d.a.name][:, n_ws:] = self.a_sol[-1][:, None]
        self.warm_start_multipliers(start_node, n_ws, repeat_last=repeat_last)
        self.last_node = i_node

    @time_fn("update_solver")
    def update_solver(self) -> None:
        """Push everything to the device (solver.py:344-353)."""
        if not getattr(self, "_device_warm_started", False):
            self.update_states()
            self.update_inputs()
        self._device_warm_started = False
        self.update_parameters()
        self.update_ref()
        self.set_ref_terminal()
        if self._lead:                        # batched setup_initial_feet_pos, after params and x0 are resident
            self.handle.pin_stance_feet(self._pin_i_node)

    @time_fn("init_solver")
    def init(self, i_node, q: np.ndarray, v: np.ndarray, base_ref: np.ndarray, base_ref_e: np.ndarray,
             joint_ref: np.ndarray, step_height, cnt_sequence: np.ndarray,
             cnt_locations: Optional[np.ndarray] = None, swing_peak: Optional[np.ndarray] = None):
        """Set the solver up for the current state and optimisation node (solver.py:356-394)."""
        self.setup_reference(base_ref, base_ref_e, joint_ref, step_height)
        self.setup_initial_state(q, v)
        self.init_contacts_parameters()
        self.setup_cnt_status(cnt_sequence, swing_peak)
        if self.restrict_cnt:
            assert cnt_locations is not None, "Contact plan not provided"
            self.setup_contact_loc(cnt_locations)
        self.setup_initial_feet_pos(i_node)
        if self.config_opt.warm_start_sol:
            if self._lead:
                # batched: problems may sit at different nodes; shift once a previous solution exists,
                # otherwise start from the cold guess (see BatchedQuadrupedSolver.cold_guess)
                if self._has_solution:
                    self.warm_start_solver(i_node, repeat_last=False)
                else:
                    self._fill_cold_guess(q, v)
                    self.last_node = np.broadcast_to(np.asarray(i_node, dtype=np.int64), self._lead).copy()
            elif i_node > 0:
                self.warm_start_solver(i_node, repeat_last=False)
            elif not self._has_solution:
                self._fill_cold_guess(q, v)          # no-op with the reference's "zero" first guess
        self.update_solver()
        if self.print_info:
            self.print_contact_constraints()

    @time_fn("solve")
    def solve(self, unpack: bool = True) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        """Solve and unpack (solver.py:397-429): q [N+1,18], v [N+1,18], a [N,18], f [N,4,3], dt [N]
        (each with the leading batch dimension when batched).  `unpack=False` leaves the solution on the device
        (the batched plan -> torque step reads it there); `fetch_solution()` unpacks it later."""
        super().solve(print_stats=self.print_info, print_time=self.print_info)
        self._has_solution = True
        if not unpack:
            return None
        return self.fetch_solution()

    def fetch_solution(self):
        self.parse_sol()
        d = self.dyn
        T = lambda a: np.swapaxes(a, -1, -2)
        self.q_sol_euler[:] = T(self.states[d.q.name])
        self.v_sol_euler[:] = T(self.states[d.v.name])
        self.h_sol[:] = T(self.states[d.h.name])
        self.a_sol[:] = T(self.inputs[d.a.name])
        self.f_sol[:] = np.stack([T(self.inputs[foot.f.name]) for foot in d.feet], axis=-2)
        if self.enable_time_opt:                 # solver.py:423-427
            self.dt_node_sol[:] = self.inputs["dt"].reshape(self.dt_node_sol.shape)
        else:
            self.dt_node_sol[:] = self.dt_nodes
        self._has_solution = True
        return self.q_sol_euler, self.v_sol_euler, self.a_sol, self.f_sol, self.dt_node_sol

    def print_timings(self):
        print_timings(self.timings)


class BatchedQuadrupedSolver(QuadrupedAcadosSolver):
    """B independent OCPs behind the same front-end: every argument of `init` gains a leading batch
    dimension (`i_node [B]`, `q [B,18]`, `cnt_sequence [B,4,N+1]` ...), `solve` returns
    `q [B,N+1,18]` etc. and `status [B]` is in `self.status`.  Before the first solve the iterate
    holds x0 at every node with the weight shared by the stance feet (`cold_guess = "hold"`); set
    `cold_guess = "zero"` for the reference's all-zero first guess."""

    cold_guess = "hold"

    def __init__(self, batch: int, path_urdf: Optional[str], feet_frame_names: List[str], config_opt: MPCOptConfig,
                 config_cost: MPCCostConfig, height_offset: float = 0., print_info: bool = False,
                 compute_timings: bool = True, device: int = 0):
        super().__init__(path_urdf, feet_frame_names, config_opt, config_cost, height_offset, print_info,
                         compute_timings, batch=int(batch), device=device)

    def init_planned(self, i_node, q: np.ndarray, v: np.ndarray, v_des: np.ndarray, w_yaw: np.ndarray,
                     tracking: Optional[np.ndarray], planner, nom_height: float, joint_ref: np.ndarray,
                     step_height: float, advance_steps: int = 0, sim_dt: float = 1.0e-3):
        """`init` with the planning step on the device (nmpc_plan_tick; nmpc_plan_tick_raibert when the contact
        restriction is on and `planner` is a RaiberContactPlanner): the contact /
        swing windows at every problem's node, the base references (mpc.py:210-272) and the per-node expansion
        (solver.py:153-275) are computed by the GPU from i_node, q, the command and the tracked pose; x0 with its
        momentum, the shift warm start (or the first cold guess) and the stance-foot pinning follow, all resident."""
        i_node = np.broadcast_to(np.asarray(i_node, dtype=np.int64), self._lead).copy()
        gait_seq, peak_seq = planner.cycle_tables()
        h = self.handle
        qv = np.concatenate((np.asarray(q, dtype=float), np.asarray(v, dtype=float)), axis=-1)
        if self.restrict_cnt:
            # contact_planner="raibert" (mpc.py:76-90, 332-341): footholds and the contact-restricted references too
            if not hasattr(planner, "offset_hip_b") or getattr(planner, "cache_cnt", False):
                raise NotImplementedError("device planning with contact restriction: RaiberContactPlanner, cache_cnt=False")
            N1 = self.config_opt.n_nodes + 1
            make_seq = (planner.switch_cnt == 1).astype(np.int8)
            t_cnt = np.array([round(j * planner.dt_nodes, 3) for j in range(N1)])
            self.last_cnt_locations = h.plan_tick_raibert(
                i_node, gait_seq, peak_seq, make_seq, qv, v_des, w_yaw, t_cnt, planner.offset_hip_b,
                planner.config_gait.stance_ratio, planner.config_gait.nominal_period, planner.foot_size,
                planner.height_offset, self.config_cost.cnt_radius, nom_height, self.height_offset, joint_ref,
                step_height, self.config_cost.W_foot_pos_constr_stab, self.config_opt.opt_peak, planner.GRAVITY,
                want_locations=getattr(self, "want_cnt_locations", False))
        else:
            h.plan_tick(i_node, gait_seq, peak_seq, q, v_des, w_yaw, tracking, nom_height, self.height_offset,
                        self.config_opt.time_horizon, joint_ref, step_height, self.config_cost.W_foot_pos_constr_stab,
                        self.config_opt.opt_peak, advance_steps, sim_dt)
        h.set_x0(qv)
        if self.config_opt.warm_start_sol and self._has_solution:
            h.shift_warm_start((i_node - self.last_node).astype(np.int32), False)
        else:
            h.cold_guess()
        h.pin_stance_feet(i_node.astype(np.int32))
        self.last_node = i_node

    def plan_torques(self, sim_dt: float = 1.0e-3, step0: int = 0, n_steps: Optional[int] = None, Kp: float = 20.,
                     Kd: float = 1.75, qv_meas: Optional[np.ndarray] = None) -> dict:
        """"Get torques" for the whole batch, on the device, from the solution of the last `solve`:
        what `LocomotionMPC._adopt_plan` + `_compute_torques_ff` + `_compute_pd_torques` do per sim step on
        the host (mpc.py:371-414, 535-599) and the DAgger label `action = (tau + Kd v)/Kp + q`
        (Behavior_Cloning/utils/Rollout_MPC.py:298).  Plan steps `step0 .. step0+n_steps-1` of the
        `round(time_horizon / sim_dt)`-sample plan; `qv_meas [B,n_steps,36]` are measured Euler states
        (None: evaluate on the plan itself).  Returns q_plan, v_plan [B,n,18], tau_ff, tau, action [B,n,12]."""
        n_interp = round(self.config_opt.time_horizon / sim_dt)
        return self.handle.plan_torques(n_interp, step0, n_steps, Kp, Kd, qv_meas)

