"""TEST INFRASTRUCTURE: an adapter that lets the reference-facing front-end
(`QuadrupedAcadosSolver` / `LocomotionMPC`) run on the CPU oracle, so that the host logic can be
tested without a GPU and so that GPU runs of the front-end have a checker.  Never imported by the
package; the product front-end only opens the CUDA library."""
import ctypes as C

import numpy as np

from oracle import oracle as O


def _to_oracle_opts(opts) -> "O.NmpcOpts":
    o = O.NmpcOpts()
    assert C.sizeof(o) == C.sizeof(opts)
    C.memmove(C.byref(o), C.byref(opts), C.sizeof(o))
    return o


class OracleHandle:
    """Method-for-method stand-in for `core.NmpcSolver` on top of `oracle.OracleSolver`."""

    total_mass = None

    def __init__(self, batch, opts, device=0):
        self.B = int(batch)
        self.opts = opts.copy()
        self.N = int(opts.n_nodes)
        self._s = O.OracleSolver(self.B, _to_oracle_opts(opts))
        dims = np.zeros(8, dtype=np.int32)
        self.total_mass = 15.017 if OracleHandle.total_mass is None else OracleHandle.total_mass
        self.launch_count = 0

    def set_opts(self, opts=None, **kw):
        o = (opts or self.opts).copy()
        for k, v in kw.items():
            setattr(o, k, v)
        self.opts = o
        self._s.set_opts(_to_oracle_opts(o))

    def solve(self, status=None, sync=True):
        return self._s.solve(1)

    def get_solution(self, duals=True):
        x, u, pi, lam = self._s.get_solution()
        if self._s.topt:          # product layout: 30 inputs, 42 constraint rows; dt and its bounds through get_dt
            u = np.ascontiguousarray(u[..., :30])
            lam = np.concatenate([lam[..., :42], lam[..., 43:85]], axis=-1)
        return x, u, pi, lam

    def get_dt(self, duals=False):
        if not self._s.topt:
            dt = np.full((self.B, self.N), float(self.opts.dt_nodes))
            return (dt, np.zeros((self.B, self.N, 2))) if duals else dt
        _, u, _, lam = self._s.get_solution()
        dt = np.ascontiguousarray(u[..., 30])
        return (dt, np.stack([lam[:, :-1, 42], lam[:, :-1, 85]], axis=-1)) if duals else dt

    def set_dt(self, dt=None):
        if not self._s.topt:
            return
        _, u, _, _ = self._s.get_solution()
        u[..., 30] = float(self.opts.dt_nodes) if dt is None else np.asarray(dt, float).reshape(self.B, self.N)
        self._s.set_guess(None, u)

    def set_guess(self, x=None, u=None):
        if u is not None and self._s.topt:
            u = np.asarray(u, float).reshape(self.B, self.N, -1)
            if u.shape[-1] == 30:         # a new 30-wide input guess restarts the node durations at dt_nodes
                u = np.concatenate([u, np.full((self.B, self.N, 1), float(self.opts.dt_nodes))], axis=-1)
        self._s.set_guess(x, u)

    def kinematics(self, qv):
        return O.kinematics(qv)

    def inverse_dynamics(self, qva, f):
        return O.inverse_dynamics(qva, f)

    def plan_torques(self, n_interp, step0=0, n_steps=None, kp=20.0, kd=1.75, qv_meas=None):
        x, u, _, _ = self.get_solution()
        dt = self.get_dt() if self._s.topt else float(self.opts.dt_nodes)
        return O.plan_torques(x, u, dt, n_interp, step0, n_steps, kp, kd, qv_meas)

    def plan_tick(self, i_node, gait_seq, peak_seq, q, v_des, w_yaw, tracking, nom_height, height_offset, t_horizon,
                  joint_ref, step_height, p_gain, opt_peak=True, advance_steps=0, sim_dt=1.0e-3):
        """Checker for nmpc_plan_tick: the host planners (problem.py restatements of contact_planner.py:121-149 and
        mpc.py:210-272), the reference's per-step increment of the tracked pose (mpc.py:204-208) and the host packing
        of `init`."""
        from iterative_learning_nmpc_b200 import problem as P
        from iterative_learning_nmpc_b200.mpc_controller.config.quadruped.utils import get_quadruped_config
        from iterative_learning_nmpc_b200.mpc_controller.utils.transform import rpy_to_matrix
        B, N1 = self.B, self.N + 1
        if tracking is not None:
            self._tracking = np.array(tracking, dtype=float).reshape(B, 12)
        v_des = np.broadcast_to(np.asarray(v_des, dtype=float), (B, 3))
        w_yaw = np.broadcast_to(np.asarray(w_yaw, dtype=float), (B,))
        for _ in range(int(advance_steps)):          # increment_base_ref_position, step by step, problem by problem
            for b in range(B):
                R_WB = rpy_to_matrix(self._tracking[b, 3:6][::-1])
                vg = np.round(R_WB @ v_des[b], 1)
                self._tracking[b, :2] += vg[:2] * sim_dt
                self._tracking[b, 3] += w_yaw[b] * sim_dt
        tracking = self._tracking
        gs, ps = np.asarray(gait_seq), np.asarray(peak_seq)
        npc = gs.shape[1]
        node = np.asarray(i_node, dtype=np.int64).reshape(B)
        idx = (node[:, None] % npc + np.arange(N1)[None, :]) % npc
        cnt = np.moveaxis(gs[:, idx], 0, 1)
        peak = np.moveaxis(ps[:, idx], 0, 1)
        ref, ref_e = P.base_ref_vel_tracking(q, v_des, w_yaw, tracking, nom_height, height_offset, t_horizon)
        _, _, cost = get_quadruped_config("trot", "go2")
        cost.W_foot_pos_constr_stab = np.asarray(p_gain, dtype=float)
        yref, yref_e = P.pack_reference(ref, ref_e, joint_ref, step_height, self.N)
        self._s.set_yref(yref, yref_e)
        self.set_params(P.pack_params(cnt, cost, height_offset, peak, False, None, opt_peak))
        self.last_plan = dict(base_ref=ref, base_ref_e=ref_e, cnt=cnt, peak=peak)

    def plan_tick_raibert(self, i_node, gait_seq, peak_seq, make_seq, qv, v_des, w_yaw, t_cnt, offset_hip_b,
                          stance_ratio, nominal_period, foot_size, planner_height_offset, cnt_radius, nom_height,
                          height_offset, joint_ref, step_height, p_gain, opt_peak=True, gravity=9.81,
                          want_locations=False):
        """Checker for nmpc_plan_tick_raibert: Raibert footholds (contact_planner.py:265-322, cache_cnt=False) and the
        contact-restricted references (mpc.py:274-315) restated with plain loops from the tables the device receives,
        then the host packing of `init`."""
        from iterative_learning_nmpc_b200 import problem as P
        from iterative_learning_nmpc_b200.mpc_controller.config.quadruped.utils import get_quadruped_config
        B, N1 = self.B, self.N + 1
        gs, ps, ms = np.asarray(gait_seq), np.asarray(peak_seq), np.asarray(make_seq)
        npc = gs.shape[1]
        node = np.asarray(i_node, dtype=np.int64).reshape(B)
        qv = np.asarray(qv, dtype=float).reshape(B, 36)
        v_des = np.broadcast_to(np.asarray(v_des, dtype=float), (B, 3))
        w_yaw = np.broadcast_to(np.asarray(w_yaw, dtype=float), (B,))
        off = np.asarray(offset_hip_b, dtype=float).reshape(4, 3)
        _, com, _ = O.kinematics(qv)
        idx = (node[:, None] % npc + np.arange(N1)[None, :]) % npc
        cnt, peak, make = (np.moveaxis(t[:, idx], 0, 1) for t in (gs, ps, ms))
        loc, ref, ref_e = np.zeros((B, 4, N1, 3)), np.zeros((B, 12)), np.zeros((B, 12))
        for b in range(B):
            yaw, vt = qv[b, 3], v_des[b, :2]
            Rz = np.array([[np.cos(yaw), -np.sin(yaw)], [np.sin(yaw), np.cos(yaw)]])
            k = 0.5 * np.sqrt((com[b, 2] - planner_height_offset) / gravity)
            ang = np.array([k * vt[1] * w_yaw[b], -k * vt[0] * w_yaw[b]])
            for f in range(4):
                for j in range(N1):
                    if make[b, f, j] == 1 and t_cnt[j] >= 0:
                        hip = com[b, :2] + Rz @ off[f, :2] + vt * t_cnt[j] * (1 + stance_ratio[f])
                        xy = hip + 0.1 * (vt - qv[b, 18:20]) + 0.5 * vt * (nominal_period * stance_ratio[f]) + ang
                        loc[b, f, j:] = (xy[0], xy[1], foot_size)
            full = [(int((loc[b, :, j] != 0).all(-1).sum()), tuple(loc[b, :, j].reshape(-1))) for j in range(N1)]
            best = max(c for c, _ in full)
            if best > 0:
                slabs = sorted(sl for c, sl in full if c == best)
                c_first, c_last = (np.array(sl).reshape(4, 3).mean(0) for sl in (slabs[0], slabs[-1]))
            else:
                c_first, c_last = loc[b, :, 0].mean(0), loc[b, :, -1].mean(0)
            ref[b, :2] = 0.35 * c_first[:2] + 0.65 * c_last[:2]
            ref_e[b, :2] = c_last[:2]
            ref[b, 2] = ref_e[b, 2] = nom_height + height_offset
        _, _, cost = get_quadruped_config("trot", "go2")
        cost.W_foot_pos_constr_stab = np.asarray(p_gain, dtype=float)
        cost.cnt_radius = float(cnt_radius)
        yref, yref_e = P.pack_reference(ref, ref_e, joint_ref, step_height, self.N, loc)
        self._s.set_yref(yref, yref_e)
        self.set_params(P.pack_params(cnt, cost, height_offset, peak, True, loc, opt_peak))
        self.last_plan = dict(base_ref=ref, base_ref_e=ref_e, cnt=cnt, peak=peak, loc=loc)
        return loc if want_locations else None

    def get_tracking(self):
        return self._tracking.copy()

    def set_params(self, p):
        self._p = np.array(p, dtype=float).reshape(self.B, self.N + 1, 4, 11)
        self._s.set_params(p)

    def set_x0(self, qv):
        qv = np.asarray(qv, dtype=float).reshape(self.B, 36)
        self._x0 = np.concatenate([qv, O.kinematics(qv)[2]], axis=1)
        self._s.set_x0(qv)

    def cold_guess(self):
        """nmpc_cold_guess_kernel on the host: x0 held, the weight shared by the stance feet of every node (from the
        planner's mask, before the stance feet of node 0 are pinned)."""
        B, N = self.B, self.N
        x = np.repeat(self._x0[:, None, :], N + 1, axis=1)
        u = np.zeros((B, N, 4 * 3 + 18))
        act = self._p[:, :N, :, 0]
        n_act = act.sum(axis=2, keepdims=True)
        fz = act * (O.model_mass() * 9.81) / np.where(n_act > 1.0, n_act, 1.0)
        u.reshape(B, N, 30)[:, :, 18:].reshape(B, N, 4, 3)[..., 2] = fz
        self.set_guess(x, u)

    def set_profiling(self, on=True):
        pass

    def kernel_times(self):
        return {}

    def synchronize(self):
        pass

    def __getattr__(self, name):          # set_x0, set_yref, set_weights, set_params, set_guess, ...
        return getattr(self._s, name)


def use_oracle_backend(monkeypatch):
    from iterative_learning_nmpc_b200.mpc_controller.utils import acados_helper as H
    monkeypatch.setattr(H.AcadosSolverHelper, "_make_handle",
                        lambda self: OracleHandle(self.batch, self._opts, self.device))
