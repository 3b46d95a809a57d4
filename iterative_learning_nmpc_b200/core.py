"""`NmpcSolver`: thin Python object over the C ABI handle (one batch of B OCPs on one GPU)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib as L
from ._lib import NG, NP, NU, NX, NY, NYE, NZ, NmpcOpts, NmpcStats, check, default_opts, ptr


def _f64(a, shape):
    if hasattr(a, "data_ptr"):           # torch tensor (host or device): must already be fp64
        import torch
        assert a.dtype == torch.float64, "fp64 tensors required"
        return a.contiguous().reshape(shape)
    return np.ascontiguousarray(a, dtype=np.float64).reshape(shape)


def _i32(a, shape):
    if hasattr(a, "data_ptr"):
        import torch
        assert a.dtype == torch.int32
        return a.contiguous().reshape(shape)
    return np.ascontiguousarray(np.broadcast_to(np.asarray(a), shape), dtype=np.int32)


def _order_after_torch(solver, *arrays):
    """Stream contract of the C ABI (include/nmpc_b200.h): the handle works on its own non-blocking stream, so a
    DEVICE buffer handed in must be complete before the call.  For CUDA tensors this is arranged here: the work queued
    on torch's current stream is awaited once per call (a host-side wait on an event; outputs need nothing, every
    getter synchronises the handle's stream before it returns)."""
    for a in arrays:
        if a is not None and hasattr(a, "is_cuda") and a.is_cuda:
            import torch
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(a.device))
            ev.synchronize()
            return


class NmpcSolver:
    """Batch of B quadruped OCPs resident on one B200.  Methods mirror include/nmpc_b200.h; array
    arguments may be numpy arrays (host) or fp64 torch tensors (host or CUDA).  A solver object may be shared by the
    two threads of `LocomotionMPC` (solve worker + control loop): the library serialises the solve path and the
    utility path (kinematics / inverse dynamics) with separate locks, streams and staging buffers."""

    def __init__(self, batch: int, opts: Optional[NmpcOpts] = None, device: int = 0):
        self.lib = L.load_library()
        self.opts = (opts or default_opts()).copy()
        self.B, self.N, self.device = int(batch), int(self.opts.n_nodes), int(device)
        err = C.c_int32(0)
        self._h = C.c_void_p(self.lib.nmpc_create(C.byref(self.opts), self.B, self.device, C.byref(err)))
        if not self._h:
            check(err.value or -1, "nmpc_create")

    def close(self):
        if getattr(self, "_h", None):
            self.lib.nmpc_destroy(self._h)
            self._h = None

    __del__ = close

    # -- options ------------------------------------------------------------------------
    def set_opts(self, opts: Optional[NmpcOpts] = None, **kw):
        o = (opts or self.opts).copy()
        for k, v in kw.items():
            if not hasattr(o, k):
                raise AttributeError(k)
            setattr(o, k, v)
        check(self.lib.nmpc_set_opts(self._h, C.byref(o)), "nmpc_set_opts")
        self.opts = o

    # -- problem data -------------------------------------------------------------------
    def set_x0(self, qv):
        a = _f64(qv, (self.B, 36))
        _order_after_torch(self, a)
        check(self.lib.nmpc_set_x0(self._h, ptr(a)), "nmpc_set_x0")
        self.synchronize_if_host(a)

    def set_x0_full(self, x0):
        a = _f64(x0, (self.B, NX))
        _order_after_torch(self, a)
        check(self.lib.nmpc_set_x0_full(self._h, ptr(a)), "nmpc_set_x0_full")
        self.synchronize_if_host(a)

    def set_yref(self, yref, yref_e):
        a = _f64(yref, (self.B, self.N, NY))
        b = _f64(yref_e, (self.B, NYE))
        _order_after_torch(self, a, b)
        check(self.lib.nmpc_set_yref(self._h, ptr(a), ptr(b)), "nmpc_set_yref")
        self.synchronize_if_host(a)

    def set_weights(self, W, W_e):
        W = np.asarray(W, dtype=np.float64) if not hasattr(W, "data_ptr") else W
        per = int(W.ndim == 2)
        a = _f64(W, (self.B, NY) if per else (NY,))
        b = _f64(W_e, (self.B, NYE) if per else (NYE,))
        _order_after_torch(self, a, b)
        check(self.lib.nmpc_set_weights(self._h, ptr(a), ptr(b), per), "nmpc_set_weights")
        self.synchronize_if_host(a)

    def set_params(self, p):
        a = _f64(p, (self.B, self.N + 1, NP))
        _order_after_torch(self, a)
        check(self.lib.nmpc_set_params(self._h, ptr(a)), "nmpc_set_params")
        self.synchronize_if_host(a)

    def init_tick(self, base_ref, base_ref_e=None, joint_ref=None, step_height=0.0, cnt_sequence=None,
                  swing_peak=None, cnt_locations=None, p_gain=None, cnt_radius: float = 0.0,
                  height_offset: float = 0.0, restrict_cnt: bool = False):
        """`QuadrupedAcadosSolver.init` with the expansion on the device (nmpc_init_tick): base_ref(_e) [B,12],
        joint_ref [B,12] or [12], step_height [B] or scalar, cnt_sequence / swing_peak int8 [B,4,N+1],
        cnt_locations [B,4,N+1,3] or None, p_gain [4]."""
        B, N1 = self.B, self.N + 1
        ref = _f64(base_ref, (B, 12))
        ref_e = None if base_ref_e is None else _f64(base_ref_e, (B, 12))
        jr = np.ascontiguousarray(joint_ref, dtype=np.float64)
        jpp = int(jr.size == B * 12 and B > 1 or jr.ndim == 2)
        jr = jr.reshape((B, 12) if jpp else (12,))
        sh = np.ascontiguousarray(np.atleast_1d(step_height), dtype=np.float64)
        hpp = int(sh.size == B and (B > 1 or np.ndim(step_height) > 0))
        sh = sh.reshape((B,) if hpp else (1,))
        if restrict_cnt and cnt_locations is None:
            raise ValueError("Contact plan not provided")
        cnt = np.ascontiguousarray(cnt_sequence, dtype=np.int8).reshape(B, 4, N1)
        peak = None if swing_peak is None else np.ascontiguousarray(swing_peak, dtype=np.int8).reshape(B, 4, N1)
        loc = None if cnt_locations is None else _f64(cnt_locations, (B, 4, N1, 3))
        gain = np.ascontiguousarray(p_gain, dtype=np.float64).reshape(4)
        _order_after_torch(self, ref, ref_e, loc)
        check(self.lib.nmpc_init_tick(self._h, ptr(ref), ptr(ref_e), ptr(jr), jpp, ptr(sh), hpp, ptr(cnt), ptr(peak),
                                      ptr(loc), ptr(gain), float(cnt_radius), float(height_offset), int(bool(restrict_cnt))),
              "nmpc_init_tick")

    def plan_tick(self, i_node, gait_seq, peak_seq, q, v_des, w_yaw, tracking, nom_height: float,
                  height_offset: float, t_horizon: float, joint_ref, step_height: float, p_gain, opt_peak: bool = True,
                  advance_steps: int = 0, sim_dt: float = 1.0e-3):
        """Planning + `init` on the device for the cyclic contact planner (nmpc_plan_tick): i_node [B], the gait
        cycle tables int8 [4, nodes_per_cycle], q [B,18], v_des [B,3], w_yaw [B]; tracking [B,12] replaces the
        resident reference pose (None keeps it), which is first advanced by `advance_steps` sim steps."""
        B = self.B
        node = _i32(i_node, (B,))
        gs = np.ascontiguousarray(gait_seq, dtype=np.int8)
        ps = np.ascontiguousarray(peak_seq, dtype=np.int8)
        assert gs.shape == ps.shape and gs.shape[0] == 4
        qq, vd, wy = _f64(q, (B, 18)), _f64(v_des, (B, 3)), _f64(w_yaw, (B,))
        tr = None if tracking is None else _f64(tracking, (B, 12))
        jr = np.ascontiguousarray(joint_ref, dtype=np.float64).reshape(12)
        gain = np.ascontiguousarray(p_gain, dtype=np.float64).reshape(4)
        _order_after_torch(self, qq, vd, wy, tr)
        check(self.lib.nmpc_plan_tick(self._h, ptr(node), ptr(gs), ptr(ps), int(gs.shape[1]), ptr(qq), ptr(vd), ptr(wy),
                                      ptr(tr), int(advance_steps), float(sim_dt), float(nom_height), float(height_offset),
                                      float(t_horizon), ptr(jr), float(step_height), ptr(gain), int(bool(opt_peak))),
              "nmpc_plan_tick")

    def plan_tick_raibert(self, i_node, gait_seq, peak_seq, make_seq, qv, v_des, w_yaw, t_cnt, offset_hip_b,
                          stance_ratio, nominal_period: float, foot_size: float, planner_height_offset: float,
                          cnt_radius: float, nom_height: float, height_offset: float, joint_ref, step_height: float,
                          p_gain, opt_peak: bool = True, gravity: float = 9.81, want_locations: bool = False):
        """Planning + `init` on the device for the Raibert contact planner (nmpc_plan_tick_raibert): i_node [B], the
        gait cycle tables int8 [4, nodes_per_cycle] (stance, swing, touch-down), qv [B,36], v_des [B,3], w_yaw [B],
        t_cnt [N+1] = round(j * dt_nodes, 3).  Returns the footholds [B,4,N+1,3] when `want_locations`."""
        from ._lib import NmpcRaibertParams
        B, N1 = self.B, self.N + 1
        node = _i32(i_node, (B,))
        gs, ps, ms = (np.ascontiguousarray(t, dtype=np.int8) for t in (gait_seq, peak_seq, make_seq))
        assert gs.shape == ps.shape == ms.shape and gs.shape[0] == 4
        q, vd, wy = _f64(qv, (B, 36)), _f64(v_des, (B, 3)), _f64(w_yaw, (B,))
        tc = np.ascontiguousarray(t_cnt, dtype=np.float64).reshape(N1)
        jr = np.ascontiguousarray(joint_ref, dtype=np.float64).reshape(12)
        gain = np.ascontiguousarray(p_gain, dtype=np.float64).reshape(4)
        rp = NmpcRaibertParams(nominal_period=float(nominal_period), foot_size=float(foot_size),
                               height_offset=float(planner_height_offset), gravity=float(gravity),
                               cnt_radius=float(cnt_radius))
        rp.offset_hip_b[:] = [float(x) for x in np.asarray(offset_hip_b, dtype=float).reshape(12)]
        rp.stance_ratio[:] = [float(x) for x in np.asarray(stance_ratio, dtype=float).reshape(4)]
        loc = np.zeros((B, 4, N1, 3)) if want_locations else None
        _order_after_torch(self, q, vd, wy)
        check(self.lib.nmpc_plan_tick_raibert(self._h, ptr(node), ptr(gs), ptr(ps), ptr(ms), int(gs.shape[1]), ptr(q),
                                              ptr(vd), ptr(wy), ptr(tc), rp, float(nom_height), float(height_offset),
                                              ptr(jr), float(step_height), ptr(gain), int(bool(opt_peak)), ptr(loc)),
              "nmpc_plan_tick_raibert")
        return loc

    def get_tracking(self) -> np.ndarray:
        tr = np.zeros((self.B, 12))
        check(self.lib.nmpc_get_tracking(self._h, ptr(tr)), "nmpc_get_tracking")
        return tr

    def cold_guess(self):
        check(self.lib.nmpc_cold_guess(self._h), "nmpc_cold_guess")

    def get_problem_data(self) -> dict:
        B, N = self.B, self.N
        out = dict(p=np.zeros((B, N + 1, NP)), yref=np.zeros((B, N, NY)), yref_e=np.zeros((B, NYE)), x0=np.zeros((B, NX)))
        check(self.lib.nmpc_get_problem_data(self._h, ptr(out["p"]), ptr(out["yref"]), ptr(out["yref_e"]), ptr(out["x0"])),
              "nmpc_get_problem_data")
        return out

    def pin_stance_feet(self, i_node):
        a = _i32(i_node, (self.B,))
        _order_after_torch(self, a)
        check(self.lib.nmpc_pin_stance_feet(self._h, ptr(a)), "nmpc_pin_stance_feet")
        self.synchronize_if_host(a)

    def set_guess(self, x=None, u=None):
        a = None if x is None else _f64(x, (self.B, self.N + 1, NX))
        b = None if u is None else _f64(u, (self.B, self.N, NU))
        _order_after_torch(self, a, b)
        check(self.lib.nmpc_set_guess(self._h, ptr(a), ptr(b)), "nmpc_set_guess")
        self.synchronize_if_host(a if a is not None else b)

    def shift_warm_start(self, start_node, repeat_last: bool = False):
        a = _i32(start_node, (self.B,))
        _order_after_torch(self, a)
        check(self.lib.nmpc_shift_warm_start(self._h, ptr(a), int(repeat_last)), "nmpc_shift_warm_start")
        self.synchronize_if_host(a)

    def reset_duals(self):
        check(self.lib.nmpc_reset_duals(self._h), "nmpc_reset_duals")

    def synchronize_if_host(self, a):
        # host buffers are only borrowed until the async copy has been consumed
        if isinstance(a, np.ndarray) or (hasattr(a, "is_cuda") and not a.is_cuda):
            self.synchronize()

    # -- solve --------------------------------------------------------------------------
    def solve(self, status=None, sync: bool = True):
        """Enqueue the solve.  With sync=True returns the per-problem status array."""
        if not sync:
            check(self.lib.nmpc_solve(self._h, None), "nmpc_solve")
            return None
        st = np.zeros(self.B, dtype=np.int32) if status is None else status
        check(self.lib.nmpc_solve(self._h, ptr(st)), "nmpc_solve")
        return st

    def synchronize(self):
        check(self.lib.nmpc_synchronize(self._h), "nmpc_synchronize")

    # -- results ------------------------------------------------------------------------
    def get_solution(self, duals: bool = True):
        B, N = self.B, self.N
        x, u = np.zeros((B, N + 1, NX)), np.zeros((B, N, NU))
        pi = np.zeros((B, N, NX)) if duals else None
        lam = np.zeros((B, N + 1, 2 * NG)) if duals else None
        check(self.lib.nmpc_get_solution(self._h, ptr(x), ptr(u), ptr(pi), ptr(lam)), "nmpc_get_solution")
        return x, u, pi, lam

    def get_dt(self, duals: bool = False):
        """Node durations of the solution, dt [B, N] (`dt_node_sol`, solver.py:423-427: dt_nodes everywhere unless
        enable_time_opt); with `duals` also the multipliers of dt_min <= dt <= dt_max, [B, N, 2]."""
        dt = np.zeros((self.B, self.N))
        lam = np.zeros((self.B, self.N, 2)) if duals else None
        check(self.lib.nmpc_get_dt(self._h, ptr(dt), ptr(lam)), "nmpc_get_dt")
        return (dt, lam) if duals else dt

    def set_dt(self, dt=None):
        """Node durations of the iterate (inputs["dt"], solver.py:189-190); None resets them to dt_nodes."""
        dt = None if dt is None else _f64(dt, (self.B, self.N))
        check(self.lib.nmpc_set_dt(self._h, ptr(dt)), "nmpc_set_dt")

    # -- checkpoint / resume ------------------------------------------------------------------
    def export_state(self) -> np.ndarray:
        """Iterate, stored QP solution and problem data as one uint8 array (nmpc_export_state)."""
        blob = np.zeros(int(self.lib.nmpc_state_bytes(self._h)), dtype=np.uint8)
        check(self.lib.nmpc_export_state(self._h, ptr(blob)), "nmpc_export_state")
        return blob

    def import_state(self, blob: np.ndarray):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        if blob.size != int(self.lib.nmpc_state_bytes(self._h)):
            raise ValueError("state blob of another solver shape")
        check(self.lib.nmpc_import_state(self._h, ptr(blob)), "nmpc_import_state")

    # -- batched rigid-body rollout (nmpc_rollout.cuh) ---------------------------------------
    def sim_set_state(self, qv):
        check(self.lib.nmpc_sim_set_state(self._h, ptr(_f64(qv, (self.B, 36)))), "nmpc_sim_set_state")
        self.synchronize()

    def sim_get_state(self):
        qv, div = np.zeros((self.B, 36)), np.zeros(self.B, dtype=np.int32)
        check(self.lib.nmpc_sim_get_state(self._h, ptr(qv), ptr(div)), "nmpc_sim_get_state")
        return qv, div

    def sim_steps(self, params, n_interp: int, step0: int, n_steps: int, record: bool = True):
        """`n_steps` closed-loop simulation steps of every rollout on the plan of the last solve: the controller of
        `LocomotionMPC.compute_torques_dof` and one physics step per sim step, on the device.  With `record`:
        dict(qv [B,n,36] state seen by the controller, tau [B,n,12], fc [B,n,12] contact forces)."""
        B = self.B
        out = dict(qv=np.zeros((B, n_steps, 36)), tau=np.zeros((B, n_steps, 12)), fc=np.zeros((B, n_steps, 12))) if record else {}
        check(self.lib.nmpc_sim_steps(self._h, C.byref(params), int(n_interp), int(step0), int(n_steps),
                                      ptr(out.get("qv")), ptr(out.get("tau")), ptr(out.get("fc"))), "nmpc_sim_steps")
        return out

    def get_solution_into(self, x=None, u=None, pi=None, lam=None):
        check(self.lib.nmpc_get_solution(self._h, ptr(x), ptr(u), ptr(pi), ptr(lam)), "nmpc_get_solution")

    def set_output(self, x=None, u=None):
        """Destinations (pinned host or CUDA fp64 tensors / arrays, x [B,N+1,42], u [B,N,30]) that every following
        `solve` fills with the final iterate, wave by wave, while the next wave is still being solved
        (nmpc_set_output).  Complete once the solve stream is synchronised (`get_status_into`, `synchronize`).  The
        solver keeps a reference to the buffers; `set_output()` switches the streaming off."""
        B, N = self.B, self.N
        for t, n in ((x, B * (N + 1) * NX), (u, B * N * NU)):
            if t is not None:
                numel = t.numel() if hasattr(t, "numel") else t.size
                assert numel == n, "output buffer of the wrong size"
        self._out_refs = (x, u)
        check(self.lib.nmpc_set_output(self._h, ptr(x), ptr(u)), "nmpc_set_output")

    def get_status_into(self, status):
        """Per-problem status of the last solve into an int32 array / tensor (host or device)."""
        check(self.lib.nmpc_get_status(self._h, ptr(status)), "nmpc_get_status")

    def solve_status(self) -> np.ndarray:
        st = np.zeros(self.B, dtype=np.int32)
        self.get_status_into(st)
        return st

    def pack_solution(self, out):
        """Packed result rows [B, (N+1)*42 + N*31 + 1] (x | u | dt_node | status) into a CUDA fp64 tensor, on the
        solver's stream and without synchronising: the send buffer of the gather to the dataset-writer rank."""
        assert out.is_cuda and out.is_contiguous() and out.numel() == self.B * self.sol_row
        check(self.lib.nmpc_pack_solution(self._h, ptr(out)), "nmpc_pack_solution")

    @property
    def sol_row(self) -> int:
        return (self.N + 1) * NX + self.N * (NU + 1) + 1

    def plan_torques_into(self, n_interp: int, step0: int, n_steps: int, kp: float = 20.0, kd: float = 1.75,
                          q_plan=None, v_plan=None, tau_ff=None, tau=None, action=None, qv_meas=None):
        """`plan_torques` with caller-owned outputs (host arrays or CUDA tensors; any may be None)."""
        _order_after_torch(self, qv_meas)
        check(self.lib.nmpc_plan_torques(self._h, ptr(qv_meas), int(n_interp), int(step0), int(n_steps), float(kp),
                                         float(kd), ptr(q_plan), ptr(v_plan), ptr(tau_ff), ptr(tau), ptr(action)),
              "nmpc_plan_torques")

    def get_stats(self) -> dict:
        B = self.B
        out = dict(sqp_iter=np.zeros(B, np.int32), qp_iter=np.zeros(B, np.int32),
                   qp_status=np.zeros(B, np.int32), res=np.zeros((B, 4)), qp_res=np.zeros((B, 4)),
                   cost=np.zeros(B), qp_ncond=np.zeros(B, np.int32))
        st = NmpcStats(*[out[k].ctypes.data for k in ("sqp_iter", "qp_iter", "qp_status", "res", "qp_res", "cost",
                                                      "qp_ncond")])
        check(self.lib.nmpc_get_stats(self._h, C.byref(st)), "nmpc_get_stats")
        return out

    def get_qp_step(self) -> np.ndarray:
        dz = np.zeros((self.B, self.N + 1, NZ))
        check(self.lib.nmpc_get_qp_step(self._h, ptr(dz)), "nmpc_get_qp_step")
        return dz

    def kinematics(self, qv):
        a = _f64(qv, (-1, 36))
        _order_after_torch(self, a)
        n = a.shape[0]
        feet, com, hg = np.zeros((n, 4, 3)), np.zeros((n, 3)), np.zeros((n, 6))
        check(self.lib.nmpc_kinematics(self._h, ptr(a), n, ptr(feet), ptr(com), ptr(hg)), "nmpc_kinematics")
        return feet, com, hg

    @property
    def total_mass(self) -> float:
        return float(self.lib.nmpc_model_mass())

    def inverse_dynamics(self, qva, f):
        """tau [n,12] = rnea(q, v, a)[joints] - sum_feet J^T f  (dynamics.py:136-163)."""
        a = _f64(qva, (-1, 54))
        n = a.shape[0]
        b = _f64(f, (n, 12))
        _order_after_torch(self, a, b)
        tau = np.zeros((n, 12))
        check(self.lib.nmpc_inverse_dynamics(self._h, ptr(a), ptr(b), n, ptr(tau)), "nmpc_inverse_dynamics")
        return tau

    def plan_torques(self, n_interp: int, step0: int = 0, n_steps: Optional[int] = None, kp: float = 20.0,
                     kd: float = 1.75, qv_meas=None) -> dict:
        """Batched plan -> torque from the resident solution (include/nmpc_b200.h::nmpc_plan_torques;
        reference mpc.py:371-414, 544-545, 592-599, dynamics.py:136-163, Rollout_MPC.py:298).
        Returns q_plan, v_plan [B,n_steps,18] and tau_ff, tau, action [B,n_steps,12]."""
        n_steps = n_interp - step0 if n_steps is None else int(n_steps)
        B = self.B
        out = {k: np.zeros((B, n_steps, w)) for k, w in
               (("q_plan", 18), ("v_plan", 18), ("tau_ff", 12), ("tau", 12), ("action", 12))}
        m = None if qv_meas is None else _f64(qv_meas, (B, n_steps, 36))
        _order_after_torch(self, m)
        check(self.lib.nmpc_plan_torques(self._h, None if m is None else ptr(m), int(n_interp), int(step0), n_steps,
                                         float(kp), float(kd), ptr(out["q_plan"]), ptr(out["v_plan"]),
                                         ptr(out["tau_ff"]), ptr(out["tau"]), ptr(out["action"])), "nmpc_plan_torques")
        return out

    # -- diagnostics --------------------------------------------------------------------
    def export_qp(self) -> dict:
        B, N = self.B, self.N
        out = dict(BAbt=np.zeros((B, N, NZ + 1, NX)), RSQrq=np.zeros((B, N + 1, NZ + 1, NZ)),
                   DCt=np.zeros((B, N + 1, NZ, NG)), lg=np.zeros((B, N + 1, NG)),
                   ug=np.zeros((B, N + 1, NG)), mask=np.zeros((B, N + 1, 2 * NG)))
        check(self.lib.nmpc_export_qp(self._h, *[ptr(out[k]) for k in ("BAbt", "RSQrq", "DCt", "lg", "ug", "mask")]),
              "nmpc_export_qp")
        return out

    def solve_qp(self, qp: dict, dx0) -> dict:
        """Solve externally supplied OCP-QPs (host arrays; the library repacks them on the host and refuses device
        pointers with NMPC_EINVAL).  QPs without the structure of this OCP family raise (NMPC_ENOSYS)."""
        B, N = self.B, self.N
        if any(hasattr(qp[k], "is_cuda") and qp[k].is_cuda for k in ("BAbt", "RSQrq", "DCt", "lg", "ug", "mask")):
            raise ValueError("nmpc_solve_qp takes host arrays")
        arrs = [_f64(qp[k], s) for k, s in (("BAbt", (B, N, NZ + 1, NX)), ("RSQrq", (B, N + 1, NZ + 1, NZ)),
                                             ("DCt", (B, N + 1, NZ, NG)), ("lg", (B, N + 1, NG)),
                                             ("ug", (B, N + 1, NG)), ("mask", (B, N + 1, 2 * NG)))]
        dx0 = _f64(dx0, (B, NX))
        out = dict(dz=np.zeros((B, N + 1, NZ)), pi=np.zeros((B, N, NX)), lam=np.zeros((B, N + 1, 2 * NG)),
                   t=np.zeros((B, N + 1, 2 * NG)), qp_iter=np.zeros(B, np.int32), qp_status=np.zeros(B, np.int32))
        check(self.lib.nmpc_solve_qp(self._h, *[ptr(a) for a in arrs], ptr(dx0),
                                     *[ptr(out[k]) for k in ("dz", "pi", "lam", "t", "qp_iter", "qp_status")]),
              "nmpc_solve_qp")
        return out

    @property
    def last_qp_path(self) -> str:
        return {1: "fused", -1: "refused"}.get(int(self.lib.nmpc_last_qp_path(self._h)), "?")

    def set_profiling(self, on: bool = True):
        check(self.lib.nmpc_set_profiling(self._h, int(on)), "nmpc_set_profiling")

    def kernel_times(self) -> dict:
        """{kernel: (ms, launches)} accumulated since the last call (CUDA events, handle stream)."""
        kinds = ("linearize", "unused1", "residual", "unused3", "unused4", "qp_other", "qp_fused")
        ms = np.zeros(len(kinds))
        cnt = np.zeros(len(kinds), dtype=np.int64)
        check(self.lib.nmpc_get_kernel_times(self._h, ptr(ms), ptr(cnt)), "nmpc_get_kernel_times")
        return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(kinds)}

    @property
    def stream_ptr(self) -> int:
        s = C.c_void_p(0)
        check(self.lib.nmpc_stream(self._h, C.byref(s)), "nmpc_stream")
        return int(s.value or 0)

    @property
    def launch_count(self) -> int:
        return int(self.lib.nmpc_launch_count(self._h))
