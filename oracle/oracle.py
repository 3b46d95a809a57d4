"""ctypes binding of the CPU oracle (oracle/libnmpc_oracle.so).  TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, nowhere else.
Mirrors the product's C ABI (include/nmpc_b200.h) call for call so the parity tests read
symmetrically.  PARITY UNPINNED: see the header of nmpc_oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnmpc_oracle.so")
LIB_PATH_TOPT = os.path.join(HERE, "libnmpc_oracle_topt.so")      # -DNMPC_ORACLE_TOPT build: nu = 31, ng = 43

NX, NU, NZ, NP, NY, NYE, NG = 42, 30, 72, 44, 76, 52, 42


class NmpcOpts(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32), ("max_iter", C.c_int32), ("max_qp_iter", C.c_int32),
        ("real_time_it", C.c_int32), ("warm_start_qp", C.c_int32), ("warm_start_nlp", C.c_int32),
        ("hpipm_mode", C.c_int32), ("enable_time_opt", C.c_int32),
        ("warm_start_first_qp", C.c_int32), ("cond_pred_corr", C.c_int32),
        ("dt_nodes", C.c_double), ("dt_min", C.c_double), ("dt_max", C.c_double),
        ("nlp_tol", C.c_double), ("qp_tol", C.c_double),
        ("reg_eps", C.c_double), ("reg_eps_e", C.c_double), ("mu_contact", C.c_double),
        ("time_opt_weight", C.c_double),
    ]


class NmpcStats(C.Structure):
    _fields_ = [("sqp_iter", C.c_void_p), ("qp_iter", C.c_void_p), ("qp_status", C.c_void_p),
                ("res", C.c_void_p), ("qp_res", C.c_void_p), ("cost", C.c_void_p),
                ("qp_ncond", C.c_void_p)]


def default_opts(**kw) -> NmpcOpts:
    o = NmpcOpts(n_nodes=25, max_iter=1, max_qp_iter=6, real_time_it=0, warm_start_qp=1,
                 warm_start_nlp=1, hpipm_mode=1, enable_time_opt=0, warm_start_first_qp=0, cond_pred_corr=1, dt_nodes=0.04, dt_min=0.02,
                 dt_max=0.07, nlp_tol=1e-1, qp_tol=1e-2, reg_eps=1e-6, reg_eps_e=1e-5, mu_contact=0.8,
                 time_opt_weight=1e4)
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def build(force: bool = False) -> str:
    """Compile the oracle with its Makefile (gcc)."""
    if force or not os.path.exists(LIB_PATH) or not os.path.exists(LIB_PATH_TOPT):
        subprocess.check_call(["make", "-C", HERE, "-s"] + (["-B"] if force else []))
    return LIB_PATH


_libs = {}


def lib(topt: bool = False):
    """The oracle library: the default build, or the `enable_time_opt` build (nu = 31, ng = 43)."""
    topt = bool(topt)
    if topt not in _libs:
        build()
        L = C.CDLL(LIB_PATH_TOPT if topt else LIB_PATH)
        L.oracle_create.restype = C.c_void_p
        L.oracle_create.argtypes = [C.POINTER(NmpcOpts), C.c_int32]
        for name in ("oracle_destroy", "oracle_set_opts", "oracle_set_x0", "oracle_set_x0_full",
                     "oracle_set_yref", "oracle_set_weights", "oracle_set_params",
                     "oracle_pin_stance_feet", "oracle_set_guess", "oracle_reset_duals",
                     "oracle_shift_warm_start", "oracle_solve", "oracle_get_solution",
                     "oracle_get_stats", "oracle_get_qp_step", "oracle_export_qp", "oracle_solve_qp",
                     "oracle_inverse_dynamics", "oracle_set_fast"):
            getattr(L, name).restype = C.c_int
        L.oracle_model_mass.restype = C.c_double
        _libs[topt] = L
    return _libs[topt]


def _p(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)


def kinematics(qv: np.ndarray):
    qv = _f64(qv).reshape(-1, 36)
    n = qv.shape[0]
    feet, com, hg = np.zeros((n, 4, 3)), np.zeros((n, 3)), np.zeros((n, 6))
    lib().oracle_kinematics(_p(qv), C.c_int32(n), _p(feet), _p(com), _p(hg))
    return feet, com, hg


def inverse_dynamics(qva: np.ndarray, f: np.ndarray) -> np.ndarray:
    qva = _f64(qva).reshape(-1, 54)
    n = qva.shape[0]
    f = _f64(f).reshape(n, 12)
    tau = np.zeros((n, 12))
    lib().oracle_inverse_dynamics(_p(qva), _p(f), C.c_int32(n), _p(tau))
    return tau


def plan_torques(x: np.ndarray, u: np.ndarray, dt_nodes, n_interp: int, step0: int = 0,
                 n_steps: Optional[int] = None, kp: float = 20.0, kd: float = 1.75, qv_meas=None) -> dict:
    """Checker for nmpc_plan_torques: the reference's own host pipeline, one problem at a time --
    scipy CubicHermiteSpline exactly as `LocomotionMPC.interpolate_trajectory_with_derivatives`
    (/root/reference/mpc_controller/mpc.py:388-414) calls it, `id_repeat` hold (mpc.py:110,544-545),
    id_torques (dynamics.py:136-163, through the C oracle), PD (mpc.py:592-599) and the DAgger label
    (Behavior_Cloning/utils/Rollout_MPC.py:298).  x [B,N+1,42], u [B,N,30]."""
    from scipy.interpolate import CubicHermiteSpline
    x, u = np.asarray(x, float), np.asarray(u, float)
    B, N = u.shape[0], u.shape[1]
    n_steps = n_interp - step0 if n_steps is None else int(n_steps)
    sl = slice(step0, step0 + n_steps)
    id_repeat = np.int32(np.linspace(0, 1, n_interp) * (N - 1))
    out = {k: np.zeros((B, n_steps, w)) for k, w in
           (("q_plan", 18), ("v_plan", 18), ("tau_ff", 12), ("tau", 12), ("action", 12))}
    for b in range(B):
        q_sol, v_sol, a_sol, f_sol = x[b, :, :18], x[b, :, 18:36], u[b, :, :18], u[b, :, 18:30]
        dt_b = np.full(N, dt_nodes) if np.ndim(dt_nodes) == 0 else np.asarray(dt_nodes, float).reshape(B, N)[b]
        time_traj = np.concatenate(([0.], np.cumsum(dt_b)))
        t = np.linspace(time_traj[0], time_traj[-1], n_interp + 1)
        pos = CubicHermiteSpline(time_traj, q_sol, v_sol)(t)[1:]
        acc = np.concatenate((a_sol[:1], a_sol))
        vel = CubicHermiteSpline(time_traj, v_sol, acc)(t)[1:]
        qp, vp = pos[sl], vel[sl]
        qm, vm = (qp, vp) if qv_meas is None else (np.asarray(qv_meas)[b, :, :18], np.asarray(qv_meas)[b, :, 18:])
        ap, fp = a_sol[id_repeat][sl], f_sol[id_repeat][sl]
        tff = inverse_dynamics(np.concatenate([qm, vm, ap], axis=1), fp)
        # `_compute_torques_ff` reads a_plan / f_plan[plan_step], `_step()` increments plan_step, then
        # `_compute_pd_torques` reads q_plan / v_plan[plan_step] (mpc.py:575-599): the PD target is the next sample
        nxt = np.minimum(np.arange(step0, step0 + n_steps) + 1, n_interp - 1)
        tau = tff + kp * (pos[nxt][:, 6:] - qm[:, 6:]) + kd * (vel[nxt][:, 6:] - vm[:, 6:])
        out["q_plan"][b], out["v_plan"][b], out["tau_ff"][b], out["tau"][b] = qp, vp, tff, tau
        out["action"][b] = (tau + kd * vm[:, 6:]) / kp + qm[:, 6:]
    return out


def model_mass() -> float:
    return float(lib().oracle_model_mass())


def sim_params_array(p) -> np.ndarray:
    """nmpc_sim_params (ctypes struct of the product, or a dict) as the 9 doubles the oracle takes."""
    names = ("sim_dt", "kp", "kd", "k_n", "d_n", "d_t", "mu", "ground_z", "tau_max")
    get = (lambda k: p[k]) if isinstance(p, dict) else (lambda k: getattr(p, k))
    return np.array([float(get(k)) for k in names])


def sim_terms(qv: np.ndarray):
    """Mass matrix [18,18] and bias [18] of the rollout's forward dynamics (generated code, nmpc_gen_sim.h)."""
    qv = _f64(qv).reshape(36)
    M, b = np.zeros((18, 18)), np.zeros(18)
    lib().oracle_sim_terms(_p(qv), _p(M), _p(b))
    return M, b


def sim_physics(params, qv: np.ndarray, tau: np.ndarray, n_steps: int):
    """n_steps physics steps of one robot from joint torques tau [12] (held) or [n_steps,12]: (qv_end, qv_traj, fc)."""
    qv = _f64(qv).reshape(36).copy()
    tau = _f64(tau)
    hold = int(tau.ndim == 1)
    out_qv, out_fc = np.zeros((n_steps, 36)), np.zeros((n_steps, 12))
    pa = sim_params_array(params)
    lib().oracle_sim_physics(_p(pa), _p(qv), _p(tau), C.c_int32(n_steps), C.c_int32(hold), _p(out_qv), _p(out_fc))
    return qv, out_qv, out_fc


def model_record(qv: np.ndarray) -> np.ndarray:
    qv = _f64(qv).reshape(36)
    dims = np.zeros(8, dtype=np.int32)
    lib().oracle_dims(_p(dims))
    rec = np.zeros(int(dims[6]))
    lib().oracle_model_record(_p(qv), _p(rec))
    return rec


class OracleSolver:
    """Batch of B OCPs solved on the host; same call sequence as the product's NmpcSolver."""

    def __init__(self, batch: int, opts: Optional[NmpcOpts] = None):
        self.opts = opts or default_opts()
        self.B = int(batch)
        self.N = int(self.opts.n_nodes)
        self.topt = bool(self.opts.enable_time_opt)
        self.lib = lib(self.topt)
        # dimensions of this build: the time-opt variant carries dt as a 31st input and one more constraint row
        self.NU, self.NZ, self.NG = (NU + 1, NZ + 1, NG + 1) if self.topt else (NU, NZ, NG)
        self._h = C.c_void_p(self.lib.oracle_create(C.byref(self.opts), C.c_int32(self.B)))
        if not self._h:
            raise RuntimeError("oracle_create failed")

    def __del__(self):
        if getattr(self, "_h", None):
            self.lib.oracle_destroy(self._h)
            self._h = None

    def set_fast(self, fast: bool = True):
        """Structured interior point (nmpc_oracle_fast.h) instead of the dense restatement: the timed CPU baseline."""
        assert self.lib.oracle_set_fast(self._h, C.c_int32(int(fast))) == 0

    def set_opts(self, opts: NmpcOpts):
        self.opts = opts
        assert self.lib.oracle_set_opts(self._h, C.byref(opts)) == 0

    def set_x0(self, qv):
        qv = _f64(qv).reshape(self.B, 36)
        self.lib.oracle_set_x0(self._h, _p(qv))

    def set_x0_full(self, x0):
        x0 = _f64(x0).reshape(self.B, NX)
        self.lib.oracle_set_x0_full(self._h, _p(x0))

    def set_yref(self, yref, yref_e):
        yref = _f64(yref).reshape(self.B, self.N, NY)
        yref_e = _f64(yref_e).reshape(self.B, NYE)
        self.lib.oracle_set_yref(self._h, _p(yref), _p(yref_e))

    def set_weights(self, W, W_e):
        W, W_e = _f64(W), _f64(W_e)
        per = int(W.ndim == 2)
        self.lib.oracle_set_weights(self._h, _p(W), _p(W_e), C.c_int32(per))

    def set_params(self, p):
        p = _f64(p).reshape(self.B, self.N + 1, NP)
        self.lib.oracle_set_params(self._h, _p(p))

    def pin_stance_feet(self, i_node):
        i_node = _i32(np.broadcast_to(i_node, (self.B,)))
        self.lib.oracle_pin_stance_feet(self._h, _p(i_node))

    def set_guess(self, x=None, u=None):
        x = None if x is None else _f64(x).reshape(self.B, self.N + 1, NX)
        u = None if u is None else _f64(u).reshape(self.B, self.N, self.NU)
        self.lib.oracle_set_guess(self._h, _p(x), _p(u))

    def reset_duals(self):
        self.lib.oracle_reset_duals(self._h)

    def shift_warm_start(self, start_node, repeat_last: bool = False):
        s = _i32(np.broadcast_to(start_node, (self.B,)))
        self.lib.oracle_shift_warm_start(self._h, _p(s), C.c_int32(int(repeat_last)))

    def solve(self, n_threads: int = 0) -> np.ndarray:
        status = np.zeros(self.B, dtype=np.int32)
        self.lib.oracle_solve(self._h, _p(status), C.c_int32(n_threads))
        return status

    def get_solution(self):
        B, N = self.B, self.N
        x, u = np.zeros((B, N + 1, NX)), np.zeros((B, N, self.NU))
        pi, lam = np.zeros((B, N, NX)), np.zeros((B, N + 1, 2 * self.NG))
        self.lib.oracle_get_solution(self._h, _p(x), _p(u), _p(pi), _p(lam))
        return x, u, pi, lam

    def get_stats(self) -> dict:
        B = self.B
        out = dict(sqp_iter=np.zeros(B, np.int32), qp_iter=np.zeros(B, np.int32),
                   qp_status=np.zeros(B, np.int32), res=np.zeros((B, 4)), qp_res=np.zeros((B, 4)),
                   cost=np.zeros(B), qp_ncond=np.zeros(B, np.int32))
        st = NmpcStats(*[out[k].ctypes.data for k in ("sqp_iter", "qp_iter", "qp_status", "res", "qp_res", "cost",
                                                      "qp_ncond")])
        self.lib.oracle_get_stats(self._h, C.byref(st))
        return out

    def sim_steps(self, params, qv: np.ndarray, n_interp: int, step0: int, n_steps: int):
        """Checker of nmpc_sim_steps: closed loop of every rollout on the plan this solver holds."""
        qv = _f64(qv).reshape(self.B, 36).copy()
        out = dict(qv=np.zeros((self.B, n_steps, 36)), tau=np.zeros((self.B, n_steps, 12)), fc=np.zeros((self.B, n_steps, 12)))
        pa = sim_params_array(params)
        self.lib.oracle_sim_steps(self._h, _p(pa), C.c_int32(n_interp), C.c_int32(step0), C.c_int32(n_steps), _p(qv),
                                  _p(out["qv"]), _p(out["tau"]), _p(out["fc"]))
        out["qv_end"] = qv
        return out

    def get_qp_step(self) -> np.ndarray:
        dz = np.zeros((self.B, self.N + 1, self.NZ))
        self.lib.oracle_get_qp_step(self._h, _p(dz))
        return dz

    def export_qp(self) -> dict:
        B, N = self.B, self.N
        out = dict(BAbt=np.zeros((B, N, self.NZ + 1, NX)), RSQrq=np.zeros((B, N + 1, self.NZ + 1, self.NZ)),
                   DCt=np.zeros((B, N + 1, self.NZ, self.NG)), lg=np.zeros((B, N + 1, self.NG)),
                   ug=np.zeros((B, N + 1, self.NG)), mask=np.zeros((B, N + 1, 2 * self.NG)))
        self.lib.oracle_export_qp(self._h, *[_p(out[k]) for k in ("BAbt", "RSQrq", "DCt", "lg", "ug", "mask")])
        return out

    def solve_qp(self, qp: dict, dx0: np.ndarray) -> dict:
        B, N = self.B, self.N
        dx0 = _f64(dx0).reshape(B, NX)
        out = dict(dz=np.zeros((B, N + 1, self.NZ)), pi=np.zeros((B, N, NX)), lam=np.zeros((B, N + 1, 2 * self.NG)),
                   t=np.zeros((B, N + 1, 2 * self.NG)), qp_iter=np.zeros(B, np.int32),
                   qp_status=np.zeros(B, np.int32), qp_res=np.zeros((B, 4)))
        args = [_p(_f64(qp[k])) for k in ("BAbt", "RSQrq", "DCt", "lg", "ug", "mask")]
        self.lib.oracle_solve_qp(self._h, *args, _p(dx0),
                              *[_p(out[k]) for k in ("dz", "pi", "lam", "t", "qp_iter", "qp_status", "qp_res")])
        return out
