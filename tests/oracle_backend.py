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
        x, u, _, _ = self._s.get_solution()
        return O.plan_torques(x, u, float(self.opts.dt_nodes), n_interp, step0, n_steps, kp, kd, qv_meas)

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
