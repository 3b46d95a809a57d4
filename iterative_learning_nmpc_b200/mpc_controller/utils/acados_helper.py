"""Host-side stand-ins for the `contact_tamp` classes the reference's solver front-end is built on
(un-vendored submodule @ 06601f49):

    ProblemFormulation, AcadosSolverHelper   (imported at mpc_controller/utils/solver.py:7-8)
    FloatingBaseDynamics, PointContact       (imported at mpc_controller/utils/dynamics.py:4-5)

Only the surface the in-tree code uses is provided, with the same method / attribute names, so
`QuadrupedAcadosSolver` and `QuadrupedDynamics` read like the reference's.  Beneath it there is
no acados: every call ends in the C ABI of libnmpc_b200.so (include/nmpc_b200.h), whose fixed
state / input / parameter / cost layouts this module maps names onto.  The library raises if the
CUDA build or the device is missing -- there is no CPU path here.

Array conventions are the reference's: `states[name] -> [dim, N+1]`, `inputs[name] -> [dim, N]`,
`params[name] -> [dim, N+1]`, `cost_ref[name] -> [dim, N]`, `cost_ref_terminal[name] -> [dim]`
(column = node).  With `batch=B` every one of them gains a leading batch dimension.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

from ... import _lib as L
from ...core import NmpcSolver

NQ, NFEET, NP_FOOT = 18, 4, 11


@dataclass(frozen=True)
class Var:
    """A named block of the state / input / parameter / cost vector (CasADi symbol stand-in)."""
    name: str
    dim: int
    offset: int                      # in the running vector
    offset_e: Optional[int] = None   # in the terminal cost vector, if the block has a terminal twin


class ProblemFormulation:
    """`ProblemFormulation(dt, dt_min, dt_max, enable_time_opt)` (solver.py:44); collects what the
    model declares.  The formulation itself is compiled into the CUDA library."""

    def __init__(self, dt: float, dt_min: float, dt_max: float, enable_time_opt: bool = False):
        self.dt, self.dt_min, self.dt_max = float(dt), float(dt_min), float(dt_max)
        self.enable_time_opt = bool(enable_time_opt)
        self.states: Dict[str, Var] = {}
        self.inputs: Dict[str, Var] = {}
        self.params: Dict[str, Var] = {}
        self.costs: Dict[str, Var] = {}
        self.terminal_costs: Dict[str, Var] = {}

    def add_state(self, var: Var):
        self.states[var.name] = var

    def add_input(self, var: Var):
        self.inputs[var.name] = var

    def add_param(self, var: Var):
        self.params[var.name] = var

    def add_cost(self, expr: Var, terminal: bool = False):
        """dynamics.py:113-116."""
        self.costs[expr.name] = expr
        if terminal:
            if expr.offset_e is None:
                raise ValueError(f"cost '{expr.name}' has no terminal slot in this build")
            self.terminal_costs[expr.name] = expr


class PointContact:
    """Per-foot contact model: parameters, force input, force-regularisation and foot-position costs
    (attribute names used at solver.py:123-137,203-225,242-275)."""

    def __init__(self, dyn: "FloatingBaseDynamics", frame: str, mu: float = 0.8, patch_restriction: bool = False,
                 index: int = 0):
        self.dyn, self.frame_name, self.mu, self.patch_restriction = dyn, frame, float(mu), bool(patch_restriction)
        self.index = i = int(index)
        b = NP_FOOT * i
        n = f"{frame}_{dyn.name}"
        self.active = Var(f"active_{n}", 1, b + 0)
        self.plane_point = Var(f"plane_point_{n}", 3, b + 1)
        self.plane_normal = Var(f"plane_normal_{n}", 3, b + 4)
        self.p_gain = Var(f"p_gain_{n}", 1, b + 7)
        self.restrict = Var(f"restrict_{n}", 1, b + 8)
        self.range_radius = Var(f"range_radius_{n}", 1, b + 9)
        self.peak = Var(f"peak_{n}", 1, b + 10)
        self.f = Var(f"f_{n}", 3, NQ + 3 * i)
        self.f_reg = Var(f"f_reg_{n}", 3, 52 + 3 * i)
        self.pos_cost = Var(f"pos_cost_{n}", 3, 64 + 3 * i, 40 + 3 * i)

    def setup(self, problem: ProblemFormulation):
        for v in (self.active, self.plane_point, self.plane_normal, self.p_gain, self.restrict,
                  self.range_radius, self.peak):
            problem.add_param(v)
        problem.add_input(self.f)
        problem.add_cost(self.f_reg)
        problem.add_cost(self.pos_cost, terminal=True)


class FloatingBaseDynamics:
    """State x = [q | v | h], input u = [a | contact forces] of a floating-base robot in the Euler
    parametrisation (dynamics.py:24,34-38,112)."""

    def __init__(self, name: str):
        self.name = name
        self.q = Var(f"q_{name}", NQ, 0)
        self.v = Var(f"v_{name}", NQ, NQ)
        self.h = Var(f"h_{name}", 6, 2 * NQ)
        self.a = Var(f"a_{name}", NQ, 0)
        self.contacts: List[PointContact] = []
        self.exprs: Dict[str, Var] = {}

    _COST_SLOTS = {"base_cost": (12, 0, 0), "joint_cost": (24, 12, 12), "acc_cost": (12, 36, None),
                   "sw_cost": (4, 48, 36)}

    def add_contacts(self, contacts: Sequence[PointContact]):
        self.contacts = list(contacts)

    def add_expr(self, name: str, expr=None) -> Var:
        if name not in self._COST_SLOTS:
            raise NotImplementedError(f"expression '{name}' is not part of the compiled formulation")
        dim, off, off_e = self._COST_SLOTS[name]
        var = Var(name, dim, off, off_e)
        self.exprs[name] = var
        return var

    def setup(self, problem: ProblemFormulation):
        for v in (self.q, self.v, self.h):
            problem.add_state(v)
        problem.add_input(self.a)


class AcadosSolverHelper:
    """Name -> array bookkeeping in front of one batch handle of the CUDA solver.

    Replaces `contact_tamp...acados_helper.AcadosSolverHelper` (solver.py:15).  `batch=None`
    reproduces the reference's unbatched shapes on a B = 1 handle; `batch=B` adds the leading
    dimension.
    """

    def __init__(self, problem: ProblemFormulation, n_nodes: int, name: str, reg_eps: float = 1e-6,
                 reg_eps_e: float = 1e-5, batch: Optional[int] = None, device: int = 0):
        self.problem, self.n_nodes, self.name = problem, int(n_nodes), name
        self.reg_eps, self.reg_eps_e = float(reg_eps), float(reg_eps_e)
        self._lead = () if batch is None else (int(batch),)
        self.batch = 1 if batch is None else int(batch)
        self.device = int(device)
        self._opts = L.default_opts(n_nodes=self.n_nodes, dt_nodes=problem.dt, dt_min=problem.dt_min,
                                    dt_max=problem.dt_max, enable_time_opt=int(problem.enable_time_opt),
                                    reg_eps=self.reg_eps, reg_eps_e=self.reg_eps_e)
        self._handle: Optional[NmpcSolver] = None
        N, ld = self.n_nodes, self._lead
        self.states = {v.name: np.zeros(ld + (v.dim, N + 1)) for v in problem.states.values()}
        self.inputs = {v.name: np.zeros(ld + (v.dim, N)) for v in problem.inputs.values()}
        self.params = {v.name: np.zeros(ld + (v.dim, N + 1)) for v in problem.params.values()}
        self.cost_ref = {v.name: np.zeros(ld + (v.dim, N)) for v in problem.costs.values()}
        self.cost_ref_terminal = {v.name: np.zeros(ld + (v.dim,)) for v in problem.terminal_costs.values()}
        if problem.enable_time_opt:
            # the node duration: a 31st input "dt" with its own cost weight and reference (solver.py:124-125,
            # 166-167, 189-190).  It lives beside the packed 30-wide input block: nmpc_set_dt / nmpc_get_dt.
            self.inputs["dt"] = np.full(ld + (1, N), problem.dt)
            self.cost_ref["dt"] = np.full(ld + (1, N), problem.dt)
        self._W = np.zeros(ld + (L.NY,))
        self._We = np.zeros(ld + (L.NYE,))
        self.status = np.zeros(self.batch, dtype=np.int32)
        self._sol = None

    # -- construction / options (solver.py:68-79) ----------------------------------------------
    def setup(self, recompile: bool = False, use_cython: bool = False, real_time_it: bool = False,
              max_qp_iter: int = 6, hpipm_mode=1):
        """Build (if asked) and open the device solver.  `use_cython` has no equivalent: the
        binding is ctypes either way."""
        from ... import build as product_build
        product_build.build()          # recompiles only when a source is newer than the library
        self._opts.real_time_it = int(bool(real_time_it))
        self._opts.max_qp_iter = int(max_qp_iter)
        self._opts.hpipm_mode = int(hpipm_mode)
        if self._handle is None:
            self._handle = self._make_handle()
        else:
            self._handle.set_opts(self._opts)
            self._handle.reset_duals()

    def _make_handle(self):
        """Open the device solver.  (The CPU test-suite substitutes the oracle here to check this
        front-end without a GPU; the product has no other backend.)"""
        return NmpcSolver(self.batch, self._opts, device=self.device)

    @property
    def handle(self) -> NmpcSolver:
        if self._handle is None:
            raise RuntimeError("setup() has not been called")
        return self._handle

    def _set_opt(self, **kw):
        for k, v in kw.items():
            setattr(self._opts, k, v)
        if self._handle is not None:
            self._handle.set_opts(self._opts)

    def set_max_iter(self, n: int):
        self._set_opt(max_iter=int(n))

    def set_warm_start_inner_qp(self, flag):
        self._set_opt(warm_start_qp=int(flag))

    def set_warm_start_nlp(self, flag):
        self._set_opt(warm_start_nlp=int(flag))

    def set_qp_tol(self, tol: float):
        self._set_opt(qp_tol=float(tol))

    def set_nlp_tol(self, tol: float):
        self._set_opt(nlp_tol=float(tol))

    # -- data templates / weights (solver.py:81,113-141) ---------------------------------------
    def get_data_template(self) -> Dict[str, Dict[str, np.ndarray]]:
        ld = self._lead
        return {
            "x": {v.name: np.zeros(ld + (v.dim,)) for v in self.problem.states.values()},
            "u": {v.name: np.zeros(ld + (v.dim,)) for v in self.problem.inputs.values()},
            "W": dict({v.name: np.zeros(ld + (v.dim,)) for v in self.problem.costs.values()},
                      **({"dt": np.zeros(ld + (1,))} if self.problem.enable_time_opt else {})),
            "W_e": {v.name: np.zeros(ld + (v.dim,)) for v in self.problem.terminal_costs.values()},
        }

    def _pack_vec(self, src: Dict[str, np.ndarray], table: Dict[str, Var], out: np.ndarray, terminal: bool):
        for name, var in table.items():
            if name not in src:
                continue
            off = var.offset_e if terminal else var.offset
            out[..., off:off + var.dim] = np.broadcast_to(np.asarray(src[name], dtype=float), out.shape[:-1] + (var.dim,))

    def set_cost_weight_constant(self, W: Dict[str, np.ndarray]):
        self._pack_vec(W, self.problem.costs, self._W, terminal=False)
        if self.problem.enable_time_opt and "dt" in W:
            w = np.unique(np.asarray(W["dt"], dtype=float))
            if w.size != 1:
                raise NotImplementedError("one time_opt weight per batch handle")
            self._set_opt(time_opt_weight=float(w[0]))

    def set_cost_weight_terminal(self, W_e: Dict[str, np.ndarray]):
        self._pack_vec(W_e, self.problem.terminal_costs, self._We, terminal=True)

    def update_cost_weights(self):
        self.handle.set_weights(self._W.reshape((-1, L.NY)) if self._lead else self._W,
                                self._We.reshape((-1, L.NYE)) if self._lead else self._We)

    # -- per-solve data (solver.py:192,349-353) --------------------------------------------------
    def set_initial_state(self, x: Dict[str, np.ndarray]):
        x0 = np.zeros((self.batch, L.NX))
        self._pack_vec(x, self.problem.states, x0.reshape(self._lead + (L.NX,)), terminal=False)
        self.handle.set_x0_full(x0)

    def _pack_nodes(self, src: Dict[str, np.ndarray], table: Dict[str, Var], width: int, n: int,
                    terminal: bool = False) -> np.ndarray:
        out = np.zeros((self.batch, n, width))
        view = out.reshape(self._lead + (n, width))
        for name, var in table.items():
            off = var.offset_e if terminal else var.offset
            a = np.broadcast_to(np.asarray(src[name], dtype=float), self._lead + (var.dim, n))
            view[..., off:off + var.dim] = np.moveaxis(a, -1, -2)
        return out

    def update_states(self):
        self.handle.set_guess(x=self._pack_nodes(self.states, self.problem.states, L.NX, self.n_nodes + 1))

    def update_inputs(self):
        self.handle.set_guess(u=self._pack_nodes(self.inputs, self.problem.inputs, L.NU, self.n_nodes))
        if self.problem.enable_time_opt:
            self.handle.set_dt(np.asarray(self.inputs["dt"], dtype=float).reshape(self.batch, self.n_nodes))

    def update_parameters(self):
        self.handle.set_params(self._pack_nodes(self.params, self.problem.params, L.NP, self.n_nodes + 1))

    def _pack_refs(self):
        yref = self._pack_nodes(self.cost_ref, self.problem.costs, L.NY, self.n_nodes)
        yref_e = np.zeros((self.batch, L.NYE))
        self._pack_vec(self.cost_ref_terminal, self.problem.terminal_costs,
                       yref_e.reshape(self._lead + (L.NYE,)), terminal=True)
        return yref, yref_e

    def update_ref(self):
        self.handle.set_yref(*self._pack_refs())

    def set_ref_terminal(self):
        self.handle.set_yref(*self._pack_refs())

    def warm_start_multipliers(self, start_node, n_warm_start: Optional[int] = None, repeat_last: bool = False):
        """Shift the multipliers (and the device copy of the iterate) by `start_node` nodes
        (solver.py:339).  `start_node` may be a scalar or one value per problem."""
        self.handle.shift_warm_start(np.broadcast_to(np.asarray(start_node, dtype=np.int32), (self.batch,)),
                                     repeat_last)

    # -- solve / results (solver.py:402-403) -----------------------------------------------------
    def solve(self, print_stats: bool = False, print_time: bool = False) -> np.ndarray:
        if print_time:
            self.handle.set_profiling(True)
        self.status = self.handle.solve()
        self._sol = None
        if print_stats:
            st = self.handle.get_stats()
            for b in range(self.batch):
                print(f"[{self.name}] ocp {b}: status {L.STATUS_NAMES.get(int(self.status[b]), self.status[b])}, "
                      f"sqp_iter {st['sqp_iter'][b]}, qp_iter {st['qp_iter'][b]}, "
                      f"res stat/eq/ineq/comp {st['res'][b]}")
        if print_time:
            for k, (ms, n) in self.handle.kernel_times().items():
                if n:
                    print(f"[{self.name}] {k:10s} {ms:9.3f} ms in {n} launches")
            self.handle.set_profiling(False)
        return self.status

    def parse_sol(self):
        """Copy the solution back into `states` / `inputs` (reference layout, column = node)."""
        x, u, _, _ = self.handle.get_solution(duals=False)
        self._sol = (x, u)
        ld = self._lead
        for var in self.problem.states.values():
            blk = x[:, :, var.offset:var.offset + var.dim]
            self.states[var.name] = np.ascontiguousarray(np.moveaxis(blk, 1, 2)).reshape(ld + (var.dim, self.n_nodes + 1))
        for var in self.problem.inputs.values():
            blk = u[:, :, var.offset:var.offset + var.dim]
            self.inputs[var.name] = np.ascontiguousarray(np.moveaxis(blk, 1, 2)).reshape(ld + (var.dim, self.n_nodes))
        if self.problem.enable_time_opt:
            self.inputs["dt"] = self.handle.get_dt().reshape(ld + (1, self.n_nodes))

    def get_stats(self) -> dict:
        return self.handle.get_stats()
