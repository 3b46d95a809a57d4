"""Synthetic OCP batches of BASELINE.json's configurations.

Initial states come from the reference's own sampler (the "proper nullspace projection perturbation
sampler", `Behavior_Cloning/utils/RolloutMPC_shift_phase_percentage.py:460-540`): uniform perturbations
of base height (+-0.15 m), roll / pitch / yaw (+-0.1, 0.1, 0.25 rad), joints (+-0.3 rad) and ALL 18
velocities (+-2.0), projected on the null space of the stance-foot contact Jacobian (positions and
velocities, `:488-503`), and re-drawn (up to 100 times) until every swing foot is at least 2 cm above the
ground and every stance foot within 3 cm of it (`:510-535`).  Velocity goals:
`Behavior_Cloning/examples/cfgs/bc_config.yaml:11-16`.  The nominal state is the Go2 home pose standing on
its feet at zero velocity (the synthetic model's feet touch the ground at base height
~0.265 m), so the time derivative of the contact Jacobian in the reference's constraint matrix vanishes.
The perturbation is applied in the solver's Euler coordinates (q = [p, yaw, pitch, roll, joints], v = q_dot).
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Sequence

import numpy as np

from . import problem as P
from .mpc_controller.config.quadruped.utils import get_quadruped_config
from .mpc_controller.utils.contact_planner import ContactPlanner, RaiberContactPlanner

FEET = ["FL_foot", "FR_foot", "RL_foot", "RR_foot"]
STAND_HEIGHT = 0.2648058464833031     # base height at which the home pose has its feet at z = 0
MASS = 15.017
_URDF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets", "go2_like.urdf")
_CHAINS = None


def _foot_chains():
    """Per foot: the list of (xyz, R_origin, axis or None, q index) joints from the base to the foot frame."""
    global _CHAINS
    if _CHAINS is None:
        from .codegen.robot import load_urdf, rpy_matrix
        m = load_urdf(_URDF)
        idx = {j.name: 6 + i for i, j in enumerate(m.act_joints)}
        by_child = {j.child: j for j in m.joints}
        chains = []
        for foot in FEET:
            chain, link = [], foot
            while link != m.root:
                j = by_child[link]
                chain.append((j.xyz, rpy_matrix(j.rpy), None if j.jtype == "fixed" else j.axis, idx.get(j.name)))
                link = j.parent
            chains.append(chain[::-1])
        _CHAINS = chains
    return _CHAINS


def feet_positions(q: np.ndarray) -> np.ndarray:
    """World positions of the four feet, [B,4,3], for Euler states q [B,18] (plain numpy forward kinematics of
    `assets/go2_like.urdf`, vectorised over the batch; the solver's `nmpc_kinematics` gives the same numbers)."""
    q = np.atleast_2d(np.asarray(q, dtype=float))
    B = q.shape[0]
    cy, sy, cp, sp, cr, sr = (f(q[:, i]) for i in (3, 4, 5) for f in (np.cos, np.sin))
    R = np.zeros((B, 3, 3))                                  # Rz(yaw) Ry(pitch) Rx(roll)
    R[:, 0, 0] = cy * cp; R[:, 0, 1] = cy * sp * sr - sy * cr; R[:, 0, 2] = cy * sp * cr + sy * sr
    R[:, 1, 0] = sy * cp; R[:, 1, 1] = sy * sp * sr + cy * cr; R[:, 1, 2] = sy * sp * cr - cy * sr
    R[:, 2, 0] = -sp;     R[:, 2, 1] = cp * sr;                R[:, 2, 2] = cp * cr
    out = np.zeros((B, 4, 3))
    for f, chain in enumerate(_foot_chains()):
        Rk, pk = R.copy(), q[:, :3].copy()
        for xyz, Ro, axis, iq in chain:
            pk = pk + Rk @ xyz
            Rk = Rk @ Ro
            if axis is not None:
                K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
                th = q[:, iq]
                Rj = np.eye(3)[None] + np.sin(th)[:, None, None] * K[None] + (1 - np.cos(th))[:, None, None] * (K @ K)[None]
                Rk = Rk @ Rj
        out[:, f] = pk
    return out


def sample_initial_states(rng, nominal: np.ndarray, stance: np.ndarray, perturb: float = 1.0,
                          max_attempts: int = 100) -> np.ndarray:
    """The reference's perturbation sampler, batched (module docstring).  nominal [36]; stance [B,4] bool = feet in
    contact at the current node.  Returns qv [B,36].  Problems still infeasible after `max_attempts` draws keep
    their last draw (the reference would `sys.exit`; does not occur at the ranges used)."""
    B = stance.shape[0]
    q0, eps = nominal[:18], 1e-6
    # contact Jacobian of the nominal pose, d foot / d q (central differences of the forward kinematics)
    qs = np.tile(q0, (36, 1))
    qs[np.arange(18), np.arange(18)] += eps
    qs[18 + np.arange(18), np.arange(18)] -= eps
    fp = feet_positions(qs)
    J = ((fp[:18] - fp[18:]) / (2 * eps)).transpose(1, 2, 0)            # [4, 3, 18]
    qv = np.tile(nominal, (B, 1))
    todo = np.arange(B)
    for _ in range(max_attempts):
        n = todo.size
        dq = np.zeros((n, 18))
        dq[:, 2] = rng.uniform(-0.15, 0.15, n)
        dq[:, 3:6] = rng.uniform(-1.0, 1.0, (n, 3)) * np.array([0.25, 0.1, 0.1])      # yaw, pitch, roll
        dq[:, 6:] = rng.uniform(-0.3, 0.3, (n, 12))
        dv = rng.uniform(-2.0, 2.0, (n, 18))
        dq *= perturb; dv *= perturb
        # null-space projection per distinct stance pattern (<= 16 of them)
        pat = stance[todo] @ (1 << np.arange(4))
        for code in np.unique(pat):
            rows = np.nonzero(pat == code)[0]
            feet = [f for f in range(4) if code >> f & 1]
            if not feet:
                continue
            Jc = J[feet].reshape(-1, 18)
            Pn = np.eye(18) - np.linalg.pinv(Jc) @ Jc
            dq[rows] = dq[rows] @ Pn.T
            dv[rows] = dv[rows] @ Pn.T
        cand = np.concatenate([q0 + dq, nominal[18:] + dv], axis=1)
        z = feet_positions(cand[:, :18])[:, :, 2]
        st = stance[todo]
        ok = np.all(np.where(st, np.abs(z) <= 0.03, z >= 0.02), axis=1)
        qv[todo] = cand
        todo = todo[~ok]
        if todo.size == 0:
            break
    return qv


def _restricted_base_ref(loc: np.ndarray, nom_height: float):
    """`LocomotionMPC.compute_base_ref_cnt_restricted` (mpc.py:274-315) for one problem, loc [4, N+1, 3]."""
    u = np.unique(loc, axis=1)
    nz = np.argwhere(np.all(u != np.zeros(3), axis=-1))
    count = np.bincount(nz[:, 1]) if len(nz) else np.zeros(0, dtype=int)
    if len(count) > 0:
        first = int(np.argmax(count))
        last = len(count) - int(np.argmax(count[::-1])) - 1
        c_first, c_last = u[:, first, :].mean(axis=0), u[:, last, :].mean(axis=0)
    else:
        c_first, c_last = loc[:, 0, :].mean(axis=0), loc[:, -1, :].mean(axis=0)
    ref, ref_e = np.zeros(12), np.zeros(12)
    ref[:2] = 0.35 * c_first[:2] + 0.65 * c_last[:2]
    ref_e[:2] = c_last[:2]
    ref[2] = ref_e[2] = nom_height
    return ref, ref_e


def make_batch(B: int, seed: int = 0, gaits: Sequence[str] = ("trot",), n_nodes: int = 25,
               dt_nodes: float = 0.04, perturb: float = 1.0, robot: str = "go2",
               restrict_cnt: bool = False, push: bool = False, kin=None) -> Dict[str, np.ndarray]:
    """B independent OCPs: perturbed x0, random gait phase, random velocity goal.

    restrict_cnt (BASELINE config 4, without time optimisation): Raibert footholds per problem
    (`contact_planner.py:265-322`), contact-patch restriction rows and foot-target costs switched on,
    references centred on the planned contacts; needs `kin(qv) -> (feet, com, hg)` (the solver's
    kinematics) for the hip offsets and the CoM.
    push (config 5): the base velocity carries the impulse of a 50-70 N push held for 0.2-0.4 s in a
    random horizontal direction (`Behavior_Cloning/examples/cfgs/bc_experimental.yaml:32-35`).

    Returns qv [B,36], W [B,76], W_e [B,52], yref [B,N,76], yref_e [B,52], p [B,N+1,44],
    i_node [B] int32, x_guess [B,N+1,42] (q,v repeated, h filled by the caller or zero),
    u_guess [B,N,30] (zero accelerations, weight shared by the stance feet), gait_id [B].
    """
    rng = np.random.default_rng(seed)
    N = n_nodes
    v_des = np.zeros((B, 3))
    v_des[:, 0] = rng.uniform(0.0, 0.3, B)
    v_des[:, 1] = rng.uniform(-0.1, 0.1, B)
    gait_id = rng.integers(0, len(gaits), B)
    i_node = rng.integers(0, 12, B).astype(np.int32)
    # feet in contact at the current node (node 0 of the horizon; every foot when the cycle starts, solver.py:198-199)
    stance = np.ones((B, 4), dtype=bool)
    for g, name in enumerate(gaits):
        idx = np.nonzero(gait_id == g)[0]
        if idx.size:
            gait = get_quadruped_config(name, robot, fallback_to_trot_cost=True, fix_swapped_heights=True)[0]
            planner = ContactPlanner(FEET, dt_nodes, gait)
            stance[idx] = np.stack([planner.get_contacts(int(i), 1)[:, 0] for i in i_node[idx]]).astype(bool)
    stance[i_node == 0] = True
    qv = sample_initial_states(rng, P.nominal_state(STAND_HEIGHT), stance, perturb)
    if push:
        dv = rng.uniform(50.0, 70.0, B) * rng.uniform(0.2, 0.4, B) / MASS
        ang = rng.uniform(-np.pi, np.pi, B)
        qv[:, 18] += dv * np.cos(ang)
        qv[:, 19] += dv * np.sin(ang)

    W = np.zeros((B, P.NY)); We = np.zeros((B, P.NYE))
    p = np.zeros((B, N + 1, P.NP))
    yref = np.zeros((B, N, P.NY)); yref_e = np.zeros((B, P.NYE))
    u_guess = np.zeros((B, N, P.NU))
    # compact planner outputs (what QuadrupedAcadosSolver.init receives; nmpc_init_tick expands them on the device)
    c_ref, c_ref_e = np.zeros((B, 12)), np.zeros((B, 12))
    c_cnt, c_peak = np.zeros((B, 4, N + 1), dtype=np.int8), np.zeros((B, 4, N + 1), dtype=np.int8)
    c_loc = np.zeros((B, 4, N + 1, 3)) if restrict_cnt else None
    c_step = np.zeros(B)
    p_gain, cnt_radius = None, 0.0
    if restrict_cnt:
        if kin is None:
            raise ValueError("restrict_cnt needs the solver's kinematics (kin=solver.kinematics)")
        feet0, _, _ = kin(P.nominal_state(0.0)[None])
        hip = np.array(feet0[0], dtype=float)
        hip[:, 2] = 0.0
        _, com, _ = kin(qv)
    for g, name in enumerate(gaits):
        idx = np.nonzero(gait_id == g)[0]
        if idx.size == 0:
            continue
        gait, opt, cost = get_quadruped_config(name, robot, fallback_to_trot_cost=True, fix_swapped_heights=True)
        planner = ContactPlanner(FEET, dt_nodes, gait)
        Wg, Weg = P.pack_weights(cost, restrict_cnt)
        W[idx], We[idx] = Wg, Weg
        cnt = np.stack([planner.get_contacts(int(i), N + 1) for i in i_node[idx]])
        peak = np.stack([planner.get_peaks(int(i), N + 1) for i in i_node[idx]])
        loc = None
        if restrict_cnt:       # LocomotionMPC.__init__ / optimize with contact_planner="raibert" (mpc.py:76-84, 332-341)
            rp = RaiberContactPlanner(FEET, dt_nodes, gait, hip.copy(), y_offset=0.02, x_offset=0.04,
                                      foot_size=0.0085, cache_cnt=False)
            loc = np.zeros((idx.size, 4, N + 1, 3))
            ref, ref_e = np.zeros((idx.size, 12)), np.zeros((idx.size, 12))
            for n, b in enumerate(idx):
                rp.set_state(qv[b, :3], qv[b, 18:21], qv[b, 3:6][::-1], com[b], v_des[b], 0.0)
                loc[n] = rp.get_locations(int(i_node[b]), N + 1)
                ref[n], ref_e[n] = _restricted_base_ref(loc[n], gait.nom_height)
        else:
            tracking = np.zeros((idx.size, 12))
            tracking[:, :2] = qv[idx, :2]
            tracking[:, 3] = qv[idx, 3]
            ref, ref_e = P.base_ref_vel_tracking(qv[idx, :18], v_des[idx], 0.0, tracking, gait.nom_height)
        p[idx] = P.pack_params(cnt, cost, 0.0, peak, restrict_cnt, loc)
        yr, yre = P.pack_reference(ref, ref_e, P.GO2_JOINT_HOME, gait.step_height, N, loc)
        yref[idx], yref_e[idx] = yr, yre
        c_ref[idx], c_ref_e[idx], c_cnt[idx], c_peak[idx], c_step[idx] = ref, ref_e, cnt, peak, gait.step_height
        if restrict_cnt:
            c_loc[idx] = loc
        g_gain = np.asarray(cost.W_foot_pos_constr_stab, dtype=float)
        assert p_gain is None or np.array_equal(p_gain, g_gain), "per-gait contact gains differ"
        p_gain, cnt_radius = g_gain, float(cost.cnt_radius)
        act = cnt[:, :, :N].transpose(0, 2, 1).astype(float)          # [b, k, foot]
        n_act = np.maximum(act.sum(-1, keepdims=True), 1.0)
        fz = act * (MASS * 9.81) / n_act
        u_guess[idx, :, 18 + 2::3] = fz
    x_guess = np.zeros((B, N + 1, P.NX))
    x_guess[:, :, :36] = qv[:, None, :]
    return dict(qv=qv, W=W, W_e=We, yref=yref, yref_e=yref_e, p=p, i_node=i_node, x_guess=x_guess,
                u_guess=u_guess, gait_id=gait_id.astype(np.int32), v_des=v_des,
                compact=dict(base_ref=c_ref, base_ref_e=c_ref_e, joint_ref=np.array(P.GO2_JOINT_HOME, dtype=float),
                             step_height=c_step, cnt_sequence=c_cnt, swing_peak=c_peak, cnt_locations=c_loc,
                             p_gain=p_gain, cnt_radius=cnt_radius, height_offset=0.0, restrict_cnt=bool(restrict_cnt)))


def take(batch: Dict[str, np.ndarray], idx) -> Dict[str, np.ndarray]:
    """Sub-batch (rows `idx`) of a `make_batch` result, the compact planner outputs included."""
    B = len(batch["qv"])
    pick = lambda v: v[idx] if isinstance(v, np.ndarray) and v.shape[:1] == (B,) else v
    out = {k: pick(v) for k, v in batch.items() if k != "compact"}
    per_problem = ("base_ref", "base_ref_e", "step_height", "cnt_sequence", "swing_peak", "cnt_locations")
    out["compact"] = {k: (v[idx] if k in per_problem and v is not None else v) for k, v in batch["compact"].items()}
    return out


def load_compact(solver, batch: Dict[str, np.ndarray]) -> None:
    """The same problem through the device-side `init` (product NmpcSolver only): compact planner outputs in,
    expansion into p / yref, x0 with h0, stance-foot pinning and the cold first guess all on the GPU."""
    solver.set_weights(batch["W"], batch["W_e"])
    solver.init_tick(**batch["compact"])
    solver.set_x0(batch["qv"])
    solver.cold_guess()                      # from the planner's mask, before node 0 is forced to stance for i_node = 0
    solver.pin_stance_feet(batch["i_node"])
    solver.reset_duals()


def load_into(solver, batch: Dict[str, np.ndarray], hg: Optional[np.ndarray] = None) -> None:
    """Push a batch into a solver object (product NmpcSolver or the oracle's OracleSolver)."""
    solver.set_weights(batch["W"], batch["W_e"])
    solver.set_yref(batch["yref"], batch["yref_e"])
    solver.set_params(batch["p"])
    xg = batch["x_guess"].copy()
    if hg is not None:
        xg[:, :, 36:] = hg[:, None, :]
    ug = batch["u_guess"]
    if getattr(solver, "NU", ug.shape[-1]) == ug.shape[-1] + 1:
        # the oracle's time-opt build carries dt as a 31st input column: nominal node duration (solver.py:189-190)
        ug = np.concatenate([ug, np.full(ug.shape[:2] + (1,), float(solver.opts.dt_nodes))], axis=2)
    solver.set_guess(xg, ug)
    solver.set_x0(batch["qv"])
    solver.pin_stance_feet(batch["i_node"])
    solver.reset_duals()
