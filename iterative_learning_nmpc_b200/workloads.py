"""Synthetic OCP batches of BASELINE.json's configurations.

Perturbation ranges follow the reference's data collection
(`Behavior_Cloning/utils/RolloutMPC_shift_phase_percentage.py:466-484`: base height, roll/pitch/
yaw, joints, velocities; `Behavior_Cloning/examples/cfgs/bc_config.yaml:11-16`: velocity goals).
The nominal state is the Go2 home pose standing on its feet (the synthetic model's feet are on
the ground at base height ~0.265 m).
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np

from . import problem as P
from .mpc_controller.config.quadruped.utils import get_quadruped_config
from .mpc_controller.utils.contact_planner import ContactPlanner, RaiberContactPlanner

FEET = ["FL_foot", "FR_foot", "RL_foot", "RR_foot"]
STAND_HEIGHT = 0.2648058464833031     # base height at which the home pose has its feet at z = 0
MASS = 15.017


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
    qv = np.tile(P.nominal_state(STAND_HEIGHT), (B, 1))
    s = perturb
    qv[:, 2] += rng.uniform(-0.05, 0.05, B) * s
    qv[:, 3] += rng.uniform(-0.25, 0.25, B) * s            # yaw
    qv[:, 4] += rng.uniform(-0.1, 0.1, B) * s              # pitch
    qv[:, 5] += rng.uniform(-0.1, 0.1, B) * s              # roll
    qv[:, 6:18] += rng.uniform(-0.3, 0.3, (B, 12)) * s
    qv[:, 18:24] += rng.uniform(-0.5, 0.5, (B, 6)) * s     # base twist
    qv[:, 24:36] += rng.uniform(-2.0, 2.0, (B, 12)) * s    # joint velocities
    if push:
        dv = rng.uniform(50.0, 70.0, B) * rng.uniform(0.2, 0.4, B) / MASS
        ang = rng.uniform(-np.pi, np.pi, B)
        qv[:, 18] += dv * np.cos(ang)
        qv[:, 19] += dv * np.sin(ang)
    v_des = np.zeros((B, 3))
    v_des[:, 0] = rng.uniform(0.0, 0.3, B)
    v_des[:, 1] = rng.uniform(-0.1, 0.1, B)
    gait_id = rng.integers(0, len(gaits), B)
    i_node = rng.integers(0, 12, B).astype(np.int32)

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
    solver.set_guess(xg, batch["u_guess"])
    solver.set_x0(batch["qv"])
    solver.pin_stance_feet(batch["i_node"])
    solver.reset_duals()
