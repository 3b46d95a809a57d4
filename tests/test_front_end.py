"""Reference-facing front-end (`QuadrupedAcadosSolver`, `BatchedQuadrupedSolver`, `LocomotionMPC`;
reference: mpc_controller/utils/solver.py, mpc_controller/mpc.py) on the CPU oracle backend:
name -> slice mapping, per-tick packing, warm-start shift, batched == unbatched, open-loop cadence.
The same scenarios run against the CUDA library in test_gpu_front_end.py."""
import numpy as np
import pytest

from iterative_learning_nmpc_b200 import problem as P
from iterative_learning_nmpc_b200 import workloads as WL
from iterative_learning_nmpc_b200.mpc_controller.config.quadruped.utils import get_quadruped_config
from iterative_learning_nmpc_b200.mpc_controller.utils.contact_planner import ContactPlanner
from oracle_backend import use_oracle_backend

FEET = WL.FEET


@pytest.fixture()
def backend(monkeypatch, oracle_mod):
    use_oracle_backend(monkeypatch)
    return oracle_mod


def _scenario(seed, i_node, gait="trot"):
    rng = np.random.default_rng(seed)
    gcfg, opt, cost = get_quadruped_config(gait, "go2")
    qv = P.nominal_state(WL.STAND_HEIGHT)
    qv[2] += rng.uniform(-0.03, 0.03)
    qv[3:6] += rng.uniform(-0.1, 0.1, 3)
    qv[6:18] += rng.uniform(-0.2, 0.2, 12)
    qv[18:] += rng.uniform(-0.5, 0.5, 18)
    planner = ContactPlanner(FEET, opt.get_dt_nodes(), gcfg)
    cnt = planner.get_contacts(i_node, opt.n_nodes + 1)
    peak = planner.get_peaks(i_node, opt.n_nodes + 1)
    base_ref = np.zeros(12); base_ref[:3] = [qv[0] + 0.1, qv[1], gcfg.nom_height]; base_ref[6] = 0.2
    base_ref_e = base_ref.copy(); base_ref_e[0] += 0.1
    return dict(gcfg=gcfg, opt=opt, cost=cost, q=qv[:18].copy(), v=qv[18:].copy(), cnt=cnt, peak=peak,
                base_ref=base_ref, base_ref_e=base_ref_e, i_node=i_node)


def _front_end(sc, batch=None):
    from iterative_learning_nmpc_b200.mpc_controller.utils.solver import BatchedQuadrupedSolver, QuadrupedAcadosSolver
    if batch is None:
        return QuadrupedAcadosSolver(None, FEET, sc["opt"], sc["cost"], print_info=False)
    return BatchedQuadrupedSolver(batch, None, FEET, sc["opt"], sc["cost"], print_info=False)


def test_init_solve_matches_packed_path(backend):
    """The front-end's name-keyed dictionaries end up in the same flat arrays as problem.py's
    vectorised packers (two independent codings of solver.py:108-275), and the solve agrees."""
    O = backend
    sc = _scenario(0, 3)
    s = _front_end(sc)
    s.dyn.update_pin(sc["q"], sc["v"])
    s.init(0, sc["q"], sc["v"], sc["base_ref"], sc["base_ref_e"], P.GO2_JOINT_HOME, sc["gcfg"].step_height,
           sc["cnt"], None, sc["peak"])
    q_sol, v_sol, a_sol, f_sol, dt_sol = s.solve()
    assert q_sol.shape == (26, 18) and v_sol.shape == (26, 18) and a_sol.shape == (25, 18)
    assert f_sol.shape == (25, 4, 3) and np.allclose(dt_sol, 0.04)

    qv = np.concatenate((sc["q"], sc["v"]))[None]
    opts = O.default_opts(max_iter=sc["opt"].max_iter, max_qp_iter=sc["opt"].max_qp_iter, warm_start_qp=1,
                          reg_eps=sc["cost"].reg_eps, reg_eps_e=sc["cost"].reg_eps_e)
    ref = O.OracleSolver(1, opts)
    W, We = P.pack_weights(sc["cost"])
    yref, yref_e = P.pack_reference(sc["base_ref"], sc["base_ref_e"], P.GO2_JOINT_HOME, sc["gcfg"].step_height, 25)
    p = P.pack_params(sc["cnt"], sc["cost"], 0.0, sc["peak"])
    ref.set_weights(W, We); ref.set_yref(yref, yref_e); ref.set_params(p)
    ref.set_x0(qv)
    # the reference pushes its `states` dict after fixing x0, and column 0 of that dict is never
    # touched by init(): the first linearisation point of node 0 is the all-zero guess
    ref.set_guess(np.zeros((1, 26, 42)), np.zeros((1, 25, 30)))
    ref.pin_stance_feet(np.array([0], dtype=np.int32))
    ref.reset_duals()
    st = ref.solve(1)
    x, u, _, _ = ref.get_solution()
    assert int(st[0]) == int(s.status[0])
    np.testing.assert_allclose(q_sol, x[0, :, :18], rtol=0, atol=1e-12)
    np.testing.assert_allclose(s.h_sol, x[0, :, 36:], rtol=0, atol=1e-12)
    np.testing.assert_allclose(a_sol, u[0, :, :18], rtol=0, atol=1e-12)
    np.testing.assert_allclose(f_sol.reshape(25, 12), u[0, :, 18:], rtol=0, atol=1e-12)


def test_states_dict_layout_and_x0(backend):
    sc = _scenario(1, 0)
    s = _front_end(sc)
    assert s.states[s.dyn.q.name].shape == (18, 26) and s.inputs[s.dyn.a.name].shape == (18, 25)
    assert s.params[s.dyn.feet[2].plane_point.name].shape == (3, 26)
    assert s.cost_ref[s.dyn.base_cost.name].shape == (12, 25)
    assert set(s.data["W"]) >= {s.dyn.base_cost.name, s.dyn.joint_cost.name, s.dyn.acc_cost.name,
                                s.dyn.swing_cost.name, s.dyn.feet[0].f_reg.name, s.dyn.feet[0].pos_cost.name}
    assert f"f_{FEET[1]}_{s.dyn.name}" in s.inputs          # the key the reference builds at solver.py:319
    s.dyn.update_pin(sc["q"], sc["v"])
    s.init(0, sc["q"], sc["v"], sc["base_ref"], None, P.GO2_JOINT_HOME, 0.05, sc["cnt"], None, sc["peak"])
    s.solve()
    np.testing.assert_allclose(s.q_sol_euler[0], sc["q"], atol=1e-9)      # x0 is a hard constraint
    np.testing.assert_allclose(s.h_sol[0], s.dyn.pin_data.hg.np, atol=1e-9)


def test_setup_initial_feet_pos_pins_stance_until_next_swing(backend):
    sc = _scenario(2, 4)
    s = _front_end(sc)
    s.dyn.update_pin(sc["q"], sc["v"])
    s.init(4, sc["q"], sc["v"], sc["base_ref"], None, P.GO2_JOINT_HOME, 0.05, sc["cnt"], None, sc["peak"])
    feet = s.dyn.get_feet_position_w()
    for i, foot in enumerate(s.dyn.feet):
        pp = s.params[foot.plane_point.name]
        if sc["cnt"][i, 0]:
            nxt = int(np.argmin(sc["cnt"][i]))
            np.testing.assert_allclose(pp[:, :nxt], np.repeat(feet[i][:, None], nxt, axis=1))
            assert np.all(pp[:, nxt:] == 0.0)
        else:
            assert np.all(pp == 0.0)


def test_warm_start_shift_host_equals_device(backend):
    """Second tick: the host-side shift of q/v/h/a/f (solver.py:290-342) and the device-side
    nmpc_shift_warm_start put the same iterate into the solver."""
    sc = _scenario(3, 0)
    s = _front_end(sc)
    s.dyn.update_pin(sc["q"], sc["v"])
    s.init(0, sc["q"], sc["v"], sc["base_ref"], None, P.GO2_JOINT_HOME, 0.05, sc["cnt"], None, sc["peak"])
    q1, v1, a1, f1, _ = (a.copy() for a in s.solve())
    x_prev, u_prev, _, _ = s.handle.get_solution()
    sc2 = _scenario(3, 2)
    s.dyn.update_pin(q1[2], v1[2])
    s.init(2, q1[2], v1[2], sc["base_ref"], None, P.GO2_JOINT_HOME, 0.05, sc2["cnt"], None, sc2["peak"])
    x_now, u_now, _, _ = s.handle.get_solution()          # iterate as pushed, before the solve
    np.testing.assert_allclose(x_now[0, 1:24, :36], x_prev[0, 3:26, :36], atol=0)
    np.testing.assert_allclose(u_now[0, :23], u_prev[0, 2:], atol=0)
    assert np.all(u_now[0, 23:, 18:] == 0.0)
    assert s.last_node == 2
    s.solve()
    assert int(s.status[0]) in (0, 2)


def test_batched_equals_unbatched(backend):
    scs = [_scenario(10 + b, i) for b, i in enumerate((0, 5, 9))]
    sols = []
    for sc in scs:
        s = _front_end(sc)
        s.dyn.update_pin(sc["q"], sc["v"])
        s.init(0, sc["q"], sc["v"], sc["base_ref"], sc["base_ref_e"], P.GO2_JOINT_HOME, 0.05, sc["cnt"], None, sc["peak"])
        sols.append([a.copy() for a in s.solve()] + [int(s.status[0])])
    sb = _front_end(scs[0], batch=3)
    sb.cold_guess = "zero"
    stack = lambda k: np.stack([sc[k] for sc in scs])
    sb.init(np.zeros(3, dtype=int), stack("q"), stack("v"), stack("base_ref"), stack("base_ref_e"), P.GO2_JOINT_HOME,
            0.05, stack("cnt"), None, stack("peak"))
    qb, vb, ab, fb, dtb = sb.solve()
    assert qb.shape == (3, 26, 18) and fb.shape == (3, 25, 4, 3) and dtb.shape == (3, 25)
    for b in range(3):
        np.testing.assert_allclose(qb[b], sols[b][0], atol=1e-12)
        np.testing.assert_allclose(ab[b], sols[b][2], atol=1e-12)
        np.testing.assert_allclose(fb[b], sols[b][3], atol=1e-12)
        assert int(sb.status[b]) == sols[b][5]


def test_batched_device_warm_start_per_problem_nodes(backend):
    sc = _scenario(20, 0)
    sb = _front_end(sc, batch=2)
    q = np.stack([sc["q"], sc["q"]]); v = np.stack([sc["v"], sc["v"]])
    cnt = np.stack([sc["cnt"], sc["cnt"]]); peak = np.stack([sc["peak"], sc["peak"]])
    sb.init(np.array([0, 0]), q, v, sc["base_ref"][None].repeat(2, 0), None, P.GO2_JOINT_HOME, 0.05, cnt, None, peak)
    sb.solve()
    x_prev, u_prev, _, _ = sb.handle.get_solution()
    sb.init(np.array([1, 3]), q, v, sc["base_ref"][None].repeat(2, 0), None, P.GO2_JOINT_HOME, 0.05, cnt, None, peak)
    x_now, u_now, _, _ = sb.handle.get_solution()
    np.testing.assert_allclose(x_now[0, 1:25, :36], x_prev[0, 2:26, :36], atol=0)
    np.testing.assert_allclose(x_now[1, 1:23, :36], x_prev[1, 4:26, :36], atol=0)
    np.testing.assert_allclose(u_now[1, :22], u_prev[1, 3:], atol=0)


def test_locomotion_mpc_open_loop_and_first_solve_settings(backend):
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    mpc = LocomotionMPC(None, FEET, "go2", gait_name="trot", print_info=False, solve_async=False)
    mpc.set_command(np.array([0.15, 0.0, 0.0]), 0.0)
    q_mj = np.zeros(19); q_mj[2] = WL.STAND_HEIGHT; q_mj[3] = 1.0; q_mj[7:] = P.GO2_JOINT_HOME
    traj = mpc.open_loop(q_mj, np.zeros(18), 0.0855)
    assert traj.shape == (86, 19) and np.isfinite(traj).all()
    assert mpc.sim_step == 86 and mpc.current_opt_node == 2 and not mpc.first_solve
    assert len(mpc.solver.timings["solve"]) == 3                      # replans at steps 0, 40, 80
    assert mpc.solver._opts.max_iter == mpc.config_opt.max_iter        # restored after the first solve
    assert abs(mpc.solver._opts.nlp_tol - mpc.config_opt.nlp_tol) < 1e-15
    assert np.abs(traj[:, 2] - WL.STAND_HEIGHT).max() < 0.08           # stays near the nominal height
    assert traj[-1, 0] > 0.0                                           # and moves forward


def test_compute_torques_dof_sync(backend):
    from types import SimpleNamespace
    from iterative_learning_nmpc_b200.mpc_controller.mpc import LocomotionMPC
    mpc = LocomotionMPC(None, FEET, "go2", print_info=False, solve_async=False)
    q_mj = np.zeros(19); q_mj[2] = WL.STAND_HEIGHT; q_mj[3] = 1.0; q_mj[7:] = P.GO2_JOINT_HOME
    data = SimpleNamespace(time=0.0, qpos=q_mj, qvel=np.zeros(18))
    mpc.compute_torques_dof(data)                 # first call solves synchronously and adopts the plan
    assert not mpc.first_solve and mpc.torques_dof.shape == (18,)
    data.time = 0.001
    mpc.compute_torques_dof(data)
    tau = mpc.torques_dof[-12:]
    assert np.isfinite(tau).all() and np.abs(tau).max() < 60.0
    # standing still on four... two stance feet of a trot: knees carry weight => non-trivial feed-forward
    assert np.abs(tau).max() > 0.5


def test_batched_locomotion_mpc_matches_unbatched_rollouts(backend):
    """`BatchedLocomotionMPC.open_loop` (B rollouts in lockstep) against B runs of `LocomotionMPC.open_loop`,
    both on the oracle backend: same visited configurations, different start states and commands."""
    from iterative_learning_nmpc_b200 import dataset as DS
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC, LocomotionMPC
    from iterative_learning_nmpc_b200.mpc_controller.utils.dynamics import QuadrupedDynamics
    B, T = 2, 0.0415
    rng = np.random.default_rng(8)
    q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
    v0 = np.zeros((B, 18))
    cmds = np.array([[0.15, 0.0, 0.0], [0.1, 0.05, 0.0]])
    single = []
    for b in range(B):
        mpc = LocomotionMPC(None, FEET, "go2", gait_name="trot", print_info=False, solve_async=False,
                            joint_ref=P.GO2_JOINT_HOME)
        mpc.solver.cold_guess = "hold"
        mpc.set_command(cmds[b], 0.0)
        q_mj, v_mj = QuadrupedDynamics.convert_to_mujoco(q0[b], v0[b])
        single.append(mpc.open_loop(q_mj, v_mj, T))
    bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME)
    bm.set_command(cmds, 0.0)
    qs, vs, taus = bm.open_loop(q0, v0, T, record_torques=True)
    assert qs.shape == (len(single[0]), B, 18) and taus.shape == (len(single[0]), B, 12)
    assert len(bm.status_log) == -(-len(single[0]) // bm.replanning_steps)    # one solve per replanning tick
    for b in range(B):
        q_mj, _ = DS.euler_to_mujoco(qs[:, b], vs[:, b])
        err = np.abs(q_mj - single[b]).max()
        assert err < 1e-6, (b, err)


def test_batched_raibert_mpc_matches_unbatched_rollouts(backend):
    """`BatchedLocomotionMPC(contact_planner="raibert")` against `LocomotionMPC(contact_planner="raibert")` run one
    rollout at a time (mpc.py:76-90, 274-315, 332-341), both on the oracle backend: the batched tick plans through
    `plan_tick_raibert` (here the checker's loop restatement), the unbatched one through RaiberContactPlanner and
    compute_base_ref_cnt_restricted as the reference does."""
    from iterative_learning_nmpc_b200 import dataset as DS
    from iterative_learning_nmpc_b200.mpc_controller.mpc import BatchedLocomotionMPC, LocomotionMPC
    from iterative_learning_nmpc_b200.mpc_controller.utils.dynamics import QuadrupedDynamics
    B, T = 2, 0.0415
    rng = np.random.default_rng(9)
    q0 = np.tile(P.nominal_state(WL.STAND_HEIGHT)[:18], (B, 1)); q0[:, 6:] += rng.uniform(-0.05, 0.05, (B, 12))
    q0[1, 3] = 0.4
    v0 = np.zeros((B, 18))
    cmds = np.array([[0.15, 0.0, 0.0], [0.1, 0.05, 0.0]])
    single = []
    for b in range(B):
        mpc = LocomotionMPC(None, FEET, "go2", gait_name="trot", print_info=False, solve_async=False,
                            joint_ref=P.GO2_JOINT_HOME, contact_planner="raibert")
        mpc.solver.cold_guess = "hold"
        mpc.set_command(cmds[b], 0.2)
        q_mj, v_mj = QuadrupedDynamics.convert_to_mujoco(q0[b], v0[b])
        single.append(mpc.open_loop(q_mj, v_mj, T))
    for on_device in (True, False):
        bm = BatchedLocomotionMPC(B, None, FEET, "go2", gait_name="trot", joint_ref=P.GO2_JOINT_HOME,
                                  contact_planner="raibert")
        bm.plan_on_device = on_device
        bm.set_command(cmds, 0.2)
        qs, vs = bm.open_loop(q0, v0, T)[:2]
        for b in range(B):
            q_mj, _ = DS.euler_to_mujoco(qs[:, b], vs[:, b])
            err = np.abs(q_mj - single[b]).max()
            assert err < 1e-6, (on_device, b, err)
