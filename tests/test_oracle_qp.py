"""Oracle OCP-QP interior point (Riccati) vs dense numpy KKT algebra (CPU only).

Stands in for golden vectors of HPIPM, which the reference does not hold: the
Riccati recursion is checked against a dense KKT solve, and converged IPM solutions against
independently computed KKT residuals.
"""
import numpy as np
import pytest

import qp_dense as QD
from iterative_learning_nmpc_b200 import workloads as WL

B = 3


@pytest.fixture(scope="module")
def qp_case(oracle_mod):
    O = oracle_mod
    batch = WL.make_batch(B, seed=11)
    _, _, hg = O.kinematics(batch["qv"])
    s = O.OracleSolver(B, O.default_opts())
    WL.load_into(s, batch, hg)
    qp = s.export_qp()
    rng = np.random.default_rng(5)
    dx0 = rng.normal(0, 1e-2, (B, 42))
    return s, qp, dx0


def test_exported_qp_shapes_and_symmetry(qp_case):
    _, qp, _ = qp_case
    N = 25
    assert qp["BAbt"].shape == (B, N, 73, 42)
    assert qp["RSQrq"].shape == (B, N + 1, 73, 72)
    assert qp["DCt"].shape == (B, N + 1, 72, 42)
    H = qp["RSQrq"][:, :, :72, :]
    np.testing.assert_allclose(H, np.swapaxes(H, -1, -2), atol=1e-12)
    # Gauss-Newton Hessians are positive definite thanks to reg_eps
    for k in (0, 7, N):
        w = np.linalg.eigvalsh(H[0, k])
        assert w.min() > 0
    # no input at the terminal node: friction rows masked off, dynamics absent
    assert (qp["mask"][:, N, :20] == 0).all()
    # node 0 carries only input constraints (x0 is fixed)
    assert (qp["mask"][:, 0, 20:42] == 0).all() and (qp["mask"][:, 0, 42 + 20:] == 0).all()


def test_riccati_equals_dense_kkt_solve(oracle_mod, qp_case):
    """With every inequality masked off, one IPM iteration is a single Riccati solve of the
    equality-constrained QP and must reproduce the dense KKT solution."""
    O = oracle_mod
    s, qp, dx0 = qp_case
    qp0 = dict(qp)
    qp0["mask"] = np.zeros_like(qp["mask"])
    s.set_opts(O.default_opts(max_qp_iter=1, qp_tol=1e-300, warm_start_qp=0))
    r = s.solve_qp(qp0, dx0)
    for b in range(B):
        st = QD.unpack(qp0, b)
        z, pi = QD.solve_equality_qp(st, dx0[b])
        scale = np.abs(z).max()
        np.testing.assert_allclose(r["dz"][b], z, atol=1e-8 * scale)
        np.testing.assert_allclose(r["pi"][b], pi, atol=1e-8 * max(np.abs(pi).max(), 1.0))


@pytest.mark.parametrize("drop_equalities,tol,iters", [(True, 1e-8, 60), (False, 1e-3, 40)])
def test_converged_ipm_satisfies_dense_kkt(oracle_mod, qp_case, drop_equalities, tol, iters):
    """KKT residuals of the returned (z, pi, lam), recomputed densely.  The stance and momentum
    rows are equalities posed as lg == ug (the way acados hands h(x) = 0 to HPIPM): their slacks
    go to zero with the residuals, so with them the IPM is only asked for 1e-3; with the pure
    inequality rows (friction pyramids) it is asked for 1e-8."""
    O = oracle_mod
    s, qp, dx0 = qp_case
    qp = {k: v.copy() for k, v in qp.items()}
    if drop_equalities:
        both = (qp["mask"][..., :42] > 0.5) & (qp["mask"][..., 42:] > 0.5)
        qp["mask"][..., :42][both] = 0
        qp["mask"][..., 42:][both] = 0
    s.set_opts(O.default_opts(max_qp_iter=iters, qp_tol=tol, warm_start_qp=0))
    r = s.solve_qp(qp, dx0)
    assert (r["qp_status"] == 0).all(), (r["qp_status"], r["qp_iter"], r["qp_res"])
    for b in range(B):
        st = QD.unpack(qp, b)
        rs, re, ri, rc = QD.kkt_residuals(st, dx0[b], r["dz"][b], r["pi"][b], r["lam"][b])
        assert rs < 10 * tol and re < 10 * tol and ri < 10 * tol and rc < 10 * tol, (rs, re, ri, rc)
        # the solver's own residual report agrees with the dense computation
        np.testing.assert_allclose(r["qp_res"][b, 0], rs, rtol=1e-3, atol=1e-9)
        # multipliers and slacks stay strictly positive
        m = qp["mask"][b] > 0.5
        assert (r["lam"][b][m] > 0).all() and (r["t"][b][m] > 0).all()


def test_iteration_cap_and_status(oracle_mod, qp_case):
    O = oracle_mod
    s, qp, dx0 = qp_case
    for cap in (0, 1, 6):
        s.set_opts(O.default_opts(max_qp_iter=cap, qp_tol=1e-9, warm_start_qp=0))
        r = s.solve_qp(qp, dx0)
        assert (r["qp_iter"] == cap).all()
        assert (r["qp_status"] == 1).all()            # HPIPM: max-iter
    # a QP that is already solved exits at iteration 0 with success: no constraints, zero gradient
    qp0 = {k: v.copy() for k, v in qp.items()}
    qp0["mask"][:] = 0
    qp0["RSQrq"][:, :, 72, :] = 0
    qp0["BAbt"][:, :, 72, :] = 0
    s.set_opts(O.default_opts(max_qp_iter=6, qp_tol=1e-9, warm_start_qp=0))
    r = s.solve_qp(qp0, np.zeros((B, 42)))
    assert (r["qp_status"] == 0).all() and (r["qp_iter"] == 0).all()
    assert np.abs(r["dz"]).max() == 0


def test_nan_is_reported_not_propagated(oracle_mod, qp_case):
    O = oracle_mod
    s, qp, dx0 = qp_case
    bad = {k: v.copy() for k, v in qp.items()}
    bad["RSQrq"][1, 3, 5, 5] = np.nan
    s.set_opts(O.default_opts(max_qp_iter=6, qp_tol=1e-2, warm_start_qp=0))
    r = s.solve_qp(bad, dx0)
    assert r["qp_status"][1] == 3
    assert r["qp_status"][0] in (0, 1) and r["qp_status"][2] in (0, 1)
    assert np.isfinite(r["dz"][0]).all() and np.isfinite(r["dz"][2]).all()


def test_structured_variant_reproduces_dense_restatement(oracle_mod):
    """oracle/nmpc_oracle_fast.h (the timed CPU baseline and the host statement of the kernel's arithmetic) against
    the dense restatement: same status / iteration counts, iterates equal to rounding on single QPs (RTI), cold and
    with the conditional predictor-corrector path; a short hot-started chain inside 1e-6 for nine problems in ten."""
    from parity_metric import chain_ok, rel_err
    from iterative_learning_nmpc_b200 import workloads as WL
    O = oracle_mod
    B = 24
    batch = WL.make_batch(B, seed=46, gaits=("trot", "pace", "bound"))
    _, _, hg = O.kinematics(batch["qv"])

    def run(fast, **opts):
        s = O.OracleSolver(B, O.default_opts(**opts))
        s.set_fast(fast)
        WL.load_into(s, batch, hg)
        st = s.solve(0)
        return st, s.get_solution(), s.get_stats()

    for opts in (dict(real_time_it=1), dict(real_time_it=1, max_qp_iter=2)):
        (s0, sol0, st0), (s1, sol1, st1) = run(False, **opts), run(True, **opts)
        np.testing.assert_array_equal(s0, s1)
        np.testing.assert_array_equal(st0["qp_iter"], st1["qp_iter"])
        for a, b in zip(sol1, sol0):
            assert rel_err(a, b) < 1e-9
    opts = dict(real_time_it=0, max_iter=3, warm_start_qp=2)
    (s0, sol0, st0), (s1, sol1, st1) = run(False, **opts), run(True, **opts)
    np.testing.assert_array_equal(s0, s1)
    np.testing.assert_array_equal(st0["qp_ncond"], st1["qp_ncond"])
    assert st0["qp_ncond"].sum() > 0
    for name, a, b in zip(("x", "u", "pi", "lam"), sol1, sol0):
        assert chain_ok("fast vs dense " + name, a, b, 1e-6)


def test_time_opt_build_tends_to_the_fixed_dt_solution_with_the_weight(oracle_mod):
    """The oracle's enable_time_opt build (nu = 31: dt_k a 31st input, weight on dt - dt_nodes, bounds) against the
    default build: as the weight grows the node durations are pinned to dt_nodes and the converged QP solution tends
    to the fixed-dt one, linearly in 1 / weight."""
    O = oracle_mod
    B = 4
    batch = WL.make_batch(B, seed=3)
    hg = O.kinematics(batch["qv"])[2]
    base = O.OracleSolver(B, O.default_opts(real_time_it=1, max_qp_iter=40, qp_tol=1e-5))
    WL.load_into(base, batch, hg)
    assert (base.solve(0) == 0).all() and (base.get_stats()["qp_status"] == 0).all()
    x0, u0, _, _ = base.get_solution()
    dev = []
    for w in (1e8, 1e10):
        s = O.OracleSolver(B, O.default_opts(real_time_it=1, enable_time_opt=1, max_qp_iter=40, qp_tol=1e-5,
                                             time_opt_weight=w))
        WL.load_into(s, batch, hg)
        assert (s.solve(0) == 0).all() and (s.get_stats()["qp_status"] == 0).all()
        x, u, _, _ = s.get_solution()
        dev.append((np.abs(x - x0).max(), np.abs(u[..., :30] - u0).max(), np.abs(u[..., 30] - 0.04).max()))
    assert dev[1][0] < 1e-3 and dev[1][2] < 1e-5
    for a, b in zip(dev[0], dev[1]):
        assert 50 < a / b < 200                      # 1 / weight


def test_time_opt_converged_qp_respects_the_dt_bounds_and_is_a_kkt_point(oracle_mod):
    O = oracle_mod
    B = 4
    batch = WL.make_batch(B, seed=3)
    hg = O.kinematics(batch["qv"])[2]
    s = O.OracleSolver(B, O.default_opts(real_time_it=1, enable_time_opt=1, max_qp_iter=60, qp_tol=1e-5))
    WL.load_into(s, batch, hg)
    assert (s.solve(0) == 0).all()
    st = s.get_stats()
    assert (st["qp_status"] == 0).all() and (st["qp_res"] < 1e-5).all()
    _, u, _, lam = s.get_solution()
    dt = u[..., 30]
    assert dt.min() >= 0.02 - 1e-6 and dt.max() <= 0.07 + 1e-6 and np.ptp(dt) > 1e-3
    # complementarity of the two dt bounds at the QP solution
    lo, up = lam[:, :-1, 42], lam[:, :-1, 85]
    assert np.abs(lo * (dt - 0.02)).max() < 1e-4 and np.abs(up * (0.07 - dt)).max() < 1e-4
