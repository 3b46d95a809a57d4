"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Tolerance (BASELINE.json north_star): <= 1e-6 relative on the primal/dual iterates
of each QP -- asserted per problem and per block (x, u, pi, lam, dz, t each normalised by that problem's own
block norm, tests/parity_metric.py; the element-wise worst case is printed) --; converged NLP solutions
within nlp_tol; identical per-problem status."""
import numpy as np
import pytest

from iterative_learning_nmpc_b200 import workloads as WL

pytestmark = pytest.mark.gpu

REL = 1e-6
N = 25


def _gpu():
    from iterative_learning_nmpc_b200._lib import default_opts
    from iterative_learning_nmpc_b200.core import NmpcSolver
    return NmpcSolver, default_opts


from parity_metric import chain_ok, rel_err, report  # noqa: E402  per problem, per block (tests/parity_metric.py)


def _pair(O, B, seed, gaits=("trot",), perturb=1.0, workload=None, **opts):
    NmpcSolver, default_opts = _gpu()
    batch = WL.make_batch(B, seed=seed, gaits=gaits, perturb=perturb, **(workload or {}))
    _, _, hg = O.kinematics(batch["qv"])
    so = O.OracleSolver(B, O.default_opts(**opts))
    sg = NmpcSolver(B, default_opts(**opts))
    WL.load_into(so, batch, hg)
    WL.load_into(sg, batch, hg)
    return so, sg, batch


def test_kinematics_parity(oracle_mod):
    NmpcSolver, default_opts = _gpu()
    batch = WL.make_batch(257, seed=2)
    fo, co, ho = oracle_mod.kinematics(batch["qv"])
    sg = NmpcSolver(1, default_opts())
    fg, cg, hg = sg.kinematics(batch["qv"])
    assert rel_err(fg, fo) < 1e-13 and rel_err(cg, co) < 1e-13 and rel_err(hg, ho) < 1e-12


def test_linearisation_parity(oracle_mod):
    """Every block of the exported OCP-QP (dynamics, Gauss-Newton Hessian/gradient, constraint
    Jacobians, bounds, masks) for three gaits."""
    so, sg, _ = _pair(oracle_mod, 12, seed=4, gaits=("trot", "pace", "bound"))
    qo, qg = so.export_qp(), sg.export_qp()
    np.testing.assert_array_equal(qg["mask"], qo["mask"])
    for k in ("BAbt", "RSQrq", "DCt", "lg", "ug"):
        assert rel_err(qg[k], qo[k]) < 1e-11, k


@pytest.mark.parametrize("cap", [0, 1, 2, 3, 6, 8])
def test_qp_iterates_parity(oracle_mod, cap):
    """Primal/dual iterate of the IPM after `cap` iterations, on identical QP data."""
    O = oracle_mod
    B = 6
    so, sg, _ = _pair(O, B, seed=5, max_qp_iter=cap, qp_tol=1e-9, warm_start_qp=0)
    qp = so.export_qp()
    dx0 = np.random.default_rng(1).normal(0, 1e-2, (B, 42))
    ro = so.solve_qp(qp, dx0)
    rg = sg.solve_qp(qp, dx0)
    assert sg.last_qp_path == "fused"          # the product's kernel, not the generic dense fallback
    np.testing.assert_array_equal(rg["qp_iter"], ro["qp_iter"])
    np.testing.assert_array_equal(rg["qp_status"], ro["qp_status"])
    m = qp["mask"] > 0.5
    for k in ("dz", "pi"):
        assert report(f"cap{cap} {k}", rg[k], ro[k]) < REL, k
    for k in ("lam", "t"):            # rows without a bound carry no multiplier (lam = 0, t = 1 by convention)
        assert report(f"cap{cap} {k}", np.where(m, rg[k], 0.0), np.where(m, ro[k], 0.0)) < REL, k
    # the primal step split into its blocks: inputs a, f and states q, v, h, each against its own norm
    for name, sl in (("a", slice(0, 18)), ("f", slice(18, 30)), ("q", slice(30, 48)), ("v", slice(48, 66)), ("h", slice(66, 72))):
        assert report(f"cap{cap} dz[{name}]", rg["dz"][..., sl], ro["dz"][..., sl]) < REL, name


def test_qp_without_inequalities_is_one_riccati_solve(oracle_mod):
    O = oracle_mod
    B = 4
    so, sg, _ = _pair(O, B, seed=6, max_qp_iter=1, qp_tol=1e-300, warm_start_qp=0)
    qp = so.export_qp()
    qp["mask"][:] = 0
    dx0 = np.random.default_rng(2).normal(0, 1e-2, (B, 42))
    ro, rg = so.solve_qp(qp, dx0), sg.solve_qp(qp, dx0)
    assert rel_err(rg["dz"], ro["dz"]) < 1e-9 and rel_err(rg["pi"], ro["pi"]) < 1e-9


def test_rti_solve_parity(oracle_mod):
    so, sg, batch = _pair(oracle_mod, 33, seed=7, real_time_it=1, warm_start_qp=0)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    for name, a, b in zip(("x", "u", "pi", "lam"), sg.get_solution(), so.get_solution()):
        assert report("rti " + name, a, b) < REL, name
    xg = sg.get_solution()[0]
    so_s, sg_s = so.get_stats(), sg.get_stats()
    for k in ("sqp_iter", "qp_iter", "qp_status", "qp_ncond"):
        np.testing.assert_array_equal(sg_s[k], so_s[k])
    assert rel_err(sg_s["cost"], so_s["cost"]) < 1e-10
    np.testing.assert_array_equal(xg[:, 0, :36], batch["qv"])


def test_full_sqp_parity_and_convergence(oracle_mod):
    opts = dict(max_iter=40, max_qp_iter=20, qp_tol=1e-3, nlp_tol=1e-1, warm_start_qp=0)
    so, sg, _ = _pair(oracle_mod, 8, seed=3, perturb=0.5, **opts)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    assert (st_g == 0).all()
    so_s, sg_s = so.get_stats(), sg.get_stats()
    np.testing.assert_array_equal(sg_s["sqp_iter"], so_s["sqp_iter"])
    assert (sg_s["res"] < 1e-1).all()                       # converged within nlp_tol
    xo, uo, _, _ = so.get_solution()
    xg, ug, _, _ = sg.get_solution()
    # many SQP iterations compound rounding differences: NLP solutions agree well inside nlp_tol
    assert rel_err(xg, xo) < 1e-5 and rel_err(ug, uo) < 1e-5


@pytest.mark.parametrize("warm", [1, 2])
def test_sqp_few_iterations_iterate_parity(oracle_mod, warm):
    """Three SQP iterations with the default 6-iteration IPM cap and the reference's QP warm start
    (warm_start_qp = True -> HPIPM level 1: the first QP of the solve is cold, acados warm_start_first_qp = 0,
    the next ones start from the previous primal step), and with HPIPM's hot start (level 2: lam, t kept and
    clipped at 0.1).  The reference's first-tick setting is 15 iterations (mpc.py:464-473)."""
    opts = dict(max_iter=3, max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-2, warm_start_qp=warm)
    so, sg, _ = _pair(oracle_mod, 10, seed=8, gaits=("trot", "pace"), **opts)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    for name, a, b in zip(("x", "u", "pi", "lam"), sg.get_solution(), so.get_solution()):
        assert chain_ok(f"sqp3 warm{warm} {name}", a, b, REL), name          # three chained QPs
    assert rel_err(sg.get_stats()["res"], so.get_stats()["res"]) < 1e-5


def test_warm_start_first_qp_option(oracle_mod):
    """acados `warm_start_first_qp`: with the flag the first QP of a solve starts from the stored QP iterate
    (here: the previous solve's), without it the stored iterate is ignored."""
    O = oracle_mod
    opts = dict(real_time_it=1, warm_start_qp=2)
    sols = {}
    for flag in (0, 1):
        so, sg, _ = _pair(O, 6, seed=15, warm_start_first_qp=flag, **opts)
        for _ in range(2):
            st_o, st_g = so.solve(0), sg.solve()
            np.testing.assert_array_equal(st_g, st_o)
        for name, a, b in zip(("x", "u", "pi", "lam"), sg.get_solution(), so.get_solution()):
            assert chain_ok(f"warm_start_first_qp={flag} {name}", a, b, REL), name
        sols[flag] = sg.get_solution()[0]
    assert np.abs(sols[0] - sols[1]).max() > 0


def test_conditional_predictor_corrector_parity(oracle_mod):
    """HPIPM cond_pred_corr: a batch in which the corrector direction is rejected (complementarity predicted above
    twice the affine prediction) and the centring direction is taken instead, in the oracle and on the device in the
    same IPM iterations; and the same batch with the option off."""
    O = oracle_mod
    opts = dict(real_time_it=0, max_iter=3, max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-1, warm_start_qp=2)
    for cond in (1, 0):
        so, sg, _ = _pair(O, 24, seed=46, gaits=("trot", "pace", "bound"), cond_pred_corr=cond, **opts)
        st_o, st_g = so.solve(0), sg.solve()
        np.testing.assert_array_equal(st_g, st_o)
        so_s, sg_s = so.get_stats(), sg.get_stats()
        np.testing.assert_array_equal(sg_s["qp_ncond"], so_s["qp_ncond"])
        assert (so_s["qp_ncond"].sum() > 0) == bool(cond)
        for name, a, b in zip(("x", "u", "pi", "lam"), sg.get_solution(), so.get_solution()):
            assert chain_ok(f"cond={cond} {name}", a, b, REL), name


def test_config3_options_convergence_and_early_exit(oracle_mod):
    """BASELINE config 3's exact options (mixed gaits, full SQP max_iter = 15, max_qp_iter = 6, qp_tol = 1e-2,
    nlp_tol = 1e-1) with HPIPM's hot start, the setting under which the 6-iteration IPM converges:
    part of the batch converges inside nlp_tol with different iteration counts, so the per-problem done flags and
    the early exit of finished CTAs are exercised.  Same status and iteration count per problem as the oracle;
    converged problems agree well inside nlp_tol (up to 15 chained QPs, 1e-6 each, compounding through a
    non-contractive iteration: 1e-4)."""
    O = oracle_mod
    B = 96
    opts = dict(real_time_it=0, max_iter=15, max_qp_iter=6, qp_tol=1e-2, nlp_tol=1e-1, warm_start_qp=2)
    so, sg, _ = _pair(O, B, seed=2, gaits=("trot", "pace", "bound"), **opts)
    st_o, st_g = so.solve(0), sg.solve()
    so_s, sg_s = so.get_stats(), sg.get_stats()
    print("config3 status histogram (oracle):", np.bincount(st_o, minlength=5).tolist(),
          "sqp_iter histogram:", np.bincount(so_s["sqp_iter"], minlength=16).tolist())
    same = (st_g == st_o) & (sg_s["sqp_iter"] == so_s["sqp_iter"])
    print("config3 problems with a different outcome:", np.nonzero(~same)[0].tolist(), st_g[~same].tolist(), st_o[~same].tolist())
    # the problems that do not converge run 15 undamped iterations, some of them into a QP failure: which of the two
    # they end in (MAXITER / QP failure) is decided by amplified rounding; every converged problem must coincide
    assert same.mean() >= 0.9 and same[st_o == 0].all() and (st_o[st_g == 0] == 0).all()
    conv = same & (st_o == 0)
    assert conv.sum() >= B // 4 and len(np.unique(so_s["sqp_iter"][conv])) >= 3
    assert (sg_s["res"][st_g == 0] < 1e-1).all()            # converged means: all four residual norms inside nlp_tol
    xo, uo, _, _ = so.get_solution()
    xg, ug, _, _ = sg.get_solution()
    assert chain_ok("config3 x (converged)", xg[conv], xo[conv], 1e-4, frac=1.0)
    assert chain_ok("config3 u (converged)", ug[conv], uo[conv], 1e-4, frac=1.0)


def test_status_codes_parity(oracle_mod):
    """MAXITER when the tolerance is unreachable, NaN confined to the poisoned problem."""
    O = oracle_mod
    B = 5
    so, sg, batch = _pair(O, B, seed=9, max_iter=2, nlp_tol=1e-12, warm_start_qp=0)
    _, _, hg = O.kinematics(batch["qv"])
    xg = batch["x_guess"].copy()
    xg[:, :, 36:] = hg[:, None, :]
    xg[3, 4, 9] = np.nan
    so.set_guess(xg, None)
    sg.set_guess(xg, None)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    assert st_g[3] == 1 and (np.delete(st_g, 3) == 2).all()
    x, u, _, _ = sg.get_solution()
    keep = [0, 1, 2, 4]
    assert np.isfinite(x[keep]).all() and np.isfinite(u[keep]).all()


@pytest.mark.parametrize("repeat_last", [False, True])
def test_shift_warm_start_parity(oracle_mod, repeat_last):
    O = oracle_mod
    NmpcSolver, default_opts = _gpu()
    B = 6
    rng = np.random.default_rng(0)
    x, u = rng.normal(size=(B, N + 1, 42)), rng.normal(size=(B, N, 30))
    so, sg = O.OracleSolver(B, O.default_opts()), NmpcSolver(B, default_opts())
    start = np.array([0, 1, 2, 5, 24, 25], dtype=np.int32)
    for s in (so, sg):
        s.set_guess(x, u)
        s.shift_warm_start(start, repeat_last)
    for a, b in zip(sg.get_solution(), so.get_solution()):
        np.testing.assert_array_equal(a, b)


def test_receding_horizon_ticks_parity(oracle_mod):
    """Three consecutive replanning ticks: solve, shift by one node, new x0 from the plan
    (LocomotionMPC.open_loop semantics, mpc.py:416-462), with QP and NLP warm starts."""
    O = oracle_mod
    B = 4
    so, sg, batch = _pair(O, B, seed=12, real_time_it=1, warm_start_qp=1)
    for tick in range(3):
        st_o, st_g = so.solve(0), sg.solve()
        np.testing.assert_array_equal(st_g, st_o)
        xo, uo, _, _ = so.get_solution()
        xg, ug, _, _ = sg.get_solution()
        assert chain_ok(f"tick{tick} x", xg, xo, REL, frac=1.0 if tick == 0 else 0.75), tick
        assert chain_ok(f"tick{tick} u", ug, uo, REL, frac=1.0 if tick == 0 else 0.75), tick
        nxt = xo[:, 1, :].copy()
        for s in (so, sg):
            s.shift_warm_start(np.ones(B, np.int32), False)
            s.set_x0_full(nxt)


def test_batch_is_independent_and_chunking_invariant(oracle_mod, monkeypatch):
    """Size-independent property used at full size: a problem's solution does not depend on its
    position in the batch nor on the workspace chunking."""
    NmpcSolver, default_opts = _gpu()
    B = 96
    batch = WL.make_batch(B, seed=13)
    _, _, hg = oracle_mod.kinematics(batch["qv"])

    def run(bt, hg_, chunk=None):
        if chunk:
            monkeypatch.setenv("NMPC_CHUNK", str(chunk))
        else:
            monkeypatch.delenv("NMPC_CHUNK", raising=False)
        s = NmpcSolver(len(bt["qv"]), default_opts(real_time_it=1, warm_start_qp=0))
        WL.load_into(s, bt, hg_)
        st = s.solve()
        x, u, _, _ = s.get_solution()
        return st, x, u

    st0, x0, u0 = run(batch, hg)
    st1, x1, u1 = run(batch, hg, chunk=40)           # 40 + 40 + 16
    np.testing.assert_array_equal(st0, st1)
    np.testing.assert_array_equal(x0, x1)
    np.testing.assert_array_equal(u0, u1)
    perm = np.random.default_rng(0).permutation(B)
    pb = {k: (v[perm] if isinstance(v, np.ndarray) and v.shape[:1] == (B,) else v) for k, v in batch.items()}
    st2, x2, u2 = run(pb, hg[perm])
    np.testing.assert_array_equal(st0[perm], st2)
    np.testing.assert_array_equal(x0[perm], x2)
    np.testing.assert_array_equal(u0[perm], u2)


def test_wave_launch_plan_is_invisible_in_the_results(oracle_mod, monkeypatch):
    """The fused QP kernel goes out wave by wave (full waves of 4 CTAs per SM, the rest with the residency capped through
    the dynamic shared memory size; run_qp_fused): a batch of two full waves and a partial one, also cut by the chunking,
    gives bit-identical results to one launch per chunk (NMPC_WAVE_PLAN=0), with and without time optimisation, and the
    capped launches (3, 2 and 1 CTAs per SM) run."""
    import torch
    NmpcSolver, default_opts = _gpu()
    n_sm = torch.cuda.get_device_properties(0).multi_processor_count
    cases = [(8 * n_sm + 2 * n_sm + 5, dict(real_time_it=1), None),          # 2 full waves + 3 per SM
             (4 * n_sm + n_sm + 9, dict(real_time_it=1), None),              # 1 full wave + 2 per SM
             (4 * n_sm + 17, dict(real_time_it=0, max_iter=2, enable_time_opt=1), None),      # + 1 per SM, time-opt kernel
             (700, dict(real_time_it=1), 333)]                               # chunks of 333 | 333 | 34
    for B, o, chunk in cases:
        batch = WL.make_batch(B, seed=31)
        _, _, hg = oracle_mod.kinematics(batch["qv"])
        out = []
        for plan in ("1", "0"):
            monkeypatch.setenv("NMPC_WAVE_PLAN", plan)
            if chunk:
                monkeypatch.setenv("NMPC_CHUNK", str(chunk))
            else:
                monkeypatch.delenv("NMPC_CHUNK", raising=False)
            s = NmpcSolver(B, default_opts(**o))
            WL.load_into(s, batch, hg)
            st = s.solve()
            x, u, pi, lam = s.get_solution()
            out.append((st, x, u, pi, lam))
            del s
        for a, b in zip(out[0], out[1]):
            np.testing.assert_array_equal(a, b)
        assert np.isfinite(out[0][1]).all() and (out[0][0] != 1).all()
    monkeypatch.delenv("NMPC_WAVE_PLAN", raising=False)
    monkeypatch.delenv("NMPC_CHUNK", raising=False)


def test_output_streaming_delivers_the_final_iterate(oracle_mod, monkeypatch):
    """nmpc_set_output: x / u copied wave by wave on the copy stream during the solve equal what nmpc_get_solution
    returns afterwards -- RTI and full SQP (copied in the last iteration), batches of several waves, chunked workspace,
    pinned host and device destinations; switching it off stops the copies."""
    import torch
    NmpcSolver, default_opts = _gpu()
    n_sm = torch.cuda.get_device_properties(0).multi_processor_count
    for B, o, chunk, on_device in ((4 * n_sm + 2 * n_sm + 3, dict(real_time_it=1), None, False),
                                   (4 * n_sm + 40, dict(real_time_it=0, max_iter=3), None, True),
                                   (500, dict(real_time_it=1), 210, False)):
        if chunk:
            monkeypatch.setenv("NMPC_CHUNK", str(chunk))
        else:
            monkeypatch.delenv("NMPC_CHUNK", raising=False)
        batch = WL.make_batch(B, seed=41)
        _, _, hg = oracle_mod.kinematics(batch["qv"])
        s = NmpcSolver(B, default_opts(**o))
        kw = dict(dtype=torch.float64, device="cuda") if on_device else dict(dtype=torch.float64, pin_memory=True)
        ox = torch.full((B, 26, 42), float("nan"), **kw)
        ou = torch.full((B, 25, 30), float("nan"), **kw)
        s.set_output(ox, ou)
        for rep in range(2):                          # the second solve starts from the first one's iterate
            if rep == 0:
                WL.load_into(s, batch, hg)
            st = s.solve()
            x, u, _, _ = s.get_solution()
            np.testing.assert_array_equal(ox.cpu().numpy(), x)
            np.testing.assert_array_equal(ou.cpu().numpy(), u)
            assert np.isfinite(x).all() and (st != 1).all()
        s.set_output()
        ox.fill_(7.0)
        s.solve()
        s.synchronize()
        assert (ox == 7.0).all()
        del s
    monkeypatch.delenv("NMPC_CHUNK", raising=False)
    # max_iter = 0: no QP runs, the registered buffers still receive the (unchanged) iterate
    B = 9
    batch = WL.make_batch(B, seed=42)
    _, _, hg = oracle_mod.kinematics(batch["qv"])
    s = NmpcSolver(B, default_opts(real_time_it=0, max_iter=0))
    WL.load_into(s, batch, hg)
    ox = torch.full((B, 26, 42), float("nan"), dtype=torch.float64, pin_memory=True)
    s.set_output(ox, None)
    s.solve()
    np.testing.assert_array_equal(ox.numpy(), s.get_solution()[0])


def test_full_size_batch_properties(oracle_mod):
    """BASELINE config 2 at full size (1024 OCPs, SQP_RTI): every problem succeeds, the measured
    state is imposed exactly, the QP residuals the device reports are finite, and a sample of the
    batch matches the oracle."""
    NmpcSolver, default_opts = _gpu()
    O = oracle_mod
    B = 1024
    batch = WL.make_batch(B, seed=1)
    sg = NmpcSolver(B, default_opts(real_time_it=1, warm_start_qp=0))
    _, _, hg = sg.kinematics(batch["qv"])
    WL.load_into(sg, batch, hg)
    st = sg.solve()
    assert (st == 0).all()
    x, u, _, _ = sg.get_solution()
    np.testing.assert_array_equal(x[:, 0, :36], batch["qv"])
    stats = sg.get_stats()
    assert np.isfinite(stats["qp_res"]).all() and (stats["qp_iter"] <= 6).all()
    idx = np.arange(0, B, 64)
    sub = {k: (v[idx] if isinstance(v, np.ndarray) and v.shape[:1] == (B,) else v) for k, v in batch.items()}
    so = O.OracleSolver(len(idx), O.default_opts(real_time_it=1, warm_start_qp=0))
    WL.load_into(so, sub, hg[idx])
    so.solve(0)
    xo, uo, _, _ = so.get_solution()
    assert rel_err(x[idx], xo) < REL and rel_err(u[idx], uo) < REL


def test_device_pointers_are_accepted(oracle_mod):
    """The C ABI takes host or device pointers (torch tensors only provide the memory)."""
    import torch
    NmpcSolver, default_opts = _gpu()
    B = 8
    batch = WL.make_batch(B, seed=14)
    _, _, hg = oracle_mod.kinematics(batch["qv"])
    s_h = NmpcSolver(B, default_opts(real_time_it=1, warm_start_qp=0))
    s_d = NmpcSolver(B, default_opts(real_time_it=1, warm_start_qp=0))
    WL.load_into(s_h, batch, hg)
    dev = {k: (torch.from_numpy(np.ascontiguousarray(v)).cuda() if isinstance(v, np.ndarray) else v)
           for k, v in batch.items()}
    xg = batch["x_guess"].copy()
    xg[:, :, 36:] = hg[:, None, :]
    s_d.set_weights(dev["W"], dev["W_e"])
    s_d.set_yref(dev["yref"], dev["yref_e"])
    s_d.set_params(dev["p"])
    s_d.set_guess(torch.from_numpy(xg).cuda(), dev["u_guess"])
    s_d.set_x0(dev["qv"])
    s_d.pin_stance_feet(dev["i_node"])
    s_d.reset_duals()
    st_d = torch.zeros(B, dtype=torch.int32, device="cuda")
    s_d.solve(status=st_d)
    st_h = s_h.solve()
    np.testing.assert_array_equal(st_d.cpu().numpy(), st_h)
    xd = torch.zeros(B, N + 1, 42, dtype=torch.float64, device="cuda")
    ud = torch.zeros(B, N, 30, dtype=torch.float64, device="cuda")
    s_d.get_solution_into(x=xd, u=ud)
    xh, uh, _, _ = s_h.get_solution()
    np.testing.assert_array_equal(xd.cpu().numpy(), xh)
    np.testing.assert_array_equal(ud.cpu().numpy(), uh)


@pytest.mark.parametrize("cfg", ["config4_restricted_contacts", "config5_push"])
def test_config4_and_5_workloads_parity(oracle_mod, cfg):
    """BASELINE configs 4 (contact-patch restriction with Raibert footholds + opt_peak; the dense QP blocks are
    compared without time optimisation, the nu = 31 solve follows in test_config4_as_written) and 5 (push-perturbed
    DAgger tick): default mpc_opt.py options, SQP with max_iter = 1."""
    O = oracle_mod
    wl = dict(restrict_cnt=True, kin=O.kinematics) if cfg.startswith("config4") else dict(push=True)
    so, sg, batch = _pair(O, 48, seed=11, workload=wl, real_time_it=0, max_iter=1, warm_start_qp=0)
    if cfg.startswith("config4"):
        assert (batch["p"].reshape(48, N + 1, 4, 11)[..., 8] > 0.5).any()        # some touch-downs are restricted
        qo, qg = so.export_qp(), sg.export_qp()
        np.testing.assert_array_equal(qg["mask"], qo["mask"])
        assert (qo["mask"].reshape(48, N + 1, 2, 42)[:, :, 1, 32:36] > 0.5).any()  # patch rows are in the QP
        for k in ("BAbt", "RSQrq", "DCt", "lg", "ug"):
            assert rel_err(qg[k], qo[k]) < 1e-11, k
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    for name, a, b in zip(("x", "u", "pi", "lam"), sg.get_solution(), so.get_solution()):
        assert report(f"{cfg} {name}", a, b) < REL, name
    so_s, sg_s = so.get_stats(), sg.get_stats()
    for k in ("sqp_iter", "qp_iter", "qp_status"):
        np.testing.assert_array_equal(sg_s[k], so_s[k])


def test_unstructured_external_qp_is_refused(oracle_mod):
    """There is no generic dense QP solver in this build: a QP without the structure of this OCP family is refused
    with NMPC_ENOSYS by the structure check (and nothing is solved by untested kernels)."""
    from iterative_learning_nmpc_b200._lib import NmpcError
    so, sg, _ = _pair(oracle_mod, 2, seed=6, max_qp_iter=2, qp_tol=1e-9, warm_start_qp=0)
    qp = so.export_qp()
    sg.solve_qp(qp, np.zeros((2, 42)))
    assert sg.last_qp_path == "fused"
    qp["RSQrq"][0, 3, 5, 40] = qp["RSQrq"][0, 3, 40, 5] = 0.25          # couples an input with a state: not in the family
    with pytest.raises(NmpcError, match="ENOSYS"):
        sg.solve_qp(qp, np.zeros((2, 42)))
    assert sg.last_qp_path == "refused"


@pytest.mark.parametrize("measured", [False, True])
def test_plan_torques_parity(oracle_mod, measured):
    """Batched plan -> torque on the device against the reference's host pipeline (scipy Hermite splines,
    id_repeat hold, inverse dynamics, PD, DAgger action).  Floating point: 1e-10 relative."""
    O = oracle_mod
    so, sg, batch = _pair(O, 9, seed=13, real_time_it=1, warm_start_qp=0)
    sg.solve()
    xg, ug, _, _ = sg.get_solution()
    n_interp, step0, n_steps = 1000, 17, 60
    meas = None
    if measured:
        rng = np.random.default_rng(3)
        ref = O.plan_torques(xg, ug, 0.04, n_interp, step0, n_steps)
        meas = np.concatenate([ref["q_plan"], ref["v_plan"]], axis=2) + rng.normal(0, 1e-2, (9, n_steps, 36))
    want = O.plan_torques(xg, ug, 0.04, n_interp, step0, n_steps, kp=20.0, kd=1.75, qv_meas=meas)
    got = sg.plan_torques(n_interp, step0, n_steps, kp=20.0, kd=1.75, qv_meas=meas)
    for k in ("q_plan", "v_plan", "tau_ff", "tau", "action"):
        assert rel_err(got[k], want[k]) < 1e-10, (k, rel_err(got[k], want[k]))
    # whole plan, including the last sample (t = T) and the first one
    want = O.plan_torques(xg[:2], ug[:2], 0.04, 250)
    got = sg.plan_torques(250)
    for k in ("q_plan", "v_plan", "tau"):
        assert rel_err(got[k][:2], want[k]) < 1e-10, k


@pytest.mark.parametrize("restrict", [False, True])
def test_device_init_matches_host_packing(oracle_mod, restrict):
    """nmpc_init_tick / nmpc_cold_guess (QuadrupedAcadosSolver.init expanded on the device from the compact
    planner outputs) against the host packing of problem.py that the oracle consumes: p, yref, yref_e
    bit-exact, and the RTI solve from the device-built problem equal to the oracle's."""
    O = oracle_mod
    wl = dict(restrict_cnt=True, kin=O.kinematics) if restrict else {}
    so, sg, batch = _pair(O, 37, seed=21, gaits=("trot", "pace", "bound"), workload=wl, real_time_it=1, warm_start_qp=0)
    host = sg.get_problem_data()                 # after load_into + pin_stance_feet
    NmpcSolver, default_opts = _gpu()
    sd = NmpcSolver(37, default_opts(real_time_it=1, warm_start_qp=0))
    WL.load_compact(sd, batch)
    dev = sd.get_problem_data()
    for k in ("p", "yref", "yref_e", "x0"):
        np.testing.assert_array_equal(dev[k], host[k], err_msg=k)
    xd0, ud0, _, _ = sd.get_solution(duals=False)
    xh0, uh0, _, _ = sg.get_solution(duals=False)
    np.testing.assert_array_equal(xd0[:, :, :36], xh0[:, :, :36])
    assert rel_err(xd0, xh0) < 1e-14             # h0: device kinematics vs the oracle's (the host path was fed the latter)
    assert rel_err(ud0, uh0) < 1e-15             # weight / n_stance: same expression, host vs device rounding
    st_o, st_d = so.solve(0), sd.solve()
    np.testing.assert_array_equal(st_d, st_o)
    xo, uo, _, _ = so.get_solution()
    xd, ud, _, _ = sd.get_solution()
    assert rel_err(xd, xo) < REL and rel_err(ud, uo) < REL


def test_device_init_argument_errors():
    NmpcSolver, default_opts = _gpu()
    s = NmpcSolver(2, default_opts())
    with pytest.raises(ValueError):              # restriction without a contact plan (solver.py:377)
        s.init_tick(np.zeros((2, 12)), joint_ref=np.zeros(12), cnt_sequence=np.ones((2, 4, N + 1)),
                    p_gain=np.zeros(4), restrict_cnt=True)


def _split_topt(uo, lo):
    """The oracle's time-opt build carries dt as the 31st input column and its two bounds as constraint row 42:
    (u [.., 30], dt [..], lam [.., 84], lam_dt [.., 2]) in the product's layout."""
    u, dt = uo[..., :30], uo[..., 30]
    lam = np.concatenate([lo[..., :42], lo[..., 43:85]], axis=-1)
    lam_dt = np.stack([lo[..., 42], lo[..., 85]], axis=-1)[:, :-1]
    return u, dt, lam, lam_dt


@pytest.mark.parametrize("cap", [1, 2, 3, 6])
def test_time_opt_rti_parity_per_qp_iteration(oracle_mod, cap):
    """enable_time_opt (nu = 31: node durations dt_k with the weight `time_opt`, bounds [dt_min, dt_max]; solver.py:43-44,
    124-125, 166-167, 189-190; BASELINE config 4): one SQP_RTI step whose QP is cut after `cap` IPM iterations, so the
    primal/dual QP iterate after 1, 2, 3, 6 iterations is compared through the full step -- x, u, the dt column with
    its two bound multipliers, pi, lam -- against the oracle's dense nu = 31 build (the product eliminates dt from
    every stage first: P~ = P+ - (P+ b)(P+ b)' / d)."""
    O = oracle_mod
    B = 12
    so, sg, _ = _pair(O, B, seed=11, real_time_it=1, max_qp_iter=cap, qp_tol=1e-9, enable_time_opt=1)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    so_s, sg_s = so.get_stats(), sg.get_stats()
    np.testing.assert_array_equal(sg_s["qp_iter"], so_s["qp_iter"])
    np.testing.assert_array_equal(sg_s["qp_status"], so_s["qp_status"])
    xo, uo, pio, lo = so.get_solution()
    xg, ug, pig, lg = sg.get_solution()
    dtg, ltg = sg.get_dt(duals=True)
    uo30, dto, lo84, lto = _split_topt(uo, lo)
    assert np.abs(dto - 0.04).max() > 1e-4          # the node durations did move
    for name, a, b in (("x", xg, xo), ("u", ug, uo30), ("dt - dt_nodes", dtg - 0.04, dto - 0.04), ("pi", pig, pio),
                       ("lam", lg, lo84), ("lam_dt", ltg, lto)):
        assert report(f"time-opt cap{cap} {name}", a, b) < REL, name
    assert report("time-opt qp_res", sg_s["qp_res"], so_s["qp_res"]) < 1e-5


def test_time_opt_sqp_and_warm_start_parity(oracle_mod):
    """Time optimisation through three SQP iterations with hot-started QPs, then the shift warm start (dt of the
    warm-started nodes shifted, nominal on the tail: solver.py:324-336) and a second solve."""
    O = oracle_mod
    B = 8
    opts = dict(real_time_it=0, max_iter=3, max_qp_iter=6, warm_start_qp=2, enable_time_opt=1)
    so, sg, _ = _pair(O, B, seed=12, **opts)
    for tick in range(2):
        st_o, st_g = so.solve(0), sg.solve()
        np.testing.assert_array_equal(st_g, st_o)
        xo, uo, pio, lo = so.get_solution()
        xg, ug, pig, lg = sg.get_solution()
        dtg, ltg = sg.get_dt(duals=True)
        uo30, dto, lo84, lto = _split_topt(uo, lo)
        for name, a, b in (("x", xg, xo), ("u", ug, uo30), ("dt", dtg, dto), ("pi", pig, pio), ("lam", lg, lo84),
                           ("lam_dt", ltg, lto)):
            assert chain_ok(f"time-opt tick{tick} {name}", a, b, REL), name
        so.shift_warm_start(3)
        sg.shift_warm_start(3)
        xo, uo, _, lo = so.get_solution()
        dtg, ltg = sg.get_dt(duals=True)
        _, dto, _, lto = _split_topt(uo, lo)
        # the shifted arrays are copies of values compared above (a sub-block has a smaller norm: same absolute error)
        assert chain_ok(f"time-opt shifted dt tick{tick}", dtg, dto, 10 * REL)
        assert chain_ok(f"time-opt shifted lam_dt tick{tick}", ltg, lto, 10 * REL)
        assert np.abs(dtg[:, -3:] - 0.04).max() == 0.0


def test_time_opt_off_reports_nominal_dt():
    NmpcSolver, default_opts = _gpu()
    s = NmpcSolver(2, default_opts(real_time_it=1))
    assert np.abs(s.get_dt() - 0.04).max() == 0.0


def test_config4_as_written(oracle_mod):
    """BASELINE config 4 as written: enable_time_opt (nu = 31, dt in [0.02, 0.07]), contact-patch restriction with
    Raibert footholds, opt_peak, default mpc_opt.py options (SQP, max_iter = 1, max_qp_iter = 6)."""
    O = oracle_mod
    B = 32
    wl = dict(restrict_cnt=True, kin=O.kinematics)
    so, sg, batch = _pair(O, B, seed=13, workload=wl, real_time_it=0, max_iter=1, enable_time_opt=1, dt_min=0.02,
                          dt_max=0.07, time_opt_weight=1e4)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    xo, uo, pio, lo = so.get_solution()
    xg, ug, pig, lg = sg.get_solution()
    dtg, ltg = sg.get_dt(duals=True)
    uo30, dto, lo84, lto = _split_topt(uo, lo)
    for name, a, b in (("x", xg, xo), ("u", ug, uo30), ("dt", dtg, dto), ("pi", pig, pio), ("lam", lg, lo84),
                       ("lam_dt", ltg, lto)):
        assert report(f"config4 {name}", a, b) < REL, name
    so_s, sg_s = so.get_stats(), sg.get_stats()
    for k in ("sqp_iter", "qp_iter", "qp_status"):
        np.testing.assert_array_equal(sg_s[k], so_s[k])
    assert report("config4 nlp residuals", sg_s["res"], so_s["res"]) < 1e-6


def test_checkpoint_resume_continues_bit_identically(oracle_mod):
    """nmpc_export_state / nmpc_import_state: a second handle resumed from the blob after tick 1 (warm-started full SQP
    with hot-started QPs, time optimisation on) produces the same tick 2, bit for bit."""
    O = oracle_mod
    NmpcSolver, default_opts = _gpu()
    B = 6
    opts = dict(real_time_it=0, max_iter=2, warm_start_qp=2, enable_time_opt=1)
    batch = WL.make_batch(B, seed=19)
    _, _, hg = O.kinematics(batch["qv"])
    a = NmpcSolver(B, default_opts(**opts))
    WL.load_into(a, batch, hg)
    a.solve()
    blob = a.export_state()
    b = NmpcSolver(B, default_opts(**opts))
    b.import_state(blob)
    for s in (a, b):
        s.shift_warm_start(1)
        s.solve()
    for x, y in zip(a.get_solution(), b.get_solution()):
        np.testing.assert_array_equal(x, y)
    np.testing.assert_array_equal(a.get_dt(), b.get_dt())
    with pytest.raises(Exception):
        NmpcSolver(B + 1, default_opts(**opts)).import_state(blob)


@pytest.mark.parametrize("n_nodes,topt", [(1, 0), (2, 0), (5, 0), (5, 1), (40, 0), (64, 1)])
def test_other_horizon_lengths(oracle_mod, n_nodes, topt):
    """The horizon is a run-time parameter (1 <= N <= 64, NMPC_MAX_NODES): the rings of the passes, the shift warm start
    and the time-opt bookkeeping at the shortest and the longest horizons, two warm-started SQP iterations."""
    O = oracle_mod
    NmpcSolver, default_opts = _gpu()
    B = 5
    opts = dict(n_nodes=n_nodes, real_time_it=0, max_iter=2, warm_start_qp=1, enable_time_opt=topt)
    batch = WL.make_batch(B, seed=41, n_nodes=n_nodes, perturb=0.5)
    _, _, hg = O.kinematics(batch["qv"])
    so, sg = O.OracleSolver(B, O.default_opts(**opts)), NmpcSolver(B, default_opts(**opts))
    WL.load_into(so, batch, hg)
    WL.load_into(sg, batch, hg)
    st_o, st_g = so.solve(0), sg.solve()
    np.testing.assert_array_equal(st_g, st_o)
    xo, uo, pio, lo = so.get_solution()
    xg, ug, pig, lg = sg.get_solution()
    if topt:
        uo, dto, lo, lto = _split_topt(uo, lo)
        dtg, ltg = sg.get_dt(duals=True)
        assert chain_ok(f"N={n_nodes} dt", dtg, dto, REL) and chain_ok(f"N={n_nodes} lam_dt", ltg, lto, REL)
    for name, a, b in (("x", xg, xo), ("u", ug, uo), ("pi", pig, pio), ("lam", lg, lo)):
        assert chain_ok(f"N={n_nodes} {name}", a, b, REL), name
    if n_nodes > 2:
        so.shift_warm_start(1); sg.shift_warm_start(1)
        for a, b in zip(sg.get_solution()[:2], so.get_solution()[:2]):
            assert chain_ok(f"N={n_nodes} shifted", a, b[..., :a.shape[-1]], 10 * REL)


def test_time_opt_chunking_and_nan_containment(oracle_mod, monkeypatch):
    """Time optimisation with the QP workspace in chunks (the dt bookkeeping is indexed by the global problem, the stage
    records by the chunk) and with one poisoned problem: bit-identical to the unchunked solve, NaN confined."""
    NmpcSolver, default_opts = _gpu()
    B = 50
    batch = WL.make_batch(B, seed=23)
    _, _, hg = oracle_mod.kinematics(batch["qv"])
    opts = dict(real_time_it=0, max_iter=2, enable_time_opt=1)
    out = []
    for chunk in (None, 16):
        if chunk:
            monkeypatch.setenv("NMPC_CHUNK", str(chunk))
        else:
            monkeypatch.delenv("NMPC_CHUNK", raising=False)
        s = NmpcSolver(B, default_opts(**opts))
        WL.load_into(s, batch, hg)
        st = s.solve()
        out.append((st, s.get_solution()[0], s.get_dt()))
    np.testing.assert_array_equal(out[0][0], out[1][0])
    np.testing.assert_array_equal(out[0][1], out[1][1])
    np.testing.assert_array_equal(out[0][2], out[1][2])
    monkeypatch.delenv("NMPC_CHUNK", raising=False)
    bad = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in batch.items()}
    bad["qv"][7, 20] = np.nan
    s = NmpcSolver(B, default_opts(**opts))
    WL.load_into(s, bad, hg)
    st = s.solve()
    assert st[7] == 1 and (np.delete(st, 7) == np.delete(out[0][0], 7)).all()
    keep = np.arange(B) != 7
    np.testing.assert_array_equal(s.get_solution()[0][keep], out[0][1][keep])
    np.testing.assert_array_equal(s.get_dt()[keep], out[0][2][keep])
