"""Oracle model functions vs an independent numeric model and finite differences (CPU only).

The generated header (shared by the oracle and the CUDA product) stands in for the
CasADi-generated model code of the reference; the reference holds no golden
vectors for it, so it is pinned algorithm-independently here.
"""
import numpy as np
import pytest

from refmodel import NumericModel

REC = dict(FOOT_P=0, FOOT_J=12, FOOT_V=228, FOOT_VQ=240, COM=456, COM_J=459, H=513, H_Q=519, H_V=627, SIZE=735)
NQ = 18


@pytest.fixture(scope="module")
def num():
    return NumericModel()


def _rand_qv(rng):
    q = np.zeros(NQ)
    q[:3] = rng.uniform(-0.5, 0.5, 3) + [0, 0, 0.3]
    q[3:6] = rng.uniform(-0.6, 0.6, 3)
    q[6:] = np.tile([0.0, 0.9, -1.8], 4) + rng.uniform(-0.4, 0.4, 12)
    v = rng.uniform(-1.5, 1.5, NQ)
    return q, v


def test_record_layout(oracle_mod):
    q, v = _rand_qv(np.random.default_rng(0))
    rec = oracle_mod.model_record(np.concatenate([q, v]))
    assert rec.shape == (REC["SIZE"],)
    assert np.isfinite(rec).all()


@pytest.mark.parametrize("seed", range(4))
def test_values_match_numeric_model(oracle_mod, num, seed):
    q, v = _rand_qv(np.random.default_rng(seed))
    rec = oracle_mod.model_record(np.concatenate([q, v]))
    feet = rec[REC["FOOT_P"]:REC["FOOT_P"] + 12].reshape(4, 3)
    np.testing.assert_allclose(feet, num.feet(q), atol=1e-12)
    np.testing.assert_allclose(rec[REC["COM"]:REC["COM"] + 3], num.com(q), atol=1e-12)
    np.testing.assert_allclose(rec[REC["H"]:REC["H"] + 6], num.momentum(q, v), atol=2e-7)
    # values-only entry point agrees with the record
    f0, c0, h0 = oracle_mod.kinematics(np.concatenate([q, v]))
    np.testing.assert_allclose(f0[0], feet, atol=1e-13)
    np.testing.assert_allclose(c0[0], rec[REC["COM"]:REC["COM"] + 3], atol=1e-13)
    np.testing.assert_allclose(h0[0], rec[REC["H"]:REC["H"] + 6], atol=1e-12)


def _fd(fun, x, eps=1e-6):
    f0 = fun(x)
    J = np.zeros((f0.size, x.size))
    for j in range(x.size):
        xp, xm = x.copy(), x.copy()
        xp[j] += eps
        xm[j] -= eps
        J[:, j] = (fun(xp) - fun(xm)) / (2 * eps)
    return J


@pytest.mark.parametrize("seed", range(3))
def test_jacobians_match_finite_differences(oracle_mod, num, seed):
    q, v = _rand_qv(np.random.default_rng(100 + seed))
    rec = oracle_mod.model_record(np.concatenate([q, v]))

    def r(qq, vv):
        return oracle_mod.model_record(np.concatenate([qq, vv]))

    FJ = rec[REC["FOOT_J"]:REC["FOOT_J"] + 12 * NQ].reshape(12, NQ)
    np.testing.assert_allclose(FJ, _fd(lambda x: num.feet(x).ravel(), q), atol=1e-8)
    # foot velocity = J v and its q-derivative
    FV = rec[REC["FOOT_V"]:REC["FOOT_V"] + 12]
    np.testing.assert_allclose(FV, FJ @ v, atol=1e-12)
    FVQ = rec[REC["FOOT_VQ"]:REC["FOOT_VQ"] + 12 * NQ].reshape(12, NQ)
    np.testing.assert_allclose(FVQ, _fd(lambda x: r(x, v)[REC["FOOT_V"]:REC["FOOT_V"] + 12], q), atol=1e-7)
    CJ = rec[REC["COM_J"]:REC["COM_J"] + 3 * NQ].reshape(3, NQ)
    np.testing.assert_allclose(CJ, _fd(num.com, q), atol=1e-8)
    HQ = rec[REC["H_Q"]:REC["H_Q"] + 6 * NQ].reshape(6, NQ)
    HV = rec[REC["H_V"]:REC["H_V"] + 6 * NQ].reshape(6, NQ)
    np.testing.assert_allclose(HQ, _fd(lambda x: r(x, v)[REC["H"]:REC["H"] + 6], q), atol=1e-6)
    np.testing.assert_allclose(HV, _fd(lambda x: r(q, x)[REC["H"]:REC["H"] + 6], v), atol=1e-7)
    # momentum is linear in v: h = H_V v ; linear part = m * com velocity
    np.testing.assert_allclose(HV @ v, rec[REC["H"]:REC["H"] + 6], atol=1e-10)
    np.testing.assert_allclose(rec[REC["H"]:REC["H"] + 3], num.mass * (CJ @ v), atol=1e-10)


def test_total_mass_and_stand_height(oracle_mod, num):
    from iterative_learning_nmpc_b200 import workloads as WL
    assert abs(num.mass - WL.MASS) < 1e-9
    qv = WL.P.nominal_state(WL.STAND_HEIGHT)
    feet, _, hg = oracle_mod.kinematics(qv)
    np.testing.assert_allclose(feet[0, :, 2], 0.0, atol=1e-9)     # feet on the ground
    np.testing.assert_allclose(hg, 0.0, atol=1e-14)


@pytest.mark.parametrize("seed", range(3))
def test_inverse_dynamics_matches_newton_euler(oracle_mod, num, seed):
    """Generated Lagrangian inverse dynamics (stand-in for pin.rnea - J^T f, dynamics.py:136-163)
    vs a numeric Newton-Euler pass with finite-difference link accelerations."""
    rng = np.random.default_rng(100 + seed)
    q, v = _rand_qv(rng)
    a = rng.uniform(-5, 5, NQ)
    f = rng.uniform(-30, 60, (4, 3))
    tau = oracle_mod.inverse_dynamics(np.concatenate([q, v, a]), f.reshape(-1))[0]
    ref = num.joint_torques(q, v, a, f)
    assert np.abs(tau - ref).max() < 2e-4 * max(1.0, np.abs(ref).max()), (tau, ref)


def test_inverse_dynamics_static_is_gravity_minus_contacts(oracle_mod, num):
    """At rest with zero forces the torques are pure gravity terms; adding the forces that carry the
    weight changes them by -J^T f exactly."""
    q = np.zeros(NQ); q[2] = 0.3; q[6:] = np.tile([0.0, 0.9, -1.8], 4)
    z = np.zeros(NQ)
    rec = oracle_mod.model_record(np.concatenate([q, z]))
    J = rec[REC["FOOT_J"]:REC["FOOT_J"] + 12 * NQ].reshape(4, 3, NQ)
    f = np.tile([0.0, 0.0, num.mass * 9.81 / 4], (4, 1))
    t0 = oracle_mod.inverse_dynamics(np.concatenate([q, z, z]), np.zeros(12))[0]
    t1 = oracle_mod.inverse_dynamics(np.concatenate([q, z, z]), f.reshape(-1))[0]
    expect = -np.einsum("fij,fi->j", J, f)[6:]
    assert np.allclose(t1 - t0, expect, atol=1e-12)


def test_plan_torques_restatement_properties(oracle_mod):
    """oracle.plan_torques (checker of nmpc_plan_torques): Hermite interpolation reproduces the node values
    at the node times, the inputs are held per `id_repeat`, and the PD term tracks the next plan sample."""
    O = oracle_mod
    rng = np.random.default_rng(0)
    B, N = 3, 25
    x = rng.normal(0, 0.3, (B, N + 1, 42))
    u = rng.normal(0, 1.0, (B, N, 30))
    n_interp = 10 * N
    r = O.plan_torques(x, u, 0.04, n_interp, kp=20.0, kd=1.75)
    for k in range(1, N + 1):                     # sample 10k-1 of the plan is t = t_k
        np.testing.assert_allclose(r["q_plan"][:, 10 * k - 1], x[:, k, :18], atol=1e-12)
        np.testing.assert_allclose(r["v_plan"][:, 10 * k - 1], x[:, k, 18:36], atol=1e-12)
    # on the plan itself the PD term is the step to the NEXT plan sample (compute_torques_ff reads plan_step, _step()
    # increments it, _compute_pd_torques reads the incremented one: mpc.py:575-599); it vanishes on the last sample
    nxt = np.minimum(np.arange(n_interp) + 1, n_interp - 1)
    pd = 20.0 * (r["q_plan"][:, nxt, 6:] - r["q_plan"][:, :, 6:]) + 1.75 * (r["v_plan"][:, nxt, 6:] - r["v_plan"][:, :, 6:])
    np.testing.assert_allclose(r["tau"], r["tau_ff"] + pd, atol=1e-10)
    np.testing.assert_allclose(r["tau"][:, -1], r["tau_ff"][:, -1], atol=1e-12)
    np.testing.assert_allclose(r["action"], (r["tau"] + 1.75 * r["v_plan"][:, :, 6:]) / 20.0 + r["q_plan"][:, :, 6:],
                               atol=1e-12)
    idr = np.int32(np.linspace(0, 1, n_interp) * (N - 1))
    qva = np.concatenate([r["q_plan"][0], r["v_plan"][0], u[0, idr, :18]], axis=1)
    np.testing.assert_allclose(r["tau_ff"][0], O.inverse_dynamics(qva, u[0, idr, 18:]), atol=1e-12)
    sub = O.plan_torques(x, u, 0.04, n_interp, step0=40, n_steps=7)
    np.testing.assert_array_equal(sub["q_plan"], r["q_plan"][:, 40:47])
