"""Batched rigid-body rollout: physical checks of the forward dynamics and the contact model on the
CPU restatement (oracle/nmpc_oracle.c, oracle_sim_*), which the GPU kernel is then held to (tests/test_gpu_sim.py)."""
import numpy as np

from iterative_learning_nmpc_b200 import problem as P

SP = dict(sim_dt=1e-3, kp=20.0, kd=1.5, k_n=2.0e4, d_n=60.0, d_t=60.0, mu=0.8, ground_z=0.0, tau_max=0.0)


def test_forward_dynamics_terms_match_the_generated_inverse_dynamics(oracle_mod):
    """M(q) a + bias(q, v) on the joint rows equals the generated rnea(q, v, a) (the code LocomotionMPC's feed-forward
    torques come from); M is symmetric positive definite with the total mass on the translation block."""
    O = oracle_mod
    rng = np.random.default_rng(0)
    for _ in range(5):
        qv = P.nominal_state(0.3)
        qv[:18] += rng.normal(0, 0.3, 18); qv[18:] = rng.normal(0, 1.0, 18)
        M, b = O.sim_terms(qv)
        assert np.abs(M - M.T).max() == 0.0 and np.linalg.eigvalsh(M).min() > 1e-4
        assert np.allclose(M[:3, :3], O.model_mass() * np.eye(3), atol=1e-12)
        a = rng.normal(0, 2.0, 18)
        tau = O.inverse_dynamics(np.concatenate([qv, a])[None], np.zeros((1, 12)))[0]
        assert np.abs((M @ a + b)[6:] - tau).max() < 1e-10


def test_free_fall(oracle_mod):
    """No contact, no torque: the base falls with g (semi-implicit Euler: z_n = z_0 - g dt^2 n (n + 1) / 2) and the
    centroidal linear momentum follows m g t."""
    O = oracle_mod
    sp = dict(SP, ground_z=-100.0, kp=0.0, kd=0.0)
    n = 300
    qe, traj, fc = O.sim_physics(sp, P.nominal_state(0.5), np.zeros(12), n)
    assert np.abs(fc).max() == 0.0
    _, _, hg = O.kinematics(qe[None])
    assert abs(hg[0, 2] + O.model_mass() * 9.81 * n * 1e-3) < 1e-9 * 50 and np.abs(hg[0, :2]).max() < 1e-9
    com0 = O.kinematics(P.nominal_state(0.5)[None])[1][0]
    com1 = O.kinematics(qe[None])[1][0]
    assert abs((com1[2] - com0[2]) + 9.81 * 1e-6 * n * (n + 1) / 2) < 1e-6


def test_standing_on_the_ground_carries_the_weight(oracle_mod):
    """PD on the home pose with gravity-compensating feed-forward torques: the robot settles on the compliant ground,
    the four normal forces add up to m g and the feet do not slide."""
    O = oracle_mod
    q0 = P.nominal_state(0.30)
    feet = O.kinematics(q0[None])[0][0]
    q0[2] -= feet[:, 2].mean() - 0.001            # feet 1 mm above the ground
    qv = q0.copy()
    mass = O.model_mass()
    for _ in range(1500):                         # PD recomputed every step on the measured state
        f_ff = np.tile([0.0, 0.0, mass * 9.81 / 4], 4)
        tau = O.inverse_dynamics(np.concatenate([qv, np.zeros(18)])[None], f_ff[None])[0]
        tau += 40.0 * (q0[6:18] - qv[6:18]) - 3.0 * qv[24:36]
        qv, _, fc = O.sim_physics(SP, qv, tau, 1)
    fn = fc[0].reshape(4, 3)[:, 2]
    assert abs(fn.sum() - mass * 9.81) < 0.02 * mass * 9.81 and (fn > 0.1 * mass * 9.81 / 4).all()
    assert np.abs(qv[18:]).max() < 2e-2 and abs(qv[2] - q0[2]) < 0.01 and np.abs(qv[3:6]).max() < 0.05


def test_friction_cone_caps_the_tangential_force(oracle_mod):
    O = oracle_mod
    q0 = P.nominal_state(0.30)
    feet = O.kinematics(q0[None])[0][0]
    q0[2] -= feet[:, 2].mean() + 0.002            # 2 mm into the ground
    q0[18] = 3.0                                  # sliding fast along x
    _, _, fc = O.sim_physics(SP, q0, np.zeros(12), 1)
    f = fc[0].reshape(4, 3)
    assert (f[:, 2] > 0).all()
    assert np.allclose(np.hypot(f[:, 0], f[:, 1]), SP["mu"] * f[:, 2], rtol=1e-12) and (f[:, 0] < 0).all()
