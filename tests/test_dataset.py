"""Dataset writer (reference formats: Behavior_Cloning/utils/Rollout_MPC.py:183-298, database.py:156-203)."""
import numpy as np
import pytest

from iterative_learning_nmpc_b200 import dataset as DS
from iterative_learning_nmpc_b200.mpc_controller.utils.dynamics import QuadrupedDynamics


def _fake(n=7, seed=0):
    rng = np.random.default_rng(seed)
    q = rng.normal(0, 0.2, (n, 18)); v = rng.normal(0, 0.5, (n, 18))
    tau = rng.normal(0, 3.0, (n, 12)); feet = rng.normal(0, 0.3, (n, 4, 3))
    cnt = rng.integers(0, 2, (n, 4))
    return q, v, tau, feet, cnt


def test_record_layout_matches_the_reference_recorder():
    q, v, tau, feet, cnt = _fake()
    t = 0.001 * np.arange(1, 8)
    r = DS.rollout_record(t, q, v, tau, feet, cnt, [0.2, 0.0, 0.0], kp=20.0, kd=1.5)
    assert set(r) == set(DS.KEYS)
    assert r["state"].shape == (7, 44) and r["action"].shape == (7, 12) and r["q"].shape == (7, 19)
    for i in range(7):                                  # against the unbatched conversion of the front-end
        q_mj, v_mj = QuadrupedDynamics.convert_to_mujoco(q[i], v[i])
        np.testing.assert_allclose(r["q"][i], q_mj, atol=1e-15)
        np.testing.assert_allclose(r["v"][i], v_mj, atol=1e-15)
        bwf = np.concatenate([(q_mj[:3] - feet[i, f])[:2] for f in range(4)])
        np.testing.assert_allclose(r["base_wrt_feet"][i], bwf, atol=1e-15)
        np.testing.assert_allclose(r["state"][i], np.concatenate([[0.0], v_mj, q_mj[2:], bwf]), atol=1e-15)
        # the recorder reads ctrl in actuator order FR FL RR RL and maps it back to FL FR RL RR (Rollout_MPC.py:270-276)
        c = r["ctrl"][i]
        back = np.concatenate([c[3:6], c[0:3], c[9:12], c[6:9]])
        np.testing.assert_array_equal(back, tau[i])
        np.testing.assert_allclose(r["action"][i], (back + 1.5 * v_mj[6:]) / 20.0 + q_mj[7:], atol=1e-15)
    np.testing.assert_array_equal(r["contact_vec"], cnt)
    np.testing.assert_array_equal(r["vc_goals"], np.tile([0.2, 0.0, 0.0], (7, 1)))
    assert r["cc_goals"].shape == (7, 8)


def test_npz_round_trip_and_file_names(tmp_path):
    q, v, tau, feet, cnt = _fake(5, 1)
    r = DS.rollout_record(0.001 * np.arange(5), q, v, tau, feet, cnt, [0.1, 0.05, 0.0])
    p1 = DS.save_rollout(r, str(tmp_path), nominal=True, timestamp="20240101_000000")
    p2 = DS.save_rollout(r, str(tmp_path), nominal=False, replanning_point=3, nth_traj=12)
    assert p1.endswith("traj_nominal_20240101_000000.npz") and p2.endswith("traj_3_12.npz")
    back = np.load(p2)
    assert sorted(back.files) == sorted(DS.KEYS)
    for k in DS.KEYS:
        np.testing.assert_array_equal(back[k], r[k])
    bad = dict(r); bad.pop("state")
    with pytest.raises(KeyError):
        DS.save_rollout(bad, str(tmp_path))


def test_hdf5_needs_h5py(tmp_path):
    pytest.importorskip("numpy")
    try:
        import h5py  # noqa: F401
    except ImportError:
        with pytest.raises(RuntimeError):
            DS.save_hdf5(str(tmp_path / "db.h5"), [])


def test_euler_to_mujoco_vectorised_matches_the_per_sample_maps():
    """The vectorised conversion against `QuadrupedDynamics.convert_to_mujoco` sample by sample, including rotations
    beyond 120 degrees (the other branches of the quaternion extraction)."""
    from iterative_learning_nmpc_b200 import dataset as DS
    from iterative_learning_nmpc_b200.mpc_controller.utils.dynamics import QuadrupedDynamics as QD
    rng = np.random.default_rng(0)
    q = rng.normal(0, 0.5, (200, 18)); v = rng.normal(0, 1.0, (200, 18))
    q[:50, 3:6] = rng.uniform(-3.1, 3.1, (50, 3)); q[:50, 4] = rng.uniform(-1.4, 1.4, 50)
    qm, vm = DS.euler_to_mujoco(q, v)
    for i in range(200):
        a, b = QD.convert_to_mujoco(q[i], v[i])
        assert np.abs(qm[i] - a).max() < 1e-12 and np.abs(vm[i] - b).max() < 1e-12
