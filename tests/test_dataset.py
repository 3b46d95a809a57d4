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


def _records(n_rollouts=3, seed=0):
    rng = np.random.default_rng(seed)
    recs = []
    for i in range(n_rollouts):
        T = 5 + 3 * i
        recs.append(dict(state=rng.normal(size=(T, 44)), action=rng.normal(size=(T, 12)),
                         vc_goals=rng.normal(size=(T, 5)), cc_goals=rng.normal(size=(T, 8)),
                         time=0.001 * np.arange(T)))
    return recs


def test_hdf5_database_file_round_trips_through_the_independent_reader(tmp_path):
    """`save_hdf5` without h5py: the built-in writer's file parsed by tests/hdf5_reader.py (superblock -> symbol table
    -> object headers -> raw data, every signature / version / size checked) gives back the Database datasets of
    database.py:156-203 bit for bit."""
    from hdf5_reader import read_hdf5
    recs = _records()
    path = DS.save_hdf5(str(tmp_path / "db.h5"), recs)
    try:
        import h5py
        with h5py.File(path, "r") as hf:
            back = {k: hf[k][:] for k in hf.keys()}
    except ImportError:
        back = read_hdf5(path)
    assert sorted(back) == ["actions", "cc_goals", "states", "traj_ids", "traj_times", "vc_goals"]
    np.testing.assert_array_equal(back["states"], np.concatenate([r["state"] for r in recs]))
    np.testing.assert_array_equal(back["actions"], np.concatenate([r["action"] for r in recs]))
    np.testing.assert_array_equal(back["vc_goals"], np.concatenate([r["vc_goals"] for r in recs]))
    np.testing.assert_array_equal(back["cc_goals"], np.concatenate([r["cc_goals"] for r in recs]))
    np.testing.assert_array_equal(back["traj_times"], np.concatenate([r["time"] for r in recs]))
    assert back["traj_ids"].dtype == np.int64 and back["states"].dtype == np.float64
    np.testing.assert_array_equal(back["traj_ids"], np.repeat([0, 1, 2], [5, 8, 11]))
    with pytest.raises(ValueError):
        DS.save_hdf5(str(tmp_path / "empty.h5"), [])


def test_hdf5_min_writer_layout(tmp_path):
    """The built-in writer on its own: header constants, alignment, special values, the 8-dataset limit; the reader
    rejects a corrupted file instead of returning data."""
    from hdf5_reader import Hdf5FormatError, read_hdf5
    from iterative_learning_nmpc_b200.hdf5_min import write_hdf5
    d = {"a": np.array([[0.0, -0.0, np.inf], [np.nan, 1e-310, -1.5e300]]), "b": np.array([-2 ** 63, 2 ** 63 - 1, 0]),
         "zz": np.zeros((0, 7)), "m": np.arange(24, dtype=np.float32).reshape(2, 3, 4), "s": np.float64(3.5)}
    path = write_hdf5(str(tmp_path / "t.h5"), d)
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and int.from_bytes(raw[40:48], "little") == len(raw)
    back = read_hdf5(path)
    assert list(back) == sorted(d)
    assert back["a"].tobytes() == d["a"].tobytes()                # bit patterns: signed zero, NaN, subnormal
    np.testing.assert_array_equal(back["b"], d["b"])
    assert back["zz"].shape == (0, 7) and back["m"].shape == (2, 3, 4) and back["m"].dtype == np.float64
    np.testing.assert_array_equal(back["m"], d["m"])
    assert back["s"].shape == (1,) and back["s"][0] == 3.5
    with pytest.raises(ValueError):
        write_hdf5(str(tmp_path / "many.h5"), {f"d{i}": np.zeros(2) for i in range(9)})
    with pytest.raises(ValueError):
        write_hdf5(str(tmp_path / "none.h5"), {})
    for off, val in ((0, b"X"), (8, b"\x02"), (len(raw) - 1, None)):
        bad = bytearray(raw)
        if val is None:
            bad = bad[:-1]                                       # truncated: end-of-file address no longer matches
        else:
            bad[off:off + 1] = val
        (tmp_path / "bad.h5").write_bytes(bytes(bad))
        with pytest.raises(Hdf5FormatError):
            read_hdf5(str(tmp_path / "bad.h5"))


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
