"""nmpc_kinematics_static -- the allocation-free entry point (include/nmpc_b200.h) -- against nmpc_kinematics, through
the C ABI.  The file name makes pytest collect it last: the entry point is an addition outside the reference's
interface, and a failure here must not hide the parity tests under `pytest -x`.  It passes alone in a fresh process
(the first call on the device does the set-up) and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-12          # the tolerance of test_kinematics_parity; the two entry points run the same kernel, so 0 is expected


def _inputs(n, seed):
    from iterative_learning_nmpc_b200 import problem as P
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray(np.tile(P.nominal_state(0.3), (n, 1)) + rng.normal(0, 0.3, (n, 36)))


def _static(lib, qv, n=None, stream=None):
    from iterative_learning_nmpc_b200._lib import ptr
    n = len(qv) if n is None else n
    feet, com, hg = np.full((len(qv), 4, 3), np.nan), np.full((len(qv), 3), np.nan), np.full((len(qv), 6), np.nan)
    rc = lib.nmpc_kinematics_static(ptr(qv), n, ptr(feet), ptr(com), ptr(hg), stream)
    return rc, feet, com, hg


def test_static_kinematics_matches_the_allocating_entry_point_twice_in_a_row():
    from iterative_learning_nmpc_b200._lib import default_opts, load_library
    from iterative_learning_nmpc_b200.core import NmpcSolver
    lib = load_library()
    h = NmpcSolver(4, default_opts())
    n_max = 4096                                     # NMPC_STATIC_MAX_N
    # two calls in a row on the shared array: a full one, then a smaller one with other inputs (stale contents of the
    # first call lie behind it and must not show), then the full size again
    for k, n in enumerate((n_max, 1000, n_max)):
        qv = _inputs(n, 100 + k)
        rc, feet, com, hg = _static(lib, qv)
        assert rc == 0, rc
        feet_r, com_r, hg_r = h.kinematics(qv)
        for a, b in ((feet, feet_r), (com, com_r), (hg, hg_r)):
            assert np.isfinite(a).all()
            assert np.abs(a - b).max() <= TOL * max(1.0, np.abs(b).max())
        assert np.abs(com_r).max() > 0.1
    # NULL outputs are skipped, as in nmpc_kinematics
    from iterative_learning_nmpc_b200._lib import ptr
    com = np.zeros((5, 3))
    assert lib.nmpc_kinematics_static(ptr(_inputs(5, 7)), 5, None, ptr(com), None, None) == 0
    assert np.abs(com - h.kinematics(_inputs(5, 7))[1]).max() <= TOL


def test_static_kinematics_rejects_what_does_not_fit(capfd):
    from iterative_learning_nmpc_b200._lib import load_library, ptr
    lib = load_library()
    qv = _inputs(8, 1)
    out = np.zeros((8, 3))
    assert lib.nmpc_kinematics_static(ptr(qv), 4097, None, ptr(out), None, None) == -5        # NMPC_E2BIG
    assert "NMPC_STATIC_MAX_N" in capfd.readouterr().err
    assert lib.nmpc_kinematics_static(ptr(qv), 0, None, ptr(out), None, None) == -1           # NMPC_EINVAL
    assert lib.nmpc_kinematics_static(None, 8, None, ptr(out), None, None) == -1
    assert not out.any()                             # nothing was written by the refused calls


def test_static_kinematics_on_a_caller_stream_with_device_buffers():
    """Device inputs and outputs on the caller's own stream: the serving-loop use."""
    import torch
    from iterative_learning_nmpc_b200._lib import default_opts, load_library
    from iterative_learning_nmpc_b200.core import NmpcSolver
    lib = load_library()
    h = NmpcSolver(4, default_opts())
    qv = _inputs(777, 3)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        d_qv = torch.from_numpy(qv).cuda()
        d_feet = torch.zeros(777, 12, dtype=torch.float64, device="cuda")
        d_hg = torch.zeros(777, 6, dtype=torch.float64, device="cuda")
    st.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.nmpc_kinematics_static(C.c_void_p(d_qv.data_ptr()), 777, C.c_void_p(d_feet.data_ptr()), None,
                                    C.c_void_p(d_hg.data_ptr()), C.c_void_p(st.cuda_stream))
    assert rc == 0 and torch.cuda.memory_allocated() == before
    feet_r, _, hg_r = h.kinematics(qv)
    assert np.abs(d_feet.cpu().numpy().reshape(777, 4, 3) - feet_r).max() <= TOL
    assert np.abs(d_hg.cpu().numpy() - hg_r).max() <= TOL * max(1.0, np.abs(hg_r).max())
