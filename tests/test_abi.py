"""The C-ABI library loads without a GPU and exports every symbol include/nmpc_b200.h declares.
No compute call is made here (that is the `-m gpu` suite)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "nmpc_b200.h")


@pytest.fixture(scope="module")
def lib():
    from iterative_learning_nmpc_b200 import build as product_build
    from iterative_learning_nmpc_b200 import _lib
    product_build.build()
    return _lib.load_library()


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nmpc_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/nmpc_b200.h but not exported"


def test_binding_table_covers_header():
    from iterative_learning_nmpc_b200 import _lib
    assert set(_declared()) == set(_lib.EXPORTED_SYMBOLS)


def test_opts_struct_layout_matches_header():
    from iterative_learning_nmpc_b200._lib import NmpcOpts
    # 10 x int32 followed by 9 x double, no padding
    assert C.sizeof(NmpcOpts) == 10 * 4 + 9 * 8
    assert NmpcOpts.time_opt_weight.offset == 40 + 8 * 8
    assert NmpcOpts.warm_start_first_qp.offset == 32 and NmpcOpts.cond_pred_corr.offset == 36
    assert NmpcOpts.dt_nodes.offset == 40 and NmpcOpts.mu_contact.offset == 40 + 7 * 8


def test_version_and_argument_errors(lib):
    assert b"sm_100a" in lib.nmpc_version()
    err = C.c_int32(0)
    assert not lib.nmpc_create(None, 1, 0, C.byref(err)) and err.value == -1       # EINVAL
    from iterative_learning_nmpc_b200._lib import default_opts
    o = default_opts(enable_time_opt=1, dt_min=0.05, dt_max=0.03)                  # inconsistent dt bounds
    assert not lib.nmpc_create(C.byref(o), 1, 0, C.byref(err)) and err.value == -1  # EINVAL
    assert lib.nmpc_solve(None, None) == -1
    assert lib.nmpc_launch_count(None) == 0


def test_no_cpu_fallback_without_a_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from iterative_learning_nmpc_b200._lib import NmpcError
    from iterative_learning_nmpc_b200.core import NmpcSolver
    with pytest.raises(NmpcError, match="ENODEV"):
        NmpcSolver(2)


def test_product_does_not_import_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pkg = os.path.join(ROOT, "iterative_learning_nmpc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")) and f != "nmpc_gen_model.h":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "from oracle" not in txt and "import oracle" not in txt and "libnmpc_oracle" not in txt, f
