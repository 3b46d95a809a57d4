"""ctypes binding of libnmpc_b200.so (include/nmpc_b200.h).  No fallback: if the library is
missing or no B200-class device is present, construction raises."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NMPC_LIB") or os.path.join(PKG, "libnmpc_b200.so")   # NMPC_LIB: debug builds

NX, NU, NZ, NP, NY, NYE, NG = 42, 30, 72, 44, 76, 52, 42
STATUS_NAMES = {0: "SUCCESS", 1: "NAN", 2: "MAXITER", 3: "MINSTEP", 4: "QP_FAILURE"}


class NmpcOpts(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32), ("max_iter", C.c_int32), ("max_qp_iter", C.c_int32),
        ("real_time_it", C.c_int32), ("warm_start_qp", C.c_int32), ("warm_start_nlp", C.c_int32),
        ("hpipm_mode", C.c_int32), ("enable_time_opt", C.c_int32),
        ("warm_start_first_qp", C.c_int32), ("cond_pred_corr", C.c_int32),
        ("dt_nodes", C.c_double), ("dt_min", C.c_double), ("dt_max", C.c_double),
        ("nlp_tol", C.c_double), ("qp_tol", C.c_double),
        ("reg_eps", C.c_double), ("reg_eps_e", C.c_double), ("mu_contact", C.c_double),
        ("time_opt_weight", C.c_double),
    ]

    def copy(self) -> "NmpcOpts":
        o = NmpcOpts()
        C.memmove(C.byref(o), C.byref(self), C.sizeof(NmpcOpts))
        return o


class NmpcSimParams(C.Structure):
    """nmpc_sim_params: batched rigid-body rollout (include/nmpc_b200.h)."""
    _fields_ = [("sim_dt", C.c_double), ("kp", C.c_double), ("kd", C.c_double), ("k_n", C.c_double),
                ("d_n", C.c_double), ("d_t", C.c_double), ("mu", C.c_double), ("ground_z", C.c_double),
                ("tau_max", C.c_double)]


class NmpcRaibertParams(C.Structure):
    """nmpc_raibert_params: RaiberContactPlanner on the device (include/nmpc_b200.h)."""
    _fields_ = [("offset_hip_b", C.c_double * 12), ("stance_ratio", C.c_double * 4), ("nominal_period", C.c_double),
                ("foot_size", C.c_double), ("height_offset", C.c_double), ("gravity", C.c_double),
                ("cnt_radius", C.c_double)]


def default_sim_params(**kw) -> NmpcSimParams:
    p = NmpcSimParams(sim_dt=1.0e-3, kp=20.0, kd=1.5, k_n=2.0e4, d_n=60.0, d_t=60.0, mu=0.8, ground_z=0.0, tau_max=0.0)
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


class NmpcStats(C.Structure):
    _fields_ = [("sqp_iter", C.c_void_p), ("qp_iter", C.c_void_p), ("qp_status", C.c_void_p),
                ("res", C.c_void_p), ("qp_res", C.c_void_p), ("cost", C.c_void_p),
                ("qp_ncond", C.c_void_p)]


def default_opts(**kw) -> NmpcOpts:
    o = NmpcOpts(n_nodes=25, max_iter=1, max_qp_iter=6, real_time_it=0, warm_start_qp=1,
                 warm_start_nlp=1, hpipm_mode=1, enable_time_opt=0, warm_start_first_qp=0, cond_pred_corr=1, dt_nodes=0.04, dt_min=0.02,
                 dt_max=0.07, nlp_tol=1e-1, qp_tol=1e-2, reg_eps=1e-6, reg_eps_e=1e-5, mu_contact=0.8,
                 time_opt_weight=1e4)
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


class NmpcError(RuntimeError):
    pass


_lib = None

_SIGS = {
    "nmpc_destroy": (None, [C.c_void_p]),
    "nmpc_set_opts": (C.c_int, [C.c_void_p, C.POINTER(NmpcOpts)]),
    "nmpc_get_dims": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "nmpc_stream": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "nmpc_set_x0": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_set_x0_full": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_set_yref": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_set_weights": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]),
    "nmpc_set_params": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_pin_stance_feet": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_set_guess": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_set_dt": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_get_dt": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_shift_warm_start": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "nmpc_reset_duals": (C.c_int, [C.c_void_p]),
    "nmpc_solve": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_synchronize": (C.c_int, [C.c_void_p]),
    "nmpc_get_solution": (C.c_int, [C.c_void_p] + [C.c_void_p] * 4),
    "nmpc_get_stats": (C.c_int, [C.c_void_p, C.POINTER(NmpcStats)]),
    "nmpc_get_status": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_set_output": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_pack_solution": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_get_qp_step": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_kinematics": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32] + [C.c_void_p] * 3),
    "nmpc_kinematics_static": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 4),
    "nmpc_inverse_dynamics": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "nmpc_init_tick": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int32]),
    "nmpc_plan_tick": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_double,
                                 C.c_void_p, C.c_double, C.c_void_p, C.c_int32]),
    "nmpc_plan_tick_raibert": (C.c_int, [C.c_void_p] * 5 + [C.c_int32] + [C.c_void_p] * 4
                               + [C.POINTER(NmpcRaibertParams), C.c_double, C.c_double, C.c_void_p, C.c_double,
                                  C.c_void_p, C.c_int32, C.c_void_p]),
    "nmpc_get_tracking": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_cold_guess": (C.c_int, [C.c_void_p]),
    "nmpc_get_problem_data": (C.c_int, [C.c_void_p] + [C.c_void_p] * 4),
    "nmpc_plan_torques": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double]
                          + [C.c_void_p] * 5),
    "nmpc_state_bytes": (C.c_int64, [C.c_void_p]),
    "nmpc_export_state": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_import_state": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_sim_set_state": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nmpc_sim_get_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_sim_steps": (C.c_int, [C.c_void_p, C.POINTER(NmpcSimParams), C.c_int32, C.c_int32, C.c_int32,
                                 C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_export_qp": (C.c_int, [C.c_void_p] + [C.c_void_p] * 6),
    "nmpc_solve_qp": (C.c_int, [C.c_void_p] + [C.c_void_p] * 13),
    "nmpc_launch_count": (C.c_int64, [C.c_void_p]),
    "nmpc_set_profiling": (C.c_int, [C.c_void_p, C.c_int32]),
    "nmpc_get_kernel_times": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "nmpc_last_qp_path": (C.c_int, [C.c_void_p]),
    "nmpc_version": (C.c_char_p, []),
    "nmpc_model_mass": (C.c_double, []),
}
EXPORTED_SYMBOLS = ["nmpc_create"] + list(_SIGS)


def load_library():
    """Load libnmpc_b200.so; raises NmpcError if it has not been built (no CPU fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NmpcError(f"{LIB_PATH} not found: run `python -m iterative_learning_nmpc_b200.build` "
                            "(nvcc, sm_100a). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        lib.nmpc_create.restype = C.c_void_p
        lib.nmpc_create.argtypes = [C.POINTER(NmpcOpts), C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def ptr(a):
    """Raw address of a numpy array (C-contiguous), a torch tensor (contiguous) or None."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return C.c_void_p(a.ctypes.data)
    if hasattr(a, "data_ptr"):
        if not a.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


def check(rc: int, what: str) -> None:
    if rc != 0:
        kind = {-1: "EINVAL", -2: "ENOSYS (option not available in this build)", -3: "ENODEV", -4: "ENOMEM",
                -5: "E2BIG (beyond a fixed capacity)"}.get(
            rc, f"cudaError {rc}" if rc > 0 else str(rc))
        raise NmpcError(f"{what} failed: {kind}")
