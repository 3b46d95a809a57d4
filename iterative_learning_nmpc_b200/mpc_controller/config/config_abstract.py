"""Configuration dataclasses of the MPC: `GaitConfig`, `MPCOptConfig`, `MPCCostConfig`.

API mirror of the reference's `mpc_controller/config/config_abstract.py:9-146` (same field
names, defaults, validation and helper methods) with one difference: `HPIPM_MODE` is defined
here instead of being imported from the missing `contact_tamp` submodule
(`config_abstract.py:5`).  Validation is done in `__post_init__` by `_check_fields`, which plays
the role of the reference's `@typechecked` decorator without depending on `typeguard`.
"""
from __future__ import annotations

import dataclasses
import enum
from dataclasses import dataclass
from typing import Tuple

import numpy as np


class HPIPM_MODE(enum.IntEnum):
    """HPIPM parameter presets (acados `qp_solver_hpipm_mode`).  Values match
    `NMPC_HPIPM_*` in include/nmpc_b200.h; only `speed` is implemented on the device."""
    speed_abs = 0
    speed = 1
    balance = 2
    robust = 3

    def acados_name(self) -> str:
        return self.name.upper()


_KINDS = {
    "str": (str,), "float": (float, int, np.floating, np.integer), "int": (int, np.integer),
    "bool": (bool, np.bool_), "np.ndarray": (np.ndarray,), "HPIPM_MODE": (HPIPM_MODE,),
}


def _check_fields(obj) -> None:
    """Runtime type check of dataclass fields (what `typeguard.typechecked` does upstream)."""
    for f in dataclasses.fields(obj):
        kinds = _KINDS.get(f.type if isinstance(f.type, str) else getattr(f.type, "__name__", ""))
        val = getattr(obj, f.name)
        if kinds is not None and not isinstance(val, kinds):
            raise TypeError(f"{type(obj).__name__}.{f.name}: expected {f.type}, got {type(val).__name__}")
        if kinds == _KINDS["int"] and isinstance(val, bool):
            raise TypeError(f"{type(obj).__name__}.{f.name}: expected int, got bool")


@dataclass
class GaitConfig:
    gait_name: str            # name of the gait
    nominal_period: float     # gait period (s)
    stance_ratio: np.ndarray  # per foot, fraction of the period in stance, in [0,1]
    phase_offset: np.ndarray  # per foot, phase of touch-down, in [0,1]
    nom_height: float         # nominal base height (m)
    step_height: float        # swing apex height (m)

    def __post_init__(self):
        _check_fields(self)
        assert all(0 <= r <= 1 for r in self.stance_ratio), "stance_ratio should be in [0,1]"
        assert all(0 <= r <= 1 for r in self.phase_offset), "phase_offset should be in [0,1]"


@dataclass
class MPCOptConfig:
    time_horizon: float                 # horizon (s)
    n_nodes: int                        # shooting intervals
    replanning_freq: int                # Hz
    Kp: float                           # joint PD gains of the torque controller
    Kd: float
    recompile: bool                     # regenerate + rebuild the model code
    max_iter: int                       # SQP iterations
    max_qp_iter: int                    # IPM iterations per QP
    real_time_it: bool = False          # SQP_RTI
    enable_time_opt: bool = False       # optimise node durations
    opt_dt_scale: Tuple[float, float] = (0.5, 1.75)   # dt bounds as ratios of the nominal dt
    enable_impact_dyn: bool = False
    opt_peak: bool = True               # swing-apex ("peak") cost active
    warm_start_sol: bool = True         # shift the previous solution into the iterate
    warm_start_nlp: bool = True
    warm_start_qp: bool = True
    hpipm_mode: HPIPM_MODE = HPIPM_MODE.speed
    use_cython: bool = False
    torque_limit: bool = True
    mu: float = 0.7
    nlp_tol: float = 1.e-1              # SQP exit tolerance
    qp_tol: float = 1.e-2               # IPM exit tolerance

    def __post_init__(self):
        _check_fields(self)
        assert len(self.opt_dt_scale) == 2, "opt_dt_scale must be of shape 2"
        assert self.mu > 0, "Friction coefficient must be positive"

    def get_dt_nodes(self) -> float:
        """Nominal node duration."""
        return round(self.time_horizon / self.n_nodes, 4)

    def get_dt_bounds(self) -> Tuple[float, float]:
        """(dt_min, dt_max) for time optimisation."""
        dt = self.get_dt_nodes()
        lo, hi = self.opt_dt_scale
        return round(dt * lo, 4), round(dt * hi, 4)


@dataclass
class MPCCostConfig:
    robot_name: str
    gait_name: str
    W_e_base: np.ndarray        # terminal base weights [pos(3) ypr(3) lin vel(3) ang vel(3)]
    W_base: np.ndarray          # running base weights, same layout
    W_joint: np.ndarray         # running joint weights [pos(12) vel(12)]
    W_e_joint: np.ndarray       # terminal joint weights
    W_acc: np.ndarray           # joint acceleration weights (12)
    W_swing: np.ndarray         # swing-height weights, per foot
    W_eeff_ori: np.ndarray      # unused upstream as well
    W_cnt_f_reg: np.ndarray     # force regularisation per foot [[x,y,z],...]
    W_foot_pos_constr_stab: np.ndarray   # per-foot contact stabilisation gain (p_gain parameter)
    W_foot_displacement: np.ndarray      # foot target penalty when contacts are restricted
    cnt_radius: float           # contact patch radius when restricted
    time_opt: np.ndarray        # weight of dt - dt_nominal
    reg_eps: float              # Levenberg-Marquardt term, running
    reg_eps_e: float            # Levenberg-Marquardt term, terminal

    def __post_init__(self):
        _check_fields(self)
        assert len(self.W_e_base) == 12, "W_e_base must be of shape 12"
        assert len(self.W_base) == 12, "W_base must be of shape 12"
        assert len(self.W_acc) == 12, "W_acc must be of shape 12"
        assert (len(self.W_swing) == len(self.W_cnt_f_reg)
                and len(self.W_swing) == len(self.W_foot_pos_constr_stab)), \
            "W_swing and W_foot should have the same length."
        for i, w in enumerate(self.W_cnt_f_reg):
            assert len(w) == 3, f"W_foot[{i}] must be of shape 3"
