"""Default solver options, reference `config/quadruped/mpc_opt.py:7-27`."""
from dataclasses import dataclass

from ..config_abstract import HPIPM_MODE, MPCOptConfig  # noqa: F401  (re-export)


@dataclass
class MPCQuadrupedCyclic(MPCOptConfig):
    time_horizon: float = 1.
    n_nodes: int = 25
    replanning_freq: int = 25
    Kp: float = 20.
    Kd: float = 1.75
    recompile: bool = False
    max_iter: int = 1
    max_qp_iter: int = 6
