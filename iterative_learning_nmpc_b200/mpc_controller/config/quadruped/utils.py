"""`get_quadruped_config(gait_name, robot_name)`, reference `config/quadruped/utils.py:7-16`."""
from typing import Tuple

from ..config_abstract import GaitConfig, MPCCostConfig, MPCOptConfig
from .mpc_cost import CostConfigFactory
from .mpc_gait import GaitConfigFactory
from .mpc_opt import MPCQuadrupedCyclic


def get_quadruped_config(gait_name: str, robot_name: str, *, fallback_to_trot_cost: bool = False,
                         fix_swapped_heights: bool = False
                         ) -> Tuple[GaitConfig, MPCOptConfig, MPCCostConfig]:
    gait = GaitConfigFactory.get(gait_name, fix_swapped_heights=fix_swapped_heights)
    cost = CostConfigFactory.get(robot_name, gait_name, fallback_to_trot=fallback_to_trot_cost)
    return gait, MPCQuadrupedCyclic(), cost
