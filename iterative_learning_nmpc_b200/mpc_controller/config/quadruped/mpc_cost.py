"""Cost weights for Go2.  Values from the reference's `config/quadruped/mpc_cost.py:15-144`.

The reference only ships `trot` and `slow_trot` costs and raises ValueError for every other gait
(`mpc_cost.py:142-144`).  `CostConfigFactory.get(..., fallback_to_trot=True)` reuses the trot
weights for the other gaits; the mixed-gait benchmark configuration needs that.
"""
from dataclasses import replace

import numpy as np

from ..config_abstract import MPCCostConfig

N_FEET = 4
HIP_SHOULDER_ELBOW_SCALE = [15., 5., 1.]
_LEG = np.tile(HIP_SHOULDER_ELBOW_SCALE, N_FEET)          # 12 joint scales


def _joint(pos_scale, vel_value, scale):
    return np.concatenate((pos_scale, np.full(12, vel_value))) * scale


def Go2TrotCost() -> MPCCostConfig:
    return MPCCostConfig(
        robot_name="Go2", gait_name="trot",
        W_base=np.array([1e3, 3e3, 1e2, 5e2, 5e2, 5e2, 5e2, 1e1, 1e0, 1e0, 2e1, 1e1]),
        W_e_base=np.array([1e1, 1e1, 1e3, 1e1, 1e2, 1e2, 5e2, 5e2, 1e3, 1e1, 1e2, 1e2]),
        W_joint=_joint(_LEG, 0.03, 5.),
        W_e_joint=_joint(_LEG, 0.1, 1.),
        W_acc=_LEG * 5.0e-4,
        W_swing=np.full(N_FEET, 2e4),
        W_eeff_ori=np.full(N_FEET, 1.),
        W_cnt_f_reg=np.tile([0.01, 0.01, 0.05], (N_FEET, 1)),
        W_foot_pos_constr_stab=np.full(N_FEET, 5e1),
        W_foot_displacement=np.array([1e3]),
        cnt_radius=0.015,
        time_opt=np.array([1.0e4]),
        reg_eps=1.0e-6, reg_eps_e=1.0e-5,
    )


_W_SLOW = np.array([0., 0., 5e3, 0., 3e3, 3e3, 0., 0., 1e1, 1e0, 1e2, 2e2])


def Go2SlowTrotCost() -> MPCCostConfig:
    return MPCCostConfig(
        robot_name="Go2", gait_name="slow_trot",
        W_base=_W_SLOW * 7., W_e_base=_W_SLOW * 10.,
        W_joint=_joint(_LEG, 0., 0.1),
        W_e_joint=_joint(_LEG, 0., 0.),
        W_acc=np.tile([7., 3., 1.], N_FEET) * 1.e-2,
        W_swing=np.full(N_FEET, 5e5),
        W_eeff_ori=np.full(N_FEET, 0.),
        W_cnt_f_reg=np.tile([1.2, 1.2, 0.9], (N_FEET, 1)),
        W_foot_pos_constr_stab=np.full(N_FEET, 5e1),
        W_foot_displacement=np.array([1e6]),
        cnt_radius=0.005,
        time_opt=np.array([1.0e4]),
        reg_eps=1.0e-6, reg_eps_e=1.0e-5,
    )


class CostConfigFactory:
    AVAILABLE_COST = {
        (cfg.robot_name.lower(), cfg.gait_name.lower()): cfg
        for cfg in (Go2TrotCost(), Go2SlowTrotCost())
    }

    @staticmethod
    def get(robot_name: str, gait_name: str, fallback_to_trot: bool = False) -> MPCCostConfig:
        key = (robot_name.lower(), gait_name.lower())
        cfg = CostConfigFactory.AVAILABLE_COST.get(key)
        if cfg is None and fallback_to_trot:
            base = CostConfigFactory.AVAILABLE_COST.get((robot_name.lower(), "trot"))
            if base is not None:
                cfg = replace(base, gait_name=gait_name)
        if cfg is None:
            raise ValueError(f"Cost config: {gait_name} for {robot_name} not available.")
        return cfg
