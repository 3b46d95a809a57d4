"""Quadruped gaits.  Values from the reference's `config/quadruped/mpc_gait.py:15-66`.

`pace` and `bound` ship upstream with `nom_height=0.05, step_height=0.32`, evidently swapped
(SURVEY.md section 5); the same numbers are kept here so the API returns what the reference
returns, and `GaitConfigFactory.get(name, fix_swapped_heights=True)` gives the corrected pair
used by the mixed-gait benchmark configuration.
"""
from dataclasses import dataclass, field, replace

import numpy as np

from ..config_abstract import GaitConfig

#        name        period  stance  offsets (FL, FR, RL, RR)  nom_h  step_h
_TABLE = [
    ("trot",      0.5, 0.5,  (0.5, 0.0, 0.0, 0.5),   0.30, 0.05),
    ("slow_trot", 1.0, 0.63, (0.5, 0.0, 0.0, 0.5),   0.32, 0.065),
    ("jump",      50., 0.4,  (0.0, 0.0, 0.0, 0.0),   0.30, 0.05),
    ("crawl",     1.0, 0.75, (0.0, 0.25, 0.5, 0.75), 0.30, 0.05),
    ("pace",      0.5, 0.6,  (0.0, 0.5, 0.5, 0.0),   0.05, 0.32),
    ("bound",     0.5, 0.6,  (0.0, 0.5, 0.5, 0.0),   0.05, 0.32),
]


@dataclass
class QuadrupedGaitConfig(GaitConfig):
    n_eeff: int = 4

    def __post_init__(self):
        super().__post_init__()
        assert len(self.stance_ratio) == self.n_eeff, f"stance_ratio must be of length {self.n_eeff}"
        assert len(self.phase_offset) == self.n_eeff, f"phase_offset must be of length {self.n_eeff}"


def _make(name, period, stance, offsets, nom_h, step_h):
    def build():
        return QuadrupedGaitConfig(
            gait_name=name, nominal_period=float(period),
            stance_ratio=np.full(4, stance), phase_offset=np.array(offsets, dtype=float),
            nom_height=nom_h, step_height=step_h)
    return build


_BUILDERS = {row[0]: _make(*row) for row in _TABLE}

# class-like callables with the reference's names
QuadrupedTrot = _BUILDERS["trot"]
QuadrupedSlowTrot = _BUILDERS["slow_trot"]
QuadrupedJump = _BUILDERS["jump"]
QuadrupedCrawl = _BUILDERS["crawl"]
QuadrupedPace = _BUILDERS["pace"]
QuadrupedBound = _BUILDERS["bound"]


class GaitConfigFactory:
    AVAILABLE_GAITS = {name: build() for name, build in _BUILDERS.items()}

    @staticmethod
    def get(gait_name: str, fix_swapped_heights: bool = False) -> GaitConfig:
        cfg = GaitConfigFactory.AVAILABLE_GAITS.get(gait_name.lower())
        if cfg is None:
            raise ValueError(f"{gait_name} not available.")
        if fix_swapped_heights and cfg.nom_height < cfg.step_height:
            cfg = replace(cfg, nom_height=cfg.step_height, step_height=cfg.nom_height)
        return cfg
