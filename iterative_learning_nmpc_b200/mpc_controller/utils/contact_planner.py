"""Gait / contact planners: periodic contact masks, swing ("peak") masks, Raibert footholds.

Same classes and call signatures as the reference's `utils/contact_planner.py`
(`GaitPlanner` :9-166, `ContactPlanner` :168-178, `RaiberContactPlanner` :180-322,
`CustomContactPlanner` :324-384, `ContactPlannerAcyclic` :386-428) including its quirks that
shape the OCP data: `nodes_per_cycle = round(period / dt)` with Python's round-half-even (trot at
dt = 0.04: 12 nodes, an effective 0.48 s cycle), `ceil` on the phase->node conversion and break
phases rounded to 2 decimals.  Windows are taken by modular indexing instead of tiling.
`cycle_tables()` additionally exposes the per-cycle tables so a batch of problems at different
phases can be gathered in one indexing operation (used by the batched front-end).
"""
from math import ceil
from typing import List, Tuple

import numpy as np

from ..config.config_abstract import GaitConfig
from .transform import rpy_to_matrix


class GaitPlanner:
    def __init__(self, feet_frame_names: List[str], dt_nodes: float, config_gait: GaitConfig):
        self.feet_frame_names = feet_frame_names
        self.n_foot = len(feet_frame_names)
        self.dt_nodes = dt_nodes
        self.config_gait = config_gait
        self.cnt_intervals = {foot: [] for foot in feet_frame_names}
        self.swing_intervals = {foot: [] for foot in feet_frame_names}
        self.nodes_per_cycle = round(config_gait.nominal_period / dt_nodes)
        shape = (self.n_foot, self.nodes_per_cycle)
        self.gait_sequence = np.zeros(shape, dtype=np.int8)
        self.peak_swing = np.zeros(shape, dtype=np.int8)
        self.switch_cnt = np.zeros(shape, dtype=np.int8)
        self._init_gait_cycle()
        self._init_peak_cycle()

    # -- phase queries ------------------------------------------------------------------
    def _is_in_cnt_phase(self, foot: str, phase: float) -> bool:
        return any(lo <= phase < hi for lo, hi in self.cnt_intervals.get(foot, []))

    def _is_in_cnt(self, foot: str, i_node: int) -> bool:
        phase = round((i_node % self.nodes_per_cycle) / self.nodes_per_cycle, 3)
        return self._is_in_cnt_phase(foot, phase)

    # -- tables -------------------------------------------------------------------------
    def _init_gait_cycle(self):
        npc = self.nodes_per_cycle
        make = self.config_gait.phase_offset
        brk = ((self.config_gait.phase_offset + self.config_gait.stance_ratio) % 1.).round(2)
        self.switch_phase = np.unique(np.concatenate((brk, make)))
        for i, (foot, mk, bk) in enumerate(zip(self.feet_frame_names, make, brk)):
            first, last = ceil(mk * npc), ceil(bk * npc)
            if mk < bk:      # one stance interval inside the cycle
                self.gait_sequence[i, first:last] = 1
                self.cnt_intervals[foot].append((mk, bk))
                for lo, hi in ((bk, 1.), (0., mk)):
                    if lo != hi:
                        self.swing_intervals[foot].append((lo, hi))
            else:            # stance wraps around the end of the cycle
                self.gait_sequence[i, first:] = 1
                self.gait_sequence[i, :last] = 1
                for lo, hi in ((mk, 1.), (0., bk)):
                    if lo != hi:
                        self.cnt_intervals[foot].append((lo, hi))
                self.swing_intervals[foot].append((bk, mk))
            self.switch_cnt[i, first] = 1
            self.switch_cnt[i, last] = -1

    def _init_peak_cycle(self):
        self.peak_swing[:, :] = 1 - self.gait_sequence

    def cycle_tables(self) -> Tuple[np.ndarray, np.ndarray]:
        """(gait_sequence, peak_swing), each [n_foot, nodes_per_cycle]."""
        return self.gait_sequence, self.peak_swing

    def _window(self, table: np.ndarray, i_node: int, n_nodes: int) -> np.ndarray:
        idx = (i_node % self.nodes_per_cycle + np.arange(n_nodes)) % self.nodes_per_cycle
        return table[:, idx]

    def get_contacts(self, i_node: int, n_nodes) -> np.ndarray:
        """[n_foot, n_nodes] contact mask (1 stance, 0 swing) starting at node i_node."""
        return self._window(self.gait_sequence, i_node, n_nodes)

    def get_peaks(self, i_node: int, n_nodes) -> np.ndarray:
        """[n_foot, n_nodes] swing mask starting at node i_node."""
        return self._window(self.peak_swing, i_node, n_nodes)

    def _window_batch(self, table: np.ndarray, i_node, n_nodes: int) -> np.ndarray:
        i_node = np.asarray(i_node, dtype=np.int64).reshape(-1)
        idx = (i_node[:, None] % self.nodes_per_cycle + np.arange(n_nodes)[None, :]) % self.nodes_per_cycle
        return np.ascontiguousarray(np.moveaxis(table[:, idx], 0, 1))           # [B, n_foot, n_nodes]

    def get_contacts_batch(self, i_node, n_nodes: int) -> np.ndarray:
        """`get_contacts` for one optimisation node per problem: [B, n_foot, n_nodes]."""
        return self._window_batch(self.gait_sequence, i_node, n_nodes)

    def get_peaks_batch(self, i_node, n_nodes: int) -> np.ndarray:
        return self._window_batch(self.peak_swing, i_node, n_nodes)

    def get_make_break_contacts(self, i_node: int, n_nodes) -> Tuple[np.ndarray, np.ndarray]:
        mk = (self.switch_cnt == 1).astype(self.switch_cnt.dtype)
        bk = (self.switch_cnt == -1).astype(self.switch_cnt.dtype)
        return self._window(mk, i_node, n_nodes), self._window(bk, i_node, n_nodes)


class ContactPlanner(GaitPlanner):
    def get_locations(self, i_node: int, n_nodes: int) -> np.ndarray:
        return None


class RaiberContactPlanner(ContactPlanner):
    GRAVITY = 9.81
    V_TRACKING = 0.05

    def __init__(self, feet_frame_names, dt_nodes, config_gait, offset_hip_b: np.ndarray,
                 x_offset: float = 0., y_offset: float = 0., foot_size: float = 0.,
                 height_offset: float = 0., cache_cnt: bool = True):
        super().__init__(feet_frame_names, dt_nodes, config_gait)
        self.foot_size = foot_size
        self.cache_cnt = cache_cnt
        self.offset_hip_b = offset_hip_b
        if self.n_foot == 4:
            self.offset_hip_b[:, 0] += np.array([1, 1, -1, -1]) * x_offset
            self.offset_hip_b[:, 1] += np.array([1, -1, 1, -1]) * y_offset
        self.default_cnt_w = np.zeros(3)
        self.height_offset = height_offset
        self.planed_cnt = {i: {} for i in range(self.n_foot)}
        self.pos = self.v_w = self.euler_rpy = self.com_xyz = self.v_des = self.w_yaw = None

    def set_state(self, pos, v_w, euler_rpy, com_xyz, v_des=np.zeros(3), w_yaw: float = 0.):
        self.pos, self.v_w, self.euler_rpy = pos, v_w, euler_rpy
        self.com_xyz, self.v_des, self.w_yaw = np.array(com_xyz), np.array(v_des), w_yaw

    def remove_cnt_before(self, i_node: int):
        self.planed_cnt = {f: {n: c for n, c in d.items() if n >= i_node}
                           for f, d in self.planed_cnt.items()}

    def get_locations(self, i_node: int, n_nodes: int) -> np.ndarray:
        """Raibert footholds [n_foot, n_nodes, 3]; zero where no touch-down is planned."""
        loc = np.zeros((self.n_foot, n_nodes, 3))
        mk, _ = self.get_make_break_contacts(i_node, n_nodes)
        com_xy, com_z = self.com_xyz[:2], self.com_xyz[-1] - self.height_offset
        vtrack = self.v_des[:2]
        R_yaw = rpy_to_matrix(np.array([0., 0., self.euler_rpy[2]]))
        ang = np.cross(0.5 * np.sqrt(com_z / self.GRAVITY) * np.append(vtrack, 0.), [0., 0., self.w_yaw])
        for i_foot, i_mk in np.argwhere(mk == 1):
            node = i_node + i_mk
            if self.cache_cnt and node in self.planed_cnt[i_foot]:
                loc[i_foot, i_mk:] = self.planed_cnt[i_foot][node]
                continue
            t_cnt = round(i_mk * self.dt_nodes, 3)
            if t_cnt < 0:
                continue
            ratio = self.config_gait.stance_ratio[i_foot]
            t_stance = self.config_gait.nominal_period * ratio
            hip = com_xy + (R_yaw @ self.offset_hip_b[i_foot])[:2] + vtrack * t_cnt * (1 + ratio)
            p = np.zeros(3)
            p[:2] = hip + 0.1 * (vtrack - self.v_w[:2]) + 0.5 * vtrack * t_stance + ang[:2]
            p[2] = self.foot_size
            loc[i_foot, i_mk:] = p
            if self.cache_cnt:
                self.planed_cnt[i_foot][node] = p
        return loc


class CustomContactPlanner(ContactPlanner):
    def __init__(self, feet_frame_names, dt_nodes, config_gait):
        super().__init__(feet_frame_names, dt_nodes, config_gait)
        self._contact_locations = None
        self.contact_locations_full = None
        self._contact_sequence = None
        self.contact_sequence_full = None
        self.peak_swing_full = None
        self.n_full = 0
        self.n_repeat = 3

    def set_contact_locations(self, contact_locations: np.ndarray) -> None:
        """contact_locations [n_cycles, n_foot, 3]: target of each gait cycle."""
        if contact_locations.ndim != 3 or contact_locations.shape[1:] != (self.n_foot, 3):
            raise ValueError(f"contact_locations: incorrect shape ({contact_locations.shape}).")
        self._contact_locations = contact_locations
        ext = np.concatenate((contact_locations,
                              np.repeat(contact_locations[-1:], self.n_repeat, axis=0)), axis=0)
        self.contact_locations_full = np.repeat(ext, self.nodes_per_cycle, axis=0).transpose(1, 0, 2)
        self.n_full = self.contact_locations_full.shape[1]

    def get_locations(self, i_node, n_nodes):
        end = i_node + n_nodes
        if end < self.n_full:
            return self.contact_locations_full[:, i_node:end, :].copy()
        return self.contact_locations_full[:, -n_nodes:, :].copy()

    def set_periodic_sequence(self, cnt_sequence: np.ndarray) -> None:
        if cnt_sequence.shape != self.gait_sequence.shape:
            raise ValueError(f"Invalid cnt_sequence shape, should be of shape {self.gait_sequence.shape}.")
        self.gait_sequence = cnt_sequence.copy()
        self.contact_sequence_full = None
        self._init_peak_cycle()

    def get_contacts(self, i_node, n_nodes) -> np.ndarray:
        if self.contact_sequence_full is not None and n_nodes + i_node <= len(self.contact_locations_full):
            return self.contact_sequence_full[:, i_node:i_node + n_nodes]
        return super().get_contacts(i_node, n_nodes)


class ContactPlannerAcyclic:
    """Explicit (non-periodic) contact sequence; the last column is held."""

    def __init__(self):
        self.n_nodes_seq = 0
        self.cnt_sequence = None
        self.center_sequence = None
        self.rot_patch_sequence = None
        self.patch_size_sequence = None

    def set_sequence(self, cnt_sequence: np.ndarray) -> None:
        self.cnt_sequence = cnt_sequence
        self.n_nodes_seq = cnt_sequence.shape[-1]

    def set_center_rot_size(self, cnt_center, cnt_rot, cnt_size) -> None:
        self.center_sequence, self.rot_patch_sequence, self.patch_size_sequence = cnt_center, cnt_rot, cnt_size

    def _held_index(self, i_node: int, n_nodes: int) -> np.ndarray:
        return np.minimum(np.arange(i_node, i_node + n_nodes), self.n_nodes_seq - 1)

    def get_sequence(self, i_node: int, n_nodes: int) -> np.ndarray:
        if self.cnt_sequence is None:
            raise ValueError("Set contact sequence first")
        return self.cnt_sequence[..., self._held_index(i_node, n_nodes)]

    def get_peak(self, i_node: int, n_nodes: int) -> np.ndarray:
        return 1 - self.get_sequence(i_node, n_nodes)

    def get_center_rot_size_patch(self, i_node: int, n_nodes: int):
        if self.center_sequence is None or self.rot_patch_sequence is None or self.patch_size_sequence is None:
            raise ValueError("Set patch data first")
        idx = self._held_index(i_node, n_nodes)
        return (self.center_sequence[:, idx], self.rot_patch_sequence[:, idx], self.patch_size_sequence[:, idx])
