"""`QuadrupedDynamics`: the OCP model front-end (reference: mpc_controller/utils/dynamics.py).

The reference builds CasADi expressions on a symbolic Pinocchio model and lets acados generate C
from them.  Here the same model is generated offline from the URDF into straight-line CUDA
(`codegen/`, `csrc/nmpc_gen_model.h`), so this class only declares the named blocks the solver
front-end addresses, converts MuJoCo <-> Euler states, and asks the device for the numeric
kinematics (`update_pin`) that the reference asks Pinocchio for.
"""
from __future__ import annotations

import os
from types import SimpleNamespace
from typing import List, Optional, Tuple

import numpy as np

from .acados_helper import FloatingBaseDynamics, PointContact, ProblemFormulation
from .transform import (euler_derivative_to_local_angular, local_angular_to_euler_derivative, matrix_to_quat,
                        matrix_to_rpy, quat_to_matrix, rpy_to_matrix)

_PKG = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
DEFAULT_URDF = os.path.join(_PKG, "assets", "go2_like.urdf")


class _Momentum:
    """`pin_data.hg` stand-in: `.np` is the 6-vector [linear | angular about the CoM]."""

    def __init__(self):
        self.np = np.zeros(6)


class QuadrupedDynamics(FloatingBaseDynamics):

    def __init__(self, urdf_path: Optional[str], feet_frame_names: List[str], cnt_patch_restriction: bool = False,
                 mu_contact: float = 0.7):
        from ...codegen.robot import load_urdf
        self.urdf_path = urdf_path or DEFAULT_URDF
        self._model = load_urdf(self.urdf_path)
        if self._model.nq != 18 or len(feet_frame_names) != 4:
            raise NotImplementedError("this build is generated for a 12-joint quadruped with four feet")
        super().__init__(self._model.name)
        self.nu = self._model.nj
        self.feet_frame_names = list(feet_frame_names)
        self.feet_frame_id = list(range(len(feet_frame_names)))
        self.mu_contact = float(mu_contact)
        self.feet = [PointContact(dyn=self, frame=f, mu=mu_contact, patch_restriction=cnt_patch_restriction, index=i)
                     for i, f in enumerate(feet_frame_names)]
        self.add_contacts(self.feet)
        self.base_cost = self.add_expr(name="base_cost")      # [q[:3], q[3:6], v[:6]]   (dynamics.py:121-124)
        self.joint_cost = self.add_expr(name="joint_cost")    # [q[6:], v[6:]]           (:126-127)
        self.acc_cost = self.add_expr(name="acc_cost")        # a[6:]                    (:129-130)
        self.swing_cost = self.add_expr(name="sw_cost")       # peak_i * foot_i.z        (:132-134)
        # numeric state (Pinocchio data stand-in), refreshed by update_pin
        self._data = SimpleNamespace(hg=_Momentum(), com=np.zeros(3), feet=np.zeros((4, 3)),
                                     q=np.zeros(18), v=np.zeros(18))
        self._kin = None                                       # bound by the solver: qv[n,36] -> feet, com, hg
        self._rnea = None                                      # bound by the solver: inverse dynamics on device
        self.referenceConfigurations = {"home": np.concatenate((np.zeros(6), np.tile([0.0, 0.9, -1.8], 4)))}

    # -- Pinocchio-facing surface -----------------------------------------------------------
    @property
    def pin_model(self):
        """Stand-in exposing what `LocomotionMPC` reads: nq = nv = 18 (the model is loaded with a
        translation + ZYX-Euler base, which is why `update_pin` takes the Euler state), and
        referenceConfigurations."""
        return SimpleNamespace(nq=self._model.nq, nv=self._model.nq, name=self._model.name,
                               referenceConfigurations=self.referenceConfigurations)

    @property
    def pin_data(self):
        return self._data

    def bind_device(self, kinematics, inverse_dynamics=None) -> None:
        self._kin, self._rnea = kinematics, inverse_dynamics

    def update_pin(self, q: np.ndarray, v: np.ndarray):
        """framesForwardKinematics + computeCentroidalMomentum (dynamics.py:48-50), evaluated by
        the generated model code on the device (include/nmpc_b200.h: nmpc_kinematics)."""
        if self._kin is None:
            raise RuntimeError("QuadrupedDynamics is not bound to a solver handle yet")
        qv = np.concatenate((np.asarray(q, dtype=float)[-18:], np.asarray(v, dtype=float)[-18:]))[None]
        feet, com, hg = self._kin(qv)
        d = self._data
        d.feet, d.com, d.hg.np, d.q, d.v = feet[0], com[0], hg[0], qv[0, :18].copy(), qv[0, 18:].copy()

    def get_feet_position_w(self) -> np.ndarray:
        return self._data.feet.copy()

    # -- MuJoCo <-> Euler state (dynamics.py:53-99) ---------------------------------------------
    @staticmethod
    def convert_from_mujoco(q_mj: np.ndarray, v_mj: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """[p, quat(wxyz), joints], [v world, w local, joint vel] -> Euler state and Euler rates."""
        q_mj, v_mj = np.asarray(q_mj, dtype=float), np.asarray(v_mj, dtype=float)
        q = np.zeros(len(q_mj) - 1)
        q[:3] = q_mj[:3]
        w, x, y, z = q_mj[3:7]
        q[3:6] = matrix_to_rpy(quat_to_matrix(np.array([x, y, z, w])))[::-1]
        q[6:] = q_mj[7:]
        v = v_mj.copy()
        v[3:6] = local_angular_to_euler_derivative(q[3:6], v_mj[3:6])
        return q, v

    @staticmethod
    def convert_to_mujoco(q: np.ndarray, v: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        q, v = np.asarray(q, dtype=float), np.asarray(v, dtype=float)
        q_mj = np.zeros(len(q) + 1)
        x, y, z, w = matrix_to_quat(rpy_to_matrix(q[3:6][::-1]))
        q_mj[:3] = q[:3]
        q_mj[3:7] = (w, x, y, z)
        q_mj[7:] = q[6:]
        v_mj = v.copy()
        v_mj[3:6] = euler_derivative_to_local_angular(q[3:6], v[3:6])
        return q_mj, v_mj

    # -- OCP declaration (dynamics.py:109-116) ------------------------------------------------------
    def setup(self, problem: ProblemFormulation):
        for f in self.feet:
            f.setup(problem)
        super().setup(problem)
        problem.add_cost(self.base_cost, terminal=True)
        problem.add_cost(self.joint_cost, terminal=True)
        problem.add_cost(self.acc_cost)
        problem.add_cost(self.swing_cost, terminal=True)

    def get_hg(self):
        return self.h

    # -- feed-forward torques (dynamics.py:136-163) -----------------------------------------------
    def id_torques(self, q_plan: np.ndarray, v_plan: np.ndarray, a_plan: np.ndarray, f_plan: np.ndarray) -> np.ndarray:
        """RNEA(q, v, a)[joints] - sum_feet J_foot^T f, on the device (nmpc_inverse_dynamics)."""
        if self._rnea is None:
            raise RuntimeError("QuadrupedDynamics is not bound to a solver handle yet")
        qva = np.concatenate((np.asarray(q_plan, float)[-18:], np.asarray(v_plan, float)[-18:],
                              np.asarray(a_plan, float)[-18:]))[None]
        return self._rnea(qva, np.asarray(f_plan, dtype=float).reshape(1, 12))[0]
