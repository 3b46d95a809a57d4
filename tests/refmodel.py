"""Independent NUMERIC whole-body model used only by the tests to check the generated model code.

Plain numpy forward kinematics over the URDF tree (no symbolic graph, no generated code): link
poses by matrix recursion, link velocities by finite differences of the poses along q_dot = v.
Conventions are those of the reference's Euler state (`mpc_controller/utils/dynamics.py:53-71`,
`mpc_controller/utils/transform.py:72-86`): q = [p_WB, yaw, pitch, roll, joints],
R_WB = Rz(yaw) Ry(pitch) Rx(roll), v = q_dot.
"""
import os

import numpy as np

from iterative_learning_nmpc_b200.codegen.robot import load_urdf, rpy_matrix

URDF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                    "iterative_learning_nmpc_b200", "assets", "go2_like.urdf")
FEET = ["FL_foot", "FR_foot", "RL_foot", "RR_foot"]


def _axis_angle(axis, th):
    a = np.asarray(axis, float)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


class NumericModel:
    def __init__(self, urdf=URDF):
        self.m = load_urdf(urdf)
        self.idx = {j.name: 6 + i for i, j in enumerate(self.m.act_joints)}
        self.mass = self.m.total_mass()

    def poses(self, q):
        yaw, pitch, roll = q[3], q[4], q[5]
        R = rpy_matrix([roll, pitch, yaw])
        out = {self.m.root: (R, np.array(q[:3], float))}

        def visit(parent):
            for j in self.m.joints:
                if j.parent != parent:
                    continue
                Rp, pp = out[parent]
                Ro = Rp @ rpy_matrix(j.rpy)
                po = pp + Rp @ j.xyz
                if j.jtype != "fixed":
                    Ro = Ro @ _axis_angle(j.axis, q[self.idx[j.name]])
                out[j.child] = (Ro, po)
                visit(j.child)

        visit(self.m.root)
        return out

    def feet(self, q):
        P = self.poses(q)
        return np.stack([P[f][1] for f in FEET])

    def com(self, q):
        P = self.poses(q)
        c = np.zeros(3)
        for name, link in self.m.links.items():
            if link.mass > 0:
                R, p = P[name]
                c += link.mass * (p + R @ link.com)
        return c / self.mass

    def momentum(self, q, v, eps=1e-6):
        """Centroidal momentum [linear; angular about the CoM] in world axes, with link twists
        obtained by central differences of the link poses along q(t) = q + t v."""
        q, v = np.asarray(q, float), np.asarray(v, float)
        Pp, Pm, P0 = self.poses(q + eps * v), self.poses(q - eps * v), self.poses(q)
        com = self.com(q)
        hl, ha = np.zeros(3), np.zeros(3)
        for name, link in self.m.links.items():
            if link.mass <= 0:
                continue
            R, p = P0[name]
            c = p + R @ link.com
            cd = ((Pp[name][1] + Pp[name][0] @ link.com) - (Pm[name][1] + Pm[name][0] @ link.com)) / (2 * eps)
            Rd = (Pp[name][0] - Pm[name][0]) / (2 * eps)
            Wx = Rd @ R.T                      # [w]_x
            w = np.array([Wx[2, 1], Wx[0, 2], Wx[1, 0]])
            hl += link.mass * cd
            ha += R @ link.inertia @ R.T @ w + np.cross(c - com, link.mass * cd)
        return np.concatenate([hl, ha])
