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

    # -- Newton-Euler joint torques with finite-difference link accelerations (independent of the
    #    Lagrangian derivation the code generator uses) -------------------------------------------
    def _omega(self, q, v, name, h=1e-6):
        Rp, Rm, R0 = self.poses(q + h * v)[name][0], self.poses(q - h * v)[name][0], self.poses(q)[name][0]
        Wx = ((Rp - Rm) / (2 * h)) @ R0.T
        return np.array([Wx[2, 1], Wx[0, 2], Wx[1, 0]])

    def joint_torques(self, q, v, a, f_feet, e=1e-4, g=9.81):
        """tau[12] = inverse dynamics of the joints minus the contact forces f_feet [4,3] (world)."""
        q, v, a = (np.asarray(x, float) for x in (q, v, a))
        traj = lambda t: (q + t * v + 0.5 * t * t * a, v + t * a)
        P0 = self.poses(q)
        Pp, Pm = self.poses(traj(e)[0]), self.poses(traj(-e)[0])
        wrench = {}                       # link -> (force, moment about the link's CoM), world axes
        for name, link in self.m.links.items():
            if link.mass <= 0:
                continue
            R, p = P0[name]
            c = lambda P: P[name][1] + P[name][0] @ link.com
            cdd = (c(Pp) - 2 * c(P0) + c(Pm)) / (e * e)
            w0 = self._omega(q, v, name)
            wd = (self._omega(*traj(e), name) - self._omega(*traj(-e), name)) / (2 * e)
            Iw = R @ link.inertia @ R.T
            wrench[name] = (link.mass * (cdd + np.array([0, 0, g])), Iw @ wd + np.cross(w0, Iw @ w0), c(P0))
        children = {}
        for j in self.m.joints:
            children.setdefault(j.parent, []).append(j)

        def subtree(link):
            out = [link]
            for j in children.get(link, []):
                out += subtree(j.child)
            return out

        tau = np.zeros(12)
        for j in self.m.act_joints:
            Rp, pp = P0[j.parent]
            pj = pp + Rp @ j.xyz
            axis_w = Rp @ rpy_matrix(j.rpy) @ j.axis
            m = np.zeros(3)
            for name in subtree(j.child):
                if name in wrench:
                    F, Nn, c = wrench[name]
                    m += Nn + np.cross(c - pj, F)
                if name in FEET:
                    m -= np.cross(P0[name][1] - pj, np.asarray(f_feet, float)[FEET.index(name)])
            tau[self.idx[j.name] - 6] = axis_w @ m
        return tau
