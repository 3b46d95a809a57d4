"""URDF reader and symbolic whole-body kinematics / centroidal momentum of a floating-base robot.

Stands in for `loadModelImpl` / `toSymModel` / `FloatingBaseDynamics` / `PointContact.get_position`
of the missing `contact_tamp` submodule (call sites: `mpc_controller/utils/dynamics.py:19-38`) and
for the Pinocchio calls `framesForwardKinematics`, `computeCentroidalMomentum`, `centerOfMass`
(`dynamics.py:49-50`, `mpc.py:339`).

Coordinates follow the reference's Euler state (`dynamics.py:53-71`, `transform.py:72-86`):
    q = [p_WB (3), yaw, pitch, roll, joints (nj)]            R_WB = Rz(yaw) Ry(pitch) Rx(roll)
    v = [p_dot world (3), yaw_dot, pitch_dot, roll_dot, joint velocities]   (so q_dot == v)
"""
from __future__ import annotations

import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .expr import E, Graph, cos, sin


# ------------------------------------------------------------------------------------------
# URDF
# ------------------------------------------------------------------------------------------
@dataclass
class Link:
    name: str
    mass: float = 0.0
    com: np.ndarray = field(default_factory=lambda: np.zeros(3))
    inertia: np.ndarray = field(default_factory=lambda: np.zeros((3, 3)))  # about com, link frame
    inertia_rpy: np.ndarray = field(default_factory=lambda: np.zeros(3))


@dataclass
class Joint:
    name: str
    jtype: str
    parent: str
    child: str
    xyz: np.ndarray
    rpy: np.ndarray
    axis: np.ndarray
    lower: float = -np.inf
    upper: float = np.inf


@dataclass
class RobotModel:
    name: str
    links: Dict[str, Link]
    joints: List[Joint]
    root: str
    # actuated joints in Pinocchio order (depth-first in URDF order of appearance per parent)
    act_joints: List[Joint] = field(default_factory=list)

    @property
    def nj(self) -> int:
        return len(self.act_joints)

    @property
    def nq(self) -> int:
        return 6 + self.nj

    def total_mass(self) -> float:
        return float(sum(l.mass for l in self.links.values()))


def _vec(s: Optional[str], default=(0.0, 0.0, 0.0)) -> np.ndarray:
    if s is None:
        return np.array(default, dtype=float)
    return np.array([float(x) for x in s.split()], dtype=float)


def rpy_matrix(rpy: Sequence[float]) -> np.ndarray:
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def load_urdf(path: str) -> RobotModel:
    root = ET.parse(path).getroot()
    links: Dict[str, Link] = {}
    for le in root.findall("link"):
        link = Link(le.get("name"))
        ie = le.find("inertial")
        if ie is not None:
            o = ie.find("origin")
            if o is not None:
                link.com = _vec(o.get("xyz"))
                link.inertia_rpy = _vec(o.get("rpy"))
            link.mass = float(ie.find("mass").get("value"))
            t = ie.find("inertia")
            ixx, ixy, ixz = float(t.get("ixx")), float(t.get("ixy", 0)), float(t.get("ixz", 0))
            iyy, iyz, izz = float(t.get("iyy")), float(t.get("iyz", 0)), float(t.get("izz"))
            I = np.array([[ixx, ixy, ixz], [ixy, iyy, iyz], [ixz, iyz, izz]])
            Ri = rpy_matrix(link.inertia_rpy)
            link.inertia = Ri @ I @ Ri.T
        links[link.name] = link
    joints: List[Joint] = []
    children = set()
    for je in root.findall("joint"):
        o = je.find("origin")
        ax = je.find("axis")
        lim = je.find("limit")
        j = Joint(
            name=je.get("name"), jtype=je.get("type"),
            parent=je.find("parent").get("link"), child=je.find("child").get("link"),
            xyz=_vec(o.get("xyz") if o is not None else None),
            rpy=_vec(o.get("rpy") if o is not None else None),
            axis=_vec(ax.get("xyz") if ax is not None else None, (1.0, 0.0, 0.0)),
        )
        if lim is not None:
            j.lower = float(lim.get("lower", -np.inf))
            j.upper = float(lim.get("upper", np.inf))
        n = np.linalg.norm(j.axis)
        if n > 0:
            j.axis = j.axis / n
        joints.append(j)
        children.add(j.child)
    roots = [n for n in links if n not in children]
    if len(roots) != 1:
        raise ValueError(f"URDF must have exactly one root link, found {roots}")
    model = RobotModel(root.get("name", "robot"), links, joints, roots[0])

    # depth-first order == order in which Pinocchio's URDF parser adds joints
    def visit(link_name: str) -> None:
        for j in joints:
            if j.parent == link_name:
                if j.jtype in ("revolute", "continuous"):
                    model.act_joints.append(j)
                elif j.jtype != "fixed":
                    raise ValueError(f"unsupported joint type {j.jtype} ({j.name})")
                visit(j.child)

    visit(model.root)
    return model


# ------------------------------------------------------------------------------------------
# small symbolic linear algebra on E
# ------------------------------------------------------------------------------------------
def _matmul(A, B):
    n, k, m = len(A), len(B), len(B[0])
    return [[sum((A[i][l] * B[l][j] for l in range(1, k)), A[i][0] * B[0][j]) for j in range(m)]
            for i in range(n)]


def _matvec(A, x):
    return [sum((A[i][l] * x[l] for l in range(1, len(x))), A[i][0] * x[0]) for i in range(len(A))]


def _transpose(A):
    return [[A[j][i] for j in range(len(A))] for i in range(len(A[0]))]


def _cross(a, b):
    return [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]


def _const_mat(g: Graph, M: np.ndarray):
    return [[E(g, g.const(M[i, j])) for j in range(M.shape[1])] for i in range(M.shape[0])]


def _const_vec(g: Graph, v: np.ndarray):
    return [E(g, g.const(x)) for x in v]


def _axis_rot(g: Graph, axis: np.ndarray, c: E, s: E):
    """Rodrigues rotation matrix about a constant unit axis with symbolic cos/sin."""
    x, y, z = axis
    K = np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]])
    K2 = K @ K
    I = np.eye(3)
    one_c = 1.0 - c
    return [[E(g, g.const(I[i, j])) + s * float(K[i, j]) + one_c * float(K2[i, j])
             for j in range(3)] for i in range(3)]


# ------------------------------------------------------------------------------------------
# symbolic model
# ------------------------------------------------------------------------------------------
@dataclass
class SymbolicKinematics:
    g: Graph
    q: List[int]
    v: List[int]
    foot_p: List[List[int]]          # [foot][3]
    foot_J: List[List[List[int]]]    # [foot][3][nq]   d foot_p / d q
    foot_v: List[List[int]]          # [foot][3]       J(q) v
    foot_vq: List[List[List[int]]]   # [foot][3][nq]   d foot_v / d q
    com: List[int]                   # [3]
    com_J: List[List[int]]           # [3][nq]
    h: List[int]                     # [6] centroidal momentum (linear, angular about CoM), world-aligned
    h_q: List[List[int]]             # [6][nq]
    h_v: List[List[int]]             # [6][nq]
    mass: float
    a: List[int] = field(default_factory=list)     # acceleration symbols (Euler-rate derivatives + joints)
    tau: List[int] = field(default_factory=list)   # [nq] generalised forces of inverse dynamics M a + C v + g
    M: List[List[int]] = field(default_factory=list)   # [nq][nq] mass matrix d2T/dv2 (forward dynamics of the batched rollout)
    bias: List[int] = field(default_factory=list)  # [nq] C v + g = tau at a = 0


GRAVITY = 9.81


def build_symbolic(model: RobotModel, feet_frames: Sequence[str], with_dynamics: bool = False) -> SymbolicKinematics:
    g = Graph()
    nq = model.nq
    q = [g.sym(f"q{i}") for i in range(nq)]
    v = [g.sym(f"v{i}") for i in range(nq)]
    qe = [E(g, n) for n in q]
    ve = [E(g, n) for n in v]

    # base
    yaw, pitch, roll = qe[3], qe[4], qe[5]
    cy, sy, cp, sp, cr, sr = cos(yaw), sin(yaw), cos(pitch), sin(pitch), cos(roll), sin(roll)
    Rz = [[cy, -sy, 0.0 * cy], [sy, cy, 0.0 * cy], [0.0 * cy, 0.0 * cy, 1.0 + 0.0 * cy]]
    Ry = [[cp, 0.0 * cp, sp], [0.0 * cp, 1.0 + 0.0 * cp, 0.0 * cp], [-sp, 0.0 * cp, cp]]
    Rx = [[1.0 + 0.0 * cr, 0.0 * cr, 0.0 * cr], [0.0 * cr, cr, -sr], [0.0 * cr, sr, cr]]
    R_WB = _matmul(_matmul(Rz, Ry), Rx)
    p_WB = [qe[0], qe[1], qe[2]]
    # local angular velocity from Euler rates (transform.py:80-86, euler_derivative_to_local_angular)
    yd, pd, rd = ve[3], ve[4], ve[5]
    w_local = [-sp * yd + rd, cp * sr * yd + cr * pd, cr * cp * yd - sr * pd]
    w_base = _matvec(R_WB, w_local)

    link_R: Dict[str, list] = {model.root: R_WB}
    link_p: Dict[str, list] = {model.root: p_WB}
    link_w: Dict[str, list] = {model.root: w_base}
    act_index = {j.name: 6 + i for i, j in enumerate(model.act_joints)}

    def visit(parent: str) -> None:
        for j in model.joints:
            if j.parent != parent:
                continue
            Rp, pp, wp = link_R[parent], link_p[parent], link_w[parent]
            Ro = _matmul(Rp, _const_mat(g, rpy_matrix(j.rpy)))
            po = [a + b for a, b in zip(pp, _matvec(Rp, _const_vec(g, j.xyz)))]
            if j.jtype == "fixed":
                link_R[j.child], link_p[j.child], link_w[j.child] = Ro, po, wp
            else:
                iq = act_index[j.name]
                Rj = _axis_rot(g, j.axis, cos(qe[iq]), sin(qe[iq]))
                link_R[j.child] = _matmul(Ro, Rj)
                link_p[j.child] = po
                axis_w = _matvec(Ro, _const_vec(g, j.axis))
                link_w[j.child] = [a + b * ve[iq] for a, b in zip(wp, axis_w)]
            visit(j.child)

    visit(model.root)

    def nodes(vec):
        return [e.n for e in vec]

    # feet
    foot_p, foot_J, foot_v, foot_vq = [], [], [], []
    for fname in feet_frames:
        if fname not in link_p:
            raise ValueError(f"frame {fname} not in URDF")
        p = nodes(link_p[fname])
        J = g.jacobian(p, q)
        vel = g.jvp(p, q, v)
        vq = g.jacobian(vel, q)
        foot_p.append(p)
        foot_J.append(J)
        foot_v.append(vel)
        foot_vq.append(vq)

    # centre of mass and centroidal momentum
    mass = model.total_mass()
    com_e = [E(g, g.zero)] * 3
    link_c: Dict[str, list] = {}
    for name, link in model.links.items():
        if link.mass <= 0.0:
            continue
        c = [a + b for a, b in zip(link_p[name], _matvec(link_R[name], _const_vec(g, link.com)))]
        link_c[name] = c
        com_e = [a + link.mass * b for a, b in zip(com_e, c)]
    com_e = [a * (1.0 / mass) for a in com_e]
    com = nodes(com_e)
    com_J = g.jacobian(com, q)

    h_lin = [E(g, g.zero)] * 3
    h_ang = [E(g, g.zero)] * 3
    kinetic = E(g, g.zero)
    potential = E(g, g.zero)
    for name, link in model.links.items():
        if link.mass <= 0.0:
            continue
        c = link_c[name]
        cd = [E(g, n) for n in g.jvp(nodes(c), q, v)]
        h_lin = [a + link.mass * b for a, b in zip(h_lin, cd)]
        R = link_R[name]
        w = link_w[name]
        # R I R^T w
        Iw = _matvec(R, _matvec(_const_mat(g, link.inertia), _matvec(_transpose(R), w)))
        if with_dynamics:
            kinetic = kinetic + (0.5 * link.mass) * (cd[0] * cd[0] + cd[1] * cd[1] + cd[2] * cd[2]) \
                + 0.5 * (w[0] * Iw[0] + w[1] * Iw[1] + w[2] * Iw[2])
            potential = potential + (link.mass * GRAVITY) * c[2]
        r = [a - b for a, b in zip(c, com_e)]
        mv = [link.mass * x for x in cd]
        orb = _cross(r, mv)
        h_ang = [a + b + c_ for a, b, c_ in zip(h_ang, Iw, orb)]
    h = nodes(h_lin) + nodes(h_ang)
    h_q = g.jacobian(h, q)
    h_v = g.jacobian(h, v)
    sk = SymbolicKinematics(g, q, v, foot_p, foot_J, foot_v, foot_vq, com, com_J, h, h_q, h_v, mass)
    if with_dynamics:
        # Inverse dynamics from the Lagrangian L = T - V in the generalised coordinates (q, v = q_dot):
        #   tau = d/dt (dT/dv) - dT/dq + dV/dq,   d/dt (dT/dv) = (d2T/dv dq) v + (d2T/dv2) a
        # which equals Pinocchio's rnea(q, v, a) on the joint rows (dynamics.py:156); the base rows are
        # the generalised forces conjugate to the Euler coordinates.
        acc = [g.sym(f"a{i}") for i in range(nq)]
        mom = g.grad(kinetic.n, v)
        mom_dot = [g.add(x, y) for x, y in zip(g.jvp(mom, q, v), g.jvp(mom, v, acc))]
        dT_dq = g.grad(kinetic.n, q)
        dV_dq = g.grad(potential.n, q)
        sk.a = acc
        sk.tau = [g.add(g.sub(md, tq), vq) for md, tq, vq in zip(mom_dot, dT_dq, dV_dq)]
        sk.M = g.jacobian(mom, v)
        sk.bias = [g.add(g.sub(md, tq), vq) for md, tq, vq in zip(g.jvp(mom, q, v), dT_dq, dV_dq)]
    return sk
