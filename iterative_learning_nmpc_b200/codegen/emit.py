"""Emit the generated model header (straight-line C usable from gcc and nvcc).

Replaces the CasADi-generated C that acados links for the reference's dynamics / cost /
constraint functions.  One header, two consumers:
`csrc/` (CUDA device code, the product) and `oracle/` (plain C, the checker, which includes the
same file by path: there is exactly one copy in the tree).

usage: python -m iterative_learning_nmpc_b200.codegen.emit [--urdf PATH] [--feet FL_foot ...]
"""
from __future__ import annotations

import argparse
import hashlib
import os
from typing import List, Tuple

from .robot import build_symbolic, load_urdf

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
REPO = os.path.dirname(PKG)
DEFAULT_URDF = os.path.join(PKG, "assets", "go2_like.urdf")
DEFAULT_FEET = ["FL_foot", "FR_foot", "RL_foot", "RR_foot"]


def record_layout(nq: int, nfeet: int):
    """Offsets (in doubles) of the kinematics record written by nmpc_model_kin."""
    off = {}
    cur = 0
    for name, size in [
        ("FOOT_P", nfeet * 3),
        ("FOOT_J", nfeet * 3 * nq),
        ("FOOT_V", nfeet * 3),
        ("FOOT_VQ", nfeet * 3 * nq),
        ("COM", 3),
        ("COM_J", 3 * nq),
        ("H", 6),
        ("H_Q", 6 * nq),
        ("H_V", 6 * nq),
    ]:
        off[name] = cur
        cur += size
    off["SIZE"] = cur
    return off


def generate(urdf: str = DEFAULT_URDF, feet: List[str] = None) -> Tuple[str, dict]:
    feet = feet or DEFAULT_FEET
    model = load_urdf(urdf)
    sk = build_symbolic(model, feet, with_dynamics=True)
    g = sk.g
    nq = model.nq
    nf = len(feet)
    off = record_layout(nq, nf)

    sym_expr = {f"q{i}": f"q[{i}]" for i in range(nq)}
    sym_expr.update({f"v{i}": f"v[{i}]" for i in range(nq)})

    # The record is emitted as NPARTS independent straight-line functions (each with its own
    # common-subexpression pool): ptxas falls back to a 32-register, spill-everything allocation
    # on a single 15k-instruction basic block, and separate parts can run on separate warps.
    parts: List[List[Tuple[str, int]]] = []

    def slot(base: str, idx: int) -> str:
        return f"NMPC_REC({off[base] + idx})"

    for f in range(nf):                       # parts 0..nf-1: one per foot
        outs: List[Tuple[str, int]] = []
        for i in range(3):
            outs.append((slot("FOOT_P", f * 3 + i), sk.foot_p[f][i]))
            outs.append((slot("FOOT_V", f * 3 + i), sk.foot_v[f][i]))
            for j in range(nq):
                outs.append((slot("FOOT_J", (f * 3 + i) * nq + j), sk.foot_J[f][i][j]))
                outs.append((slot("FOOT_VQ", (f * 3 + i) * nq + j), sk.foot_vq[f][i][j]))
        parts.append(outs)
    outs = []                                  # part nf: centre of mass + linear momentum rows
    for i in range(3):
        outs.append((slot("COM", i), sk.com[i]))
        for j in range(nq):
            outs.append((slot("COM_J", i * nq + j), sk.com_J[i][j]))
    for i in range(3):
        outs.append((slot("H", i), sk.h[i]))
        for j in range(nq):
            outs.append((slot("H_Q", i * nq + j), sk.h_q[i][j]))
            outs.append((slot("H_V", i * nq + j), sk.h_v[i][j]))
    parts.append(outs)
    for i in range(3, 6):                      # parts nf+1..nf+3: one angular-momentum row each
        outs = [(slot("H", i), sk.h[i])]
        for j in range(nq):
            outs.append((slot("H_Q", i * nq + j), sk.h_q[i][j]))
            outs.append((slot("H_V", i * nq + j), sk.h_v[i][j]))
        parts.append(outs)
    bodies = [g.emit_body(o, sym_expr) for o in parts]
    stats_full = {"flops": sum(b[1]["flops"] for b in bodies), "trig": sum(b[1]["trig"] for b in bodies),
                  "nodes": sum(b[1]["nodes"] for b in bodies),
                  "part_flops": [b[1]["flops"] for b in bodies]}

    outs0: List[Tuple[str, int]] = []
    for f in range(nf):
        for i in range(3):
            outs0.append((f"feet[{f * 3 + i}]", sk.foot_p[f][i]))
    for i in range(3):
        outs0.append((f"com[{i}]", sk.com[i]))
    for i in range(6):
        outs0.append((f"h[{i}]", sk.h[i]))
    body0, stats0 = g.emit_body(outs0, sym_expr)

    # feed-forward joint torques: inverse dynamics minus the contact forces mapped through the foot
    # Jacobians (dynamics.py:136-163: pin.rnea(q, v, a)[-nu:] - sum_feet f_ee @ J_ee[:3, -nu:])
    sym_expr.update({f"a{i}": f"a[{i}]" for i in range(nq)})
    fsym = [g.sym(f"f{i}") for i in range(3 * nf)]
    sym_expr.update({f"f{i}": f"f[{i}]" for i in range(3 * nf)})
    outs_tau: List[Tuple[str, int]] = []
    for j in range(6, nq):
        t = sk.tau[j]
        for ft in range(nf):
            for i in range(3):
                t = g.sub(t, g.mul(sk.foot_J[ft][i][j], fsym[3 * ft + i]))
        outs_tau.append((f"tau[{j - 6}]", t))
    body_tau, stats_tau = g.emit_body(outs_tau, sym_expr)

    with open(urdf, "rb") as fh:
        digest = hashlib.sha256(fh.read()).hexdigest()[:16]

    lim_lo = ", ".join(repr(float(j.lower)) for j in model.act_joints)
    lim_hi = ", ".join(repr(float(j.upper)) for j in model.act_joints)
    hdr = []
    hdr.append("/* GENERATED by iterative_learning_nmpc_b200/codegen/emit.py -- do not edit.\n"
               f" * model: {model.name}  urdf sha256[:16]={digest}  feet={','.join(feet)}\n"
               f" * full record: {stats_full['flops']} flops, {stats_full['trig']} sincos;"
               f" values-only: {stats0['flops']} flops.\n"
               " * Stands in for the CasADi-generated C of the reference's acados model\n"
               " * (SURVEY.md 2.2 E2/E7; call sites mpc_controller/utils/dynamics.py:19-38,48-50).\n"
               " * Includer defines NMPC_HD (function qualifiers) and NMPC_REC(i) (lvalue of record slot i).\n"
               " */\n")
    hdr.append("#ifndef NMPC_GEN_MODEL_H\n#define NMPC_GEN_MODEL_H\n")
    hdr.append(f"#define NMPC_NQ {nq}\n#define NMPC_NJ {model.nj}\n#define NMPC_NFEET {nf}\n")
    hdr.append(f"#define NMPC_MASS {sk.mass!r}\n")
    for k, val in off.items():
        hdr.append(f"#define NMPC_REC_{k} {val}\n")
    hdr.append(f"#define NMPC_MODEL_FLOPS {stats_full['flops']}\n")
    hdr.append(f"#define NMPC_JOINT_LOWER {{{lim_lo}}}\n#define NMPC_JOINT_UPPER {{{lim_hi}}}\n")
    hdr.append("#ifndef NMPC_HD\n#define NMPC_HD static\n#endif\n")
    hdr.append(f"#define NMPC_MODEL_PARTS {len(bodies)}\n")
    hdr.append("\n/* full kinematics record (feet, CoM, centroidal momentum and their Jacobians) in\n"
               " * NMPC_MODEL_PARTS independent parts; part p writes a disjoint set of record slots */\n")
    hdr.append("#ifdef NMPC_REC\n")
    for i, (body, st) in enumerate(bodies):
        hdr.append(f"/* part {i}: {st['flops']} flops */\n")
        hdr.append(f"NMPC_HD void nmpc_model_part{i}(const double* __restrict__ q, const double* __restrict__ v,"
                   " double* __restrict__ rec)\n{\n")
        hdr.append(body)
        hdr.append("}\n")
    hdr.append("NMPC_HD void nmpc_model_part(int part, const double* __restrict__ q, const double* __restrict__ v,"
               " double* __restrict__ rec)\n{\n    switch (part) {\n")
    for i in range(len(bodies)):
        hdr.append(f"    case {i}: nmpc_model_part{i}(q, v, rec); break;\n")
    hdr.append("    default: break;\n    }\n}\n")
    hdr.append("NMPC_HD void nmpc_model_kin(const double* __restrict__ q, const double* __restrict__ v,"
               " double* __restrict__ rec)\n{\n")
    hdr.append(f"    for (int part = 0; part < {len(bodies)}; part++) nmpc_model_part(part, q, v, rec);\n")
    hdr.append("}\n#endif\n")
    hdr.append("\n/* values only: feet positions [nfeet*3], CoM [3], centroidal momentum [6] */\n")
    hdr.append("NMPC_HD void nmpc_model_kin0(const double* __restrict__ q, const double* __restrict__ v,"
               " double* __restrict__ feet, double* __restrict__ com, double* __restrict__ h)\n{\n")
    hdr.append(body0)
    hdr.append("}\n")
    hdr.append(f"\n/* joint torques tau[{model.nj}] = rnea(q, v, a)[joints] - sum_feet J_foot^T f: {stats_tau['flops']} flops */\n")
    hdr.append("#ifdef NMPC_WITH_RNEA\n")
    hdr.append("NMPC_HD void nmpc_model_rnea(const double* __restrict__ q, const double* __restrict__ v,"
               " const double* __restrict__ a, const double* __restrict__ f, double* __restrict__ tau)\n{\n")
    hdr.append(body_tau)
    hdr.append("}\n#endif\n#endif /* NMPC_GEN_MODEL_H */\n")
    meta = {"nq": nq, "nfeet": nf, "mass": sk.mass, "layout": off,
            "stats_full": stats_full, "stats0": stats0, "stats_tau": stats_tau, "model": model, "sk": sk}
    return "".join(hdr), meta


def generate_sim(urdf: str = DEFAULT_URDF, feet: List[str] = None) -> Tuple[str, dict]:
    """Straight-line code of the forward-dynamics terms for the batched rigid-body rollout (nmpc_rollout.cuh):
    mass matrix M(q) (lower triangle, packed by rows), bias C(q, v) v + g(q), and the feet (positions, velocities,
    position Jacobians) -- everything in the Euler coordinates of the solver, so that M a + bias = S' tau + sum J' f."""
    feet = feet or DEFAULT_FEET
    model = load_urdf(urdf)
    sk = build_symbolic(model, feet, with_dynamics=True)
    g, nq, nf = sk.g, model.nq, len(feet)
    sym_expr = {f"q{i}": f"q[{i}]" for i in range(nq)}
    sym_expr.update({f"v{i}": f"v[{i}]" for i in range(nq)})
    outs_m = [(f"M[{i * (i + 1) // 2 + j}]", sk.M[i][j]) for i in range(nq) for j in range(i + 1)]
    outs_b = [(f"bias[{i}]", sk.bias[i]) for i in range(nq)]
    outs_f: List[Tuple[str, int]] = []
    for f in range(nf):
        for i in range(3):
            outs_f.append((f"fp[{f * 3 + i}]", sk.foot_p[f][i]))
            outs_f.append((f"fv[{f * 3 + i}]", sk.foot_v[f][i]))
            for j in range(nq):
                outs_f.append((f"fJ[{(f * 3 + i) * nq + j}]", sk.foot_J[f][i][j]))
    body_m, st_m = g.emit_body(outs_m, sym_expr)
    body_b, st_b = g.emit_body(outs_b, sym_expr)
    body_f, st_f = g.emit_body(outs_f, sym_expr)
    hdr = ["/* GENERATED by iterative_learning_nmpc_b200/codegen/emit.py --sim -- do not edit.\n"
           f" * model: {model.name}; forward-dynamics terms of the batched rigid-body rollout (nmpc_rollout.cuh).\n */\n",
           "#ifndef NMPC_GEN_SIM_H\n#define NMPC_GEN_SIM_H\n#ifndef NMPC_HD\n#define NMPC_HD static\n#endif\n",
           f"/* mass matrix, lower triangle packed by rows M[i (i + 1) / 2 + j]: {st_m['flops']} flops */\n",
           "NMPC_HD void nmpc_sim_mass(const double* __restrict__ q, double* __restrict__ M)\n{\n", body_m, "}\n",
           f"/* bias C(q, v) v + g(q): {st_b['flops']} flops */\n",
           "NMPC_HD void nmpc_sim_bias(const double* __restrict__ q, const double* __restrict__ v, double* __restrict__ bias)\n{\n",
           body_b, "}\n",
           f"/* feet: positions fp[12], velocities fv[12], position Jacobians fJ[12][{nq}]: {st_f['flops']} flops */\n",
           "NMPC_HD void nmpc_sim_feet(const double* __restrict__ q, const double* __restrict__ v, double* __restrict__ fp,"
           " double* __restrict__ fv, double* __restrict__ fJ)\n{\n", body_f, "}\n#endif /* NMPC_GEN_SIM_H */\n"]
    return "".join(hdr), {"mass": st_m, "bias": st_b, "feet": st_f}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sim", action="store_true", help="emit csrc/nmpc_gen_sim.h (forward-dynamics terms) instead")
    ap.add_argument("--urdf", default=DEFAULT_URDF)
    ap.add_argument("--feet", nargs="*", default=DEFAULT_FEET)
    ap.add_argument("--out", nargs="*", default=[
        os.path.join(PKG, "csrc", "nmpc_gen_model.h"),
    ])
    args = ap.parse_args()
    if args.sim:
        text, meta = generate_sim(args.urdf, args.feet)
        path = os.path.join(PKG, "csrc", "nmpc_gen_sim.h")
        with open(path, "w") as f:
            f.write(text)
        print(f"wrote {path}: {len(text)} bytes, {meta}")
        return
    text, meta = generate(args.urdf, args.feet)
    for path in args.out:
        with open(path, "w") as f:
            f.write(text)
        print(f"wrote {path}: {len(text)} bytes, {meta['stats_full']}")


if __name__ == "__main__":
    main()
