"""Minimal symbolic expression graph with hash-consing, algebraic AD and C/CUDA emission.

This replaces what the reference gets from CasADi 3.6.7 (`environment.yml:25`; used through
`contact_tamp` at `mpc_controller/utils/dynamics.py:7`): symbolic expressions of the robot
kinematics, their derivatives, and generated straight-line C.  It is our own design: a flat
node table (op, a, b) with structural hashing (common sub-expression elimination is therefore
free), forward/reverse differentiation by graph traversal, and an emitter that writes one
`const double` per live node in topological order.
"""
from __future__ import annotations

import os

import math
from typing import Dict, Iterable, List, Sequence, Tuple

CONST, SYM, ADD, SUB, MUL, DIV, NEG, SIN, COS, SQRT = range(10)
_UNARY = (NEG, SIN, COS, SQRT)
_BINARY = (ADD, SUB, MUL, DIV)


class Graph:
    """Hash-consed expression DAG.  Nodes are integers; use `E` for operator sugar."""

    def __init__(self) -> None:
        self.op: List[int] = []
        self.a: List[int] = []
        self.b: List[int] = []
        self.val: List[float] = []          # constants
        self.name: List[str] = []           # symbols
        self._cache: Dict[Tuple, int] = {}
        self.zero = self.const(0.0)
        self.one = self.const(1.0)

    # ---- node construction -------------------------------------------------------------
    def _new(self, op: int, a: int = -1, b: int = -1, val: float = 0.0, name: str = "") -> int:
        key = (op, a, b, val, name)
        idx = self._cache.get(key)
        if idx is None:
            idx = len(self.op)
            self.op.append(op)
            self.a.append(a)
            self.b.append(b)
            self.val.append(val)
            self.name.append(name)
            self._cache[key] = idx
        return idx

    def const(self, v: float) -> int:
        v = float(v)
        if v == 0.0:
            v = 0.0  # merge -0.0
        return self._new(CONST, val=v)

    def sym(self, name: str) -> int:
        return self._new(SYM, name=name)

    def is_const(self, n: int) -> bool:
        return self.op[n] == CONST

    def cval(self, n: int) -> float:
        return self.val[n]

    def add(self, x: int, y: int) -> int:
        if self.is_const(x) and self.is_const(y):
            return self.const(self.val[x] + self.val[y])
        if x == self.zero:
            return y
        if y == self.zero:
            return x
        if self.op[y] == NEG:
            return self.sub(x, self.a[y])
        if self.op[x] == NEG:
            return self.sub(y, self.a[x])
        if x > y:
            x, y = y, x  # canonical order (commutative)
        return self._new(ADD, x, y)

    def sub(self, x: int, y: int) -> int:
        if self.is_const(x) and self.is_const(y):
            return self.const(self.val[x] - self.val[y])
        if y == self.zero:
            return x
        if x == self.zero:
            return self.neg(y)
        if x == y:
            return self.zero
        if self.op[y] == NEG:
            return self.add(x, self.a[y])
        return self._new(SUB, x, y)

    def mul(self, x: int, y: int) -> int:
        if self.is_const(x) and self.is_const(y):
            return self.const(self.val[x] * self.val[y])
        if x == self.zero or y == self.zero:
            return self.zero
        if x == self.one:
            return y
        if y == self.one:
            return x
        if self.is_const(x) and self.val[x] == -1.0:
            return self.neg(y)
        if self.is_const(y) and self.val[y] == -1.0:
            return self.neg(x)
        if self.op[x] == NEG and self.op[y] == NEG:
            return self.mul(self.a[x], self.a[y])
        if self.op[x] == NEG:
            return self.neg(self.mul(self.a[x], y))
        if self.op[y] == NEG:
            return self.neg(self.mul(x, self.a[y]))
        if x > y:
            x, y = y, x
        return self._new(MUL, x, y)

    def div(self, x: int, y: int) -> int:
        if self.is_const(y):
            return self.mul(x, self.const(1.0 / self.val[y]))
        if x == self.zero:
            return self.zero
        return self._new(DIV, x, y)

    def neg(self, x: int) -> int:
        if self.is_const(x):
            return self.const(-self.val[x])
        if self.op[x] == NEG:
            return self.a[x]
        if self.op[x] == SUB:
            return self.sub(self.b[x], self.a[x])
        return self._new(NEG, x)

    def sin(self, x: int) -> int:
        if self.is_const(x):
            return self.const(math.sin(self.val[x]))
        return self._new(SIN, x)

    def cos(self, x: int) -> int:
        if self.is_const(x):
            return self.const(math.cos(self.val[x]))
        return self._new(COS, x)

    def sqrt(self, x: int) -> int:
        if self.is_const(x):
            return self.const(math.sqrt(self.val[x]))
        return self._new(SQRT, x)

    # ---- traversal ---------------------------------------------------------------------
    def topo(self, outputs: Iterable[int]) -> List[int]:
        """Nodes reachable from outputs, children first.  Node ids are already topological
        (a child is always created before its parent), so sorting the reachable set suffices."""
        seen = set()
        stack = list(outputs)
        while stack:
            n = stack.pop()
            if n in seen:
                continue
            seen.add(n)
            o = self.op[n]
            if o in _BINARY:
                stack.append(self.a[n])
                stack.append(self.b[n])
            elif o in _UNARY:
                stack.append(self.a[n])
        return sorted(seen)

    def dfs_order(self, outputs: Iterable[int]) -> List[int]:
        """Depth-first post-order from the outputs, in output order: the operands of an output are computed
        right before it instead of in creation order (shorter live ranges in the emitted straight-line code)."""
        seen = set()
        order: List[int] = []
        for root in outputs:
            stack = [(root, False)]
            while stack:
                n, done = stack.pop()
                if done:
                    order.append(n)
                    continue
                if n in seen:
                    continue
                seen.add(n)
                stack.append((n, True))
                o = self.op[n]
                if o in _BINARY:
                    stack.append((self.b[n], False))
                    stack.append((self.a[n], False))
                elif o in _UNARY:
                    stack.append((self.a[n], False))
        return order

    # ---- differentiation ---------------------------------------------------------------
    def jvp(self, outputs: Sequence[int], wrt: Sequence[int], tangents: Sequence[int]) -> List[int]:
        """Forward mode: d(outputs)/d(wrt) . tangents, tangents being expressions."""
        order = self.topo(outputs)
        dot: Dict[int, int] = {w: t for w, t in zip(wrt, tangents)}
        z = self.zero
        for n in order:
            if n in dot:
                continue
            o = self.op[n]
            if o in (CONST, SYM):
                dot[n] = z
                continue
            a = self.a[n]
            da = dot[a]
            if o == ADD:
                dot[n] = self.add(da, dot[self.b[n]])
            elif o == SUB:
                dot[n] = self.sub(da, dot[self.b[n]])
            elif o == MUL:
                b = self.b[n]
                dot[n] = self.add(self.mul(da, b), self.mul(a, dot[b]))
            elif o == DIV:
                b = self.b[n]
                # d(a/b) = (da - n*db)/b
                dot[n] = self.div(self.sub(da, self.mul(n, dot[b])), b)
            elif o == NEG:
                dot[n] = self.neg(da)
            elif o == SIN:
                dot[n] = self.mul(self.cos(a), da)
            elif o == COS:
                dot[n] = self.neg(self.mul(self.sin(a), da))
            elif o == SQRT:
                dot[n] = self.div(da, self.mul(self.const(2.0), n))
        return [dot[o] for o in outputs]

    def grad(self, output: int, wrt: Sequence[int]) -> List[int]:
        """Reverse mode gradient of one scalar output w.r.t. the symbols in `wrt`."""
        order = self.topo([output])
        bar: Dict[int, int] = {output: self.one}
        z = self.zero
        for n in reversed(order):
            g = bar.get(n)
            if g is None or g == z:
                continue
            o = self.op[n]
            if o in (CONST, SYM):
                continue
            a = self.a[n]

            def acc(node: int, contrib: int) -> None:
                prev = bar.get(node)
                bar[node] = contrib if prev is None else self.add(prev, contrib)

            if o == ADD:
                acc(a, g)
                acc(self.b[n], g)
            elif o == SUB:
                acc(a, g)
                acc(self.b[n], self.neg(g))
            elif o == MUL:
                b = self.b[n]
                acc(a, self.mul(g, b))
                acc(b, self.mul(g, a))
            elif o == DIV:
                b = self.b[n]
                t = self.div(g, b)
                acc(a, t)
                acc(b, self.neg(self.mul(t, n)))
            elif o == NEG:
                acc(a, self.neg(g))
            elif o == SIN:
                acc(a, self.mul(g, self.cos(a)))
            elif o == COS:
                acc(a, self.neg(self.mul(g, self.sin(a))))
            elif o == SQRT:
                acc(a, self.div(g, self.mul(self.const(2.0), n)))
        return [bar.get(w, z) for w in wrt]

    def jacobian(self, outputs: Sequence[int], wrt: Sequence[int]) -> List[List[int]]:
        """Dense Jacobian (list of rows) by one reverse sweep per output."""
        return [self.grad(o, wrt) for o in outputs]

    # ---- numeric evaluation (used by the tests; never on the product path) -------------
    def evaluate(self, outputs: Sequence[int], env: Dict[str, float]) -> List[float]:
        order = self.topo(outputs)
        v: Dict[int, float] = {}
        for n in order:
            o = self.op[n]
            if o == CONST:
                v[n] = self.val[n]
            elif o == SYM:
                v[n] = env[self.name[n]]
            elif o == ADD:
                v[n] = v[self.a[n]] + v[self.b[n]]
            elif o == SUB:
                v[n] = v[self.a[n]] - v[self.b[n]]
            elif o == MUL:
                v[n] = v[self.a[n]] * v[self.b[n]]
            elif o == DIV:
                v[n] = v[self.a[n]] / v[self.b[n]]
            elif o == NEG:
                v[n] = -v[self.a[n]]
            elif o == SIN:
                v[n] = math.sin(v[self.a[n]])
            elif o == COS:
                v[n] = math.cos(v[self.a[n]])
            elif o == SQRT:
                v[n] = math.sqrt(v[self.a[n]])
        return [v[o] for o in outputs]

    # ---- emission ----------------------------------------------------------------------
    def emit_body(self, outputs: Sequence[Tuple[str, int]], sym_expr: Dict[str, str],
                  indent: str = "    ") -> Tuple[str, Dict[str, int]]:
        """Emit straight-line code computing every (lvalue, node) in `outputs`.

        `sym_expr` maps symbol names to C rvalues (e.g. "q[3]").  Returns (code, stats).
        sin/cos of the same argument are emitted as one `sincos` call.
        """
        # Depth-first order (operands right before their first use) instead of creation order: 37 % fewer spill
        # stores / 39 % fewer reloads in the sm_100a build of the model parts, inverse-dynamics frame 10.6 -> 2.1 KB
        # (heavier-operand-first variants were no better).  NMPC_EMIT_ORDER=topo gives the old order.
        if os.environ.get("NMPC_EMIT_ORDER", "dfs") == "dfs":
            nodes = self.dfs_order([n for _, n in outputs])
        else:
            nodes = self.topo([n for _, n in outputs])
        lines: List[str] = []
        ref: Dict[int, str] = {}
        # An output is stored as soon as its node has been computed, not at the end of the body: with all stores at
        # the end every output value (735 record slots per stage) stayed live across the whole function and ptxas
        # spilled 20-40 KB per thread.
        stores: Dict[int, List[str]] = {}
        for lv, n in outputs:
            stores.setdefault(n, []).append(lv)
        pending: List[Tuple[str, int]] = []

        def flush(n: int) -> None:
            for lv in stores.get(n, ()):
                lines.append(f"{indent}{lv} = {ref[n]};")
        n_flops = 0
        n_trig = 0
        # pair sin/cos
        trig_args: Dict[int, List[int]] = {}
        for n in nodes:
            if self.op[n] in (SIN, COS):
                trig_args.setdefault(self.a[n], []).append(n)
        emitted_trig = set()
        for n in nodes:
            o = self.op[n]
            if o == CONST:
                ref[n] = repr(self.val[n]) if self.val[n] >= 0 else f"({self.val[n]!r})"
                flush(n)
                continue
            if o == SYM:
                ref[n] = sym_expr[self.name[n]]
                flush(n)
                continue
            if o == NEG:
                ref[n] = f"(-{ref[self.a[n]]})"
                flush(n)
                continue
            t = f"t{n}"
            if o in (SIN, COS):
                arg = self.a[n]
                group = trig_args[arg]
                if len(group) == 2:
                    if arg not in emitted_trig:
                        s_node = group[0] if self.op[group[0]] == SIN else group[1]
                        c_node = group[0] if self.op[group[0]] == COS else group[1]
                        lines.append(f"{indent}double t{s_node}, t{c_node}; "
                                     f"sincos({ref[arg]}, &t{s_node}, &t{c_node});")
                        emitted_trig.add(arg)
                        n_trig += 1
                    ref[n] = t
                    flush(n)
                    continue
                fn = "sin" if o == SIN else "cos"
                lines.append(f"{indent}const double {t} = {fn}({ref[arg]});")
                n_trig += 1
                ref[n] = t
                flush(n)
                continue
            if o == SQRT:
                lines.append(f"{indent}const double {t} = sqrt({ref[self.a[n]]});")
                n_flops += 1
                ref[n] = t
                flush(n)
                continue
            sym = {ADD: "+", SUB: "-", MUL: "*", DIV: "/"}[o]
            lines.append(f"{indent}const double {t} = {ref[self.a[n]]} {sym} {ref[self.b[n]]};")
            n_flops += 1
            ref[n] = t
            flush(n)
        return "\n".join(lines) + "\n", {"flops": n_flops, "trig": n_trig, "nodes": len(nodes)}


class E:
    """Operator-overloading handle on a Graph node."""
    __slots__ = ("g", "n")

    def __init__(self, g: Graph, n: int) -> None:
        self.g = g
        self.n = n

    @staticmethod
    def _lift(g: Graph, x) -> int:
        return x.n if isinstance(x, E) else g.const(x)

    def __add__(self, o):
        return E(self.g, self.g.add(self.n, self._lift(self.g, o)))

    __radd__ = __add__

    def __sub__(self, o):
        return E(self.g, self.g.sub(self.n, self._lift(self.g, o)))

    def __rsub__(self, o):
        return E(self.g, self.g.sub(self._lift(self.g, o), self.n))

    def __mul__(self, o):
        return E(self.g, self.g.mul(self.n, self._lift(self.g, o)))

    __rmul__ = __mul__

    def __truediv__(self, o):
        return E(self.g, self.g.div(self.n, self._lift(self.g, o)))

    def __rtruediv__(self, o):
        return E(self.g, self.g.div(self._lift(self.g, o), self.n))

    def __neg__(self):
        return E(self.g, self.g.neg(self.n))


def sin(x: E) -> E:
    return E(x.g, x.g.sin(x.n))


def cos(x: E) -> E:
    return E(x.g, x.g.cos(x.n))


def sqrt(x: E) -> E:
    return E(x.g, x.g.sqrt(x.n))
