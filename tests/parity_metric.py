"""Error metric of the parity tests (BASELINE.json north_star: "<= 1e-6 relative on the primal/dual iterates of
each QP").

`rel_err(a, b)` is evaluated PER PROBLEM and PER BLOCK: the arrays handed in are one block of the iterate (x, u, pi,
lam, dz, t, ...) with the problem index leading; each problem's error max|a_b - b_b| is normalised by THAT problem's
own block norm max|b_b|, and the worst problem is returned.  A block that is identically zero in the checker (e.g. pi
before the first IPM iteration) is compared absolutely.  `worst_elementwise` reports, for the logs, the largest
element-wise relative deviation |a - b| / |b| over the entries whose magnitude is at least `rel_floor` times the
block norm of their problem (entries below that are rounding noise of sums of O(block norm) terms).
"""
import numpy as np

TINY = 1e-300


def _per_problem(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.ndim == 0:
        a, b = a.reshape(1, 1), b.reshape(1, 1)
    elif a.ndim == 1:                      # one scalar per problem
        a, b = a[:, None], b[:, None]
    n = a.shape[0]
    if a.size == 0:
        return np.zeros((n, 0)), np.zeros((n, 0))
    return a.reshape(n, -1), b.reshape(n, -1)


def rel_err_per_problem(a, b):
    a, b = _per_problem(a, b)
    if a.size == 0:
        return np.zeros(a.shape[0])
    d = np.abs(a - b).max(axis=1)
    d = np.where(np.isnan(a).any(axis=1) != np.isnan(b).any(axis=1), np.inf, d)
    nrm = np.abs(b).max(axis=1)
    return np.where(nrm > TINY, d / np.maximum(nrm, TINY), d)


def rel_err(a, b):
    """Worst problem's block-relative error (see the module docstring)."""
    e = rel_err_per_problem(a, b)
    return float(e.max()) if e.size else 0.0


def worst_elementwise(a, b, rel_floor=1e-6):
    a, b = _per_problem(a, b)
    if a.size == 0:
        return 0.0
    nrm = np.abs(b).max(axis=1, keepdims=True)
    sig = np.abs(b) >= rel_floor * np.maximum(nrm, TINY)
    if not sig.any():
        return 0.0
    return float((np.abs(a - b)[sig] / np.abs(b)[sig]).max())


def chain_ok(name, a, b, tol, frac=0.9):
    """Chained QPs (several SQP iterations / ticks with undamped full steps): the per-QP deviation of 1e-12..1e-9 is
    amplified from one QP to the next, without bound for the problems whose iteration is not contractive (the CPU
    oracle does the same to a 1e-9 change of its inputs, and its dense and structured variants drift apart the same
    way: tests/test_oracle_qp.py).  So a chain is held to: at least `frac` of the problems inside `tol` per block,
    every problem finite; the worst problem is printed."""
    e = rel_err_per_problem(a, b)
    ok = float((e < tol).mean()) if e.size else 1.0
    print(f"parity {name}: {100 * ok:.1f} % of {e.size} problems inside {tol:.0e}, median {np.median(e) if e.size else 0.0:.3e}, "
          f"worst problem {int(e.argmax()) if e.size else -1} block-rel {e.max() if e.size else 0.0:.3e}")
    return ok >= frac and bool(np.isfinite(e).all())


def report(name, a, b):
    """One log line per block: worst problem (block-relative) and worst significant element."""
    e = rel_err_per_problem(a, b)
    print(f"parity {name}: worst problem {int(e.argmax()) if e.size else -1} block-rel {e.max() if e.size else 0.0:.3e}, "
          f"worst element-wise rel {worst_elementwise(a, b):.3e}")
    return float(e.max()) if e.size else 0.0
