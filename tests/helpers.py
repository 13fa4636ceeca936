"""Shared test helpers: golden fixtures -> tableaux, random LP families."""
import json
import os
from fractions import Fraction

import numpy as np

import oracle_lib as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_dumps():
    with open(os.path.join(GOLD, "frontend_dumps.json")) as f:
        return json.load(f)


def load_dumps_dual():
    """The same models through the reference's LPStandardize(model, dual = 1) (Source/simplex.c:177-178)."""
    with open(os.path.join(GOLD, "frontend_dumps_dual.json")) as f:
        return json.load(f)


def load_kat():
    with open(os.path.join(GOLD, "rational_kat.json")) as f:
        return json.load(f)


def _frac(t):
    return Fraction(t[0], t[1])


def tableau_from_dump(d, method):
    """Reference dump (CreateSMatrix output) -> canonical tableau of Fractions.

    Rows lacking an exact unit column get one trailing artificial column each, in row order
    (the reference's AddBigM, Source/simplex.c:79-82, is an empty stub; its identity
    heuristic at Source/matrix.c:59-78 is re-checked exactly here).
    Returns (T rows of Fractions incl. objective rows, basis 1-based, n_art, n).
    """
    m, n0 = d["m"], d["n"]
    rows = [[_frac(x) for x in r] for r in d["rows"]]
    costs = [_frac(x) for x in d["costs"]]
    basis = [0] * m
    for i in range(m):
        name = d["basis"][i]
        if name is None:
            continue
        j = d["names"].index(name)
        col = [rows[r][j + 1] for r in range(m)]
        if col[i] == 1 and all(col[r] == 0 for r in range(m) if r != i) and costs[j] == 0:
            basis[i] = j + 1
    # a column may be claimed by one row only
    lack = [i for i in range(m) if basis[i] == 0]
    n_art = len(lack)
    if method == O.PLAIN:
        assert n_art == 0, "plain method needs an identity basis"
    n = n0 + n_art
    T = []
    for i in range(m):
        r = rows[i] + [Fraction(0)] * n_art
        if i in lack:
            a = n0 + 1 + lack.index(i)
            r[a] = Fraction(1)
            basis[i] = a
        T.append(r)
    T.append([Fraction(0)] + costs + [Fraction(0)] * n_art)
    if method != O.PLAIN:
        om = [Fraction(0)] * (n + 1)
        for a in range(n0 + 1, n + 1):
            om[a] = Fraction(-1)
        for i in lack:
            om = [x + y for x, y in zip(om, T[i])]
        T.append(om)
    return T, basis, n_art, n


def frac_pairs(T):
    return [[(x.numerator, x.denominator) for x in r] for r in T]


def frac_f64(T):
    return np.array([[x.numerator / x.denominator for x in r] for r in T], dtype=np.float64)


def random_mixed_lp(rng, m, ns, p_ge=0.25, p_eq=0.15, neg_b=0.2, dense=True):
    """Random LP  max c.x, A x rel b  with mixed relations (needs artificials)."""
    A = rng.uniform(-0.5, 1.5, size=(m, ns))
    if not dense:
        A *= rng.random((m, ns)) < 0.6
    x0 = rng.uniform(0.0, 2.0, size=ns)
    rel = np.full(m, -2, np.int32)
    u = rng.random(m)
    rel[u < p_ge] = 2
    rel[(u >= p_ge) & (u < p_ge + p_eq)] = 3
    ax = A @ x0
    b = np.where(rel == -2, ax + rng.uniform(0.1, 2.0, m), np.where(rel == 2, ax - rng.uniform(0.1, 2.0, m), ax))
    flip = rng.random(m) < neg_b
    A[flip] *= -1
    b[flip] *= -1
    rel[flip & (rel != 3)] *= -1
    c = rng.uniform(-1.0, 1.0, size=ns)
    return A, b, c, rel


def random_dual_lp(rng, m, ns, p_ge=0.5, infeasible=False):
    """max c.x with c <= 0, rows  a.x <= b  where the '>=' rows were flipped (so some b_i < 0): a slack
    basis that is dual feasible but not primal feasible -- the start the dual simplex wants
    (LPStandardize(model, dual = 1), Source/simplex.c:177-178).  Returns (T, basis, (A, b, c, rel)) with
    the (A, b, c, rel) form for the primal Big-M / two-phase solvers."""
    A = rng.uniform(0.0, 1.5, size=(m, ns))
    A[rng.random((m, ns)) < 0.2] = 0.0
    x0 = rng.uniform(0.0, 2.0, size=ns)
    ge = rng.random(m) < p_ge
    ax = A @ x0
    b = np.where(ge, ax - rng.uniform(0.1, 1.0, m), ax + rng.uniform(0.1, 2.0, m))
    b = np.where(ge, np.maximum(b, 0.05), b)                  # keep '>=' right-hand sides positive
    c = -rng.uniform(0.1, 1.0, size=ns)
    rel = np.where(ge, 2, -2).astype(np.int32)
    if infeasible:
        A[0] = np.abs(A[0]); rel[0] = -2; b[0] = -1.0         # a.x <= -1 with a >= 0, x >= 0
    n = ns + m
    T = np.zeros((m + 1, n + 1))
    sgn = np.where(rel == 2, -1.0, 1.0)                       # '>=' rows times -1 -> '<='
    T[:m, 0] = sgn * b
    T[:m, 1:ns + 1] = sgn[:, None] * A
    T[:m, ns + 1:] = np.eye(m)
    T[m, 1:ns + 1] = c
    basis = (ns + 1 + np.arange(m)).astype(np.int32)
    return T, basis, (A, b, c, rel)
