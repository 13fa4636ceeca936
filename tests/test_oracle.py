"""CPU tests of the oracle itself (no GPU).

What is pinned by the reference (tests/golden/*.json, produced by oracle/gen_golden.py
from the reference's own unmodified code):
  * rational arithmetic known answers  (Source/numOprts.c:37-129)
  * the initial tableau of each model  (Source/simplex.c:91-260, Source/matrix.c:19-91,
    SURVEY.md App. B)
What is NOT pinned by the reference -- it has no pivot loop (Source/simplex.c:57-63) -- and
is therefore checked against hand-derived sequences and against
invariants: pivot sequences, statuses, iteration counts.
"""
from fractions import Fraction

import numpy as np
import pytest

import helpers as H
import oracle_lib as O


# --------------------------------------------------------------------- App. A KATs
@pytest.mark.parametrize("kat", [k for k in H.load_kat() if not k["has_m"]],
                         ids=lambda k: f"{k['op']}:{k['a']}:{k['b']}")
def test_rational_kat_matches_reference(kat):
    L = O.lib()
    a = O.rat_parse(kat["a"])
    if kat["op"] in ("parse", "dec"):
        r = a
    else:
        b = O.rat_parse(kat["b"])
        r = getattr(L, "lpo_rat_" + kat["op"])(a, b)
    assert r.valid == kat["valid"]
    if kat["valid"]:
        assert (r.num, r.den) == (kat["num"], kat["den"])     # same representation, sign on den kept
        assert L.lpo_rat_to_double(r) == kat["dec"]


def test_rational_cmp_handles_negative_denominator():
    L = O.lib()
    assert L.lpo_rat_cmp(O.rat(1, -2), O.rat(-1, 3)) < 0
    assert L.lpo_rat_cmp(O.rat(1, -2), O.rat(-1, 2)) == 0
    assert L.lpo_rat_cmp(O.rat(3, 4), O.rat(2, 3)) > 0


# ------------------------------------------------------- App. B tableaux -> App. C runs
def _run_exact(name, method):
    d = H.load_dumps()[name]
    T, basis, n_art, n = H.tableau_from_dump(d, method)
    return d, T, basis, n_art, n, O.solve_rat(H.frac_pairs(T), basis, method=method, n_art=n_art)


def _run_f64(name, method):
    d = H.load_dumps()[name]
    T, basis, n_art, n = H.tableau_from_dump(d, method)
    return O.solve_f64(H.frac_f64(T), basis, method=method, n_art=n_art)


def test_reference_rejects_shipped_testdata():
    # Source/testdata.txt:2 says "ma:"; WriteIn only accepts max/min (Source/dataReader.c:455)
    assert H.load_dumps()["testdata_asshipped.txt"] == {"parsed": 0}


def test_testdata_front_end_dump_is_appendix_b1():
    d = H.load_dumps()["testdata_fixed.txt"]
    assert (d["m"], d["n"], d["valid"]) == (2, 4, 1)
    assert d["names"] == ["x1", "x2", "x3", "x4"] and d["basis"] == ["x3", "x4"]
    assert d["constant"][:2] == [19, 3]
    assert [r[0][:2] for r in d["rows"]] == [[51, 14], [1, 3]]


def test_testdata_plain_exact_sequence():
    d, T, basis, n_art, n, r = _run_exact("testdata_fixed.txt", O.PLAIN)
    assert r["status"] == O.OPTIMAL and r["iters"] == 3
    assert r["log"].tolist() == [[1, 1], [2, 0], [4, 1]]
    assert r["basis"].tolist() == [2, 4]
    assert Fraction(*r["T"][0][0]) == Fraction(17, 3) and Fraction(*r["T"][1][0]) == 6
    assert -Fraction(*r["T"][2][0]) + Fraction(19, 3) == 12


@pytest.mark.parametrize("method", [O.BIGM, O.TWOPHASE])
def test_testdata_forced_artificials_same_optimum(method):
    # config 2: the same LP pushed through Big-M / two-phase with an artificial on every row
    d = dict(H.load_dumps()["testdata_fixed.txt"])
    d["basis"] = [None, None]
    T, basis, n_art, n = H.tableau_from_dump(d, method)
    assert n_art == 2
    r = O.solve_rat(H.frac_pairs(T), basis, method=method, n_art=n_art)
    assert r["status"] == O.OPTIMAL
    assert -Fraction(*r["T"][2][0]) == Fraction(17, 3)
    f = O.solve_f64(H.frac_f64(T), basis, method=method, n_art=n_art)
    assert f["status"] == O.OPTIMAL and f["log"].tolist() == r["log"].tolist()
    assert abs(f["z"] - 17 / 3) <= 1e-9 * (17 / 3)


def test_min_surplus_bigm_and_twophase():
    d, T, basis, n_art, n, r = _run_exact("min_surplus.txt", O.BIGM)
    assert n_art == 1 and basis == [5, 4]
    assert r["status"] == O.OPTIMAL and r["log"].tolist() == [[1, 0]]
    assert r["basis"].tolist() == [1, 4] and r["z"] == -8.0          # max(-z) = -8  =>  min z = 8
    _, _, _, _, _, r2 = _run_exact("min_surplus.txt", O.TWOPHASE)
    assert r2["status"] == O.OPTIMAL and r2["log"].tolist() == [[1, 0]] and r2["phase1_iters"] == 1
    assert r2["z"] == -8.0


def test_infeasible_and_unbounded_statuses():
    for method in (O.BIGM, O.TWOPHASE):
        *_, r = _run_exact("infeasible.txt", method)
        assert r["status"] == O.INFEASIBLE
        assert _run_f64("infeasible.txt", method)["status"] == O.INFEASIBLE
    *_, r = _run_exact("infeasible.txt", O.BIGM)
    assert r["log"].tolist() == [[1, 1]] and r["zM"] == -2.0          # App. C: M-part of z = -2
    *_, u = _run_exact("unbounded.txt", O.PLAIN)
    assert u["status"] == O.UNBOUNDED and u["log"].tolist() == [[1, 0]] and u["iters"] == 1
    assert _run_f64("unbounded.txt", O.PLAIN)["status"] == O.UNBOUNDED


def test_identity_heuristic_defect_is_not_inherited():
    # Source/matrix.c:70 accepts column (1, 1/2) as a unit column; the exact re-check rejects it
    d = H.load_dumps()["eq_fraction_col.txt"]
    assert d["basis"] == ["x1", None]
    T, basis, n_art, n = H.tableau_from_dump(d, O.BIGM)
    assert n_art == 2 and basis == [3, 4]
    r = O.solve_rat(H.frac_pairs(T), basis, method=O.BIGM, n_art=n_art)
    assert r["status"] == O.OPTIMAL
    x = {int(b): Fraction(*r["T"][i][0]) for i, b in enumerate(r["basis"])}
    assert x[1] + x[2] == 4 and Fraction(1, 2) * x[1] + 3 * x[2] == 6


@pytest.mark.parametrize("name,method", [("three_by_three.txt", O.PLAIN), ("unr_inverted.txt", O.PLAIN),
                                         ("mixed_three.txt", O.BIGM), ("mixed_three.txt", O.TWOPHASE),
                                         ("min_equality.txt", O.BIGM), ("min_equality.txt", O.TWOPHASE)])
def test_exact_and_f64_agree_on_model_family(name, method):
    *_, r = _run_exact(name, method)
    f = _run_f64(name, method)
    # unr_inverted.txt is unbounded (x3 unrestricted, 2*x3 in a min objective); the rest are optimal
    expect = O.UNBOUNDED if name == "unr_inverted.txt" else O.OPTIMAL
    assert r["status"] == f["status"] == expect
    assert r["log"].tolist() == f["log"].tolist()
    assert r["basis"].tolist() == f["basis"].tolist()
    assert abs(r["z"] - f["z"]) <= 1e-9 * max(1.0, abs(r["z"]))


def test_mixed_three_known_optimum():
    # textbook: min 4x1+x2, 3x1+x2=3, 4x1+3x2>=6, x1+2x2<=4  ->  x=(2/5, 9/5), z = 17/5
    for method in (O.BIGM, O.TWOPHASE):
        *_, r = _run_exact("mixed_three.txt", method)
        assert r["z"] == pytest.approx(-17 / 5, rel=1e-15)


def test_rational_overflow_is_reported():
    big = 3037000500
    T = [[(1, 1), (big, 1), (1, 1), (0, 1)], [(1, 1), (big, big - 1), (0, 1), (1, 1)], [(0, 1), (1, 1), (0, 1), (0, 1)]]
    r = O.solve_rat(T, [2, 3], method=O.PLAIN)
    assert r["status"] in (O.OVERFLOW, O.OPTIMAL)


# ------------------------------------------------------------------ fp64 invariants
def _check_optimal(T0, res, n_art=0, tol=1e-7):
    """Primal feasibility + objective consistency of an OPTIMAL result against the initial tableau."""
    R = T0.shape[0] - len(res["basis"])
    m = T0.shape[0] - R
    x = res["x"]
    A, b = T0[:m, 1:], T0[:m, 0]
    assert np.all(x >= -tol)
    assert np.allclose(A @ x, b, atol=tol * max(1.0, np.abs(b).max()))
    c = T0[m, 1:]
    assert res["z"] == pytest.approx(float(c @ x) - T0[m, 0], rel=1e-7, abs=1e-7)
    if n_art:
        assert np.all(np.abs(x[-n_art:]) <= tol)


@pytest.mark.parametrize("m,ns", [(4, 4), (8, 12), (32, 32), (64, 64), (40, 7)])
def test_synth_feasible_solves_to_optimal(m, ns):
    T0, basis = O.synth_feasible(1234, 5, m, ns)
    assert np.all(T0[:m, 0] > 0) and np.all(T0[m, 1:ns + 1] > 0)
    res = O.solve_f64(T0, basis)
    assert res["status"] == O.OPTIMAL and res["iters"] > 0
    _check_optimal(T0, res)
    # threads do not change a bit
    res4 = O.solve_f64(T0, basis, threads=4)
    assert res4["log"].tolist() == res["log"].tolist() and np.array_equal(res4["T"], res["T"])


def test_synth_is_counter_based():
    T1, _ = O.synth_feasible(1234, 7, 6, 5)
    T2, _ = O.synth_feasible(1234, 7, 6, 5)
    T3, _ = O.synth_feasible(1234, 8, 6, 5)
    assert np.array_equal(T1, T2) and not np.array_equal(T1, T3)
    u = O.lib().lpo_uniform(1234, 7, 0, 0.1, 1.1)
    assert T1[0, 1] == u and 0.1 <= u < 1.1


@pytest.mark.parametrize("seed", range(12))
def test_bigm_and_twophase_agree_on_random_mixed_lps(seed):
    rng = np.random.default_rng(seed)
    m, ns = int(rng.integers(3, 12)), int(rng.integers(3, 12))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    T0, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
    rb = O.solve_f64(T0, basis, method=O.BIGM, n_art=n_art, max_iter=2000)
    rt = O.solve_f64(T0, basis, method=O.TWOPHASE, n_art=n_art, max_iter=2000)
    assert rb["status"] == rt["status"]
    assert rb["status"] in (O.OPTIMAL, O.UNBOUNDED)     # constructed feasible
    if rb["status"] == O.OPTIMAL:
        _check_optimal(T0, rb, n_art)
        _check_optimal(T0, rt, n_art)
        assert rb["z"] == pytest.approx(rt["z"], rel=1e-8, abs=1e-8)


def test_build_tableau_layout_and_priceout():
    A = np.array([[1.0, 1.0], [1.0, 3.0], [2.0, -1.0]])
    b = np.array([4.0, 9.0, -3.0])
    c = np.array([-2.0, -3.0])
    rel = np.array([2, -2, 3], np.int32)
    T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
    # columns: x1 x2 | surplus(row0) slack(row1) | art(row0) art(row2)
    assert T.shape == (5, 7) and n_art == 2 and basis.tolist() == [5, 4, 6]
    assert T[0].tolist() == [4, 1, 1, -1, 0, 1, 0]
    assert T[1].tolist() == [9, 1, 3, 0, 1, 0, 0]
    assert T[2].tolist() == [3, -2, 1, 0, 0, 0, 1]            # "=" row negated, relation kept
    assert T[3].tolist() == [0, -2, -3, 0, 0, 0, 0]
    assert T[4].tolist() == [7, -1, 2, -1, 0, 0, 0]           # sum of artificial rows, 0 on artificials


def test_iteration_cap_and_log_prefix():
    T0, basis = O.synth_feasible(1234, 0, 16, 16)
    full = O.solve_f64(T0, basis)
    assert full["iters"] >= 4
    part = O.solve_f64(T0, basis, max_iter=3)
    assert part["status"] == O.ITER_LIMIT and part["iters"] == 3
    assert part["log"].tolist() == full["log"][:3].tolist()
    exact = O.solve_f64(T0, basis, max_iter=full["iters"])
    assert exact["status"] == O.OPTIMAL


def test_single_pivot_matches_definition():
    rng = np.random.default_rng(0)
    T = rng.uniform(-1, 1, size=(6, 9))
    r, k = 2, 4
    out = O.pivot(T, r, k)
    prow = T[r] / T[r, k]
    for i in range(6):
        if i == r:
            assert np.array_equal(out[i], prow)
        else:
            # one rounding: exact product + sum in Fractions, rounded once by float()
            ref = np.array([float(Fraction(-T[i, k]) * Fraction(prow[j]) + Fraction(T[i, j])) for j in range(9)])
            assert np.array_equal(out[i], ref)
    assert out[r, k] == 1.0 and np.all(out[np.arange(6) != r, k] == 0.0)


def test_batched_matches_single():
    T, basis = O.synth_feasible_batch(1234, 0, 64, 8, 8)
    rb = O.solve_batched_f64(T, basis, threads=4, log_cap=64)
    for p in (0, 17, 63):
        rs = O.solve_f64(T[p], basis[p])
        assert rb["status"][p] == rs["status"] and rb["iters"][p] == rs["iters"]
        assert rb["z"][p] == rs["z"] and np.array_equal(rb["x"][p], rs["x"])
        assert rb["log"][p, :rs["iters"]].tolist() == rs["log"].tolist()


# ------------------------------------------------------------------ dual simplex
@pytest.mark.parametrize("seed", range(12))
def test_dual_simplex_reaches_the_primal_optimum(seed):
    """Dual simplex from a dual-feasible slack basis (b of mixed sign) vs Big-M on the same LP."""
    rng = np.random.default_rng(900 + seed)
    m, ns = int(rng.integers(3, 25)), int(rng.integers(3, 25))
    T, basis, (A, b, c, rel) = H.random_dual_lp(rng, m, ns)
    dual = O.solve_f64(T, basis, method=O.DUAL)
    Tb, bb, n_art = O.build_tableau(A, b, c, rel, R=2)
    prim = O.solve_f64(Tb, bb, method=O.BIGM, n_art=n_art)
    assert dual["status"] == prim["status"] == O.OPTIMAL
    assert dual["z"] == pytest.approx(prim["z"], rel=1e-9, abs=1e-9)
    assert np.all(dual["T"][:m, 0] >= -1e-9) and np.all(dual["T"][m, 1:] <= 1e-9)      # primal and dual feasible


def test_dual_simplex_statuses_and_rule():
    rng = np.random.default_rng(5)
    T, basis, _ = H.random_dual_lp(rng, 6, 5, infeasible=True)
    out = O.solve_f64(T, basis, method=O.DUAL)
    assert out["status"] == O.INFEASIBLE
    # textbook example: min 2x1 + 3x2 + 4x3, x1 + 2x2 + x3 >= 3, 2x1 - x2 + 3x3 >= 4  ->  z = 28/5
    T = np.array([[-3.0, -1, -2, -1, 1, 0], [-4.0, -2, 1, -3, 0, 1], [0.0, -2, -3, -4, 0, 0]])
    out = O.solve_f64(T, np.array([4, 5], np.int32), method=O.DUAL, log_cap=8)
    assert out["status"] == O.OPTIMAL and out["log"].tolist() == [[1, 1], [2, 0]]
    assert -out["z"] == pytest.approx(28 / 5) and out["x"][:3] == pytest.approx([11 / 5, 2 / 5, 0])
    # already optimal (b >= 0): zero pivots
    T2, b2 = O.synth_feasible(1, 0, 4, 4)
    T2[4, 1:] = -np.abs(T2[4, 1:])
    out = O.solve_f64(T2, b2, method=O.DUAL)
    assert out["status"] == O.OPTIMAL and out["iters"] == 0
    # iteration cap
    T, basis, _ = H.random_dual_lp(np.random.default_rng(3), 12, 10)
    full = O.solve_f64(T, basis, method=O.DUAL)
    if full["iters"] > 1:
        capped = O.solve_f64(T, basis, method=O.DUAL, max_iter=1)
        assert capped["status"] == O.ITER_LIMIT and capped["iters"] == 1


# ------------------------------------------------------------------ independent solver (scipy / HiGHS)
@pytest.mark.parametrize("seed", range(20))
def test_oracle_optimum_agrees_with_scipy_highs(seed):
    """The pivot PATH is this repo's rule (unpinned by the reference), but the END of the path is not a
    matter of convention: status and optimal objective must agree with an independent LP solver."""
    scipy_opt = pytest.importorskip("scipy.optimize")
    rng = np.random.default_rng(4000 + seed)
    m, ns = int(rng.integers(3, 30)), int(rng.integers(3, 30))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns, dense=bool(seed % 2))
    if seed % 5 == 4:                       # make some instances infeasible: x1 + ... <= -1 with A >= 0
        A[0] = np.abs(A[0]); b[0] = -1.0; rel[0] = -2
    le, ge, eq = rel == -2, rel == 2, rel == 3
    res = scipy_opt.linprog(-c, A_ub=np.vstack([A[le], -A[ge]]) if (le | ge).any() else None,
                            b_ub=np.concatenate([b[le], -b[ge]]) if (le | ge).any() else None,
                            A_eq=A[eq] if eq.any() else None, b_eq=b[eq] if eq.any() else None,
                            bounds=(0, None), method="highs")
    want = {0: O.OPTIMAL, 2: O.INFEASIBLE, 3: O.UNBOUNDED}.get(res.status)
    for method in (O.BIGM, O.TWOPHASE):
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
        got = O.solve_f64(T, basis, method=method, n_art=n_art)
        if want is None:
            continue
        assert got["status"] == want, (method, O.STATUS_NAMES[got["status"]], res.message)
        if want == O.OPTIMAL:
            assert got["z"] == pytest.approx(-res.fun, rel=1e-7, abs=1e-7)
            x = got["x"][:ns]
            assert np.all(x >= -1e-9) and c @ x == pytest.approx(-res.fun, rel=1e-7, abs=1e-7)


@pytest.mark.parametrize("seed", range(6))
def test_oracle_synthetic_family_agrees_with_scipy_highs(seed):
    scipy_opt = pytest.importorskip("scipy.optimize")
    m, ns = 24 + seed, 20 + 2 * seed
    T, basis = O.synth_feasible(1234, seed, m, ns)
    got = O.solve_f64(T, basis)
    res = scipy_opt.linprog(-T[m, 1:ns + 1], A_ub=T[:m, 1:ns + 1], b_ub=T[:m, 0], bounds=(0, None), method="highs")
    assert got["status"] == O.OPTIMAL and res.status == 0
    assert got["z"] == pytest.approx(-res.fun, rel=1e-9)


# ------------------------------------------------------------------ pricing option: Bland's rule
def _beale():
    """Beale's cycling example: max 3/4 x1 - 20 x2 + 1/2 x3 - 6 x4,  1/4 x1 - 8 x2 - x3 + 9 x4 <= 0,
    1/2 x1 - 12 x2 - 1/2 x3 + 3 x4 <= 0,  x3 <= 1; optimum 5/4.  Every entry is a dyadic fraction, so fp64 is exact."""
    A = np.array([[0.25, -8.0, -1.0, 9.0], [0.5, -12.0, -0.5, 3.0], [0.0, 0.0, 1.0, 0.0]])
    b = np.array([0.0, 0.0, 1.0])
    c = np.array([0.75, -20.0, 0.5, -6.0])
    T, basis, n_art = O.build_tableau(A, b, c, np.full(3, -2, np.int32), R=1)
    assert n_art == 0
    return T, basis


def test_bland_pricing_breaks_beales_cycle():
    T, basis = _beale()
    dantzig = O.solve_f64(T, basis, max_iter=600, log_cap=600)
    assert dantzig["status"] == O.ITER_LIMIT                      # largest reduced cost + lowest-row ties: the 6-pivot cycle
    assert dantzig["log"][:6].tolist() == dantzig["log"][6:12].tolist()
    bland = O.solve_f64(T, basis, max_iter=600, log_cap=600, pricing=O.BLAND)
    assert bland["status"] == O.OPTIMAL and bland["z"] == 1.25 and bland["iters"] < 20


@pytest.mark.parametrize("seed", range(6))
def test_bland_pricing_reaches_the_same_optimum(seed):
    rng = np.random.default_rng(900 + seed)
    A, b, c, rel = H.random_mixed_lp(rng, 12, 10)
    for method, R in ((O.BIGM, 2), (O.TWOPHASE, 2)):
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=R)
        d = O.solve_f64(T, basis, method=method, n_art=n_art, max_iter=5000)
        bl = O.solve_f64(T, basis, method=method, n_art=n_art, max_iter=5000, pricing=O.BLAND, log_cap=64)
        assert bl["status"] == d["status"]
        if d["status"] == O.OPTIMAL:
            assert abs(bl["z"] - d["z"]) <= 1e-9 * max(1.0, abs(d["z"]))
    # the first Bland pivot enters the LOWEST column with a positive reduced cost
    T, basis = O.synth_feasible(1234, seed, 10, 8)
    bl = O.solve_f64(T, basis, pricing=O.BLAND, log_cap=8)
    assert bl["log"][0][0] == 1 + int(np.argmax(T[10, 1:] > 1e-9))


# ------------------------------------------------------------------ pricing option: Devex reference weights
def _devex_replay(T0, basis0, log, method=O.PLAIN, n_art=0, eps=1e-9):
    """numpy restatement of the Devex rule: replays the oracle's pivot log and checks that every entering column is the arg-max
    of sigma_j^2 / w_j under the weight recurrence of devex_update() (lp_oracle.c)."""
    T = np.array(T0, dtype=np.float64)
    R = 1 if method == O.PLAIN else 2
    m, n = T.shape[0] - R, T.shape[1] - 1
    w = np.ones(n + 1)
    for k, r in log:
        sig = T[m]
        elig = sig[1:] > eps
        assert elig[k - 1]
        score = np.where(elig, (sig[1:] * sig[1:]) / w[1:], -1.0)
        assert int(np.argmax(score)) + 1 == k
        T = O.pivot(T, r, k)
        pr = T[r]
        upd = (pr * pr) * w[k]
        keep = w[k]
        w = np.maximum(w, upd)
        w[k] = 1.0
        del keep
    return T


@pytest.mark.parametrize("seed", range(4))
def test_devex_pricing_follows_its_recurrence_and_reaches_the_optimum(seed):
    T, basis = O.synth_feasible(77, seed, 24, 30)
    d = O.solve_f64(T, basis)
    dx = O.solve_f64(T, basis, pricing=O.DEVEX)
    assert d["status"] == dx["status"] == O.OPTIMAL
    assert abs(d["z"] - dx["z"]) <= 1e-9 * max(1.0, abs(d["z"]))
    Tend = _devex_replay(T, basis, dx["log"].tolist())
    assert np.array_equal(Tend, dx["T"])
    assert dx["log"].tolist() != d["log"].tolist()[:len(dx["log"])] or dx["iters"] == d["iters"]


@pytest.mark.parametrize("seed", range(6))
def test_devex_pricing_bigm_twophase_same_optimum(seed):
    rng = np.random.default_rng(1900 + seed)
    A, b, c, rel = H.random_mixed_lp(rng, 12, 10)
    for method, R in ((O.BIGM, 2), (O.TWOPHASE, 2)):
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=R)
        d = O.solve_f64(T, basis, method=method, n_art=n_art, max_iter=5000)
        dx = O.solve_f64(T, basis, method=method, n_art=n_art, max_iter=5000, pricing=O.DEVEX)
        assert dx["status"] == d["status"]
        if d["status"] == O.OPTIMAL:
            assert abs(dx["z"] - d["z"]) <= 1e-7 * max(1.0, abs(d["z"]))


def test_devex_needs_fewer_pivots_on_badly_scaled_columns():
    """The largest-coefficient rule depends on the scale of each column, Devex (approximately) does not: with the structural
    columns of the synthetic family scaled by powers of two between 2^-6 and 2^6 it needs well under half the pivots."""
    tot_d = tot_x = 0
    for lp in range(6):
        T, basis = O.synth_feasible(4321, lp, 96, 128)
        T[:, 1:129] *= 2.0 ** np.random.default_rng(lp).integers(-6, 7, size=128)
        d, x = O.solve_f64(T, basis), O.solve_f64(T, basis, pricing=O.DEVEX)
        assert d["status"] == x["status"] == O.OPTIMAL and abs(d["z"] - x["z"]) <= 1e-9 * abs(d["z"])
        tot_d += d["iters"]
        tot_x += x["iters"]
    assert 2 * tot_x < tot_d


def test_devex_rejected_for_dual():
    T, basis = O.synth_feasible(1, 0, 4, 4)
    res = O.Result()
    opts = O.Opts(1e-9, 10, O.DUAL, 0, 1, O.DEVEX)
    import ctypes as C
    rc = O.lib().lpo_solve_f64(4, 8, 1, O._dp(T.copy()), 9, O._ip(basis.copy()), C.byref(opts), None, 0, C.byref(res))
    assert rc == -1
