"""GPU parity tests: the CUDA path, called through the C ABI (liblpgpu.so), against the CPU oracle
on the same inputs.  Bar: bit-exact pivot sequence, basis, iteration count, status AND tableau
(the method has no floating-point sums: per-element div / fma and exact comparisons only);
objective and x therefore agree to 0 ulp, well inside the 1e-9 relative tolerance north_star states.
"""
import numpy as np
import pytest

import helpers as H
import oracle_lib as O

import linearprogramming_b200 as lp
from linearprogramming_b200 import capi

pytestmark = pytest.mark.gpu

REL_TOL = 1e-9      # north_star's stated tolerance for objective / solution values (fp64)


@pytest.fixture(scope="module")
def ctx():
    c = lp.Context(0)
    yield c
    c.close()


def _same_bits(a, b):
    return np.array_equal(a, b) and np.array_equal(np.signbit(a), np.signbit(b))


def _assert_dense_equal(got, ref, tableau=True):
    assert got["status"] == ref["status"], (capi.STATUS_NAMES[got["status"]], O.STATUS_NAMES[ref["status"]])
    assert got["iters"] == ref["iters"]
    assert got["log"].tolist() == ref["log"].tolist()
    assert np.array_equal(got["basis"], ref["basis"])
    assert abs(got["z"] - ref["z"]) <= REL_TOL * max(1.0, abs(ref["z"]))
    assert np.allclose(got["x"], ref["x"], rtol=REL_TOL, atol=0.0)
    if tableau:
        assert np.array_equal(got["T"], ref["T"])        # bit-exact (+0 == -0 tolerated here)
        assert got["z"] == ref["z"] and np.array_equal(got["x"], ref["x"])


# ------------------------------------------------------------------ config 1: testdata.txt family
def test_cfg1_testdata_pivot_sequence(ctx):
    d = H.load_dumps()["testdata_fixed.txt"]
    T, basis, n_art, n = H.tableau_from_dump(d, O.PLAIN)
    T0 = H.frac_f64(T)
    exact = O.solve_rat(H.frac_pairs(T), basis)
    ref = O.solve_f64(T0, basis, log_cap=16)
    got = lp.solve_dense(ctx, T0, basis, log_cap=16)
    _assert_dense_equal(got, ref)
    assert got["log"].tolist() == exact["log"].tolist() == [[1, 1], [2, 0], [4, 1]]
    assert got["basis"].tolist() == [2, 4] and got["status"] == lp.OPTIMAL
    assert abs(got["z"] + 19 / 3 - 12) <= REL_TOL * 12          # + stripped constant 19/3 (Source/matrix.c:24)


# ------------------------------------------------------------------ config 2: Big-M / two-phase / statuses
@pytest.mark.parametrize("name", ["testdata_fixed.txt", "min_surplus.txt", "min_equality.txt", "mixed_three.txt",
                                  "eq_fraction_col.txt", "infeasible.txt", "three_by_three.txt"])
@pytest.mark.parametrize("method", [O.BIGM, O.TWOPHASE])
def test_cfg2_model_family_bigm_twophase(ctx, name, method):
    d = dict(H.load_dumps()[name])
    if name in ("testdata_fixed.txt", "three_by_three.txt"):
        d["basis"] = [None] * d["m"]                # force an artificial on every row
    T, basis, n_art, n = H.tableau_from_dump(d, method)
    T0 = H.frac_f64(T)
    exact = O.solve_rat(H.frac_pairs(T), basis, method=method, n_art=n_art)
    ref = O.solve_f64(T0, basis, method=method, n_art=n_art, log_cap=64)
    got = lp.solve_dense(ctx, T0, basis, method=method, n_art=n_art, log_cap=64)
    _assert_dense_equal(got, ref)
    assert got["status"] == exact["status"] and got["log"].tolist() == exact["log"].tolist()
    assert got["phase1_iters"] == ref["phase1_iters"]


def test_cfg2_unbounded(ctx):
    for name in ("unbounded.txt", "unr_inverted.txt"):
        d = H.load_dumps()[name]
        T, basis, n_art, n = H.tableau_from_dump(d, O.PLAIN)
        ref = O.solve_f64(H.frac_f64(T), basis, log_cap=16)
        got = lp.solve_dense(ctx, H.frac_f64(T), basis, log_cap=16)
        _assert_dense_equal(got, ref)
        assert got["status"] == lp.UNBOUNDED


@pytest.mark.parametrize("seed", range(10))
@pytest.mark.parametrize("method", [O.BIGM, O.TWOPHASE])
def test_random_mixed_lps_dense(ctx, seed, method):
    rng = np.random.default_rng(seed)
    m, ns = int(rng.integers(3, 40)), int(rng.integers(3, 40))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    if seed % 3 == 2:        # make some infeasible: contradict the first row
        A = np.vstack([A, A[0]]); b = np.append(b, b[0] - 5.0 if rel[0] != -2 else b[0] + 5.0)
        rel = np.append(rel, -2 if rel[0] != -2 else 2).astype(np.int32)
    T0, basis, n_art = lp.build_tableau(A, b, c, rel, R=2)
    ref = O.solve_f64(T0, basis, method=method, n_art=n_art, max_iter=5000, log_cap=5000)
    got = lp.solve_dense(ctx, T0, basis, method=method, n_art=n_art, max_iter=5000, log_cap=5000)
    _assert_dense_equal(got, ref)


# ------------------------------------------------------------------ the elimination pass alone
@pytest.mark.parametrize("shape", [(5, 7), (33, 65), (100, 511), (257, 1030), (64, 2049)])
@pytest.mark.parametrize("variant", [0, 1, 2, 0x200, 0x201, 0x202, 0x400, 0x401])
def test_single_pivot_pass_bitwise(ctx, shape, variant):
    variant |= capi.RANK1
    rows, cols = shape
    rng = np.random.default_rng(rows * 1000 + cols)
    T0 = rng.uniform(-2, 2, size=(rows, cols))
    T0[rng.random((rows, cols)) < 0.1] = 0.0
    m, n = rows - 1, cols - 1
    r, k = int(rng.integers(0, m)), int(rng.integers(1, n + 1))
    T0[r, k] = 1.7
    ref = O.pivot(T0, r, k)
    d = lp.DenseTableau(ctx, m, n, 1)
    d.reset(variant=variant)
    d.upload(T0, np.arange(1, m + 1, dtype=np.int32))
    d.pivot(r, k)
    out = d.download(want_log=False)
    d.close()
    assert _same_bits(out["T"], ref)
    assert out["basis"][r] == k


# ------------------------------------------------------------------ large path on synthetic LPs
@pytest.mark.parametrize("m,ns", [(32, 32), (64, 64), (200, 120), (300, 500), (512, 512)])
@pytest.mark.parametrize("variant", [0, 1, 2])
def test_dense_synth_full_solve(ctx, m, ns, variant):
    variant |= capi.RANK1                    # the three rank-1 elimination kernels
    T0, basis = O.synth_feasible(1234, 3, m, ns)
    ref = O.solve_f64(T0, basis, threads=4, log_cap=4096)
    got = lp.solve_dense(ctx, T0, basis, log_cap=4096, variant=variant)
    _assert_dense_equal(got, ref)
    assert got["status"] == lp.OPTIMAL


def test_dense_full_solve_2048x4096_bitwise(ctx):
    """A few thousand pivots to the optimum of a 2048 x 4096 tableau: pivot log, basis and the final
    tableau still equal the oracle's bit for bit (no drift: the method has no floating-point sums)."""
    m = ns = 2048
    T, basis = O.synth_feasible(1234, 1, m, ns)
    ref = O.solve_f64(T, basis, threads=16, log_cap=16384)
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.synth(1234, 1, ns)
    d.reset(log_cap=16384)
    res = d.run(2**31 - 1)
    out = d.download()
    d.close()
    assert res["status"] == ref["status"] == lp.OPTIMAL and res["iters"] == ref["iters"] > 1000
    assert out["log"][:res["iters"]].tolist() == ref["log"].tolist() and np.array_equal(out["basis"], ref["basis"])
    assert np.array_equal(out["T"], ref["T"]) and res["z"] == ref["z"]


def test_dense_device_synth_matches_host_and_windows(ctx):
    m, ns = 300, 212
    n = m + ns
    T0, basis = O.synth_feasible(77, 5, m, ns)
    d = lp.DenseTableau(ctx, m, n, 1)
    d.synth(77, 5, ns)
    d.reset(log_cap=4096, check_every=7)
    first = d.download()
    assert np.array_equal(first["T"], T0) and np.array_equal(first["basis"], basis)   # generator parity
    # run in windows of 25 pivots: state carries over, result identical to a single run
    total = 0
    while True:
        r = d.run(25)
        assert r["iters"] - total <= 25
        total = r["iters"]
        if r["status"] != lp.RUNNING:
            break
        part = O.solve_f64(T0, basis, max_iter=total)
        assert np.array_equal(d.download()["T"], part["T"])
    ref = O.solve_f64(T0, basis, log_cap=4096)
    out = d.download()
    assert r["status"] == ref["status"] == lp.OPTIMAL and r["iters"] == ref["iters"]
    assert out["log"][:r["iters"]].tolist() == ref["log"].tolist()
    assert np.array_equal(out["T"], ref["T"]) and np.array_equal(out["x"], ref["x"])
    d.close()


def test_dense_iteration_limit(ctx):
    T0, basis = O.synth_feasible(1234, 0, 64, 64)
    full = O.solve_f64(T0, basis)
    ref = O.solve_f64(T0, basis, max_iter=5, log_cap=16)
    got = lp.solve_dense(ctx, T0, basis, max_iter=5, log_cap=16)
    _assert_dense_equal(got, ref)
    assert got["status"] == lp.ITER_LIMIT and got["iters"] == 5
    exact = lp.solve_dense(ctx, T0, basis, max_iter=full["iters"])
    assert exact["status"] == lp.OPTIMAL and exact["iters"] == full["iters"]


def test_dense_timed_run_reports_kernel_time(ctx):
    m, ns = 512, 512
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.synth(1234, 0, ns)
    d.reset(time_kernels=1, variant=capi.RANK1)
    r = d.run(20)
    assert r["update_launches"] == 20 and r["update_ms"] > 0 and r["select_ms"] > 0
    d.reset(time_kernels=1, variant=_blocked(8))          # blocked path: one pass per 8 pivots
    d.synth(1234, 0, ns)
    r = d.run(20)
    assert r["iters"] == 20 and r["update_launches"] == 3 and r["update_ms"] > 0 and r["select_ms"] > 0
    d.close()


# ------------------------------------------------------------------ blocked path (delayed rank-1 applications)
def _blocked(K=0):
    # pivots per block (0 = the default, 32), PIPELINED blocks whatever the size (the default picks them from 0.4 GB on);
    # K >= 100: K - 100 pivots per block, blocks applied in place.  Odd K: the kernels that are not the default (persistent
    # step kernel, DMMA pass), so that every K-parametrised test covers both sets.
    alt = (capi.STEP_PERSISTENT | capi.DMMA_PASS) if K % 2 == 1 else (capi.DFMA_PASS if K % 4 == 2 else 0)
    return ((K % 100) << 16) | (capi.INPLACE if K >= 100 else capi.PIPELINED) | alt


def test_log_cap_shrinks_between_calls_no_overflow(ctx):
    """the device log buffer only grows, the ACTIVE limit is this call's log_cap.  Same shape solved
    twice on one context with a decreasing log_cap; the second pivot_log is guard padded and must keep its guard,
    n_logged = min(iters, log_cap), and log_cap = 0 with a non-NULL pivot_log writes nothing."""
    import ctypes as C
    T0, basis0 = O.synth_feasible(1234, 3, 96, 96)
    ref = O.solve_f64(T0, basis0, log_cap=4096)
    assert ref["iters"] > 12
    first = lp.solve_dense(ctx, T0, basis0, log_cap=1000)
    assert first["n_logged"] == ref["iters"] and first["log"].tolist() == ref["log"].tolist()
    for cap in (10, 0):
        T = np.array(T0, order="C", copy=True)
        basis = np.array(basis0, np.int32, copy=True)
        guard = np.full((cap + 64, 2), -77, np.int32)
        o = capi.default_opts(log_cap=cap)
        res = capi.Result()
        x = np.zeros(T.shape[1] - 1)
        rc = capi.lib().lpgpu_solve_dense(ctx.handle, 96, 192, 1, T.ctypes.data_as(C.c_void_p), T.shape[1], basis.ctypes.data_as(C.c_void_p),
                                          C.byref(o), guard.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p), C.byref(res))
        assert rc == 0 and res.iters == ref["iters"] and res.n_logged == min(ref["iters"], cap)
        assert guard[:cap].tolist() == ref["log"][:cap].tolist()
        assert (guard[cap:] == -77).all(), "pivot_log written beyond 2 * log_cap ints"
        assert np.array_equal(T, ref["T"])
    # batched handle API: the stride and the copy length follow this run's log_cap, not the largest one seen
    Tb, bb = O.synth_feasible_batch(1234, 0, 64, 16, 16)
    refb = O.solve_batched_f64(Tb, bb, log_cap=64)
    bt = lp.BatchedLPs(ctx, 64, 16, 32)
    bt.upload(Tb, bb)
    for cap in (64, 4):
        o = capi.default_opts(log_cap=cap)
        rc = capi.lib().lpgpu_batched_run(bt._h, C.byref(o), None)
        assert rc == 0
        guard = np.full((64 * cap + 64, 2), -77, np.int32)
        rc = capi.lib().lpgpu_batched_download(bt._h, None, None, None, None, None, guard.ctypes.data_as(C.c_void_p), None)
        assert rc == 0
        assert np.array_equal(guard[:64 * cap].reshape(64, cap, 2), refb["log"][:, :cap, :])
        assert (guard[64 * cap:] == -77).all(), "batched pivot_log written beyond n_lp * 2 * log_cap ints"
    bt.close()


@pytest.mark.parametrize("m,ns", [(32, 32), (64, 64), (200, 120), (300, 500), (512, 512)])
@pytest.mark.parametrize("K", [0, 1, 2, 5, 32])
def test_blocked_synth_full_solve_bitwise(ctx, m, ns, K):
    """Up to K pivots decided on lazily updated slices and applied in one pass: pivot log, basis, statuses and
    the tableau equal the oracle's (and so the rank-1 path's) bit for bit."""
    T0, basis = O.synth_feasible(1234, 3, m, ns)
    ref = O.solve_f64(T0, basis, threads=4, log_cap=4096)
    got = lp.solve_dense(ctx, T0, basis, log_cap=4096, variant=_blocked(K))
    _assert_dense_equal(got, ref)
    assert got["status"] == lp.OPTIMAL


@pytest.mark.parametrize("seed", range(10))
@pytest.mark.parametrize("K", [0, 3, 103, 100])
def test_blocked_bigm_random_mixed(ctx, seed, K):
    rng = np.random.default_rng(seed)
    m, ns = int(rng.integers(3, 40)), int(rng.integers(3, 40))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    if seed % 3 == 2:        # make some infeasible: contradict the first row
        A = np.vstack([A, A[0]]); b = np.append(b, b[0] - 5.0 if rel[0] != -2 else b[0] + 5.0)
        rel = np.append(rel, -2 if rel[0] != -2 else 2).astype(np.int32)
    T0, basis, n_art = lp.build_tableau(A, b, c, rel, R=2)
    ref = O.solve_f64(T0, basis, method=O.BIGM, n_art=n_art, max_iter=5000, log_cap=5000)
    got = lp.solve_dense(ctx, T0, basis, method=O.BIGM, n_art=n_art, max_iter=5000, log_cap=5000, variant=_blocked(K))
    _assert_dense_equal(got, ref)


@pytest.mark.parametrize("seed", range(8))
@pytest.mark.parametrize("K", [0, 3, 103, 100])
def test_blocked_twophase_windows_and_driveout(ctx, seed, K):
    """Two-phase through the blocked path: phase one and the main phase in blocks, the drive-out pivots between
    them through the rank-1 decision; run in windows of 5 pivots with a state read every 3.  Some LPs carry a
    duplicated equality row (an artificial stays basic at level zero after phase one), some are infeasible."""
    rng = np.random.default_rng(4000 + seed)
    m, ns = int(rng.integers(6, 40)), int(rng.integers(6, 40))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    if seed % 4 == 1:        # infeasible: contradict the first row
        A = np.vstack([A, A[0]]); b = np.append(b, b[0] - 5.0 if rel[0] != -2 else b[0] + 5.0)
        rel = np.append(rel, -2 if rel[0] != -2 else 2).astype(np.int32)
    if seed % 4 in (2, 3):   # redundant equality: twice the same row
        i = int(rng.integers(0, m))
        rel = rel.copy(); rel[i] = 3
        A = np.vstack([A, A[i]]); b = np.append(b, b[i]); rel = np.append(rel, 3).astype(np.int32)
    T0, basis, n_art = lp.build_tableau(A, b, c, rel, R=2)
    ref = O.solve_f64(T0, basis, method=O.TWOPHASE, n_art=n_art, max_iter=5000, log_cap=5000)
    d = lp.DenseTableau(ctx, T0.shape[0] - 2, T0.shape[1] - 1, 2)
    d.upload(T0, basis)
    d.reset(method=O.TWOPHASE, n_art=n_art, max_iter=5000, log_cap=5000, check_every=3, variant=_blocked(K))
    total = 0
    while True:
        r = d.run(5)
        assert r["iters"] - total <= 5
        total = r["iters"]
        if r["status"] != lp.RUNNING:
            break
    out = d.download()
    d.close()
    assert r["status"] == ref["status"] and r["iters"] == ref["iters"] and r["phase1_iters"] == ref["phase1_iters"]
    assert out["log"][:r["iters"]].tolist() == ref["log"].tolist() and np.array_equal(out["basis"], ref["basis"])
    assert np.array_equal(out["T"], ref["T"]) and r["z"] == ref["z"]


def test_blocked_statuses_windows_and_limits(ctx):
    # unbounded
    for name in ("unbounded.txt", "unr_inverted.txt"):
        d = H.load_dumps()[name]
        T, basis, n_art, n = H.tableau_from_dump(d, O.PLAIN)
        ref = O.solve_f64(H.frac_f64(T), basis, log_cap=16)
        got = lp.solve_dense(ctx, H.frac_f64(T), basis, log_cap=16, variant=_blocked())
        _assert_dense_equal(got, ref)
        assert got["status"] == lp.UNBOUNDED
    # iteration limit, and exactly the needed number of pivots
    T0, basis = O.synth_feasible(1234, 0, 64, 64)
    full = O.solve_f64(T0, basis)
    ref = O.solve_f64(T0, basis, max_iter=5, log_cap=16)
    got = lp.solve_dense(ctx, T0, basis, max_iter=5, log_cap=16, variant=_blocked())
    _assert_dense_equal(got, ref)
    assert got["status"] == lp.ITER_LIMIT and got["iters"] == 5
    exact = lp.solve_dense(ctx, T0, basis, max_iter=full["iters"], variant=_blocked())
    assert exact["status"] == lp.OPTIMAL and exact["iters"] == full["iters"]
    # windows of 25 pivots with blocks of 7 and a state read every 10: the tableau is up to date after every call
    m, ns = 300, 212
    T0, basis = O.synth_feasible(77, 5, m, ns)
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.synth(77, 5, ns)
    d.reset(log_cap=4096, check_every=10, variant=_blocked(7))
    total = 0
    while True:
        r = d.run(25)
        assert r["iters"] - total <= 25
        total = r["iters"]
        if r["status"] != lp.RUNNING:
            break
        part = O.solve_f64(T0, basis, max_iter=total)
        assert np.array_equal(d.download()["T"], part["T"])
    ref = O.solve_f64(T0, basis, log_cap=4096)
    out = d.download()
    assert r["status"] == ref["status"] == lp.OPTIMAL and r["iters"] == ref["iters"]
    assert out["log"][:r["iters"]].tolist() == ref["log"].tolist()
    assert np.array_equal(out["T"], ref["T"]) and np.array_equal(out["x"], ref["x"])
    d.close()


def test_blocked_equals_rank1_path_midsize_bitwise(ctx):
    """1024 x 4096, 300 pivots: blocked (K = 32 and 13) and rank-1 runs leave the same bits; then the two paths
    alternate on one tableau object (reset between them) and still agree."""
    m, ns, Pn = 1024, 3072, 300
    outs = []
    for variant in (capi.RANK1, _blocked(), _blocked(13)):
        d = lp.DenseTableau(ctx, m, m + ns, 1)
        d.synth(1234, 3, ns)
        d.reset(log_cap=Pn, variant=variant)
        r = d.run(Pn)
        outs.append((r, d.download()))
        d.close()
    for r, o in outs[1:]:
        assert r["iters"] == outs[0][0]["iters"] == Pn and r["z"] == outs[0][0]["z"]
        assert np.array_equal(o["log"], outs[0][1]["log"]) and np.array_equal(o["basis"], outs[0][1]["basis"])
        assert np.array_equal(o["T"], outs[0][1]["T"])


@pytest.mark.parametrize("m,ns", [(64, 64), (300, 500), (700, 1500)])
@pytest.mark.parametrize("K", [0, 1, 5, 13])
def test_pipelined_blocks_equal_inplace_blocks_and_oracle(ctx, m, ns, K):
    """Default = pipelined blocks (the pass of block b out of place on a second stream while block b+1 is decided
    through the slices of both blocks); INPLACE = steps, pass, steps on one stream.  Same bits, and the oracle's."""
    T0, basis = O.synth_feasible(99, 2, m, ns)
    ref = O.solve_f64(T0, basis, threads=4, log_cap=8192)
    a = lp.solve_dense(ctx, T0, basis, log_cap=8192, variant=_blocked(K))
    b = lp.solve_dense(ctx, T0, basis, log_cap=8192, variant=_blocked(K) | capi.INPLACE)
    _assert_dense_equal(a, ref)
    _assert_dense_equal(b, ref)
    assert _same_bits(a["T"], b["T"])
    # the step kernel as a cluster of 8 CTAs (default: 16 where the device places such clusters), and the DMMA pass on ordinary loads / the scalar-DFMA pass in halves / the persistent step
    # kernel (defaults: the ring pass, one step launch per pivot)
    for extra in (capi.STEP_CL8, capi.STEP_CL8 | capi.INPLACE,
                  capi.DMMA_PASS, capi.DMMA_PASS | capi.INPLACE, capi.STEP_PERSISTENT, capi.STEP_PERSISTENT | capi.DMMA_PASS,
                  capi.DFMA_PASS, capi.DFMA_PASS | capi.INPLACE):
        c = lp.solve_dense(ctx, T0, basis, log_cap=8192, variant=_blocked(K) | extra)
        _assert_dense_equal(c, ref)


def test_pipelined_blocks_windows_of_odd_sizes(ctx):
    """Run calls of 1, 2, 33, 64, 7, ... pivots on one tableau object: every call drains the pipeline (the tableau
    is up to date when it returns, whichever of the two buffers holds it) and the next call starts a fresh block."""
    m, ns = 300, 212
    T0, basis = O.synth_feasible(78, 6, m, ns)
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.upload(T0, basis)
    d.reset(log_cap=4096, check_every=5, variant=_blocked(4))
    total, sizes, q = 0, [1, 2, 33, 64, 7, 9, 100, 3], 0
    while True:
        r = d.run(sizes[q % len(sizes)])
        q += 1
        total = r["iters"]
        if r["status"] != lp.RUNNING:
            break
        part = O.solve_f64(T0, basis, max_iter=total)
        assert np.array_equal(d.download()["T"], part["T"])
    ref = O.solve_f64(T0, basis, log_cap=4096)
    out = d.download()
    assert r["status"] == ref["status"] == lp.OPTIMAL and r["iters"] == ref["iters"]
    assert out["log"][:r["iters"]].tolist() == ref["log"].tolist() and np.array_equal(out["T"], ref["T"])
    d.close()


def test_pipelined_blocks_stay_pending_across_run_calls(ctx):
    """A pipelined run call that ends RUNNING leaves its open block and the pass in flight pending (z comes from the eagerly
    kept objective row); the next call continues the block.  Many short calls without a download in between, then a forced
    pivot / a reset to the rank-1 path in the middle (both settle the tableau first): the oracle's bits."""
    m, ns = 300, 212
    T0, basis = O.synth_feasible(79, 4, m, ns)
    ref = O.solve_f64(T0, basis, log_cap=4096)
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.upload(T0, basis)
    d.reset(log_cap=4096, check_every=4, variant=_blocked(5))
    sizes, q, total = [3, 1, 4, 1, 5, 9, 2, 6], 0, 0
    while True:
        r = d.run(sizes[q % len(sizes)])
        q += 1
        total = r["iters"]
        if r["status"] != lp.RUNNING:
            break
        assert r["z"] == O.solve_f64(T0, basis, max_iter=total)["z"]       # from objcur while blocks are pending
    out = d.download()
    assert r["status"] == ref["status"] and r["iters"] == ref["iters"] and r["z"] == ref["z"]
    assert out["log"][:r["iters"]].tolist() == ref["log"].tolist() and np.array_equal(out["T"], ref["T"])
    # pending blocks, then a switch of path without a download: reset() must settle first
    d.upload(T0, basis)
    d.reset(log_cap=4096, variant=_blocked(7))
    d.run(10)
    part = O.solve_f64(T0, basis, max_iter=10)
    d.reset(log_cap=4096, variant=capi.RANK1)
    r2 = d.run(10**6)
    ref2 = O.solve_f64(part["T"], part["basis"], log_cap=4096)
    assert r2["status"] == ref2["status"] and r2["iters"] == ref2["iters"]
    assert np.array_equal(d.download()["T"], ref2["T"])
    d.close()


def test_dense_upload_async_stream_of_problems(ctx):
    """lpgpu_dense_upload_async: the next problem travels into one tableau object (pinned buffers, own copy stream) while
    another object pivots; runs, downloads and further uploads on an object are ordered after its asynchronous upload."""
    m, ns = 256, 192
    lps = [O.synth_feasible(500 + q, q, m, ns) for q in range(5)]
    refs = [O.solve_f64(T, b, log_cap=4096) for T, b in lps]
    objs = [lp.DenseTableau(ctx, m, m + ns, 1), lp.DenseTableau(ctx, m, m + ns, 1)]
    bufs = [(ctx.pinned((m + 1, m + ns + 1)), ctx.pinned((m,), np.int32)) for _ in range(2)]
    bufs[0][0][...] = lps[0][0]; bufs[0][1][...] = lps[0][1]
    objs[0].upload_async(*bufs[0])
    for q in range(5):
        X = objs[q % 2]
        if q + 1 < 5:
            Tn, bn = bufs[(q + 1) % 2]
            Tn[...] = lps[q + 1][0]; bn[...] = lps[q + 1][1]
            objs[(q + 1) % 2].upload_async(Tn, bn)
        X.reset(log_cap=4096, check_every=16, variant=capi.PIPELINED)
        r = X.run(10 if q == 2 else 10**6)            # problem 2 is left RUNNING with blocks pending: the next upload settles it
        out = X.download()
        ref = refs[q] if q != 2 else O.solve_f64(*lps[2], max_iter=10, log_cap=4096)
        assert r["iters"] == ref["iters"] and out["log"][:r["iters"]].tolist() == ref["log"].tolist()
        assert np.array_equal(out["T"], ref["T"]) and np.array_equal(out["basis"], ref["basis"])
    # an asynchronous upload followed directly by a blocking one and by a device synth: the last writer wins
    objs[0].upload_async(*bufs[0])
    objs[0].upload(*lps[3])
    objs[0].reset(log_cap=4096)
    r = objs[0].run(10**6)
    assert r["iters"] == refs[3]["iters"] and np.array_equal(objs[0].download()["T"], refs[3]["T"])
    for o in objs:
        o.close()


def test_pipelined_blocks_full_grid_2048x8192_bitwise(ctx):
    """Large enough for the pass to run as a full persistent grid (the case in which it leaves the step kernels'
    SMs free): 400 pivots, pipelined and in place, against the rank-1 path and the oracle."""
    m, ns, Pn = 2048, 6144, 400
    outs = []
    for variant in (capi.RANK1, 0, capi.PIPELINED, capi.INPLACE, _blocked(24), capi.STEP_CL8 | capi.PIPELINED, capi.PIPELINED | capi.STEP_PERSISTENT | capi.DMMA_PASS,
                    capi.PIPELINED | capi.DFMA_PASS, capi.INPLACE | capi.DFMA_PASS):
        d = lp.DenseTableau(ctx, m, m + ns, 1)
        d.synth(1234, 7, ns)
        d.reset(log_cap=Pn, variant=variant)
        r = d.run(150)
        r = d.run(Pn - 150)
        outs.append((r, d.download()))
        d.close()
    T0, basis0 = lp.synth_host(1234, 7, m, ns)
    ref = O.solve_f64(T0, basis0, max_iter=Pn, threads=16, log_cap=Pn)
    for r, o in outs:
        assert r["iters"] == Pn and r["z"] == ref["z"]
        assert o["log"][:Pn].tolist() == ref["log"].tolist() and np.array_equal(o["basis"], ref["basis"])
        assert np.array_equal(o["T"], ref["T"])


@pytest.mark.parametrize("method", [O.BIGM, O.TWOPHASE])
def test_pipelined_blocks_two_objective_rows_full_grid_bitwise(ctx, method):
    """Big-M and two-phase (R = 2) on a mixed-relation LP large enough for the pass to run as a full persistent grid next
    to the step cluster: the first 500 pivots, pipelined and in place, against the oracle (log, basis, tableau)."""
    rng = np.random.default_rng(2024)
    A, b, c, rel = H.random_mixed_lp(rng, 2048, 5000)
    T0, basis, n_art = lp.build_tableau(A, b, c, rel, R=2)
    Pn = 500
    ref = O.solve_f64(T0, basis, method=method, n_art=n_art, max_iter=Pn, threads=16, log_cap=Pn)
    for variant in (capi.PIPELINED, capi.INPLACE, capi.PIPELINED | capi.STEP_PERSISTENT | capi.DMMA_PASS, capi.PIPELINED | capi.DFMA_PASS):
        got = lp.solve_dense(ctx, T0, basis, method=method, n_art=n_art, max_iter=Pn, log_cap=Pn, variant=variant)
        _assert_dense_equal(got, ref)
    assert ref["iters"] == Pn


# ------------------------------------------------------------------ pricing option: Bland's rule
@pytest.mark.parametrize("variant", [capi.RANK1, capi.INPLACE, capi.PIPELINED, (3 << 16) | capi.PIPELINED, capi.STEP_CL8 | capi.INPLACE,
                                     capi.PIPELINED | capi.STEP_PERSISTENT])
def test_bland_pricing_dense_bitwise(ctx, variant):
    """variant bit BLAND: the lowest eligible column enters.  Rank-1 path, blocks in place and pipelined: the oracle's
    (pricing = Bland) pivot log, basis and tableau, for the plain method, Big-M and two-phase; Beale's cycling example
    runs into the iteration limit under the default rule and is solved under Bland."""
    T0, basis = O.synth_feasible(321, 2, 150, 120)
    ref = O.solve_f64(T0, basis, pricing=O.BLAND, log_cap=8192, threads=4)
    got = lp.solve_dense(ctx, T0, basis, log_cap=8192, variant=variant | capi.BLAND)
    _assert_dense_equal(got, ref)
    assert got["log"].tolist() != O.solve_f64(T0, basis, log_cap=8192)["log"].tolist()      # a different path than the default rule
    rng = np.random.default_rng(31)
    for seed in range(4):
        A, b, c, rel = H.random_mixed_lp(rng, 20 + seed, 16)
        for method in (O.BIGM, O.TWOPHASE):
            T, bs, n_art = lp.build_tableau(A, b, c, rel, R=2)
            ref = O.solve_f64(T, bs, method=method, n_art=n_art, max_iter=5000, log_cap=5000, pricing=O.BLAND)
            got = lp.solve_dense(ctx, T, bs, method=method, n_art=n_art, max_iter=5000, log_cap=5000, variant=variant | capi.BLAND)
            _assert_dense_equal(got, ref)
    A = np.array([[0.25, -8.0, -1.0, 9.0], [0.5, -12.0, -0.5, 3.0], [0.0, 0.0, 1.0, 0.0]])
    T, bs, _ = lp.build_tableau(A, np.array([0.0, 0.0, 1.0]), np.array([0.75, -20.0, 0.5, -6.0]), np.full(3, -2, np.int32), R=1)
    cyc = lp.solve_dense(ctx, T, bs, max_iter=120, log_cap=120, variant=variant)
    assert cyc["status"] == lp.ITER_LIMIT
    ok = lp.solve_dense(ctx, T, bs, max_iter=120, log_cap=120, variant=variant | capi.BLAND)
    assert ok["status"] == lp.OPTIMAL and ok["z"] == 1.25
    _assert_dense_equal(ok, O.solve_f64(T, bs, max_iter=120, log_cap=120, pricing=O.BLAND))


def test_bland_pricing_batched_and_refusals(ctx):
    T, basis = O.synth_feasible_batch(77, 0, 300, 24, 20)
    ref = O.solve_batched_f64(T, basis, threads=4, pricing=O.BLAND)
    got = lp.solve_batched(ctx, T, basis, variant=capi.BLAND)          # routed to the shared-memory kernel (decide())
    _assert_batched_equal(got, ref)
    assert not np.array_equal(ref["iters"], O.solve_batched_f64(T, basis, threads=4)["iters"])
    rng = np.random.default_rng(5)
    Td, bd, _ = H.random_dual_lp(rng, 12, 10)
    d = lp.DenseTableau(ctx, 12, Td.shape[1] - 1, 1)
    with pytest.raises(Exception):
        d.reset(method=lp.DUAL, variant=capi.BLAND)                    # primal methods only
    d.close()


# ------------------------------------------------------------------ pricing option: Devex reference weights
@pytest.mark.parametrize("variant", [capi.RANK1, capi.INPLACE, capi.PIPELINED, (3 << 16) | capi.PIPELINED, capi.STEP_CL8 | capi.INPLACE,
                                     capi.PIPELINED | capi.STEP_PERSISTENT])
def test_devex_pricing_dense_bitwise(ctx, variant):
    """variant bit DEVEX: sigma_j^2 / w_j decides, the weights follow every pivot's scaled row.  Rank-1 path, blocks in place
    and pipelined (persistent step kernel and per-pivot launches): the oracle's (pricing = Devex) pivot log, basis and
    tableau, for the plain method (columns scaled by powers of two, where the rule matters), Big-M and two-phase incl.
    the drive-out pivots; a second run call on the same object continues with the weights it left."""
    T0, basis = O.synth_feasible(321, 2, 150, 120)
    T0[:, 1:121] *= 2.0 ** np.random.default_rng(9).integers(-6, 7, size=120)
    ref = O.solve_f64(T0, basis, pricing=O.DEVEX, log_cap=8192, threads=4)
    got = lp.solve_dense(ctx, T0, basis, log_cap=8192, variant=variant | capi.DEVEX)
    _assert_dense_equal(got, ref)
    assert ref["status"] == O.OPTIMAL and ref["iters"] > 40
    assert got["log"].tolist() != O.solve_f64(T0, basis, log_cap=8192)["log"].tolist()[:ref["iters"]]
    rng = np.random.default_rng(31)
    for seed in range(4):
        A, b, c, rel = H.random_mixed_lp(rng, 20 + seed, 16)
        for method in (O.BIGM, O.TWOPHASE):
            T, bs, n_art = lp.build_tableau(A, b, c, rel, R=2)
            ref2 = O.solve_f64(T, bs, method=method, n_art=n_art, max_iter=5000, log_cap=5000, pricing=O.DEVEX)
            got2 = lp.solve_dense(ctx, T, bs, method=method, n_art=n_art, max_iter=5000, log_cap=5000, variant=variant | capi.DEVEX)
            _assert_dense_equal(got2, ref2)
    # windows: the weights live in the tableau object between run calls
    d = lp.DenseTableau(ctx, 150, 270, 1)
    d.upload(T0, basis)
    d.reset(log_cap=8192, variant=variant | capi.DEVEX)
    for _ in range(400):
        if d.run(7)["status"] != lp.RUNNING:
            break
    out = d.download()
    d.close()
    assert np.array_equal(out["T"], ref["T"]) and out["log"][:ref["iters"]].tolist() == ref["log"].tolist()


def test_devex_pricing_batched_and_refusals(ctx):
    T, basis = O.synth_feasible_batch(77, 0, 300, 24, 20)
    T[:, :, 1:21] *= 2.0 ** np.random.default_rng(3).integers(-5, 6, size=(300, 1, 20))
    ref = O.solve_batched_f64(T, basis, threads=4, pricing=O.DEVEX)
    got = lp.solve_batched(ctx, T, basis, variant=capi.DEVEX)          # the shared-memory kernel keeps the weights
    _assert_batched_equal(got, ref)
    assert not np.array_equal(ref["iters"], O.solve_batched_f64(T, basis, threads=4)["iters"])
    rng = np.random.default_rng(6)
    A, b, c, rel = H.random_mixed_lp(rng, 14, 12)
    for method in (O.BIGM, O.TWOPHASE):
        Tm, bm, n_art = lp.build_tableau(A, b, c, rel, R=2)
        Tb = np.repeat(Tm[None], 40, axis=0).copy()
        Tb[:, :-2, 0] *= 1.0 + 0.01 * np.arange(40)[:, None]
        bb = np.repeat(np.asarray(bm, np.int32)[None], 40, axis=0).copy()
        refb = O.solve_batched_f64(Tb, bb, method=method, n_art=n_art, max_iter=2000, threads=4, pricing=O.DEVEX)
        gotb = lp.solve_batched(ctx, Tb, bb, method=method, n_art=n_art, max_iter=2000, variant=capi.DEVEX)
        _assert_batched_equal(gotb, refb)
    Td, bd, _ = H.random_dual_lp(rng, 12, 10)
    d = lp.DenseTableau(ctx, 12, Td.shape[1] - 1, 1)
    with pytest.raises(Exception):
        d.reset(method=lp.DUAL, variant=capi.DEVEX)                    # primal methods only
    with pytest.raises(Exception):
        d.reset(variant=capi.DEVEX | capi.BLAND)                       # one rule at a time
    d.close()


# ------------------------------------------------------------------ dual simplex
@pytest.mark.parametrize("seed", range(8))
def test_dual_simplex_dense_bitwise(ctx, seed):
    rng = np.random.default_rng(900 + seed)
    m, ns = int(rng.integers(3, 60)), int(rng.integers(3, 60))
    T, basis, _ = H.random_dual_lp(rng, m, ns, infeasible=(seed == 7))
    ref = O.solve_f64(T, basis, method=O.DUAL, log_cap=512)
    got = lp.solve_dense(ctx, T, basis, method=lp.DUAL, log_cap=512)
    _assert_dense_equal(got, ref)
    assert got["status"] == (lp.INFEASIBLE if seed == 7 else lp.OPTIMAL)
    capped = lp.solve_dense(ctx, T, basis, method=lp.DUAL, log_cap=512, max_iter=1)
    _assert_dense_equal(capped, O.solve_f64(T, basis, method=O.DUAL, log_cap=512, max_iter=1))


def test_dual_simplex_midsize_and_batched(ctx):
    rng = np.random.default_rng(77)
    T, basis, _ = H.random_dual_lp(rng, 300, 200)
    ref = O.solve_f64(T, basis, method=O.DUAL, log_cap=4096)
    got = lp.solve_dense(ctx, T, basis, method=lp.DUAL, log_cap=4096)
    _assert_dense_equal(got, ref)
    assert ref["status"] == O.OPTIMAL and ref["iters"] > 20
    _assert_dense_equal(lp.solve_dense(ctx, T, basis, method=lp.DUAL, log_cap=4096, variant=capi.PIPELINED), ref)      # pipelined blocks
    # batched: the shared-memory kernel takes the dual method through the same decide()
    Ts, Bs = zip(*[H.random_dual_lp(rng, 10, 12)[:2] for _ in range(200)])
    Tb, Bb = np.stack(Ts), np.stack(Bs)
    refb = O.solve_batched_f64(Tb, Bb, method=O.DUAL, threads=4)
    gotb = lp.solve_batched(ctx, Tb, Bb, method=lp.DUAL)
    _assert_batched_equal(gotb, refb)


@pytest.mark.parametrize("variant", [0, 4 << 16, capi.STEP_CL8, (4 << 16) | capi.STEP_CL8, capi.RANK1, capi.PIPELINED, (5 << 16) | capi.PIPELINED,
                                     capi.PIPELINED | capi.STEP_CL8, (3 << 16) | capi.PIPELINED | capi.DFMA_PASS])
def test_dual_simplex_blocked_windows_and_rank1(ctx, variant):
    """Dual simplex through the blocked path (in place and pipelined; blocks of 32 / 28, 5, 4 and 3 pivots; step kernel as a cluster
    of 16 x 512 and of 8 x 1024 threads) and through the rank-1 path, in windows of 7 pivots: all leave the oracle's bits."""
    rng = np.random.default_rng(555)
    for (m, ns, infeasible) in [(120, 90, False), (40, 70, False), (30, 30, True)]:
        T, basis, _ = H.random_dual_lp(rng, m, ns, infeasible=infeasible)
        ref = O.solve_f64(T, basis, method=O.DUAL, log_cap=4096)
        d = lp.DenseTableau(ctx, m, T.shape[1] - 1, 1)
        d.upload(T, basis)
        d.reset(method=lp.DUAL, log_cap=4096, check_every=3, variant=variant)
        total = 0
        while True:
            r = d.run(7)
            assert r["iters"] - total <= 7
            total = r["iters"]
            if r["status"] != lp.RUNNING:
                break
        out = d.download()
        d.close()
        assert r["status"] == ref["status"] == (lp.INFEASIBLE if infeasible else lp.OPTIMAL) and r["iters"] == ref["iters"]
        assert out["log"][:r["iters"]].tolist() == ref["log"].tolist() and np.array_equal(out["basis"], ref["basis"])
        assert np.array_equal(out["T"], ref["T"]) and r["z"] == ref["z"] and np.array_equal(out["x"], ref["x"])


# ------------------------------------------------------------------ batched path
def _assert_batched_equal(got, ref, final=None):
    assert np.array_equal(got["status"], ref["status"])
    assert np.array_equal(got["iters"], ref["iters"])
    assert np.array_equal(got["basis"], ref["basis"])
    assert np.array_equal(got["z"], ref["z"]) and np.array_equal(got["x"], ref["x"])
    if final is not None:
        assert np.array_equal(final, ref["T"])


@pytest.mark.parametrize("m,ns,n_lp", [(32, 32, 2000), (64, 64, 300), (8, 8, 999), (5, 9, 100), (33, 20, 101)])
def test_cfg3_batched_synth_bitwise(ctx, m, ns, n_lp):
    T, basis = O.synth_feasible_batch(1234, 10, n_lp, m, ns)
    ref = O.solve_batched_f64(T, basis, threads=8, log_cap=256)
    b = lp.BatchedLPs(ctx, n_lp, m, m + ns, 1)
    b.upload(T, basis)
    ms = b.run(log_cap=256, keep_final=True)
    assert ms > 0
    got = b.download(want_log=True, want_final=True)
    _assert_batched_equal(got, ref, got["T"])
    assert np.array_equal(got["log"], ref["log"])
    assert np.all(got["status"] == lp.OPTIMAL)
    # device generator produces the same LPs
    b.synth(1234, 10, ns)
    b.run()
    got2 = b.download()
    _assert_batched_equal(got2, ref)
    b.close()
    # one-shot host-buffer entry point
    got3 = lp.solve_batched(ctx, T, basis)
    _assert_batched_equal(got3, ref)


@pytest.mark.parametrize("m,ns,n_lp", [(32, 32, 700), (64, 64, 150), (16, 48, 200), (40, 88, 120), (2, 2, 50),
                                       (32, 31, 97), (31, 33, 97), (64, 63, 40), (7, 120, 40)])
@pytest.mark.parametrize("force_smem", [False, True])
def test_batched_register_and_smem_kernels_agree_with_oracle(ctx, m, ns, n_lp, force_smem):
    """Shapes on and around the register tiles (m <= 32, n <= 64 -> one warp per LP; m <= 64, n <= 128 ->
    four warps per LP) through both kernels: bit-exact results, pivot logs and final tableaux."""
    T, basis = O.synth_feasible_batch(77, 3, n_lp, m, ns)
    ref = O.solve_batched_f64(T, basis, threads=8, log_cap=300)
    b = lp.BatchedLPs(ctx, n_lp, m, m + ns, 1)
    b.upload(T, basis)
    b.run(log_cap=300, keep_final=True, force_smem=force_smem)
    got = b.download(want_log=True, want_final=True)
    b.close()
    _assert_batched_equal(got, ref, got["T"])
    assert np.array_equal(got["log"], ref["log"])


@pytest.mark.parametrize("force_smem", [False, True])
def test_batched_plain_unbounded_and_iteration_limit(ctx, force_smem):
    """Plain-method LPs with negative entries: some unbounded, some optimal; plus a pivot cap."""
    rng = np.random.default_rng(11)
    n_lp, m, ns = 400, 12, 10
    T = np.zeros((n_lp, m + 1, m + ns + 1))
    B = np.zeros((n_lp, m), np.int32)
    for p in range(n_lp):
        A = rng.uniform(-1.0, 1.0, size=(m, ns))
        A[rng.random((m, ns)) < 0.3] = 0.0
        bvec = rng.uniform(0.0, 3.0, m)
        bvec[rng.random(m) < 0.2] = 0.0                  # degenerate rows -> ratio ties
        c = rng.uniform(-0.5, 1.0, ns)
        T[p], B[p], _ = lp.build_tableau(A, bvec, c, np.full(m, -2, np.int32), R=1)
    for cap in (2**31 - 1, 3):
        ref = O.solve_batched_f64(T, B, max_iter=cap, threads=4, log_cap=128)
        b = lp.BatchedLPs(ctx, n_lp, m, m + ns, 1)
        b.upload(T, B)
        b.run(max_iter=cap, log_cap=128, keep_final=True, force_smem=force_smem)
        got = b.download(want_log=True, want_final=True)
        b.close()
        _assert_batched_equal(got, ref, got["T"])
        assert np.array_equal(got["log"], ref["log"])
    assert lp.UNBOUNDED in set(ref["status"].tolist()) | set(O.solve_batched_f64(T, B)["status"].tolist())


def test_cfg3_full_size_batched_bitwise(ctx):
    """BASELINE config 3 at its full size: 100 000 LPs m=32 n=64, device generator, every LP's status,
    iteration count, objective, x and basis bit-compared with the oracle (host generator)."""
    n_lp, m, ns = 100_000, 32, 32
    b = lp.BatchedLPs(ctx, n_lp, m, m + ns, 1)
    b.synth(1234, 0, ns)
    b.run()
    got = b.download()
    b.close()
    T, basis = O.synth_feasible_batch(1234, 0, n_lp, m, ns)
    ref = O.solve_batched_f64(T, basis, threads=16)
    _assert_batched_equal(got, ref)
    assert np.all(got["status"] == lp.OPTIMAL) and got["iters"].mean() > 10


def test_cfg5a_sample_batched_bitwise(ctx):
    """BASELINE config 5a shape (m=64 n=128), LP indices 600 000 .. 620 000 of the million (what rank 4 of 8
    would hold first), against the oracle."""
    n_lp, m, ns, lo = 20_000, 64, 64, 600_000
    b = lp.BatchedLPs(ctx, n_lp, m, m + ns, 1)
    b.synth(1234, lo, ns)
    b.run()
    got = b.download()
    b.close()
    T, basis = O.synth_feasible_batch(1234, lo, n_lp, m, ns)
    ref = O.solve_batched_f64(T, basis, threads=16)
    _assert_batched_equal(got, ref)


@pytest.mark.parametrize("method", [O.BIGM, O.TWOPHASE])
def test_batched_bigm_twophase(ctx, method):
    rng = np.random.default_rng(5)
    m, ns, n_lp = 6, 5, 64
    rel = np.array([2, -2, 3, -2, 2, -2], np.int32)
    Ts, Bs = [], []
    for p in range(n_lp):
        A = rng.uniform(0.1, 1.5, size=(m, ns))
        x0 = rng.uniform(0, 2, ns)
        ax = A @ x0
        b = np.where(rel == -2, ax + 1.0, np.where(rel == 2, ax - 0.5, ax))
        if p % 5 == 4:
            b[1] = -1.0 - p          # A>0, x>=0: row "<= negative" -> infeasible
        c = rng.uniform(-1, 1, ns)
        T0, basis, n_art = lp.build_tableau(A, b, c, rel, R=2)
        Ts.append(T0); Bs.append(basis)
    # rows with b<0 flip relation and change the column count: keep the common-shape ones
    shape = Ts[0].shape
    keep = [i for i in range(n_lp) if Ts[i].shape == shape]
    T = np.stack([Ts[i] for i in keep]); B = np.stack([Bs[i] for i in keep])
    n_art = 3
    ref = O.solve_batched_f64(T, B, method=method, n_art=n_art, max_iter=500, threads=4)
    got = lp.solve_batched(ctx, T, B, method=method, n_art=n_art, max_iter=500)
    _assert_batched_equal(got, ref)
    assert set(np.unique(got["status"])) <= {lp.OPTIMAL, lp.UNBOUNDED, lp.INFEASIBLE}


def _same_shape_batch(build, n_lp):
    """Stack the LPs of a generator that share the shape of the first one."""
    Ts, Bs, na = [], [], None
    for p in range(n_lp):
        T0, basis, n_art = build(p)
        if Ts and (T0.shape != Ts[0].shape or n_art != na):
            continue
        na = n_art
        Ts.append(T0); Bs.append(basis)
    return np.stack(Ts), np.stack(Bs), na


@pytest.mark.parametrize("shape", [(6, 5), (20, 24), (40, 50)])
@pytest.mark.parametrize("method", [O.BIGM, O.TWOPHASE])
@pytest.mark.parametrize("bland", [False, True])
def test_batched_register_kernels_two_objective_rows(ctx, shape, method, bland):
    """Big-M and two-phase through the register-resident kernels (batched_regx.cuh): one warp per LP (m <= 32, n <= 64) and
    four warps per LP (m <= 64, n <= 128), random mixed-relation LPs incl. infeasible ones and redundant equality rows
    (two-phase drive-out); bit-exact against the oracle, and against the shared-memory kernel (force_smem) incl. the final
    tableaux and pivot logs.  Bland's rule as well."""
    m, ns = shape
    rng = np.random.default_rng(1000 * m + ns)

    rel = np.array([(-2, 2, 3, -2)[i % 4] for i in range(m)], np.int32)       # the same relations in every LP: one shape
    rel[0] = rel[1] = 3

    def build(p):
        A = rng.uniform(-0.5, 1.5, size=(m, ns))
        x0 = rng.uniform(0.0, 2.0, size=ns)
        ax = A @ x0
        b = np.where(rel == -2, ax + rng.uniform(0.1, 2.0, m), np.where(rel == 2, ax - rng.uniform(0.1, 2.0, m), ax))
        c = rng.uniform(-1.0, 1.0, ns)
        if p % 4 == 1:      # infeasible: two equality rows that contradict each other
            A[1] = A[0]; b[1] = b[0] + 3.0
        if p % 4 == 2:      # redundant equality rows: an artificial stays basic at level zero (two-phase drive-out)
            A[1] = A[0]; b[1] = b[0]
        return O.build_tableau(A, b, c, rel, R=2)

    T, B, n_art = _same_shape_batch(build, 96)
    assert T.shape[0] >= 16 and T.shape[2] - 1 <= (64 if m <= 32 else 128)
    pricing = O.BLAND if bland else 0
    variant = capi.BLAND if bland else 0
    ref = O.solve_batched_f64(T, B, method=method, n_art=n_art, max_iter=2000, threads=8, log_cap=64, pricing=pricing)
    for force_smem in (False, True):
        b = lp.BatchedLPs(ctx, T.shape[0], m, T.shape[2] - 1, 2)
        b.upload(T, B)
        b.run(method=method, n_art=n_art, max_iter=2000, log_cap=64, keep_final=True, force_smem=force_smem, variant=variant)
        got = b.download(want_log=True, want_final=True)
        b.close()
        _assert_batched_equal(got, ref)
        assert np.array_equal(got["log"], ref["log"])
        # final tableaux: status OPTIMAL / INFEASIBLE leave a complete tableau behind
        assert np.array_equal(got["T"], ref["T"])
    assert len(set(np.unique(ref["status"]))) >= 2


@pytest.mark.parametrize("shape", [(10, 12), (30, 34), (50, 60)])
def test_batched_register_kernels_dual(ctx, shape):
    """The dual simplex through the register-resident kernels, against the oracle and the shared-memory kernel."""
    m, ns = shape
    rng = np.random.default_rng(31 * m + ns)
    pairs = [H.random_dual_lp(rng, m, ns, infeasible=(q % 7 == 3))[:2] for q in range(80)]
    T = np.stack([a for a, _ in pairs]); B = np.stack([b for _, b in pairs])
    ref = O.solve_batched_f64(T, B, method=O.DUAL, max_iter=4000, threads=8, log_cap=32)
    for force_smem in (False, True):
        b = lp.BatchedLPs(ctx, T.shape[0], m, T.shape[2] - 1, 1)
        b.upload(T, B)
        b.run(method=O.DUAL, max_iter=4000, log_cap=32, keep_final=True, force_smem=force_smem)
        got = b.download(want_log=True, want_final=True)
        b.close()
        _assert_batched_equal(got, ref)
        assert np.array_equal(got["log"], ref["log"]) and np.array_equal(got["T"], ref["T"])
    assert lp.OPTIMAL in ref["status"]


@pytest.mark.parametrize("m,ns", [(32, 32), (64, 64), (17, 40)])
def test_batched_register_kernels_bland_plain(ctx, m, ns):
    T, basis = O.synth_feasible_batch(5, 0, 128, m, ns)
    ref = O.solve_batched_f64(T, basis, threads=8, pricing=O.BLAND, log_cap=16)
    for force_smem in (False, True):
        b = lp.BatchedLPs(ctx, 128, m, m + ns, 1)
        b.upload(T, basis)
        b.run(log_cap=16, keep_final=True, force_smem=force_smem, variant=capi.BLAND)
        got = b.download(want_log=True, want_final=True)
        b.close()
        _assert_batched_equal(got, ref)
        assert np.array_equal(got["log"], ref["log"]) and np.array_equal(got["T"], ref["T"])


def test_batched_edge_cases(ctx):
    # empty batch
    out = lp.solve_batched(ctx, np.zeros((0, 4, 6)), np.zeros((0, 3), np.int32))
    assert out["status"].shape == (0,)
    # single LP, iteration cap
    T, basis = O.synth_feasible_batch(1, 0, 1, 16, 16)
    ref = O.solve_batched_f64(T, basis, max_iter=2)
    got = lp.solve_batched(ctx, T, basis, max_iter=2)
    _assert_batched_equal(got, ref)
    assert got["status"][0] == lp.ITER_LIMIT
    # shape too large for shared memory is refused, not silently rerouted
    with pytest.raises(lp.LpgpuError):
        lp.BatchedLPs(ctx, 1, 200, 400, 1)


# ------------------------------------------------------------------ full-size properties
def test_cfg4_full_size_window_properties(ctx):
    """8192 x 16384: after P pivots the basic columns are exact unit vectors, b stays >= 0, the
    objective never decreases, and the first pivots equal the oracle's (bit-exact tableau)."""
    m = ns = 8192
    n = m + ns
    d = lp.DenseTableau(ctx, m, n, 1)
    d.synth(1234, 0, ns)
    d.reset(log_cap=64)
    r1 = d.run(3)
    out = d.download()
    T0, basis0 = lp.synth_host(1234, 0, m, ns)
    ref = O.solve_f64(T0, basis0, max_iter=3, threads=8, log_cap=64)
    assert out["log"][:3].tolist() == ref["log"].tolist()
    assert np.array_equal(out["T"], ref["T"]) and np.array_equal(out["basis"], ref["basis"])
    z_prev = r1["z"]
    r2 = d.run(29)
    assert r2["iters"] == 32 and r2["z"] >= z_prev
    out = d.download()
    T = out["T"]
    assert np.all(T[:m, 0] >= 0.0)
    cols = out["basis"]
    sub = T[:, cols]                      # (m+1) x m: must be [I; 0]
    assert np.array_equal(sub[:m], np.eye(m)) and np.all(sub[m] == 0.0)
    d.close()


def test_cfg4_full_size_pipelined_equals_inplace_bitwise(ctx):
    """8192 x 16384, 200 pivots in run calls of 70 + 130 (six full blocks and a partial one, blocks left pending between the
    calls): the default (pipelined at this size) and the in-place order leave the same pivot log, basis and tableau, bit
    for bit; basic columns are exact unit vectors and b stays >= 0."""
    m = ns = 8192
    outs = []
    for variant in (0, capi.INPLACE):
        d = lp.DenseTableau(ctx, m, m + ns, 1)
        d.synth(1234, 0, ns)
        d.reset(log_cap=256, variant=variant)
        z1 = d.run(70)["z"]
        r = d.run(130)
        assert r["iters"] == 200 and r["z"] >= z1
        outs.append(d.download())
        d.close()
    a, b = outs
    assert a["log"][:200].tolist() == b["log"][:200].tolist() and np.array_equal(a["basis"], b["basis"])
    assert _same_bits(a["T"], b["T"])
    T = a["T"]
    assert np.all(T[:m, 0] >= 0.0)
    sub = T[:, a["basis"]]
    assert np.array_equal(sub[:m], np.eye(m)) and np.all(sub[m] == 0.0)


def test_cfg4_full_size_default_path_96_pivots_vs_oracle(ctx):
    """BASELINE config 4 (8192 x 16384) on the DEFAULT path (pipelined blocks at this size), compared with the ORACLE:
    96 pivots in two run calls of 40 + 56 -- three blocks of 32, the second left open and the first one's pass left
    pending across the call boundary, both tableau buffers used -- then pivot log, basis and the whole tableau
    bit for bit against O.solve_f64(max_iter = 96)."""
    import os
    m = ns = 8192
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.synth(1234, 0, ns)
    d.reset(log_cap=128)
    r1 = d.run(40)
    assert r1["iters"] == 40 and r1["status"] == lp.RUNNING
    r2 = d.run(56)
    assert r2["iters"] == 96 and r2["status"] == lp.RUNNING and r2["z"] >= r1["z"]
    out = d.download()
    d.close()
    T0, basis0 = lp.synth_host(1234, 0, m, ns)
    ref = O.solve_f64(T0, basis0, max_iter=96, threads=os.cpu_count() or 8, log_cap=128)
    assert ref["iters"] == 96
    assert out["log"][:96].tolist() == ref["log"].tolist()
    assert np.array_equal(out["basis"], ref["basis"])
    assert _same_bits(out["T"], ref["T"])
    assert r2["z"] == ref["z"]


def _golden_cfg4(name="cfg4_full_solve.json"):
    import json
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name)
    if not os.path.exists(path):
        pytest.skip(f"{path} not generated (python oracle/gen_golden_cfg4.py)")
    return json.load(open(path))


def _check_against_golden(ctx, g, variant=0, window=None):
    import hashlib
    m = g["m"]
    ns = g["n"] - m
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.synth(g["seed"], g["lp"], ns)
    d.reset(log_cap=g["iters"] + 64, variant=variant, check_every=256)
    if window:
        r = d.run(window)
        while r["status"] == lp.RUNNING:
            r = d.run(window)
    else:
        r = d.run(2**31 - 1)
    out = d.download(want_T=False)
    d.close()
    assert lp.STATUS_NAMES[r["status"]] == g["status"] and r["iters"] == g["iters"]
    log = np.ascontiguousarray(out["log"][:r["iters"]], np.int32)
    for k, h in g["log_prefix_sha256"].items():
        assert hashlib.sha256(log[:int(k)].tobytes()).hexdigest() == h, f"pivot log diverges from the oracle within the first {k} pivots"
    assert hashlib.sha256(log.tobytes()).hexdigest() == g["log_sha256"]
    assert out["basis"].tolist() == g["basis"]
    assert float(r["z"]).hex() == g["z_hex"]
    assert hashlib.sha256(np.ascontiguousarray(out["x"]).tobytes()).hexdigest() == g["x_sha256"]


def test_cfg4_whole_solve_matches_golden(ctx):
    """The WHOLE config-4 solve on the default path: status, iteration count, z (exact bits), final basis, x and the
    SHA-256 of the pivot log equal the oracle's record in tests/golden/cfg4_full_solve.json (oracle/gen_golden_cfg4.py)."""
    _check_against_golden(ctx, _golden_cfg4())


@pytest.mark.parametrize("variant,window", [(0, None), (capi.PIPELINED, 37), (capi.INPLACE, None), (capi.RANK1, None)])
def test_small_whole_solve_matches_golden(ctx, variant, window):
    """The same check at m = 512 (tests/golden/cfg4_full_solve_m512.json) for every path, incl. pipelined blocks in run
    calls of 37 pivots."""
    _check_against_golden(ctx, _golden_cfg4("cfg4_full_solve_m512.json"), variant=variant, window=window)


def test_dense_lookahead_is_dropped_across_upload_and_forced_pivot(ctx):
    """The select kernel prices one pivot ahead; an upload, a device synth or a forced pivot between run()
    calls must invalidate that lookahead."""
    m, ns = 96, 80
    Ta, ba = O.synth_feasible(5, 1, m, ns)
    Tb, bb = O.synth_feasible(5, 2, m, ns)
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.upload(Ta, ba)
    d.reset(log_cap=1024)
    assert d.run(7)["status"] == lp.RUNNING
    d.upload(Tb, bb)                       # new LP behind the lookahead's back
    d.reset(log_cap=1024)
    r = d.run(10**6)
    ref = O.solve_f64(Tb, bb, log_cap=1024)
    out = d.download()
    assert r["status"] == ref["status"] and r["iters"] == ref["iters"]
    assert out["log"][:r["iters"]].tolist() == ref["log"].tolist() and np.array_equal(out["T"], ref["T"])
    # forced pivot in the middle of a run
    d.upload(Ta, ba)
    d.reset(log_cap=1024)
    d.run(4)
    part = O.solve_f64(Ta, ba, max_iter=4)
    k = int(np.argmax(part["T"][m, 1:ns + 1] != 0) + 1)
    rr = int(np.argmax(part["T"][:m, k] > 0.05))
    d.pivot(rr, k)
    forced = O.pivot(part["T"], rr, k)
    bf = part["basis"].copy(); bf[rr] = k
    r2 = d.run(10**6)
    ref2 = O.solve_f64(forced, bf)
    assert r2["status"] == ref2["status"] and r2["iters"] - 4 == ref2["iters"]
    assert np.array_equal(d.download()["T"], ref2["T"])
    d.close()
