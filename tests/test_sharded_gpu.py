"""GPU parity tests of the column-sharded large-tableau path (lpgpu_shard_*) against the CPU oracle.

world > 1 inside ONE process on one device exercises the whole N > 1 logic (partition, candidate
exchange, arg-max with ties to the lowest global column, entering column from its owner, redundant
ratio test) with device copies as the transport; the NCCL transport is covered at world = 1 on any box
and at world = 2 (one process per GPU, torchrun) where two GPUs exist.  Bar: bit-exact pivot log,
basis, status, iteration count and tableau."""
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as H
import oracle_lib as O

import linearprogramming_b200 as lp
from linearprogramming_b200 import capi, sharding

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ctx():
    c = lp.Context(0)
    yield c
    c.close()


def _solve_sharded(ctx, T, basis, world, method=O.PLAIN, n_art=0, log_cap=4096, nccl_id=None, window=None, **kw):
    R = 1 if method == O.PLAIN else 2
    m, n = T.shape[0] - R, T.shape[1] - 1
    s = lp.ShardedTableau(ctx, m, n, R, world=world, rank=0, nccl_id=nccl_id)
    s.upload(T, basis)
    s.reset(method=method, n_art=n_art, log_cap=log_cap, **kw)
    if window is None:
        res = s.run(2**31 - 1)
    else:
        while True:
            res = s.run(window)
            if res["status"] != lp.RUNNING:
                break
    out = s.download()
    s.close()
    res.update(T=out["T"], basis=out["basis"], log=out["log"][:res["n_logged"]], x=out["x"])
    return res


def _assert_equal(got, ref):
    assert got["status"] == ref["status"], (capi.STATUS_NAMES[got["status"]], O.STATUS_NAMES[ref["status"]])
    assert got["iters"] == ref["iters"]
    assert got["log"].tolist() == ref["log"].tolist()
    assert np.array_equal(got["basis"], ref["basis"])
    assert np.array_equal(got["T"], ref["T"])
    assert got["z"] == ref["z"] and np.array_equal(got["x"], ref["x"])


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("m,ns", [(24, 20), (64, 64), (200, 120)])
def test_sharded_synth_matches_oracle(ctx, world, m, ns):
    T, basis = O.synth_feasible(1234, 5, m, ns)
    ref = O.solve_f64(T, basis, log_cap=4096)
    got = _solve_sharded(ctx, T, basis, world)
    assert ref["status"] == O.OPTIMAL
    _assert_equal(got, ref)


@pytest.mark.parametrize("world", [2, 5])
@pytest.mark.parametrize("seed", range(6))
def test_sharded_bigm_random_mixed(ctx, world, seed):
    rng = np.random.default_rng(300 + seed)
    m, ns = int(rng.integers(4, 40)), int(rng.integers(4, 40))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
    ref = O.solve_f64(T, basis, method=O.BIGM, n_art=n_art, log_cap=4096)
    got = _solve_sharded(ctx, T, basis, world, method=O.BIGM, n_art=n_art)
    _assert_equal(got, ref)
    assert got["zM"] == ref["zM"]


@pytest.mark.parametrize("name,method", [("infeasible.txt", O.BIGM), ("unbounded.txt", O.PLAIN),
                                         ("testdata_fixed.txt", O.PLAIN), ("min_surplus.txt", O.BIGM)])
def test_sharded_statuses_on_model_family(ctx, name, method):
    d = H.load_dumps()[name]
    T, basis, n_art, n = H.tableau_from_dump(d, method)
    T0 = H.frac_f64(T)
    ref = O.solve_f64(T0, basis, method=method, n_art=n_art, log_cap=64)
    for world in (1, 2):
        got = _solve_sharded(ctx, T0, basis, world, method=method, n_art=n_art, log_cap=64)
        _assert_equal(got, ref)


def test_sharded_windows_iteration_limit_and_device_synth(ctx):
    m, ns = 96, 80
    T, basis = O.synth_feasible(1234, 11, m, ns)
    ref = O.solve_f64(T, basis, log_cap=4096)
    # windows of 7 pivots, state read every 3: same result as one run
    got = _solve_sharded(ctx, T, basis, 3, window=7, check_every=3)
    _assert_equal(got, ref)
    # iteration cap
    cap = ref["iters"] // 2
    refc = O.solve_f64(T, basis, max_iter=cap, log_cap=4096)
    gotc = _solve_sharded(ctx, T, basis, 4, max_iter=cap)
    assert refc["status"] == O.ITER_LIMIT
    _assert_equal(gotc, refc)
    # device generator == host generator, block by block
    s = lp.ShardedTableau(ctx, m, ns + m, 1, world=4)
    s.synth(1234, 11, ns)
    s.reset(log_cap=4096)
    first = s.download(want_log=False)
    assert np.array_equal(first["T"], T) and np.array_equal(first["basis"], basis)
    res = s.run(2**31 - 1)
    out = s.download()
    s.close()
    assert res["status"] == ref["status"] and res["iters"] == ref["iters"]
    assert np.array_equal(out["T"], ref["T"]) and out["log"][:res["n_logged"]].tolist() == ref["log"].tolist()


def test_sharded_equals_single_gpu_path_bitwise_midsize(ctx):
    """1024 x 4096, first 300 pivots: the column-sharded run (4 and 7 ranks) leaves exactly the tableau,
    basis and pivot log of the single-tableau path (tier T4)."""
    m, ns, P = 1024, 3072, 300
    d = lp.DenseTableau(ctx, m, m + ns, 1)
    d.synth(1234, 3, ns)
    d.reset(log_cap=P)
    r1 = d.run(P)
    one = d.download()
    d.close()
    for world in (4, 7):
        s = lp.ShardedTableau(ctx, m, m + ns, 1, world=world)
        s.synth(1234, 3, ns)
        s.reset(log_cap=P)
        r2 = s.run(P)
        many = s.download()
        s.close()
        assert r2["iters"] == r1["iters"] == P and r2["status"] == r1["status"] == lp.RUNNING
        assert np.array_equal(many["log"], one["log"]) and np.array_equal(many["basis"], one["basis"])
        assert np.array_equal(many["T"], one["T"]) and r2["z"] == r1["z"]


def test_sharded_rejects_bad_shapes_and_twophase_off_the_blocked_path(ctx):
    with pytest.raises(lp.LpgpuError):
        lp.ShardedTableau(ctx, 4, 3, 1, world=4)             # more ranks than columns
    s = lp.ShardedTableau(ctx, 4, 8, 2, world=2)             # in-process ranks: the rank-1 pipeline, no two-phase there
    with pytest.raises(lp.LpgpuError):
        s.reset(method=lp.TWOPHASE, n_art=2)
    s.close()
    s = lp.ShardedTableau(ctx, 4, 8, 1, world=1, rank=0, p2p=True)
    with pytest.raises(lp.LpgpuError):
        s.reset(method=lp.DUAL)
    s.close()


def _solve_p2p_world1(ctx, T, basis, method, n_art, variant=0, window=11, check_every=4):
    R = 1 if method == O.PLAIN else 2
    s = lp.ShardedTableau(ctx, T.shape[0] - R, T.shape[1] - 1, R, world=1, rank=0, p2p=True)
    s.upload(T, basis)
    s.reset(method=method, n_art=n_art, log_cap=8192, variant=variant, check_every=check_every)
    while True:
        res = s.run(window)
        if res["status"] != lp.RUNNING:
            break
    out = s.download()
    s.close()
    res.update(T=out["T"], basis=out["basis"], log=out["log"][:res["n_logged"]], x=out["x"])
    return res


@pytest.mark.parametrize("variant", [0, 3 << 16])
@pytest.mark.parametrize("seed", range(8))
def test_sharded_p2p_world1_twophase(ctx, seed, variant):
    """Two-phase method on the column-sharded blocked path (one rank over the peer-memory transport): phase switch inside
    the step kernel, drive-out by forced steps from the host loop, main phase without the artificial columns; random
    mixed-relation LPs incl. infeasible ones and redundant equality rows (artificials left basic at level zero, so the
    drive-out has work), run calls of 11 and of 1 pivots.  Bit-exact against the oracle."""
    rng = np.random.default_rng(4100 + seed)
    m, ns = int(rng.integers(4, 40)), int(rng.integers(4, 40))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    if seed % 3 == 1:          # infeasible: a row against itself
        A = np.vstack([A, A[0]]); b = np.append(b, b[0] - 5.0 if rel[0] != -2 else b[0] + 5.0)
        rel = np.append(rel, -2 if rel[0] != -2 else 2).astype(np.int32)
    if seed % 3 == 2:          # redundant equalities: duplicated rows, their artificials stay basic at zero
        A = np.vstack([A, A[1], 2.0 * A[1]]); b = np.append(b, [b[1], 2.0 * b[1]])
        rel = np.append(rel, [3, 3]).astype(np.int32); rel[1] = 3
    T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
    ref = O.solve_f64(T, basis, method=O.TWOPHASE, n_art=n_art, log_cap=8192)
    for window in (11, 1):
        got = _solve_p2p_world1(ctx, T, basis, O.TWOPHASE, n_art, variant=variant, window=window)
        _assert_equal(got, ref)
        assert got["phase1_iters"] == ref["phase1_iters"]


@pytest.mark.parametrize("method", [O.PLAIN, O.BIGM, O.TWOPHASE])
def test_sharded_p2p_world1_bland(ctx, method):
    """Bland's rule on the column-sharded paths (blocked and rank-1 pipeline)."""
    rng = np.random.default_rng(77)
    if method == O.PLAIN:
        T, basis = O.synth_feasible(1234, 5, 60, 44)
        n_art = 0
    else:
        A, b, c, rel = H.random_mixed_lp(rng, 24, 30)
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
    ref = O.solve_f64(T, basis, method=method, n_art=n_art, log_cap=8192, pricing=O.BLAND)
    variants = [capi.BLAND, capi.BLAND | (5 << 16)] + ([capi.BLAND | capi.RANK1] if method != O.TWOPHASE else [])
    for variant in variants:
        got = _solve_p2p_world1(ctx, T, basis, method, n_art, variant=variant)
        _assert_equal(got, ref)


def test_sharded_nccl_transport_world1(ctx):
    """ncclAllGather as the exchange (one rank): the NCCL code path on a single GPU."""
    import ctypes as C
    buf = C.create_string_buffer(capi.NCCL_ID_BYTES)
    assert capi.lib().lpgpu_shard_unique_id(buf) == capi.OK
    T, basis = O.synth_feasible(1234, 2, 64, 48)
    ref = O.solve_f64(T, basis, log_cap=4096)
    got = _solve_sharded(ctx, T, basis, 1, nccl_id=bytes(buf.raw))
    _assert_equal(got, ref)


def test_sharded_p2p_transport_world1(ctx):
    """Peer-memory transport with a single rank: flags, parity buffers and windows on one GPU."""
    T, basis = O.synth_feasible(1234, 2, 64, 48)
    ref = O.solve_f64(T, basis, log_cap=4096)
    s = lp.ShardedTableau(ctx, 64, 112, 1, world=1, rank=0, p2p=True)
    s.upload(T, basis)
    s.reset(log_cap=4096, check_every=5)
    while True:
        res = s.run(7)
        if res["status"] != lp.RUNNING:
            break
    out = s.download()
    s.close()
    res.update(T=out["T"], basis=out["basis"], log=out["log"][:res["n_logged"]], x=out["x"])
    _assert_equal(res, ref)


@pytest.mark.parametrize("variant", [0, 3 << 16, 0x800])
def test_sharded_p2p_world1_blocked_and_rank1_bigm(ctx, variant):
    """Peer-memory transport, one rank: the blocked path (default, blocks of 32 and 3) and the rank-1 pipeline on
    Big-M problems incl. infeasible ones; bit-exact against the oracle."""
    for seed in range(4):
        rng = np.random.default_rng(900 + seed)
        m, ns = int(rng.integers(4, 40)), int(rng.integers(4, 40))
        A, b, c, rel = H.random_mixed_lp(rng, m, ns)
        if seed == 3:
            A = np.vstack([A, A[0]]); b = np.append(b, b[0] - 5.0 if rel[0] != -2 else b[0] + 5.0)
            rel = np.append(rel, -2 if rel[0] != -2 else 2).astype(np.int32)
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
        ref = O.solve_f64(T, basis, method=O.BIGM, n_art=n_art, log_cap=4096)
        s = lp.ShardedTableau(ctx, T.shape[0] - 2, T.shape[1] - 1, 2, world=1, rank=0, p2p=True)
        s.upload(T, basis)
        s.reset(method=O.BIGM, n_art=n_art, log_cap=4096, variant=variant, check_every=4)
        while True:
            res = s.run(11)
            if res["status"] != lp.RUNNING:
                break
        out = s.download()
        s.close()
        res.update(T=out["T"], basis=out["basis"], log=out["log"][:res["n_logged"]], x=out["x"])
        _assert_equal(res, ref)
        assert res["zM"] == ref["zM"]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_real_peers(world):
    """One process per GPU (torchrun) at world = 2 / 4 / 8, NCCL and peer-memory transports, plain / Big-M / two-phase /
    Bland, every output bit-compared with the oracle inside the script (tests/sharded_parity.py)."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29631 + world), os.path.join(ROOT, "tests", "mp_sharded_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "sharded parity ok" in p.stdout
