"""CPU tests of the multi-GPU host logic (gloo, world_size 2 and 3; no GPU work).

What runs on CPU here is the plumbing of linearprogramming_b200/sharding.py and the exchange PROTOCOL of
the column-sharded path (csrc/shard.cu), emulated with numpy + the oracle's pivot on each rank's column
block: candidates all-gathered, arg-max with ties to the lowest global column, entering column taken from
its owner, ratio test redundant on every rank.  The emulation lives in tests/ only; the product has no
CPU path.  The same partition and tie-break run on the GPU in tests/test_sharded_gpu.py."""
import os
import socket
import sys

import numpy as np
import pytest

import oracle_lib as O

from linearprogramming_b200 import sharding as S

EPS = 1e-9


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_entry, args=(fn, world, port, args), nprocs=world, join=True)


def _entry(rank, fn, world, port, args):
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.dirname(here)):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(dist, rank, world, *args)
    finally:
        dist.destroy_process_group()


# ------------------------------------------------------------------ partitions (single process)
@pytest.mark.parametrize("n,world", [(10, 1), (10, 3), (16384, 8), (65536, 8), (7, 7), (129, 4)])
def test_column_blocks_tile_the_columns(n, world):
    nxt = 1
    widths = []
    for r in range(world):
        c0, nc = S.column_block(n, world, r)
        assert c0 == nxt and nc >= 1
        nxt += nc
        widths.append(nc)
    assert nxt == n + 1 and max(widths) - min(widths) <= 1
    for k in (1, n, (n + 1) // 2):
        r = S.owner_of_column(n, world, k)
        c0, nc = S.column_block(n, world, r)
        assert c0 <= k < c0 + nc


@pytest.mark.parametrize("n_lp,world", [(0, 2), (5, 8), (100000, 8), (1000000, 4), (17, 3)])
def test_lp_ranges_tile_the_instances(n_lp, world):
    nxt = 0
    for r in range(world):
        lo, hi = S.lp_range(n_lp, world, r)
        assert lo == nxt and hi >= lo
        nxt = hi
    assert nxt == n_lp
    sizes = [S.lp_range(n_lp, world, r)[1] - S.lp_range(n_lp, world, r)[0] for r in range(world)]
    assert max(sizes) - min(sizes) <= 1


def test_bad_partitions_are_rejected():
    with pytest.raises(ValueError):
        S.column_block(4, 5, 0)
    with pytest.raises(ValueError):
        S.column_block(4, 2, 2)
    with pytest.raises(ValueError):
        S.lp_range(4, 2, 2)


# ------------------------------------------------------------------ column-sharded protocol over gloo
def _better(a, b):
    """(p, q, idx) candidates: maximise (p, q), ties to the lowest index (common.cuh cand_better)."""
    if a[2] < 0:
        return False
    if b[2] < 0:
        return True
    if a[0] != b[0]:
        return a[0] > b[0]
    if a[1] != b[1]:
        return a[1] > b[1]
    return a[2] < b[2]


def _local_candidate(obj, objM, c0, method):
    best = (0.0, 0.0, -1)
    for j in range(len(obj)):
        q = obj[j]
        if method == O.BIGM:
            p = objM[j]
            if abs(p) <= EPS:
                p = 0.0
            ok = p > 0.0 or (p == 0.0 and q > EPS)
        else:
            p, ok = 0.0, q > EPS
        t = (p, q, c0 + j)
        if ok and _better(t, best):
            best = t
    return best


def _sharded_solve_emulated(dist, rank, world, T, basis, method, n_art, max_iter=10_000):
    R = 1 if method == O.PLAIN else 2
    m, n = T.shape[0] - R, T.shape[1] - 1
    c0, nc = S.column_block(n, world, rank)
    Tl = np.ascontiguousarray(np.concatenate([T[:, :1], T[:, c0:c0 + nc], np.zeros((m + R, 1))], axis=1))
    basis = basis.copy()
    log = []
    while True:
        cand = _local_candidate(Tl[m, 1:1 + nc], Tl[m + 1, 1:1 + nc] if R == 2 else None, c0, method)
        pairs = [None] * world
        dist.all_gather_object(pairs, cand)                       # exchange A
        best, owner = (0.0, 0.0, -1), 0
        for g, t in enumerate(pairs):
            if _better(t, best):
                best, owner = t, g
        k = best[2]
        if k < 0:
            nreal = n - n_art
            bad = method == O.BIGM and any(basis[i] > nreal and Tl[i, 0] > EPS for i in range(m))
            return (O.INFEASIBLE if bad else O.OPTIMAL), log, Tl[:, :1 + nc], basis
        assert owner == S.owner_of_column(n, world, k)
        box = [Tl[:, 1 + k - c0].copy() if owner == rank else None]
        dist.broadcast_object_list(box, src=owner)                # exchange B
        col = box[0]
        r, theta = -1, None
        for i in range(m):
            if col[i] > EPS:
                t = Tl[i, 0] / col[i]
                if theta is None or t < theta:
                    r, theta = i, t
        if r < 0:
            return O.UNBOUNDED, log, Tl[:, :1 + nc], basis
        if len(log) >= max_iter:
            return O.ITER_LIMIT, log, Tl[:, :1 + nc], basis
        Tl[:, 1 + nc] = col                                       # the oracle's pivot on [b | block | column k]
        Tl = O.pivot(Tl, r, 1 + nc)
        basis[r] = k
        log.append((k, r))


def _check_sharded(dist, rank, world, case):
    import helpers as H
    if case == "synth":
        T, basis = O.synth_feasible(1234, 3, 24, 20)
        method, n_art = O.PLAIN, 0
    else:
        rng = np.random.default_rng(7)
        A, b, c, rel = H.random_mixed_lp(rng, 10, 8)
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
        method = O.BIGM
    ref = O.solve_f64(T, basis, method=method, n_art=n_art, log_cap=4096)
    status, log, Tl, bs = _sharded_solve_emulated(dist, rank, world, T, basis, method, n_art)
    n = T.shape[1] - 1
    c0, nc = S.column_block(n, world, rank)
    assert status == ref["status"]
    assert [list(p) for p in log] == ref["log"].tolist()
    assert np.array_equal(bs, ref["basis"])
    assert np.array_equal(Tl[:, 0], ref["T"][:, 0]) and np.array_equal(Tl[:, 1:], ref["T"][:, c0:c0 + nc])
    # block merge on rank 0 reproduces the single-process tableau
    full = np.full(T.shape, np.nan)
    full[:, 0] = Tl[:, 0]
    full[:, c0:c0 + nc] = Tl[:, 1:]
    merged = S.merge_column_blocks(dist, full, n)
    if rank == 0:
        assert np.array_equal(merged, ref["T"])


@pytest.mark.parametrize("world,case", [(2, "synth"), (3, "synth"), (2, "bigm")])
def test_column_sharded_protocol_matches_single_process_oracle(world, case):
    _spawn(_check_sharded, world, case)


# ------------------------------------------------------------------ blocked column-sharded protocol over gloo
def _fma(a, b, c):
    """round(a*b + c) with ONE rounding, as the device's __fma_rn (exact rational arithmetic; tiny sizes only)."""
    from fractions import Fraction
    return float(Fraction(a) * Fraction(b) + Fraction(c))


def _sharded_blocked_emulated(dist, rank, world, T, basis, method, n_art, K, max_iter=10_000):
    """shard_blk_step_kernel + blk_flush_kernel, restated: up to K pivots are decided on lazily updated slices
    (the owner's entering column through the block's earlier pivots, every rank's slice of the pivot row), the
    objective rows and b are kept eagerly, and the block is applied to the local column block in one pass."""
    R = 1 if method == O.PLAIN else 2
    m, n = T.shape[0] - R, T.shape[1] - 1
    c0, nc = S.column_block(n, world, rank)
    Tl = np.ascontiguousarray(np.concatenate([T[:, :1], T[:, c0:c0 + nc]], axis=1))     # [b | block]
    obj, bcur = Tl[m:, :].copy(), Tl[:m, 0].copy()
    basis = basis.copy()
    F, P, start, log = [], [], {}, []

    def lazy(i, j, upto):
        s0 = start.get(i, 0)
        a = P[s0 - 1][j] if s0 else Tl[i, j]
        for s in range(s0, upto):
            a = _fma(-F[s][i], P[s][j], a)
        return a

    def flush():
        nonlocal F, P, start
        t = len(F)
        if t:
            new = np.array([[lazy(i, j, t) for j in range(1 + nc)] for i in range(m + R)])
            assert np.array_equal(new[m:, :], obj) and np.array_equal(new[:m, 0], bcur)   # eager copies == the chain
            Tl[...] = new
        F, P, start = [], [], {}

    while True:
        cand = _local_candidate(obj[0, 1:1 + nc], obj[1, 1:1 + nc] if R == 2 else None, c0, method)
        pairs = [None] * world
        dist.all_gather_object(pairs, cand)                       # candidates of this pivot
        best, owner = (0.0, 0.0, -1), 0
        for g, c in enumerate(pairs):
            if _better(c, best):
                best, owner = c, g
        k, t = best[2], len(F)
        status = None
        if k < 0:
            nreal = n - n_art
            bad = method == O.BIGM and any(basis[i] > nreal and bcur[i] > EPS for i in range(m))
            status = O.INFEASIBLE if bad else O.OPTIMAL
        else:
            box = [None]
            if owner == rank:                                     # the owner's column, brought up to date lazily
                kl = 1 + k - c0
                box[0] = np.array([lazy(i, kl, t) for i in range(m)] + [obj[q, kl] for q in range(R)])
            dist.broadcast_object_list(box, src=owner)
            col = box[0]
            r, theta = -1, None
            for i in range(m):
                if col[i] > EPS:
                    q = bcur[i] / col[i]
                    if theta is None or q < theta:
                        r, theta = i, q
            if r < 0:
                status = O.UNBOUNDED
            elif len(log) >= max_iter:
                status = O.ITER_LIMIT
        if status is not None:
            flush()
            return status, log, Tl, basis
        row = np.array([lazy(r, j, t) for j in range(1 + nc)])
        pr = np.where(row == 0.0, row, row / col[r])              # div_skip_zero
        for q in range(R):
            obj[q, :] = [_fma(-col[m + q], pr[j], obj[q, j]) for j in range(1 + nc)]
        bcur[:] = [pr[0] if i == r else _fma(-col[i], pr[0], bcur[i]) for i in range(m)]
        F.append(col); P.append(pr)
        start[r] = t + 1
        basis[r] = k
        log.append((k, r))
        if len(F) >= K:
            flush()


def _check_sharded_blocked(dist, rank, world, case, K):
    import helpers as H
    if case == "synth":
        T, basis = O.synth_feasible(1234, 3, 16, 12)
        method, n_art = O.PLAIN, 0
    else:
        rng = np.random.default_rng(7)
        A, b, c, rel = H.random_mixed_lp(rng, 10, 8)
        T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
        method = O.BIGM
    ref = O.solve_f64(T, basis, method=method, n_art=n_art, log_cap=4096)
    status, log, Tl, bs = _sharded_blocked_emulated(dist, rank, world, T, basis, method, n_art, K)
    n = T.shape[1] - 1
    c0, nc = S.column_block(n, world, rank)
    assert status == ref["status"]
    assert [list(p) for p in log] == ref["log"].tolist()
    assert np.array_equal(bs, ref["basis"])
    assert np.array_equal(Tl[:, 0], ref["T"][:, 0]) and np.array_equal(Tl[:, 1:], ref["T"][:, c0:c0 + nc])


@pytest.mark.parametrize("world,case,K", [(2, "synth", 32), (2, "synth", 3), (3, "synth", 5), (2, "bigm", 4)])
def test_blocked_column_sharded_protocol_matches_oracle(world, case, K):
    _spawn(_check_sharded_blocked, world, case, K)


# ------------------------------------------------------------------ batched sharding + id exchange over gloo
def _check_batched(dist, rank, world):
    n_lp, m, ns = 37, 8, 8
    lo, hi = S.lp_range(n_lp, world, rank)
    Tb, bb = O.synth_feasible_batch(1234, lo, hi - lo, m, ns)        # LP index = global index
    mine = O.solve_batched_f64(Tb, bb)                               # stand-in for lpgpu_solve_batched
    out = S.gather_batched(dist, dict(status=mine["status"], iters=mine["iters"], z=mine["z"], x=mine["x"]))
    ident = S.exchange_nccl_id(dist)                                 # 128 bytes of host data over gloo
    assert isinstance(ident, bytes) and len(ident) == 128
    ids = [None] * world
    dist.all_gather_object(ids, ident)
    assert all(i == ids[0] for i in ids)
    if rank == 0:
        Ta, ba = O.synth_feasible_batch(1234, 0, n_lp, m, ns)
        ref = O.solve_batched_f64(Ta, ba)
        for k in ("status", "iters", "z", "x"):
            assert np.array_equal(out[k], ref[k]), k
    else:
        assert out is None


def test_batched_ranges_and_nccl_id_exchange_world2():
    _spawn(_check_batched, 2)
