"""Bit-exact checks of the column-sharded path against the CPU oracle, one process per GPU (test infrastructure).

Used by tests/mp_sharded_check.py (torchrun by hand or from tests/test_sharded_gpu.py) and by bench.py, which runs
`check_small` and `check_large_prefix` at world = N BEFORE it times anything at N > 1 and records the outcome in
`config.sharded_parity` -- the driver reaches more than one real GPU only through bench.py.
The oracle is the checker here, never the thing measured (oracle/lp_oracle.c header)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)

import oracle_lib as O  # noqa: E402


def _case(name):
    """(T, basis, method, n_art) of a named case; the same bits on every rank."""
    import helpers as H
    if name.startswith("synth"):
        _, m, ns = name.split(":")
        T, basis = O.synth_feasible(1234, 7, int(m), int(ns))
        return T, basis, O.PLAIN, 0
    kind, m, ns, seed = name.split(":")
    A, b, c, rel = H.random_mixed_lp(np.random.default_rng(int(seed)), int(m), int(ns))
    T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
    return T, basis, (O.BIGM if kind == "bigm" else O.TWOPHASE), n_art


def solve_sharded(lp, sharding, dist, ctx, rank, world, T, basis, method, n_art, transport="p2p", variant=0, window=0,
                  log_cap=1 << 16, max_pivots=2**31 - 1):
    R = 1 if method in (O.PLAIN, O.DUAL) else 2
    m, n = T.shape[0] - R, T.shape[1] - 1
    if transport == "nccl":
        s = lp.ShardedTableau(ctx, m, n, R, world=world, rank=rank, nccl_id=sharding.exchange_nccl_id(dist))
    else:
        s = lp.ShardedTableau(ctx, m, n, R, world=world, rank=rank, p2p=True)
        sharding.attach_p2p(dist, s)
    s.upload(T, basis)
    s.reset(method=method, n_art=n_art, log_cap=log_cap, variant=variant)
    if window > 0:
        done = 0
        while True:                                   # short run calls: flags keep counting across them
            res = s.run(min(window, max_pivots - done))
            done = res["iters"]
            if res["status"] != lp.RUNNING or done >= max_pivots:
                break
    else:
        res = s.run(max_pivots)
    out = s.download()
    dist.barrier()                                    # p2p: peers may still be storing flags into this rank's region
    s.close()
    return res, out


def _agree(dist, err):
    """Every rank learns every rank's verdict, so that all raise (or none does) and no collective is left half done."""
    errs = [None] * dist.get_world_size()
    dist.all_gather_object(errs, err)
    bad = [e for e in errs if e]
    if bad:
        raise AssertionError("; ".join(str(e) for e in bad))


def check_case(lp, sharding, dist, ctx, rank, world, name, transport="p2p", variant=0, window=0, threads=4):
    """One full solve, every output against the oracle.  All ranks raise AssertionError together (case in the message)."""
    T, basis, method, n_art = _case(name)
    n = T.shape[1] - 1
    pricing = O.BLAND if (variant & lp.capi.BLAND) else 0
    ref = O.solve_f64(T, basis, method=method, n_art=n_art, log_cap=1 << 16, threads=threads, pricing=pricing)
    res, out = solve_sharded(lp, sharding, dist, ctx, rank, world, T, basis, method, n_art, transport, variant, window)
    tag = f"{name} {transport} variant {variant:#x} world {world} rank {rank}"
    err = None
    c0, nc = sharding.column_block(n, world, rank)
    if res["status"] != ref["status"] or res["iters"] != ref["iters"]:
        err = f"{tag}: status/iters {res['status']}/{res['iters']} vs oracle {ref['status']}/{ref['iters']}"
    elif out["log"][:res["n_logged"]].tolist() != ref["log"].tolist():
        err = f"{tag}: pivot log differs"
    elif not (np.array_equal(out["basis"], ref["basis"]) and np.array_equal(out["x"], ref["x"]) and res["z"] == ref["z"]):
        err = f"{tag}: basis / x / z differ"
    elif not np.array_equal(out["T"][:, 0], ref["T"][:, 0]):
        err = f"{tag}: b column differs"
    elif not np.array_equal(out["T"][:, c0:c0 + nc], ref["T"][:, c0:c0 + nc]):
        err = f"{tag}: own column block differs"
    full = sharding.merge_column_blocks(dist, out["T"], n)
    if rank == 0 and err is None and not np.array_equal(full, ref["T"]):
        err = f"{tag}: merged tableau differs"
    _agree(dist, err)
    return ref["iters"]


SMALL_CASES = ("synth:300:260", "synth:1024:3072", "bigm:40:30:5")


def check_small(lp, sharding, dist, ctx, rank, world, transports=("p2p",), methods_extra=True, window=0):
    """The small sharded cases against the oracle at this world size.  Returns the number of (case, variant) runs."""
    runs = 0
    for transport in transports:
        cases = [(c, 0) for c in SMALL_CASES]
        if transport == "p2p":
            cases += [("synth:300:260", lp.capi.BLOCK_K(5)), ("synth:300:260", lp.capi.RANK1)]
            if methods_extra:
                cases += [("twophase:40:30:5", 0), ("twophase:60:40:11", 0), ("synth:300:260", lp.capi.BLAND), ("bigm:40:30:5", lp.capi.BLAND)]
        for name, variant in cases:
            check_case(lp, sharding, dist, ctx, rank, world, name, transport, variant, window=window if transport == "p2p" else 0)
            runs += 1
    return runs


def check_large_prefix(lp, sharding, dist, ctx, rank, world, m, n, pivots=64, seed=1234):
    """First `pivots` pivots of the column-sharded run on the (m x n) synthetic tableau == the single-tableau path's
    pivot log (rank 0 runs the single-GPU blocked path on the same LP; that path is pinned to the oracle by the
    -m gpu tests).  b column and z are compared as well."""
    ns = n - m
    s = lp.ShardedTableau(ctx, m, n, 1, world=world, rank=rank, p2p=True)
    sharding.attach_p2p(dist, s)
    s.synth(seed, 0, ns)
    s.reset(log_cap=pivots, check_every=256)
    res = s.run(pivots)
    out = s.download(want_T=False)
    dist.barrier()
    s.close()
    err = None
    if rank == 0:
        d = lp.DenseTableau(ctx, m, n, 1)
        d.synth(seed, 0, ns)
        d.reset(log_cap=pivots, check_every=256)
        r1 = d.run(pivots)
        o1 = d.download(want_T=False)
        d.close()
        if r1["iters"] != res["iters"] or o1["log"][:pivots].tolist() != out["log"][:pivots].tolist():
            err = f"pivot log of the {m}x{n} sharded run differs from the single-tableau path within {pivots} pivots"
        elif not (np.array_equal(o1["basis"], out["basis"]) and np.array_equal(o1["x"], out["x"]) and r1["z"] == res["z"]):
            err = f"basis / x / z of the {m}x{n} sharded run differ from the single-tableau path after {pivots} pivots"
    # every rank must agree on the replicated state
    it = [None] * world
    dist.all_gather_object(it, (res["iters"], res["status"], float(res["z"]).hex()))
    if err is None and not all(v == it[0] for v in it):
        err = f"ranks disagree on (iters, status, z): {it}"
    _agree(dist, err)
    return res["iters"]
