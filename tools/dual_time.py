"""Dual simplex on a mid-size dual-feasible LP: blocked path with the 16 x 512 and the 8 x 1024 step cluster, rank-1 path.
usage: python tools/dual_time.py [m] [ns]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
import helpers as H
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi
m = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
ns = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
rng = np.random.default_rng(77)
T, basis, _ = H.random_dual_lp(rng, m, ns)
ctx = lp.Context(0)
ref = None
for name, variant in (("cl16", 0), ("pipelined", capi.PIPELINED), ("cl8", capi.STEP_CL8), ("rank1", capi.RANK1)):
    d = lp.DenseTableau(ctx, m, T.shape[1] - 1, 1)
    d.upload(T, basis)
    d.reset(method=lp.DUAL, log_cap=1 << 16, check_every=256, variant=variant)
    d.run(32)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = d.run(1 << 20)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    log = d.download(want_T=False)["log"][:r["iters"]]
    if ref is None:
        ref = log
    print(f"{name:6s} {(r['iters'] - 32) / dt:10.1f} pivots/s  iters {r['iters']} status {r['status']}  log equal: {bool(np.array_equal(log, ref))}")
    d.close()
