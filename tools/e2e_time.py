"""Where the time of one lpgpu_solve_dense() call with host buffers goes (headline tableau)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi
m = 8192
ctx = lp.Context(0)
Th = ctx.pinned((m + 1, 2 * m + 1)); bh = ctx.pinned((m,), np.int32)
T0, b0 = lp.synth_host(1234, 0, m, m)
Th[...] = T0; bh[...] = b0
for variant in (0, capi.RANK1):
    for P in (0, 1, 32, 1000):
        ts = []
        for rep in range(3):
            bh[...] = b0
            torch.cuda.synchronize(); t0 = time.perf_counter()
            r = lp.solve_dense(ctx, Th, bh, max_iter=P, results_only=True, inplace=True, variant=variant)
            torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
        print(f"variant {variant:#x} P={P:5d}: {min(ts) * 1e3:8.2f} ms  iters {r['iters']}")
