import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import helpers as H
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi
m, ns = 8192, 8192
rng = np.random.default_rng(77)
T, basis, _ = H.random_dual_lp(rng, m, ns)
ctx = lp.Context(0)
for name, variant in (("default(pipelined)", 0), ("inplace", capi.INPLACE)):
    d = lp.DenseTableau(ctx, m, T.shape[1] - 1, 1)
    d.upload(T, basis)
    d.reset(method=lp.DUAL, check_every=256, variant=variant)
    d.run(64)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = d.run(3000)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"{name:20s} {3000 / dt:10.1f} pivots/s status {r['status']}")
    d.close()
