"""Rank-1 path vs blocked path on the headline tableau: pivots/s (device events around the whole run)
and per-kernel times.  usage: python tools/blocked_time.py [m] [pivots]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi

m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
piv = int(sys.argv[2]) if len(sys.argv) > 2 else 640
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, m, 2 * m, 1)
ref_log = None
names = (("rank1", capi.RANK1), ("blocked32", 0), ("blocked24", 24 << 16), ("blocked16", 16 << 16), ("blocked8", 8 << 16),
         ("inplace32", capi.INPLACE), ("inplace24", (24 << 16) | capi.INPLACE))
if os.environ.get("BLK_ONLY"):
    names = tuple(x for x in names if x[0] in os.environ["BLK_ONLY"].split(","))
if os.environ.get("BLK_VARIANTS"):      # name=variant,name=variant (hex allowed), always after the rank-1 reference run
    names = (("rank1", capi.RANK1),) + tuple((kv.split("=")[0], int(kv.split("=")[1], 0)) for kv in os.environ["BLK_VARIANTS"].split(","))
for name, variant in names:
    for timed in (0, 1):
        d.synth(1234, 0, m)
        d.reset(log_cap=piv, check_every=max(64, piv), variant=variant, time_kernels=timed)
        d.run(32)                                    # warm-up window
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = d.run(piv - 32)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if timed:
            print(f"{name:10s} timed: step/select {r['select_ms'] / (piv - 32) * 1e3:8.2f} us/pivot, pass {r['update_ms'] / max(1, r['update_launches']):8.4f} ms x {r['update_launches']}")
        else:
            done = max(1, r['iters'] - 32)            # the LP may reach its optimum before the window ends
            print(f"{name:10s} {done / dt:10.1f} pivots/s  ({dt / done * 1e6:7.2f} us/pivot)  iters {r['iters']}")
    log = d.download(want_T=False)["log"] if hasattr(d, "download") else None
    if ref_log is None:
        ref_log = log
    else:
        print("   log equal to rank-1:", bool((log == ref_log).all()))
d.close()
