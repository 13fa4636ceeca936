"""Scratch: solve the config-4 LP to optimality on the GPU and report iterations / time."""
import sys, time, json
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
m = ns = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
variant = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, m, m + ns, 1)
d.synth(1234, 0, ns)
d.reset(variant=variant, check_every=256)
t0 = time.perf_counter()
r = d.run(2**31 - 1)
dt = time.perf_counter() - t0
print(json.dumps({"m": m, "n": m + ns, "variant": variant, "status": lp.STATUS_NAMES[r["status"]], "iters": r["iters"], "seconds": dt,
                  "pivots_per_sec": r["iters"] / dt, "z": r["z"]}))
