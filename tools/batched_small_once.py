"""One run of the one-warp register kernel (m=32, n=64) for an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
ctx = lp.Context(0)
n_lp = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
b = lp.BatchedLPs(ctx, n_lp, 32, 64, 1)
b.synth(1234, 0, 32)
for _ in range(2):
    ms = b.run()
out = b.download(want_x=False)
print("ms", ms, "pivots", int(out["iters"].sum()))
b.close()
