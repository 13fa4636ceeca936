"""One run of the four-warp register kernel (m=64, n=128) for an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
ctx = lp.Context(0)
n_lp = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
b = lp.BatchedLPs(ctx, n_lp, 64, 128, 1)
b.synth(1234, 0, 64)
for _ in range(2):
    ms = b.run()
print("ms", ms)
b.close()
