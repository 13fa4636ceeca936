"""LPGPU_BLK_DEBUG=1 python tools/blocked_debug.py [m] [K] [extra variant bits] : phase timing of the blocked path's step kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi
m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K = int(sys.argv[2]) if len(sys.argv) > 2 else 32
extra = int(sys.argv[3], 0) if len(sys.argv) > 3 else 0
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, m, 2 * m, 1)
d.synth(1234, 0, m)
d.reset(check_every=640, variant=(K << 16) | extra)
r = d.run(640)
print(r["iters"], r["status"])
d.close()
