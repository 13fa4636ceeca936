"""A short blocked-path run on the headline tableau for an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi
m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, m, 2 * m, 1)
d.synth(1234, 0, m)
d.reset(check_every=128, variant=capi.NO_PDL | int(os.environ.get("BLK_VARIANT", "0"), 0))      # BLK_VARIANT: extra variant bits (e.g. 0x2100 = in place, DMMA pass)
r = d.run(96)
print(r["iters"], r["status"])
d.close()
