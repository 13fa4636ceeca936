"""A long run of the default path on the headline tableau (for clock / power sampling): python tools/pass_loop.py [pivots] [variant]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import linearprogramming_b200 as lp
piv = int(sys.argv[1]) if len(sys.argv) > 1 else 12000
variant = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, 8192, 16384, 1)
d.synth(1234, 0, 8192)
d.reset(check_every=1024, variant=variant)
d.run(64)
torch.cuda.synchronize(); t0 = time.perf_counter()
r = d.run(piv)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"variant {variant:#x}: {piv / dt:.1f} pivots/s over {dt:.2f} s, iters {r['iters']}")
d.close()
