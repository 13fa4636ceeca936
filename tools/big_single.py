import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import linearprogramming_b200 as lp
ctx = lp.Context(0)
m, n = 16384, 65536
d = lp.DenseTableau(ctx, m, n, 1)
d.synth(1234, 0, n - m)
for variant in (0, 0x800):
    d.synth(1234, 0, n - m)
    d.reset(check_every=256, variant=variant, time_kernels=0)
    d.run(64)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    P = 256 if variant == 0 else 64
    r = d.run(P)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(hex(variant), P / dt, "pivots/s", dt / P * 1e6, "us/pivot")
    d.reset(check_every=256, variant=variant, time_kernels=1)
    r = d.run(64)
    print("   pass ms", r["update_ms"] / r["update_launches"], "step us", r["select_ms"] / 64 * 1e3)
d.close()
