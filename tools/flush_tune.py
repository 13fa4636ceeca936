import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
m = 8192
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, m, 2 * m, 1)
for K in (32, 24):
    d.synth(1234, 0, m)
    d.reset(check_every=320, variant=(K << 16) | int(os.environ.get("V", "0"), 0), time_kernels=1)
    d.run(32)
    r = d.run(320)
    print(os.environ.get("V"), os.environ.get("LPGPU_FL_RB"), os.environ.get("LPGPU_FL_PF"), "K", K, "pass ms", round(r["update_ms"] / r["update_launches"], 4), "step us", round(r["select_ms"] / 320 * 1e3, 2))
d.close()
