"""Pass and step kernels side by side on the headline tableau: python tools/pass_compare.py [m]
(ms per pass launch, us per pivot of the decision kernels in timed mode, untimed pivots/s)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import linearprogramming_b200 as lp
from linearprogramming_b200 import capi
m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
ctx = lp.Context(0)
d = lp.DenseTableau(ctx, m, 2 * m, 1)
for name, v in (("ring", 0), ("mma", capi.DMMA_PASS), ("dfma_sp", capi.DFMA_PASS)):
    for mode, mv in (("inplace", capi.INPLACE), ("pipe+launches", capi.PIPELINED), ("pipe+persistent", capi.PIPELINED | capi.STEP_PERSISTENT)):
        for K in ((32, 16) if not os.environ.get("PC_K32") else (32,)):
            try:
                d.synth(1234, 0, m)
                d.reset(check_every=320, variant=capi.BLOCK_K(K) | v | mv, time_kernels=1)
                d.run(64)
                r = d.run(320)
                d.reset(check_every=320, variant=capi.BLOCK_K(K) | v | mv)
                d.run(64); torch.cuda.synchronize(); t0 = time.perf_counter(); d.run(960); torch.cuda.synchronize(); dt = time.perf_counter() - t0
                print(f"{name:8s} {mode:15s} K {K:2d}: pass {r['update_ms'] / max(r['update_launches'], 1):.4f} ms, steps {r['select_ms'] / 320 * 1e3:.2f} us/pivot (timed mode), untimed {960 / dt:.0f} pivots/s", flush=True)
            except Exception as e:
                print(name, mode, K, "FAILED", e, flush=True)
d.close()
