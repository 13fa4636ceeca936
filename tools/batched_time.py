import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linearprogramming_b200 as lp
ctx=lp.Context(0)
for (n_lp,m,n) in [(100000,32,64),(200000,64,128)]:
    b=lp.BatchedLPs(ctx,n_lp,m,n,1); b.synth(1234,0,n-m)
    for fs, var in ((False, 0), (True, 0)):
        if var and m > 32: continue
        for _ in range(2): b.run(force_smem=fs, variant=var)
        ms=[b.run(force_smem=fs, variant=var) for _ in range(5)]
        out=b.download(want_x=False)
        print(n_lp,m,n,"smem" if fs else ("reg2w" if var else "reg"), sorted(ms)[2], "ms", n_lp/sorted(ms)[2]*1e3, "LPs/s", out["iters"].mean(), (out["status"]==0).all())
    b.close()
