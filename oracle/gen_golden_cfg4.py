#!/usr/bin/env python3
"""Golden record of the WHOLE config-4 solve (test infrastructure).

Runs the CPU oracle (oracle/lp_oracle.c, fp64 mode, the rule of DESIGN.md section 3) on the
BASELINE config-4 LP -- one dense tableau m = 8192, n = 16384, synthetic generator seed 1234,
lp index 0 -- from pivot 0 to OPTIMAL and commits what a `-m gpu` test compares the default
CUDA path with: status, iteration count, z (exact bits), the final basis, a SHA-256 of the
pivot log and, at a few check points along the way, the SHA-256 of the log prefix.

The reference has no pivot loop (Source/simplex.c:40-67), so this pins the GPU to THIS REPO's
oracle, not to the reference ("parity unpinned"); the optimum itself is
cross-checked against an independent solver at small sizes in tests/test_oracle.py.

    python oracle/gen_golden_cfg4.py [m] [threads]      (a few minutes of CPU at m = 8192)
"""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402

SEED, LP = 1234, 0


def log_sha(log):
    return hashlib.sha256(np.ascontiguousarray(log, dtype=np.int32).tobytes()).hexdigest()


def main():
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    threads = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    ns = m
    T, basis = O.synth_feasible(SEED, LP, m, ns)
    t0 = time.perf_counter()
    r = O.solve_f64(T, basis, threads=threads, log_cap=8 * (m + ns))
    dt = time.perf_counter() - t0
    assert r["iters"] == len(r["log"]), "pivot log truncated"
    marks = [k for k in (32, 96, 1000, 5000, 6000, 10000) if k <= r["iters"]]
    rec = {
        "generator": "oracle/gen_golden_cfg4.py",
        "workload": f"dense {m}x{m + ns} fp64, synthetic seed {SEED} lp {LP}, plain method, largest-coefficient pricing",
        "m": m, "n": m + ns, "seed": SEED, "lp": LP,
        "status": O.STATUS_NAMES[r["status"]],
        "iters": int(r["iters"]),
        "z": float(r["z"]),
        "z_hex": float(r["z"]).hex(),
        "basis_sha256": hashlib.sha256(r["basis"].astype(np.int32).tobytes()).hexdigest(),
        "basis": [int(v) for v in r["basis"]],
        "log_sha256": log_sha(r["log"]),
        "log_prefix_sha256": {str(k): log_sha(r["log"][:k]) for k in marks},
        "log_head": [[int(a), int(b)] for a, b in r["log"][:8]],
        "x_sha256": hashlib.sha256(np.ascontiguousarray(r["x"]).tobytes()).hexdigest(),
        "oracle_seconds": round(dt, 1), "oracle_threads": threads,
    }
    name = "cfg4_full_solve.json" if m == 8192 else f"cfg4_full_solve_m{m}.json"
    out = os.path.join(ROOT, "tests", "golden", name)
    with open(out, "w") as f:
        json.dump(rec, f, separators=(",", ":"))
        f.write("\n")
    print(f"{out}: {rec['status']} after {rec['iters']} pivots, z = {rec['z']!r}, {dt:.1f} s on {threads} threads")


if __name__ == "__main__":
    main()
