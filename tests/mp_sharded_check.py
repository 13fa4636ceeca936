"""Run under torchrun, one process per GPU: the column-sharded path over NCCL, bit-compared with the
CPU oracle (tests/test_sharded_gpu.py::test_sharded_nccl_two_processes launches it; it can also be run
by hand: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/mp_sharded_check.py)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import oracle_lib as O  # noqa: E402

import linearprogramming_b200 as lp  # noqa: E402
from linearprogramming_b200 import sharding  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")                      # host plumbing only; the data path is our own NCCL comm
    ctx = lp.Context(local)
    window = int(os.environ.get("MP_WINDOW", "9"))
    # peer-memory transport: the blocked path (default; blocks of 32 and of 5 pivots) and the rank-1 pipeline (0x800)
    runs = [(t, 0) for t in os.environ.get("MP_TRANSPORTS", "nccl,p2p").split(",")]
    if any(t == "p2p" for t, _ in runs):
        runs += [("p2p", 5 << 16), ("p2p", 0x800)]
    for transport, variant in runs:
      for (m, ns, method) in [(64, 48, O.PLAIN), (300, 260, O.PLAIN), (40, 30, O.BIGM)]:
          if method == O.PLAIN:
              T, basis = O.synth_feasible(1234, 7, m, ns)
              n_art = 0
          else:
              import helpers as H
              A, b, c, rel = H.random_mixed_lp(np.random.default_rng(5), m, ns)
              T, basis, n_art = O.build_tableau(A, b, c, rel, R=2)
          R = 1 if method == O.PLAIN else 2
          n = T.shape[1] - 1
          ref = O.solve_f64(T, basis, method=method, n_art=n_art, log_cap=8192)
          if transport == "nccl":
              ident = sharding.exchange_nccl_id(dist)
              s = lp.ShardedTableau(ctx, m, n, R, world=world, rank=rank, nccl_id=ident)
          else:
              s = lp.ShardedTableau(ctx, m, n, R, world=world, rank=rank, p2p=True)
              sharding.attach_p2p(dist, s)
          s.upload(T, basis)
          s.reset(method=method, n_art=n_art, log_cap=8192, variant=variant)
          if transport == "p2p" and window > 0:
              while True:                                   # windows of 9 pivots: flags keep counting across run calls
                  res = s.run(window)
                  if res["status"] != lp.RUNNING:
                      break
          else:
              res = s.run(2**31 - 1)
          out = s.download()
          dist.barrier()                                   # p2p: peers may still be storing flags into this rank's region
          s.close()
          assert res["status"] == ref["status"] and res["iters"] == ref["iters"], (transport, m, res, ref["status"], ref["iters"])
          assert out["log"][:res["n_logged"]].tolist() == ref["log"].tolist(), (transport, m, out["log"][:res["n_logged"]].tolist(), ref["log"].tolist())
          assert np.array_equal(out["basis"], ref["basis"]) and np.array_equal(out["x"], ref["x"]) and res["z"] == ref["z"]
          c0, nc = sharding.column_block(n, world, rank)
          assert np.array_equal(out["T"][:, 0], ref["T"][:, 0])
          assert np.array_equal(out["T"][:, c0:c0 + nc], ref["T"][:, c0:c0 + nc])
          full = sharding.merge_column_blocks(dist, out["T"], n)
          if rank == 0:
              assert np.array_equal(full, ref["T"])
    ctx.close()
    dist.barrier()
    if rank == 0:
        print(f"sharded nccl ok: world {world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
