"""Run under torchrun, one process per GPU: the column-sharded path over NCCL and over the peer-memory transport,
bit-compared with the CPU oracle at world = WORLD_SIZE (2, 4, 8; tests/test_sharded_gpu.py launches it where the GPUs
exist; by hand: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/mp_sharded_check.py).
The checks themselves live in tests/sharded_parity.py, which bench.py runs too before it times anything at N > 1."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import sharded_parity as SP  # noqa: E402

import linearprogramming_b200 as lp  # noqa: E402
from linearprogramming_b200 import sharding  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")                      # host plumbing only; the data path is our own exchange
    ctx = lp.Context(local)
    window = int(os.environ.get("MP_WINDOW", "9"))       # p2p: run calls of 9 pivots (flags keep counting across them)
    transports = tuple(os.environ.get("MP_TRANSPORTS", "nccl,p2p").split(","))
    runs = SP.check_small(lp, sharding, dist, ctx, rank, world, transports=transports, window=window)
    if os.environ.get("MP_LARGE"):
        lm = int(os.environ["MP_LARGE"])
        SP.check_large_prefix(lp, sharding, dist, ctx, rank, world, lm, 4 * lm, pivots=64)
    ctx.close()
    dist.barrier()
    if rank == 0:
        print(f"sharded parity ok: world {world}, {runs} runs against the oracle")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
