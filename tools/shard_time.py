"""Column-sharded path, one 16384 x 65536 tableau over the ranks of a torchrun launch: us per pivot of the blocked path over the
peer-memory transport; with LPGPU_BLK_DEBUG=1 rank 0 also prints the mean time between the step kernel's phase stamps
(0->1 wait for every rank's candidate | 1->2 owner: column computed, stored, flagged | 2->3 wait for the column | 3->4 ratio
test | 4->5 row, objective rows, b | 5->6 candidate posted).
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/shard_time.py [m] [pivots]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import linearprogramming_b200 as lp
from linearprogramming_b200 import sharding

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
m = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
piv = int(sys.argv[2]) if len(sys.argv) > 2 else 640
ctx = lp.Context(int(os.environ.get("LOCAL_RANK", rank)))
sh = lp.ShardedTableau(ctx, m, 4 * m, 1, world=world, rank=rank, p2p=True)
sharding.attach_p2p(dist, sh)
sh.synth(1234, 0, 3 * m)
sh.reset(check_every=256)
sh.run(64)
torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
r = sh.run(piv)
torch.cuda.synchronize(); dt = torch.tensor([time.perf_counter() - t0], device="cuda")
dist.all_reduce(dt, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"world {world}: {piv / dt.item():.0f} pivots/s ({1e6 * dt.item() / piv:.2f} us per pivot), iters {r['iters']}", flush=True)
dist.barrier()
sh.close()
dist.barrier()
dist.destroy_process_group()
