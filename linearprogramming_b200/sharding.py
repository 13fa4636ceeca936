"""Host-side plumbing for the multi-GPU paths (one process per GPU, torch.distributed).

* batched path: LP instances are independent, so rank r solves the contiguous index range
  ``lp_range(n_lp, world, r)`` with ``lpgpu_solve_batched`` and there is no data-path collective;
  ``gather_batched`` only concatenates the per-rank result arrays on rank 0.
* column-sharded large tableau: ``column_block`` is the partition ``lpgpu_shard_create`` uses,
  ``exchange_nccl_id`` ships rank 0's NCCL id to the others over the process group (works on any
  backend, gloo included -- it is 128 bytes of host data).

The reference has no multi-device code at all; the partitioning follows the
tableau layout of Source/matrix.c:42-72 (columns are variables, rows are constraints).
"""
import ctypes as C

import numpy as np

from . import capi


def lp_range(n_lp, world, rank):
    """[lo, hi) of the LP indices rank solves: contiguous, sizes differ by at most one."""
    if world < 1 or not 0 <= rank < world or n_lp < 0:
        raise ValueError("bad (n_lp, world, rank)")
    base, extra = divmod(n_lp, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def column_block(n, world, rank):
    """(col0, ncols): global 1-based first column and width of rank's block (the C ABI's split)."""
    c0, nc = C.c_int(), C.c_int()
    rc = capi.lib().lpgpu_shard_columns(int(n), int(world), int(rank), C.byref(c0), C.byref(nc))
    if rc != capi.OK:
        raise ValueError("bad (n, world, rank)")
    return c0.value, nc.value


def owner_of_column(n, world, k):
    """Rank whose block holds global column k (1-based)."""
    for r in range(world):
        c0, nc = column_block(n, world, r)
        if c0 <= k < c0 + nc:
            return r
    raise ValueError("column out of range")


def exchange_nccl_id(dist, src=0):
    """Rank ``src`` creates the NCCL id (lpgpu_shard_unique_id), everybody returns its 128 bytes."""
    box = [None]
    if dist.get_rank() == src:
        buf = C.create_string_buffer(capi.NCCL_ID_BYTES)
        rc = capi.lib().lpgpu_shard_unique_id(buf)
        if rc != capi.OK:
            raise capi.LpgpuError(f"lpgpu_shard_unique_id failed with code {rc} (libnccl.so.2 not loadable?)")
        box[0] = bytes(buf.raw)
    dist.broadcast_object_list(box, src=src)
    return box[0]


def attach_p2p(dist, shard):
    """Peer-memory transport: all-gather the ranks' CUDA IPC handles over the process group and map them."""
    handles = [None] * dist.get_world_size()
    dist.all_gather_object(handles, shard.p2p_export())
    shard.p2p_attach(handles)
    dist.barrier()          # nobody starts posting before everybody has mapped


def gather_batched(dist, local, dst=0):
    """Concatenate per-rank result dicts (status, iters, z, x, basis) in rank order on ``dst``."""
    world = dist.get_world_size()
    parts = [None] * world if dist.get_rank() == dst else None
    dist.gather_object({k: v for k, v in local.items() if isinstance(v, np.ndarray)}, parts, dst=dst)
    if dist.get_rank() != dst:
        return None
    return {k: np.concatenate([p[k] for p in parts], axis=0) for k in parts[0]}


def merge_column_blocks(dist, T_local, n, dst=0):
    """Rank-local downloads (NaN outside the own block) -> the full tableau on ``dst``."""
    world = dist.get_world_size()
    parts = [None] * world if dist.get_rank() == dst else None
    dist.gather_object(T_local, parts, dst=dst)
    if dist.get_rank() != dst:
        return None
    T = parts[0].copy()
    for r in range(1, world):
        c0, nc = column_block(n, world, r)
        T[:, c0:c0 + nc] = parts[r][:, c0:c0 + nc]
    return T
