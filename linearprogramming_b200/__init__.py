"""B200-native dense-tableau simplex engine behind the C ABI of include/lpgpu.h.

Only what the pivot path needs lives here: csrc/ (sm_100a CUDA kernels + the C ABI),
host/ (C host side that keeps the reference's simplex.h entry points) and this thin
Python mirror used by the tests and bench.py.  No CPU fallback.
"""
from .capi import (BIGM, DUAL, INFEASIBLE, ITER_LIMIT, OPTIMAL, PLAIN, RUNNING, STATUS_NAMES, TWOPHASE, UNBOUNDED,
                   LpgpuError)
from .simplex import (BatchedLPs, Context, DenseTableau, ShardedTableau, build_tableau, solve_batched, solve_dense, synth_host)

__all__ = ["Context", "DenseTableau", "BatchedLPs", "ShardedTableau", "solve_dense", "solve_batched", "build_tableau", "synth_host",
           "PLAIN", "BIGM", "TWOPHASE", "DUAL", "OPTIMAL", "UNBOUNDED", "INFEASIBLE", "ITER_LIMIT", "RUNNING",
           "STATUS_NAMES", "LpgpuError"]
