"""Host-side mirror of the simplex entry points over the C ABI (include/lpgpu.h).

The reference's solve entry point is ``newSimplex(LPModel*)`` (Source/simplex.h:13), which
builds a SimplexMatrix and stops (Source/simplex.c:40,67).  The classes below drive what
follows that point -- the pivot loop -- on one B200 through liblpgpu.so.  Arrays are numpy
(host) buffers; device memory is owned by the opaque handles of the library.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (BIGM, DUAL, INFEASIBLE, ITER_LIMIT, OPTIMAL, PLAIN, RUNNING, STATUS_NAMES, TWOPHASE, UNBOUNDED,  # noqa: F401
                   LpgpuError, Opts, Result)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _R(method):
    return 1 if method in (PLAIN, DUAL) else 2


class Context:
    """One per process and GPU.  ``stream``: a raw cudaStream_t (int) to launch on, e.g.
    ``torch.cuda.current_stream().cuda_stream`` so that torch CUDA events bracket the kernels."""

    def __init__(self, device=0, stream=None):
        self._h = C.c_void_p()
        rc = capi.lib().lpgpu_create(C.byref(self._h), int(device), C.c_void_p(stream) if stream else None)
        if rc != capi.OK:
            self._h = C.c_void_p()
            raise LpgpuError(f"lpgpu_create(device={device}) failed with code {rc}: "
                             "no sm_100 CUDA device (there is no CPU fallback)")
        self.device = device

    @property
    def handle(self):
        return self._h

    def info(self):
        sm, cc1, cc2, mem = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
        capi.check(self._h, capi.lib().lpgpu_device_info(self._h, C.byref(sm), C.byref(mem), C.byref(cc1), C.byref(cc2)),
                   "lpgpu_device_info")
        return dict(sm_count=sm.value, hbm_bytes=mem.value, cc=(cc1.value, cc2.value))

    def launch_count(self):
        return int(capi.lib().lpgpu_launch_count(self._h))

    def sync(self):
        capi.check(self._h, capi.lib().lpgpu_sync(self._h), "lpgpu_sync")

    def pinned(self, shape, dtype=np.float64):
        """numpy array over pinned host memory (freed with the context)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = capi.lib().lpgpu_host_alloc(self._h, max(n, 16))
        if not p:
            raise LpgpuError("lpgpu_host_alloc failed")
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        buf = (C.c_char * max(n, 16)).from_address(p)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def close(self):
        if self._h:
            for p in getattr(self, "_pinned", []):
                capi.lib().lpgpu_host_free(self._h, p)
            self._pinned = []
            capi.lib().lpgpu_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _result_dict(res):
    return dict(status=res.status, iters=res.iters, phase1_iters=res.phase1_iters, phase=res.phase, z=res.z,
                zM=res.zM, n_logged=res.n_logged, update_launches=res.update_launches,
                update_ms=res.update_ms, select_ms=res.select_ms)


def solve_dense(ctx, T, basis, method=PLAIN, n_art=0, eps=1e-9, max_iter=2**31 - 1, log_cap=0, check_every=0,
                variant=0, results_only=False, inplace=False):
    """One-shot solve of a canonical tableau (host arrays in, host arrays out).

    T: (m+R) x (n+1) float64, updated on a copy; returns dict(status, iters, z, zM, T, basis, log, x, ...).
    """
    if not inplace:      # inplace: T / basis are caller-owned C-contiguous buffers (e.g. pinned)
        T = np.array(T, dtype=np.float64, order="C", copy=True)
        basis = np.array(basis, dtype=np.int32, copy=True)
    R = _R(method)
    m, n = T.shape[0] - R, T.shape[1] - 1
    if basis.shape != (m,):
        raise ValueError("basis must have one entry per constraint row")
    o = capi.default_opts(eps=eps, max_iter=int(min(max_iter, 2**31 - 1)), method=method, n_art=n_art,
                          log_cap=log_cap, check_every=check_every, variant=variant,
                          results_only=1 if results_only else 0)
    log = np.full((max(log_cap, 1), 2), -1, np.int32)
    x = np.zeros(n)
    res = Result()
    rc = capi.lib().lpgpu_solve_dense(ctx.handle, m, n, R, _ptr(T), T.shape[1], _ptr(basis), C.byref(o),
                                      _ptr(log) if log_cap else None, _ptr(x), C.byref(res))
    capi.check(ctx.handle, rc, "lpgpu_solve_dense")
    out = _result_dict(res)
    out.update(T=T, basis=basis, log=log[:res.n_logged].copy() if log_cap else log[:0], x=x)
    return out


class DenseTableau:
    """Device-resident large tableau (lpgpu_dense_*): upload or synthesise once, pivot in windows."""

    def __init__(self, ctx, m, n, R=1):
        self.ctx, self.m, self.n, self.R = ctx, m, n, R
        self._h = C.c_void_p()
        capi.check(ctx.handle, capi.lib().lpgpu_dense_create(ctx.handle, m, n, R, C.byref(self._h)), "lpgpu_dense_create")

    @property
    def ld(self):
        return int(capi.lib().lpgpu_dense_ld(self._h))

    def upload(self, T, basis):
        T = np.ascontiguousarray(T, np.float64)
        basis = np.ascontiguousarray(basis, np.int32)
        if T.shape != (self.m + self.R, self.n + 1) or basis.shape != (self.m,):
            raise ValueError("shape mismatch")
        capi.check(self.ctx.handle, capi.lib().lpgpu_dense_upload(self._h, _ptr(T), T.shape[1], _ptr(basis)),
                   "lpgpu_dense_upload")

    def upload_async(self, T, basis):
        """upload() without blocking the host: T and basis (pinned: Context.pinned) must stay untouched until the next
        call on this object returns; the copy overlaps whatever runs on OTHER tableau objects meanwhile."""
        if T.dtype != np.float64 or basis.dtype != np.int32 or not T.flags.c_contiguous or not basis.flags.c_contiguous:
            raise ValueError("upload_async needs C-contiguous float64 / int32 arrays (no hidden copies)")
        if T.shape != (self.m + self.R, self.n + 1) or basis.shape != (self.m,):
            raise ValueError("shape mismatch")
        self._async_refs = (T, basis)           # keep the buffers alive while the copy is in flight
        capi.check(self.ctx.handle, capi.lib().lpgpu_dense_upload_async(self._h, _ptr(T), T.shape[1], _ptr(basis)),
                   "lpgpu_dense_upload_async")

    def synth(self, seed, lp_index, ns):
        capi.check(self.ctx.handle, capi.lib().lpgpu_dense_synth(self._h, seed, lp_index, ns), "lpgpu_dense_synth")

    def reset(self, method=None, n_art=0, eps=1e-9, max_iter=2**31 - 1, log_cap=0, check_every=0, variant=0,
              time_kernels=0):
        if method is None:
            method = PLAIN if self.R == 1 else BIGM
        o = capi.default_opts(eps=eps, max_iter=int(min(max_iter, 2**31 - 1)), method=method, n_art=n_art,
                              log_cap=log_cap, check_every=check_every, variant=variant, time_kernels=time_kernels)
        self._log_cap = log_cap
        capi.check(self.ctx.handle, capi.lib().lpgpu_dense_reset(self._h, C.byref(o)), "lpgpu_dense_reset")

    def run(self, max_pivots):
        res = Result()
        capi.check(self.ctx.handle, capi.lib().lpgpu_dense_run(self._h, int(max_pivots), C.byref(res)), "lpgpu_dense_run")
        return _result_dict(res)

    def pivot(self, r, k):
        capi.check(self.ctx.handle, capi.lib().lpgpu_dense_pivot(self._h, int(r), int(k)), "lpgpu_dense_pivot")

    def download(self, want_T=True, want_log=True, want_x=True):
        T = np.zeros((self.m + self.R, self.n + 1)) if want_T else None
        basis = np.zeros(self.m, np.int32)
        cap = getattr(self, "_log_cap", 0)
        log = np.full((max(cap, 1), 2), -1, np.int32) if (want_log and cap) else None
        x = np.zeros(self.n) if want_x else None
        capi.check(self.ctx.handle,
                   capi.lib().lpgpu_dense_download(self._h, _ptr(T), self.n + 1, _ptr(basis), _ptr(log), _ptr(x)),
                   "lpgpu_dense_download")
        return dict(T=T, basis=basis, log=log, x=x)

    def close(self):
        if self._h:
            capi.lib().lpgpu_dense_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedTableau:
    """One large tableau column-sharded over several GPUs (lpgpu_shard_*).

    nccl_id = 128 bytes from ``sharding.exchange_nccl_id`` : this process is rank ``rank`` of ``world``
    (one process per GPU, exchanges over NCCL).  nccl_id = None : all ``world`` ranks live in this
    process on the context's device (single-GPU vehicle for the N > 1 logic).  p2p = True: one rank per
    process, exchanges are stores into peer GPU memory (``sharding.attach_p2p`` maps the peers)."""

    def __init__(self, ctx, m, n, R=1, world=1, rank=0, nccl_id=None, p2p=False):
        self.ctx, self.m, self.n, self.R, self.world, self.rank = ctx, m, n, R, world, rank
        self._h = C.c_void_p()
        self._log_cap = 0
        if p2p:
            capi.check(ctx.handle, capi.lib().lpgpu_shard_create_p2p(ctx.handle, rank, world, m, n, R, C.byref(self._h)),
                       "lpgpu_shard_create_p2p")
            return
        if nccl_id is not None and len(nccl_id) != capi.NCCL_ID_BYTES:
            raise ValueError("nccl_id must be 128 bytes")
        idbuf = C.create_string_buffer(bytes(nccl_id), capi.NCCL_ID_BYTES) if nccl_id is not None else None
        capi.check(ctx.handle, capi.lib().lpgpu_shard_create(ctx.handle, rank, world, idbuf, m, n, R, C.byref(self._h)),
                   "lpgpu_shard_create")

    def p2p_export(self):
        """This rank's 64-byte CUDA IPC handle of its exchange region."""
        buf = C.create_string_buffer(capi.IPC_HANDLE_BYTES)
        capi.check(self.ctx.handle, capi.lib().lpgpu_shard_p2p_export(self._h, buf), "lpgpu_shard_p2p_export")
        return bytes(buf.raw)

    def p2p_attach(self, handles):
        """handles: the export of every rank, in rank order."""
        blob = b"".join(handles)
        if len(blob) != self.world * capi.IPC_HANDLE_BYTES:
            raise ValueError("need one 64-byte handle per rank")
        capi.check(self.ctx.handle, capi.lib().lpgpu_shard_p2p_attach(self._h, C.create_string_buffer(blob, len(blob))),
                   "lpgpu_shard_p2p_attach")

    def upload(self, T, basis):
        T = np.ascontiguousarray(T, np.float64)
        basis = np.ascontiguousarray(basis, np.int32)
        if T.shape != (self.m + self.R, self.n + 1) or basis.shape != (self.m,):
            raise ValueError("shape mismatch")
        capi.check(self.ctx.handle, capi.lib().lpgpu_shard_upload(self._h, _ptr(T), T.shape[1], _ptr(basis)),
                   "lpgpu_shard_upload")

    def synth(self, seed, lp_index, ns):
        capi.check(self.ctx.handle, capi.lib().lpgpu_shard_synth(self._h, seed, lp_index, ns), "lpgpu_shard_synth")

    def reset(self, method=None, n_art=0, eps=1e-9, max_iter=2**31 - 1, log_cap=0, check_every=0, variant=0):
        if method is None:
            method = PLAIN if self.R == 1 else BIGM
        o = capi.default_opts(eps=eps, max_iter=int(min(max_iter, 2**31 - 1)), method=method, n_art=n_art,
                              log_cap=log_cap, check_every=check_every, variant=variant)
        self._log_cap = log_cap
        capi.check(self.ctx.handle, capi.lib().lpgpu_shard_reset(self._h, C.byref(o)), "lpgpu_shard_reset")

    def run(self, max_pivots):
        res = Result()
        capi.check(self.ctx.handle, capi.lib().lpgpu_shard_run(self._h, int(max_pivots), C.byref(res)), "lpgpu_shard_run")
        return _result_dict(res)

    def download(self, want_T=True, want_log=True, want_x=True):
        """T in the GLOBAL layout; columns owned by ranks of other processes stay NaN."""
        T = np.full((self.m + self.R, self.n + 1), np.nan) if want_T else None
        basis = np.zeros(self.m, np.int32)
        log = np.full((max(self._log_cap, 1), 2), -1, np.int32) if (want_log and self._log_cap) else None
        x = np.zeros(self.n) if want_x else None
        capi.check(self.ctx.handle,
                   capi.lib().lpgpu_shard_download(self._h, _ptr(T), self.n + 1, _ptr(basis), _ptr(log), _ptr(x)),
                   "lpgpu_shard_download")
        return dict(T=T, basis=basis, log=log, x=x)

    def close(self):
        if self._h:
            capi.lib().lpgpu_shard_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def solve_batched(ctx, T_all, basis_all, method=PLAIN, n_art=0, eps=1e-9, max_iter=2**31 - 1, log_cap=0, variant=0):
    """One-shot batched solve: n_lp tableaux (n_lp, m+R, n+1) in host memory."""
    T_all = np.ascontiguousarray(T_all, np.float64)
    basis_all = np.ascontiguousarray(basis_all, np.int32)
    n_lp, rows, ld = T_all.shape
    R = _R(method)
    m, n = rows - R, ld - 1
    status = np.full(n_lp, -1, np.int32)
    iters = np.zeros(n_lp, np.int32)
    z = np.zeros(n_lp)
    x = np.zeros((n_lp, n))
    basis_out = np.zeros((n_lp, m), np.int32)
    log = np.full((n_lp, max(log_cap, 1), 2), -1, np.int32)
    o = capi.default_opts(eps=eps, max_iter=int(min(max_iter, 2**31 - 1)), method=method, n_art=n_art, log_cap=log_cap,
                          variant=variant)
    rc = capi.lib().lpgpu_solve_batched(ctx.handle, n_lp, m, n, R, _ptr(T_all), _ptr(basis_all), C.byref(o),
                                        _ptr(status), _ptr(iters), _ptr(z), _ptr(x), _ptr(basis_out),
                                        _ptr(log) if log_cap else None)
    capi.check(ctx.handle, rc, "lpgpu_solve_batched")
    return dict(status=status, iters=iters, z=z, x=x, basis=basis_out, log=log)


class BatchedLPs:
    """Device-resident batch (lpgpu_batched_*): one LP per CTA, resident in shared memory."""

    def __init__(self, ctx, n_lp, m, n, R=1):
        self.ctx, self.n_lp, self.m, self.n, self.R = ctx, n_lp, m, n, R
        self._h = C.c_void_p()
        capi.check(ctx.handle, capi.lib().lpgpu_batched_create(ctx.handle, n_lp, m, n, R, C.byref(self._h)),
                   "lpgpu_batched_create")
        self._log_cap = 0

    def upload(self, T_all, basis_all):
        T_all = np.ascontiguousarray(T_all, np.float64)
        basis_all = np.ascontiguousarray(basis_all, np.int32)
        if T_all.shape != (self.n_lp, self.m + self.R, self.n + 1) or basis_all.shape != (self.n_lp, self.m):
            raise ValueError("shape mismatch")
        capi.check(self.ctx.handle, capi.lib().lpgpu_batched_upload(self._h, _ptr(T_all), _ptr(basis_all)),
                   "lpgpu_batched_upload")

    def synth(self, seed, lp0, ns):
        capi.check(self.ctx.handle, capi.lib().lpgpu_batched_synth(self._h, seed, lp0, ns), "lpgpu_batched_synth")

    def run(self, method=None, n_art=0, eps=1e-9, max_iter=2**31 - 1, log_cap=0, keep_final=False, timed=True,
            force_smem=False, variant=0):
        if method is None:
            method = PLAIN if self.R == 1 else BIGM
        o = capi.default_opts(eps=eps, max_iter=int(min(max_iter, 2**31 - 1)), method=method, n_art=n_art,
                              log_cap=log_cap,
                              variant=(capi.KEEP_FINAL if keep_final else 0) | (capi.FORCE_SMEM if force_smem else 0) | variant)
        self._log_cap = log_cap          # the library's stride / download length is this run's log_cap
        self._keep = keep_final
        ms = C.c_float(0.0)
        capi.check(self.ctx.handle, capi.lib().lpgpu_batched_run(self._h, C.byref(o), C.byref(ms) if timed else None),
                   "lpgpu_batched_run")
        return ms.value

    def download(self, want_x=True, want_log=False, want_final=False):
        n_lp, m, n = self.n_lp, self.m, self.n
        status = np.full(n_lp, -1, np.int32)
        iters = np.zeros(n_lp, np.int32)
        z = np.zeros(n_lp)
        x = np.zeros((n_lp, n)) if want_x else None
        basis = np.zeros((n_lp, m), np.int32)
        log = np.full((n_lp, max(self._log_cap, 1), 2), -1, np.int32) if (want_log and self._log_cap) else None
        Tf = np.zeros((n_lp, m + self.R, n + 1)) if want_final else None
        capi.check(self.ctx.handle,
                   capi.lib().lpgpu_batched_download(self._h, _ptr(status), _ptr(iters), _ptr(z), _ptr(x), _ptr(basis),
                                                     _ptr(log), _ptr(Tf)), "lpgpu_batched_download")
        return dict(status=status, iters=iters, z=z, x=x, basis=basis, log=log, T=Tf)

    def close(self):
        if self._h:
            capi.lib().lpgpu_batched_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------ host helpers (no GPU)
def build_tableau(A, b, c, rel, R):
    """Canonical tableau of  max c.x, A x rel b, x >= 0  (rel: -2 <=, 2 >=, 3 =)."""
    A = np.ascontiguousarray(A, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    c = np.ascontiguousarray(c, np.float64)
    rel = np.ascontiguousarray(rel, np.int32)
    m, ns = A.shape
    na = C.c_int(0)
    n = capi.lib().lpgpu_count_columns(m, ns, _ptr(b), _ptr(rel), C.byref(na))
    if n < 0:
        raise ValueError("bad relation codes")
    T = np.zeros((m + R, n + 1))
    basis = np.zeros(m, np.int32)
    n2 = capi.lib().lpgpu_build_tableau(m, ns, _ptr(A), ns, _ptr(b), _ptr(c), _ptr(rel), R, _ptr(T), n + 1,
                                        _ptr(basis), C.byref(na))
    if n2 != n:
        raise ValueError(f"lpgpu_build_tableau failed with {n2} (plain method needs all '<=' rows)")
    return T, basis, na.value


def synth_host(seed, lp_index, m, ns):
    n = ns + m
    T = np.zeros((m + 1, n + 1))
    basis = np.zeros(m, np.int32)
    capi.lib().lpgpu_synth_host(seed, lp_index, m, ns, _ptr(T), n + 1, _ptr(basis))
    return T, basis
