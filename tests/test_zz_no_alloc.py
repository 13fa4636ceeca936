"""lpgpu_pivot_noalloc (include/lpgpu.h): the allocation-free forced pivot on a caller-owned device tableau, against the
existing entry point lpgpu_dense_pivot -- bit for bit, as every other parity test here, on two calls in a row.

Collected last on purpose.  Passes alone in a fresh process and after the other tests:
it uses a context of its own and nothing but the C ABI (torch only supplies the caller's device buffer)."""
import ctypes as C

import numpy as np
import pytest

import linearprogramming_b200 as lp
from linearprogramming_b200 import capi



def _tableau(seed, m, n, R):
    rng = np.random.default_rng(seed)
    T = rng.uniform(-3.0, 3.0, size=(m + R, n + 1))
    T[rng.random(T.shape) < 0.2] = 0.0          # exact zeros: the signed-zero rule of the row scaling
    T[:, 0] = np.abs(T[:, 0])
    return T


def _pivot_existing(ctx, T, R, r, k):
    m, n = T.shape[0] - R, T.shape[1] - 1
    d = lp.DenseTableau(ctx, m, n, R)
    d.upload(T, np.arange(1, m + 1, dtype=np.int32))
    d.pivot(r, k)
    out = d.download(want_log=False, want_x=False)["T"]
    d.close()
    return out


def _pivot_noalloc(ctx, dev, T, ld, r, k):
    import torch
    rows, ncols = T.shape
    view = dev[: rows * ld].view(rows, ld)
    view.fill_(float("nan"))                     # padding the entry point must not touch
    view[:, :ncols] = torch.from_numpy(T)
    rc = capi.lib().lpgpu_pivot_noalloc(ctx.handle, C.c_void_p(dev.data_ptr()), ld, rows, ncols, int(r), int(k))
    capi.check(ctx.handle, rc, "lpgpu_pivot_noalloc")
    ctx.sync()
    got = view.cpu().numpy()
    assert np.isnan(got[:, ncols:]).all()
    return got[:, :ncols].copy()


@pytest.mark.gpu
def test_noalloc_pivot_matches_dense_pivot_two_calls_in_a_row():
    import torch
    torch.cuda.init()
    ctx = lp.Context(0, stream=torch.cuda.current_stream().cuda_stream)
    dev = torch.empty(4096 * 4096, dtype=torch.float64, device="cuda:0")      # the CALLER's buffer, made before the calls
    # (m, n, R, ld, r, k): the second call is smaller than the first and pivots elsewhere -- stale scratch contents
    # of the first call lie behind and beside what it uses; the third has an odd leading dimension (scalar accesses)
    cases = [(300, 700, 1, 704, 17, 333), (61, 150, 2, 151, 60, 1), (129, 257, 2, 259, 0, 257)]
    want = []
    for i, (m, n, R, ld, r, k) in enumerate(cases):
        T = _tableau(100 + i, m, n, R)
        T[r, k] = 1.75 if i != 1 else -0.625
        want.append((T, _pivot_existing(ctx, T, R, r, k)))
    ctx.sync()
    for i, (m, n, R, ld, r, k) in enumerate(cases):
        T, ref = want[i]
        got = _pivot_noalloc(ctx, dev, T, ld, r, k)
        assert np.array_equal(got.view(np.uint64), ref.view(np.uint64)), f"call {i}: tableau differs from lpgpu_dense_pivot's"
    # the same call again on the first case: the array holds the third call's contents now
    T, ref = want[0]
    m, n, R, ld, r, k = cases[0]
    assert np.array_equal(_pivot_noalloc(ctx, dev, T, ld, r, k).view(np.uint64), ref.view(np.uint64))
    ctx.close()


@pytest.mark.gpu
def test_noalloc_rejects_shapes_beyond_the_fixed_scratch():
    import torch
    ctx = lp.Context(0)
    dev = torch.zeros(1024, dtype=torch.float64, device="cuda:0")
    L, H = capi.lib(), capi._H
    p = C.c_void_p(dev.data_ptr())
    assert L.lpgpu_pivot_noalloc(ctx.handle, p, 8, H["NOALLOC_MAX_ROWS"] + 1, 8, 0, 1) == capi.EINVAL
    assert b"exceed" in L.lpgpu_last_error(ctx.handle)
    assert L.lpgpu_pivot_noalloc(ctx.handle, p, H["NOALLOC_MAX_COLS"] + 1, 4, H["NOALLOC_MAX_COLS"] + 1, 0, 1) == capi.EINVAL
    assert L.lpgpu_pivot_noalloc(ctx.handle, p, 8, 4, 8, 4, 1) == capi.EINVAL      # row out of range
    assert L.lpgpu_pivot_noalloc(ctx.handle, p, 8, 4, 8, 0, 0) == capi.EINVAL      # the b column is not a pivot column
    assert L.lpgpu_pivot_noalloc(ctx.handle, p, 4, 4, 8, 0, 1) == capi.EINVAL      # ld < ncols
    assert L.lpgpu_pivot_noalloc(ctx.handle, None, 8, 4, 8, 0, 1) == capi.EINVAL
    ctx.sync()
    assert float(dev.abs().sum()) == 0.0                                           # nothing ran
    ctx.close()


def test_noalloc_source_calls_no_allocating_routine():
    """By construction: the translation unit of the entry point names no allocating CUDA routine outside its comments."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(capi.HERE), "linearprogramming_b200", "csrc", "noalloc.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    for word in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "cudaMallocManaged", "new ", "malloc("):
        assert word not in code, word
    assert "cudaGetSymbolAddress" in code and "static __device__ __align__(256)" in code
