"""ctypes loader of liblpgpu.so -- the C ABI declared in include/lpgpu.h.

The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no
CPU fallback: if the shared object is missing, loading raises, and if there is no sm_100
device ``lpgpu_create`` returns LPGPU_ENODEV, which ``Context`` turns into an exception.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblpgpu.so")

OK, EINVAL, ENODEV, ECUDA, ENOMEM, ENCCL = 0, -1, -2, -3, -4, -5
OPTIMAL, UNBOUNDED, INFEASIBLE, ITER_LIMIT, RUNNING = 0, 1, 2, 3, 4
PLAIN, BIGM, TWOPHASE, DUAL = 0, 1, 2, 3
STATUS_NAMES = {0: "OPTIMAL", 1: "UNBOUNDED", 2: "INFEASIBLE", 3: "ITER_LIMIT", 4: "RUNNING"}
def _header_defines():
    """#define LPGPU_<NAME> <integer> of include/lpgpu.h: the header is the one place option bits are defined."""
    import re
    hdr = os.path.join(os.path.dirname(HERE), "include", "lpgpu.h")
    text = open(hdr).read()
    return {k: int(v, 0) for k, v in re.findall(r"#define\s+LPGPU_(\w+)\s+\(?(-?(?:0x[0-9a-fA-F]+|\d+))\)?", text)}


_H = _header_defines()
# lpgpu_opts.variant bits (include/lpgpu.h; DESIGN.md section 4 says what each selects)
KEEP_FINAL = _H["VARIANT_BATCH_KEEP_FINAL"]    # batched: keep the final tableaux on the device (parity tests)
FORCE_SMEM = _H["VARIANT_BATCH_SMEM"]          # batched: shared-memory kernel even where a register-resident one applies
STATIC = _H["VARIANT_UPD_STATIC"]              # dense, rank-1 path: static task stride (default: work stealing)
DIRECT = _H["VARIANT_UPD_DIRECT"]              # dense, rank-1 path: direct-load elimination kernel
NO_ALTERNATE = _H["VARIANT_NO_ALTERNATE"]      # dense, rank-1 path: do not alternate the sweep direction between pivots
NO_PDL = _H["VARIANT_NO_PDL"]                  # dense: plain stream-ordered launches (default: programmatic dependent launch)
RANK1 = _H["VARIANT_RANK1"]                    # dense: one pass over the tableau per pivot (default: one pass per block of pivots)
PIPELINED = _H["VARIANT_PIPELINED"]            # dense, blocked path: always pipeline the pass with the next block's decisions
INPLACE = _H["VARIANT_INPLACE"]                # dense, blocked path: never pipeline (default: by size)
DFMA_PASS = _H["VARIANT_DFMA_PASS"]            # dense, blocked path: the pass with scalar DFMAs (default: the ring pass, TMA ring of slabs + DMMA chain)
DMMA_PASS = _H["VARIANT_DMMA_PASS"]            # dense, blocked path: the pass with the FP64 matrix instruction as the FMA chain (default: scalar DFMAs)
STEP_PERSISTENT = _H["VARIANT_STEP_PERSISTENT"]  # dense, pipelined blocks: one persistent step launch per block (default: one cluster launch per pivot)
STEP_CL8 = _H["VARIANT_STEP_CL8"]              # dense, blocked path: step kernel as a portable 8-CTA cluster
BLAND = _H["VARIANT_BLAND"]                    # Bland's rule (primal methods)
DEVEX = _H["VARIANT_DEVEX"]                    # Devex reference-framework pricing (primal methods)


def BLOCK_K(k):
    """variant bits for `k` pivots per block (1..32)."""
    return (int(k) & 0x3f) << 16


NCCL_ID_BYTES = 128
IPC_HANDLE_BYTES = 64


class Opts(C.Structure):
    _fields_ = [("eps", C.c_double), ("max_iter", C.c_int), ("method", C.c_int), ("n_art", C.c_int),
                ("log_cap", C.c_int), ("check_every", C.c_int), ("variant", C.c_int),
                ("time_kernels", C.c_int), ("results_only", C.c_int)]


class Result(C.Structure):
    _fields_ = [("status", C.c_int), ("iters", C.c_int), ("phase1_iters", C.c_int), ("phase", C.c_int),
                ("z", C.c_double), ("zM", C.c_double), ("n_logged", C.c_int), ("update_launches", C.c_int),
                ("update_ms", C.c_float), ("select_ms", C.c_float)]


class LpgpuError(RuntimeError):
    pass


_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p

# name -> (restype, argtypes): every symbol include/lpgpu.h declares
SIGNATURES = {
    "lpgpu_default_opts": (None, [C.POINTER(Opts)]),
    "lpgpu_abi_sizeof": (C.c_size_t, [C.c_int]),
    "lpgpu_create": (C.c_int, [C.POINTER(_vp), C.c_int, _vp]),
    "lpgpu_destroy": (None, [_vp]),
    "lpgpu_last_error": (C.c_char_p, [_vp]),
    "lpgpu_device_info": (C.c_int, [_vp, _ip, C.POINTER(C.c_size_t), _ip, _ip]),
    "lpgpu_launch_count": (C.c_long, [_vp]),
    "lpgpu_sync": (C.c_int, [_vp]),
    "lpgpu_host_alloc": (_vp, [_vp, C.c_size_t]),
    "lpgpu_host_free": (None, [_vp, _vp]),
    "lpgpu_solve_dense": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp, C.c_size_t, _vp, C.POINTER(Opts), _vp, _vp,
                                    C.POINTER(Result)]),
    "lpgpu_dense_create": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]),
    "lpgpu_dense_destroy": (None, [_vp]),
    "lpgpu_dense_upload": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "lpgpu_dense_upload_async": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "lpgpu_dense_synth": (C.c_int, [_vp, C.c_uint64, C.c_uint64, C.c_int]),
    "lpgpu_dense_reset": (C.c_int, [_vp, C.POINTER(Opts)]),
    "lpgpu_dense_run": (C.c_int, [_vp, C.c_int, C.POINTER(Result)]),
    "lpgpu_dense_pivot": (C.c_int, [_vp, C.c_int, C.c_int]),
    "lpgpu_dense_download": (C.c_int, [_vp, _vp, C.c_size_t, _vp, _vp, _vp]),
    "lpgpu_dense_ld": (C.c_size_t, [_vp]),
    "lpgpu_pivot_noalloc": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int]),
    "lpgpu_solve_batched": (C.c_int, [_vp, C.c_long, C.c_int, C.c_int, C.c_int, _vp, _vp, C.POINTER(Opts), _vp, _vp,
                                      _vp, _vp, _vp, _vp]),
    "lpgpu_batched_create": (C.c_int, [_vp, C.c_long, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]),
    "lpgpu_batched_destroy": (None, [_vp]),
    "lpgpu_batched_upload": (C.c_int, [_vp, _vp, _vp]),
    "lpgpu_batched_synth": (C.c_int, [_vp, C.c_uint64, C.c_uint64, C.c_int]),
    "lpgpu_batched_run": (C.c_int, [_vp, C.POINTER(Opts), C.POINTER(C.c_float)]),
    "lpgpu_batched_download": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "lpgpu_shard_unique_id": (C.c_int, [_vp]),
    "lpgpu_shard_columns": (C.c_int, [C.c_int, C.c_int, C.c_int, _ip, _ip]),
    "lpgpu_shard_create": (C.c_int, [_vp, C.c_int, C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]),
    "lpgpu_shard_create_p2p": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]),
    "lpgpu_shard_p2p_export": (C.c_int, [_vp, _vp]),
    "lpgpu_shard_p2p_attach": (C.c_int, [_vp, _vp]),
    "lpgpu_shard_destroy": (None, [_vp]),
    "lpgpu_shard_upload": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "lpgpu_shard_synth": (C.c_int, [_vp, C.c_uint64, C.c_uint64, C.c_int]),
    "lpgpu_shard_reset": (C.c_int, [_vp, C.POINTER(Opts)]),
    "lpgpu_shard_run": (C.c_int, [_vp, C.c_int, C.POINTER(Result)]),
    "lpgpu_shard_download": (C.c_int, [_vp, _vp, C.c_size_t, _vp, _vp, _vp]),
    "lpgpu_count_columns": (C.c_int, [C.c_int, C.c_int, _vp, _vp, _ip]),
    "lpgpu_build_tableau": (C.c_int, [C.c_int, C.c_int, _vp, C.c_size_t, _vp, _vp, _vp, C.c_int, _vp, C.c_size_t, _vp,
                                      _ip]),
    "lpgpu_synth_host": (None, [C.c_uint64, C.c_uint64, C.c_int, C.c_int, _vp, C.c_size_t, _vp]),
}


def lib():
    """Load liblpgpu.so (once).  Raises LpgpuError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LpgpuError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                         "(nvcc, sm_100a).  There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        f = getattr(L, name)          # AttributeError if the ABI and the header drifted apart
        f.restype = res
        f.argtypes = args
    if L.lpgpu_abi_sizeof(0) != C.sizeof(Opts) or L.lpgpu_abi_sizeof(1) != C.sizeof(Result):
        raise LpgpuError("capi.py structs do not match include/lpgpu.h (rebuild liblpgpu.so)")
    _lib = L
    return L


def check(ctx_handle, rc, what):
    if rc != OK:
        msg = lib().lpgpu_last_error(ctx_handle) if ctx_handle else b""
        raise LpgpuError(f"{what} failed with code {rc}: {(msg or b'').decode(errors='replace')}")


def default_opts(**kw):
    o = Opts()
    lib().lpgpu_default_opts(C.byref(o))
    for k, v in kw.items():
        if v is not None:
            setattr(o, k, v)
    return o
