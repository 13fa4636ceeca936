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
KEEP_FINAL = 0x1000      # batched: keep the final tableaux on the device (parity tests)
FORCE_SMEM = 0x2000      # batched: shared-memory kernel even where the register-resident one applies
STATIC = 0x1             # dense: TMA-staged elimination with a static task stride (default: work stealing)
DIRECT = 0x2             # dense: direct-load elimination kernel instead of the TMA-staged default
RANK1 = 0x800            # dense: one pass over the tableau per pivot (default for plain / Big-M: up to K pivots decided on lazily
                         # updated slices, then applied in one pass; K = variant bits 16..21, 0 = 32)
INPLACE = 0x2000         # dense, blocked path: steps, pass, steps, ... in place on one stream
PIPELINED = 0x1000       # dense, blocked path (primal methods): pipelined blocks -- the pass of a block runs out of place on a second
                         # stream while the next block is decided through the slices of both blocks (bit-identical; needs a second
                         # tableau buffer).  Default (neither bit): pipelined from about 0.4 GB of tableau on, in place below
WHOLE_TILE_PASS = 0x4000 # dense, blocked path: the pass kernel that loads a whole tile before it computes (default: software-pipelined halves)
STEP_CL8 = 0x8000        # dense, blocked path: primal step kernel as a cluster of 8 CTAs (default: 16 where the device allows it)
BLAND = 0x400000         # dense and batched, primal methods: Bland's rule -- the LOWEST eligible column enters (default: the largest
                         # reduced cost, ties to the lowest column); with the lowest-row ratio tie-break it cannot cycle
WS_PASS = 0x100          # dense, blocked path: warp-specialised pass kernel (TMA ring; measured slower, kept as an alternative)
NO_PDL = 0x400           # dense: plain stream-ordered launches (default: programmatic dependent launch)
NO_ALTERNATE = 0x200     # dense: do not alternate the sweep direction between pivots
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
