"""ctypes binding to liblphost.so: the host C front end (model text -> tableau) and the file-level
solve that stands where the reference's newSimplex stops (Source/simplex.c:27-73)."""
import ctypes as C
import os

import numpy as np

from . import capi

LIB_PATH = os.path.join(capi.HERE, "liblphost.so")


class Rat(C.Structure):
    _fields_ = [("num", C.c_long), ("den", C.c_long)]


class Var(C.Structure):
    _fields_ = [("name", C.c_char * 8), ("relation", C.c_int), ("former", C.c_char * 8), ("latter", C.c_char * 8)]


class Model(C.Structure):
    _fields_ = [("valid", C.c_int), ("error", C.c_char * 256), ("sense", C.c_int), ("constant", Rat),
                ("m", C.c_int), ("n", C.c_int), ("names", C.POINTER(C.c_char * 8)), ("inverted", C.POINTER(C.c_int)),
                ("costs", C.POINTER(Rat)), ("rows", C.POINTER(Rat)), ("basis", C.POINTER(C.c_int)),
                ("nvars", C.c_int), ("nuser", C.c_int), ("vars", C.POINTER(Var))]


class Solution(C.Structure):
    _fields_ = [("status", C.c_int), ("iters", C.c_int), ("method", C.c_int), ("z", C.c_double),
                ("n_named", C.c_int), ("names", C.POINTER(C.c_char * 8)), ("values", C.POINTER(C.c_double)),
                ("n_log", C.c_int), ("pivot_log", C.POINTER(C.c_int))]


_lib = None


def lib():
    global _lib
    if _lib is None:
        capi.lib()                      # liblphost.so links against liblpgpu.so
        if not os.path.exists(LIB_PATH):
            raise capi.LpgpuError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        L = C.CDLL(LIB_PATH)
        L.lpf_parse_file.argtypes = [C.c_char_p, C.POINTER(Model)]
        L.lpf_parse_string.argtypes = [C.c_char_p, C.POINTER(Model)]
        L.lpf_parse_file_mode.argtypes = [C.c_char_p, C.c_int, C.POINTER(Model)]
        L.lpf_parse_string_mode.argtypes = [C.c_char_p, C.c_int, C.POINTER(Model)]
        L.lpf_free.argtypes = [C.POINTER(Model)]
        L.lpf_build_tableau.argtypes = [C.POINTER(Model), C.c_int, C.POINTER(C.POINTER(C.c_double)),
                                        C.POINTER(C.POINTER(C.c_int)), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.lpf_solve_model.argtypes = [C.c_void_p, C.POINTER(Model), C.c_int, C.POINTER(Solution), C.c_char_p, C.c_size_t]
        L.lpf_solve_file.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(Solution), C.c_char_p, C.c_size_t]
        L.lpf_free_solution.argtypes = [C.POINTER(Solution)]
        _lib = L
    return _lib


def _model_dict(m):
    n, rows = m.n, m.m
    return dict(
        sense=m.sense, m=rows, n=n, constant=(m.constant.num, m.constant.den),
        names=[m.names[j].value.decode() for j in range(n)], inverted=[m.inverted[j] for j in range(n)],
        costs=[(m.costs[j].num, m.costs[j].den) for j in range(n)],
        rows=[[(m.rows[i * (n + 1) + j].num, m.rows[i * (n + 1) + j].den) for j in range(n + 1)] for i in range(rows)],
        basis=[m.basis[i] for i in range(rows)],
        vars=[dict(name=m.vars[v].name.decode(), relation=m.vars[v].relation, former=m.vars[v].former.decode(),
                   latter=m.vars[v].latter.decode()) for v in range(m.nvars)], nuser=m.nuser)


def parse(text=None, path=None, dual=False):
    """Parse + standardise + align a model; returns a dict, raises ValueError with the front end's message.
    dual = True: the dual-simplex standard form (">=" rows negated into "<=", b keeps its sign)."""
    m = Model()
    d = 1 if dual else 0
    rc = (lib().lpf_parse_file_mode(path.encode(), d, C.byref(m)) if path
          else lib().lpf_parse_string_mode(text.encode(), d, C.byref(m)))
    if rc != 0:
        raise ValueError(m.error.decode(errors="replace"))
    out = _model_dict(m)
    lib().lpf_free(C.byref(m))
    return out


def tableau(path, method):
    """(T, basis, n_art) as the front end hands them to lpgpu_solve_dense (method 3 = dual simplex)."""
    m = Model()
    dual = method == capi.DUAL
    if lib().lpf_parse_file_mode(path.encode(), 1 if dual else 0, C.byref(m)) != 0:
        raise ValueError(m.error.decode(errors="replace"))
    Tp, bp, n, na = C.POINTER(C.c_double)(), C.POINTER(C.c_int)(), C.c_int(), C.c_int()
    rc = lib().lpf_build_tableau(C.byref(m), 0 if dual else method, C.byref(Tp), C.byref(bp), C.byref(n), C.byref(na))
    if rc != 0:
        lib().lpf_free(C.byref(m))
        raise ValueError("the plain method needs an identity basis")
    R = 1 if method in (0, capi.DUAL) else 2
    T = np.ctypeslib.as_array(Tp, shape=(m.m + R, n.value + 1)).copy()
    basis = np.ctypeslib.as_array(bp, shape=(m.m,)).astype(np.int32).copy()
    libc = C.CDLL(None)
    libc.free.argtypes = [C.c_void_p]
    libc.free(Tp); libc.free(bp)
    lib().lpf_free(C.byref(m))
    return T, basis, na.value


def solve_file(ctx, path, method=capi.BIGM):
    """Solve a model file on the GPU; returns dict(status, iters, method, z, x{name: value}, log)."""
    s = Solution()
    err = C.create_string_buffer(512)
    rc = lib().lpf_solve_file(ctx.handle, path.encode(), method, C.byref(s), err, 512)
    if rc != 0:
        raise capi.LpgpuError(err.value.decode(errors="replace"))
    out = dict(status=s.status, iters=s.iters, method=s.method, z=s.z,
               x={s.names[v].value.decode(): s.values[v] for v in range(s.n_named)},
               log=[(s.pivot_log[2 * p], s.pivot_log[2 * p + 1]) for p in range(s.n_log)])
    lib().lpf_free_solution(C.byref(s))
    return out
