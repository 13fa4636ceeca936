"""ctypes binding to the CPU oracle (oracle/liblp_oracle.so) -- test infrastructure only.

Nothing under linearprogramming_b200/ imports this module; it is the checker that the
parity tests, __graft_entry__.smoke() and bench.py's cpu_baseline leg compare against.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

OPTIMAL, UNBOUNDED, INFEASIBLE, ITER_LIMIT, RUNNING, OVERFLOW = range(6)
PLAIN, BIGM, TWOPHASE, DUAL = 0, 1, 2, 3
STATUS_NAMES = ["OPTIMAL", "UNBOUNDED", "INFEASIBLE", "ITER_LIMIT", "RUNNING", "OVERFLOW"]


DANTZIG, BLAND, DEVEX = 0, 1, 2          # lpo_opts.pricing


class Opts(C.Structure):
    _fields_ = [("eps", C.c_double), ("max_iter", C.c_int), ("method", C.c_int),
                ("n_art", C.c_int), ("threads", C.c_int), ("pricing", C.c_int)]


class Result(C.Structure):
    _fields_ = [("status", C.c_int), ("iters", C.c_int), ("phase1_iters", C.c_int),
                ("phase", C.c_int), ("z", C.c_double), ("zM", C.c_double)]


class Rat(C.Structure):
    _fields_ = [("num", C.c_long), ("den", C.c_long), ("valid", C.c_int)]


_lib = None


def build():
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, capture_output=True)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(ORACLE_DIR, "liblp_oracle.so")
    src = os.path.join(ORACLE_DIR, "lp_oracle.c")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
        build()
    L = C.CDLL(path)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    L.lpo_solve_f64.argtypes = [C.c_int, C.c_int, C.c_int, dp, C.c_size_t, ip, C.POINTER(Opts), ip,
                                C.c_int, C.POINTER(Result)]
    L.lpo_solve_f64.restype = C.c_int
    L.lpo_pivot_f64.argtypes = [C.c_int, C.c_int, dp, C.c_size_t, C.c_int, C.c_int, C.c_int]
    L.lpo_pivot_f64.restype = None
    L.lpo_extract_x.argtypes = [C.c_int, C.c_int, dp, C.c_size_t, ip, dp]
    L.lpo_extract_x.restype = None
    L.lpo_solve_batched_f64.argtypes = [C.c_long, C.c_int, C.c_int, C.c_int, dp, ip, C.POINTER(Opts),
                                        ip, ip, dp, dp, ip, C.c_int]
    L.lpo_solve_batched_f64.restype = C.c_int
    L.lpo_build_tableau.argtypes = [C.c_int, C.c_int, dp, C.c_size_t, dp, dp, ip, C.c_int, dp,
                                    C.c_size_t, ip, ip]
    L.lpo_build_tableau.restype = C.c_int
    L.lpo_count_columns.argtypes = [C.c_int, C.c_int, dp, ip, ip]
    L.lpo_count_columns.restype = C.c_int
    L.lpo_mix64.argtypes = [C.c_uint64]
    L.lpo_mix64.restype = C.c_uint64
    L.lpo_uniform.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_double, C.c_double]
    L.lpo_uniform.restype = C.c_double
    L.lpo_synth_feasible.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int, dp, C.c_size_t, ip]
    L.lpo_synth_feasible.restype = None
    for name in ("add", "sub", "mul", "div"):
        f = getattr(L, "lpo_rat_" + name)
        f.argtypes = [Rat, Rat]
        f.restype = Rat
    L.lpo_rat_neg.argtypes = [Rat]
    L.lpo_rat_neg.restype = Rat
    L.lpo_rat_to_double.argtypes = [Rat]
    L.lpo_rat_to_double.restype = C.c_double
    L.lpo_rat_cmp.argtypes = [Rat, Rat]
    L.lpo_rat_cmp.restype = C.c_int
    L.lpo_rat_parse.argtypes = [C.c_char_p]
    L.lpo_rat_parse.restype = Rat
    L.lpo_solve_rat.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(Rat), C.c_size_t, ip,
                                C.POINTER(Opts), ip, C.c_int, C.POINTER(Result)]
    L.lpo_solve_rat.restype = C.c_int
    _lib = L
    return L


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def solve_f64(T, basis, method=PLAIN, n_art=0, eps=1e-9, max_iter=1 << 30, threads=1, log_cap=None, R=None, pricing=0):
    """Solve in place on copies; returns dict(status, iters, z, zM, T, basis, log, x, phase1_iters)."""
    T = np.array(T, dtype=np.float64, order="C", copy=True)
    basis = np.array(basis, dtype=np.int32, copy=True)
    if R is None:
        R = 1 if method in (PLAIN, DUAL) else 2
    rows, ld = T.shape
    m, n = rows - R, ld - 1
    if log_cap is None:
        log_cap = 4 * (m + n) + 64
    log = np.full((log_cap, 2), -1, dtype=np.int32)
    res = Result()
    opts = Opts(eps, int(min(max_iter, 2**31 - 1)), method, n_art, threads, pricing)
    rc = lib().lpo_solve_f64(m, n, R, _dp(T), ld, _ip(basis), C.byref(opts), _ip(log), log_cap, C.byref(res))
    assert rc == 0, "oracle rejected arguments"
    x = np.zeros(n)
    lib().lpo_extract_x(m, n, _dp(T), ld, _ip(basis), _dp(x))
    return dict(status=res.status, iters=res.iters, phase1_iters=res.phase1_iters, z=res.z, zM=res.zM,
                T=T, basis=basis, log=log[:min(res.iters, log_cap)].copy(), x=x, phase=res.phase)


def solve_batched_f64(T_all, basis_all, method=PLAIN, n_art=0, eps=1e-9, max_iter=1 << 30, threads=1,
                      log_cap=0, pricing=0):
    T_all = np.array(T_all, dtype=np.float64, order="C", copy=True)
    basis_all = np.array(basis_all, dtype=np.int32, order="C", copy=True)
    n_lp, rows, ld = T_all.shape
    R = 1 if method in (PLAIN, DUAL) else 2
    m, n = rows - R, ld - 1
    status = np.zeros(n_lp, np.int32)
    iters = np.zeros(n_lp, np.int32)
    z = np.zeros(n_lp)
    x = np.zeros((n_lp, n))
    log = np.full((n_lp, max(log_cap, 1), 2), -1, np.int32)
    opts = Opts(eps, int(min(max_iter, 2**31 - 1)), method, n_art, threads, pricing)
    rc = lib().lpo_solve_batched_f64(n_lp, m, n, R, _dp(T_all), _ip(basis_all), C.byref(opts), _ip(status),
                                     _ip(iters), _dp(z), _dp(x), _ip(log) if log_cap else None, log_cap)
    assert rc == 0
    return dict(status=status, iters=iters, z=z, x=x, T=T_all, basis=basis_all, log=log)


def pivot(T, r, k, threads=1):
    T = np.array(T, dtype=np.float64, order="C", copy=True)
    rows, ld = T.shape
    lib().lpo_pivot_f64(rows, ld - 1, _dp(T), ld, r, k, threads)
    return T


def build_tableau(A, b, c, rel, R):
    A = np.ascontiguousarray(A, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    c = np.ascontiguousarray(c, np.float64)
    rel = np.ascontiguousarray(rel, np.int32)
    m, ns = A.shape
    na = C.c_int(0)
    n = lib().lpo_count_columns(m, ns, _dp(b), _ip(rel), C.byref(na))
    assert n > 0
    T = np.zeros((m + R, n + 1))
    basis = np.zeros(m, np.int32)
    n2 = lib().lpo_build_tableau(m, ns, _dp(A), ns, _dp(b), _dp(c), _ip(rel), R, _dp(T), n + 1, _ip(basis),
                                 C.byref(na))
    assert n2 == n, (n2, n)
    return T, basis, na.value


def synth_feasible(seed, lp, m, ns):
    n = ns + m
    T = np.zeros((m + 1, n + 1))
    basis = np.zeros(m, np.int32)
    lib().lpo_synth_feasible(seed, lp, m, ns, _dp(T), n + 1, _ip(basis))
    return T, basis


def synth_feasible_batch(seed, lp0, n_lp, m, ns):
    n = ns + m
    T = np.zeros((n_lp, m + 1, n + 1))
    basis = np.zeros((n_lp, m), np.int32)
    L = lib()
    for p in range(n_lp):
        L.lpo_synth_feasible(seed, lp0 + p, m, ns, _dp(T[p]), n + 1, _ip(basis[p]))
    return T, basis


def rat(num, den=1):
    return Rat(num, den, 1 if den != 0 else 0)


def rat_parse(s):
    return lib().lpo_rat_parse(s.encode())


def solve_rat(T_pairs, basis, method=PLAIN, n_art=0, max_iter=1 << 30, log_cap=256):
    """T_pairs: rows x cols list of (num, den).  Returns dict like solve_f64 with T as (num, den) tuples."""
    rows, cols = len(T_pairs), len(T_pairs[0])
    R = 1 if method in (PLAIN, DUAL) else 2
    m, n = rows - R, cols - 1
    arr = (Rat * (rows * cols))()
    for i in range(rows):
        for j in range(cols):
            nu, de = T_pairs[i][j]
            arr[i * cols + j] = Rat(nu, de, 1)
    basis = np.array(basis, np.int32, copy=True)
    log = np.full((log_cap, 2), -1, np.int32)
    res = Result()
    opts = Opts(0.0, int(min(max_iter, 2**31 - 1)), method, n_art, 1)
    rc = lib().lpo_solve_rat(m, n, R, arr, cols, _ip(basis), C.byref(opts), _ip(log), log_cap, C.byref(res))
    assert rc == 0
    T = [[(arr[i * cols + j].num, arr[i * cols + j].den) for j in range(cols)] for i in range(rows)]
    return dict(status=res.status, iters=res.iters, phase1_iters=res.phase1_iters, z=res.z, zM=res.zM, T=T,
                basis=basis, log=log[:res.iters].copy())
