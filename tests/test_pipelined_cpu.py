"""CPU restatement of the PIPELINED blocked path's bookkeeping (csrc/dense_block.cuh blk_step_kernel<R, PIPE = true>,
csrc/dense.cu close_block_pipe / drain_pipe) with exact-rational single-rounding fma: two tableau buffers, two sets of
slices / row stamps / block info alternating by block parity, lazy chains that run through the PENDING block's slices
first and then the open block's, a pass that is out of place and applied one block late.  It must reproduce the
oracle's tableau, basis and pivot log bit for bit -- the argument of DESIGN.md 4.0 (same operands, same order, only
later) checked without a GPU."""
from fractions import Fraction

import numpy as np
import pytest

import helpers as H
import oracle_lib as O

EPS = 1e-9
KMAX = 32


def _fma(a, b, c):
    return float(Fraction(a) * Fraction(b) + Fraction(c))


def _row_start(stamp, seq):
    return (stamp & 0xFF) if (stamp >> 8) == seq else 0


class Pipelined:
    def __init__(self, T, basis, K):
        self.m, self.n = T.shape[0] - 1, T.shape[1] - 1
        self.T, self.T2 = T.copy(), np.full_like(T, np.nan)          # T2's content is never read before a pass wrote it
        self.F = np.zeros((2, KMAX, self.m + 1)); self.P = np.zeros((2, KMAX, self.n + 1))
        self.stamp = np.zeros((2, self.m + 1), dtype=np.int64)
        self.info = [[0, 0], [0, 0]]                                 # [parity] = [kc, seq]
        self.obj, self.b = T[self.m].copy(), T[:self.m, 0].copy()    # kept eagerly
        self.basis, self.K, self.par, self.pend, self.seq, self.t = basis.copy(), K, 0, False, 0, 0
        self.open, self.log, self.status = False, [], None
        self.inflight = None                                         # the pass enqueued on "stream 2", applied one block late

    # combined step c: the pending block's steps first (set par ^ 1), then the open block's (set par)
    def _src(self, c, Kp):
        return (self.par ^ 1, c) if c < Kp else (self.par, c - Kp)

    def _lazy(self, i, j, Kp, t, seq):
        s0 = _row_start(self.stamp[self.par][i], seq)
        c0 = Kp + s0 if s0 else (_row_start(self.stamp[self.par ^ 1][i], self.info[self.par ^ 1][1]) if Kp else 0)
        if c0:
            q, s = self._src(c0 - 1, Kp)
            a = self.P[q][s][j]
        else:
            a = self.T[i, j]
        for c in range(c0, Kp + t):
            q, s = self._src(c, Kp)
            a = _fma(-self.F[q][s][i], self.P[q][s][j], a)
        return a

    def _pass(self, src, dst, par):
        """blk_flush_* + blk_fixrows out of place: dst = src run through the block of parity `par`."""
        kc, seq = self.info[par]
        for i in range(self.m + 1):
            s0 = _row_start(self.stamp[par][i], seq)
            for j in range(self.n + 1):
                a = self.P[par][s0 - 1][j] if s0 else src[i, j]
                for s in range(s0, kc):
                    a = _fma(-self.F[par][s][i], self.P[par][s][j], a)
                dst[i, j] = a

    def _close_block(self):
        # the pass before this one must be done before the NEXT block's steps start (they read its output): run it now
        if self.inflight:
            self._pass(*self.inflight)
            self.inflight = None
        src, dst = (self.T2, self.T) if self.pend else (self.T, self.T2)
        self.inflight = (src, dst, self.par)          # runs "concurrently" with the next block: applied at its close / at the drain
        if self.pend:
            self.T, self.T2 = self.T2, self.T         # T = the tableau before the block just closed
        self.pend, self.par, self.open = True, self.par ^ 1, False

    def _drain(self):
        if self.open:
            self._close_block()
        if self.inflight:
            self._pass(*self.inflight)
            self.inflight = None
        if self.pend:
            self.T, self.T2 = self.T2, self.T
            self.pend = False

    def run(self, budget, max_iter=10**9, lazy=False):
        m, n = self.m, self.n
        while budget > 0 and self.status is None:
            first = not self.open
            if first:
                self.t, self.seq = 0, self.seq + 1
            t, seq = self.t, self.seq
            Kp = self.info[self.par ^ 1][0] if self.pend else 0
            self.info[self.par] = [t, seq]
            self.open = True
            k = max(range(1, n + 1), key=lambda j: (self.obj[j] > EPS, self.obj[j], -j))
            if not self.obj[k] > EPS:
                self.status = O.OPTIMAL
                break
            col = np.array([self._lazy(i, k, Kp, t, seq) for i in range(m)] + [self.obj[k]])
            r, theta = -1, None
            for i in range(m):
                if col[i] > EPS:
                    q = self.b[i] / col[i]
                    if theta is None or q < theta:
                        r, theta = i, q
            if r < 0:
                self.status = O.UNBOUNDED
                break
            if len(self.log) >= max_iter:
                self.status = O.ITER_LIMIT
                break
            row = np.array([self.b[r]] + [self._lazy(r, j, Kp, t, seq) for j in range(1, n + 1)])
            pr = np.where(row == 0.0, row, row / col[r])
            self.F[self.par][t], self.P[self.par][t] = col, pr
            self.obj = np.array([_fma(-col[m], pr[j], self.obj[j]) for j in range(n + 1)])
            self.b = np.array([pr[0] if i == r else _fma(-col[i], pr[0], self.b[i]) for i in range(m)])
            self.stamp[self.par][r] = (seq << 8) | (t + 1)
            self.basis[r] = k
            self.log.append([k, r])
            self.t, budget = t + 1, budget - 1
            self.info[self.par] = [t + 1, seq]
            if self.t >= self.K:
                self._close_block()
        if not lazy or self.status is not None:
            self._drain()               # self.T is up to date (download / reset / a terminal status settle the tableau)
        elif not self.open:
            # the device's decision-only launch at the end of a run call starts a block of its own (first = 1) that stays
            # empty: the sequence number moves on without a block -- the pending block is found by ITS number (info), not seq - 1
            self.seq += 1
            self.info[self.par] = [0, self.seq]


def _lp(seed, m=10, ns=8):
    """all-<= LP with mixed-sign A and c (slack basis, plain method): a handful of pivots, rows that pivot twice"""
    rng = np.random.default_rng(seed)
    A, b, c, rel = H.random_mixed_lp(rng, m, ns, p_ge=0, p_eq=0, neg_b=0)
    T, basis, n_art = O.build_tableau(A, b, c, rel, R=1)
    assert n_art == 0
    return T, basis


@pytest.mark.parametrize("K", [1, 2, 3, 32])
@pytest.mark.parametrize("seed,sizes", [(2, [10**6]), (8, [10**6]), (8, [1, 2, 3]), (11, [2, 1]), (16, [10**6])])
def test_pipelined_blocks_restated_match_oracle(K, seed, sizes):
    T, basis = _lp(seed) if seed < 16 else O.synth_feasible(1234, 3, 16, 12)
    ref = O.solve_f64(T, basis, log_cap=4096)
    assert ref["iters"] >= 5
    p = Pipelined(T, basis, K)
    q = 0
    while p.status is None:
        p.run(sizes[q % len(sizes)])
        q += 1
        part = O.solve_f64(T, basis, max_iter=len(p.log))
        assert np.array_equal(p.T, part["T"])                  # up to date after every call, whichever buffer it is
        assert np.array_equal(p.T[p.m], p.obj) and np.array_equal(p.T[:p.m, 0], p.b)   # eager copies == the chains
    assert p.status == ref["status"] == O.OPTIMAL
    assert p.log == ref["log"].tolist() and np.array_equal(p.basis, ref["basis"])
    assert np.array_equal(p.T, ref["T"])


@pytest.mark.parametrize("K", [1, 2, 3])
def test_pipelined_blocks_restated_stay_pending_across_run_calls(K):
    """Run calls that leave the open block and the pass in flight pending (as lpgpu_dense_run does while RUNNING)."""
    T, basis = _lp(8)
    ref = O.solve_f64(T, basis, log_cap=4096)
    p = Pipelined(T, basis, K)
    sizes, q = [1, 2, 1, 3], 0
    while p.status is None:
        p.run(sizes[q % len(sizes)], lazy=True)
        q += 1
        assert p.obj[0] == O.solve_f64(T, basis, max_iter=len(p.log))["T"][p.m, 0]
    assert p.status == ref["status"] and p.log == ref["log"].tolist() and np.array_equal(p.T, ref["T"])


def test_pipelined_blocks_restated_unbounded_and_limit():
    d = H.load_dumps()["unbounded.txt"]
    T, basis, n_art, n = H.tableau_from_dump(d, O.PLAIN)
    T = H.frac_f64(T)
    ref = O.solve_f64(T, basis, log_cap=16)
    p = Pipelined(T, basis, 2)
    p.run(100)
    assert p.status == ref["status"] == O.UNBOUNDED and p.log == ref["log"].tolist() and np.array_equal(p.T, ref["T"])
    T, basis = _lp(8)
    ref = O.solve_f64(T, basis, max_iter=5, log_cap=16)
    p = Pipelined(T, basis, 2)
    p.run(100, max_iter=5)
    assert p.status == ref["status"] == O.ITER_LIMIT and p.log == ref["log"].tolist() and np.array_equal(p.T, ref["T"])
