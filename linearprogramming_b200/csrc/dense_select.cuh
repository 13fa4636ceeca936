// dense_select.cuh -- pricing, ratio test and pivot-row scaling for ONE large tableau.
//
// One CTA of 1024 threads runs the whole decision for a pivot:
//   entering column  = arg max reduced cost (> eps), ties -> lowest column
//   leaving row      = arg min b_i / a_ik over a_ik > eps, ties -> lowest row
//   pivot row        = a_rj / a_rk  -> compact buffer prow[0..ld)
//   pivot column     = a_ik         -> compact buffer pcol[0..m+R)
// and the Big-M / two-phase objective-row handling.  The reference has none of this
// (Source/simplex.c:57-63 empty cases, :79-82 empty AddBigM); the rule is this repo's and is
// restated on the CPU by the test oracle.  The arg-min/arg-max reductions are warp
// shuffles plus a block-level tree (common.cuh: block_best).
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"

namespace lpgpu {

struct DenseParams {
    double *T;          // (m+R) x ld, row-major, ld % 16 == 0, padding columns are zero
    size_t ld;
    int m, n, R;
    int *basis;         // m entries, 1-based columns
    double *prow;       // ld doubles: scaled pivot row
    double *pcol;       // colbuf: 2 x pcn doubles; half st->cur = pivot column of this pivot, the other
                        // half receives the NEXT entering column from the elimination pass (lookahead)
    size_t pcn;         // doubles per half (>= m+R, zero padded to the row-tile multiple)
    double *bnext;      // pcn doubles: b column as left by the elimination pass (lookahead)
    PivotState *st;
    int *log;           // 2*log_cap ints
    int log_cap;
    double eps;
    int max_iter;
    int method;
    int n_art;
    int bland;          // 1: Bland's rule -- every eligible column gets the same key, so the arg-max's tie-break (lowest
                        // column) picks the entering column (variant bit 0x400000, primal methods)
    double *wts;        // Devex reference weights w[0..ld) (LPGPU_VARIANT_DEVEX, primal methods), else nullptr: the entering column
                        // maximises sigma_j^2 / w_j; every pivot updates them from its scaled pivot row (block_common.cuh devex_weight)
};

// Devex weight of column j after the pivot on (r, k): pr = the scaled pivot row entry a_rj', wk = w_k before the pivot
// (the CPU checker restates it as devex_update: two rounded multiplies in this order, max, the entering column restarts at 1)
__device__ __forceinline__ double devex_weight(double w, double pr, double wk, bool entering) {
    const double u = (pr * pr) * wk;
    return entering ? 1.0 : (u > w ? u : w);
}

constexpr int SELECT_THREADS = 1024;

namespace cg = cooperative_groups;

constexpr int SELECT_CLUSTER = 8;     // CTAs per cluster (portable maximum); one cluster per launch

// Best candidate over the whole cluster: block reduction, then every CTA pushes its winner into
// every CTA's shared memory (DSMEM stores) and one cluster barrier makes them visible.
// `slot` must be used for ONE exchange per launch (no reuse hazards).
__device__ __forceinline__ Cand cluster_best(cg::cluster_group &cl, Cand c, Cand *scratch, Cand *slot) {
    c = block_best(c, scratch);
    if (threadIdx.x < SELECT_CLUSTER) {
        Cand *remote = cl.map_shared_rank(slot, threadIdx.x);
        remote[cl.block_rank()] = c;
    }
    cl.sync();
    const int lane = threadIdx.x & 31;
    Cand v = (lane < SELECT_CLUSTER) ? slot[lane] : cand_none();
    return warp_best(v);
}

__device__ __forceinline__ Cand price_row(const DenseParams &P, int row, int jmax, Cand *scratch) {
    const double *x = P.T + (size_t)row * P.ld;
    Cand c = cand_none();
    for (int j = 1 + (int)threadIdx.x; j <= jmax; j += blockDim.x) {
        const double v = x[j];
        if (v > P.eps) {
            Cand t;
            t.p = 0.0; t.q = P.bland ? 1.0 : P.wts ? (v * v) / P.wts[j] : v; t.idx = j;
            if (cand_better(t, c)) c = t;
        }
    }
    return block_best(c, scratch);
}

__device__ __forceinline__ Cand price_bigm(const DenseParams &P, Cand *scratch) {
    const double *xc = P.T + (size_t)P.m * P.ld;
    const double *xm = P.T + (size_t)(P.m + 1) * P.ld;
    Cand c = cand_none();
    for (int j = 1 + (int)threadIdx.x; j <= P.n; j += blockDim.x) {
        double p = xm[j];
        const double q = xc[j];
        if (fabs(p) <= P.eps) p = 0.0;
        if (p > 0.0 || (p == 0.0 && q > P.eps)) {
            Cand t;
            t.p = P.bland ? 0.0 : p; t.q = P.bland ? 1.0 : q; t.idx = j;
            if (P.wts) {                                 // Devex: both levels scaled by the weight (oracle price_bigm, pricing == 2)
                const double w = P.wts[j];
                t.q = (p == 0.0) ? (q * q) / w : q;
                t.p = (p * p) / w;
            }
            if (cand_better(t, c)) c = t;
        }
    }
    return block_best(c, scratch);
}

__device__ __forceinline__ Cand ratio_test(const DenseParams &P, int k, Cand *scratch) {
    Cand c = cand_none();
    for (int i = (int)threadIdx.x; i < P.m; i += blockDim.x) {
        const double a = P.T[(size_t)i * P.ld + (size_t)k];
        if (a > P.eps) {
            Cand t;
            t.p = 0.0;
            t.q = -(P.T[(size_t)i * P.ld] / a);     // exact negation: arg max -theta == arg min theta
            t.idx = i;
            if (cand_better(t, c)) c = t;
        }
    }
    return block_best(c, scratch);
}

// Dual simplex (LPGPU_DUAL): leaving row = most negative b_i < -eps, ties -> lowest row.
__device__ __forceinline__ Cand dual_leave(const DenseParams &P, Cand *scratch) {
    Cand c = cand_none();
    for (int i = (int)threadIdx.x; i < P.m; i += blockDim.x) {
        const double b = P.T[(size_t)i * P.ld];
        if (b < -P.eps) {
            Cand t;
            t.p = 0.0; t.q = -b; t.idx = i;              // exact negation: arg max -b == arg min b
            if (cand_better(t, c)) c = t;
        }
    }
    return block_best(c, scratch);
}

// ... entering column = min sigma_j / a_rj over a_rj < -eps, ties -> lowest column.
__device__ __forceinline__ Cand dual_enter(const DenseParams &P, int r, Cand *scratch) {
    const double *xr = P.T + (size_t)r * P.ld;
    const double *oc = P.T + (size_t)P.m * P.ld;
    Cand c = cand_none();
    for (int j = 1 + (int)threadIdx.x; j <= P.n; j += blockDim.x) {
        const double a = xr[j];
        if (a < -P.eps) {
            Cand t;
            t.p = 0.0; t.q = -(oc[j] / a); t.idx = j;
            if (cand_better(t, c)) c = t;
        }
    }
    return block_best(c, scratch);
}

// any artificial still basic at a positive level?
__device__ __forceinline__ bool art_basic_positive(const DenseParams &P) {
    const int nreal = P.n - P.n_art;
    int f = 0;
    for (int i = (int)threadIdx.x; i < P.m; i += blockDim.x)
        if (P.basis[i] > nreal && P.T[(size_t)i * P.ld] > P.eps) f = 1;
    return __syncthreads_or(f) != 0;
}

// Drive-out at the phase switch: lowest row with an artificial basic variable that has a
// non-artificial entry |a_ij| > eps; lowest such column.  Returns false if there is none.
__device__ __forceinline__ bool driveout_find(const DenseParams &P, Cand *scratch, int *r, int *k) {
    const int nreal = P.n - P.n_art;
    int start = 0;
    while (true) {
        Cand c = cand_none();
        for (int i = start + (int)threadIdx.x; i < P.m; i += blockDim.x)
            if (P.basis[i] > nreal) { c.p = 0.0; c.q = 0.0; c.idx = i; break; }
        c = block_best(c, scratch);                  // equal (p,q): lowest index wins
        if (c.idx < 0) return false;
        const int row = c.idx;
        const double *x = P.T + (size_t)row * P.ld;
        Cand d = cand_none();
        for (int j = 1 + (int)threadIdx.x; j <= nreal; j += blockDim.x)
            if (fabs(x[j]) > P.eps) { d.p = 0.0; d.q = 0.0; d.idx = j; break; }
        d = block_best(d, scratch);
        if (d.idx > 0) { *r = row; *k = d.idx; return true; }
        start = row + 1;
    }
}

// The decision of one pivot: block-uniform, every thread returns the same values.
// iters is the number of pivots already applied; max_iter turns a pending pivot into
// LPGPU_ITER_LIMIT.  Shared by the large path (tableau in HBM) and the batched path (tableau in
// shared memory): P.T is a generic pointer.
struct Decision {
    int status, phase, p1, r, k;
    bool have;       // a pivot on (r, k) is to be applied
};

__device__ __forceinline__ Decision decide(const DenseParams &P, Cand *scratch, int status, int phase, int iters, int p1) {
    Decision D;
    D.r = -1; D.k = 0; D.have = false;
    while (!D.have && status == LPGPU_RUNNING) {
        if (P.method == LPGPU_DUAL) {
            const Cand l = dual_leave(P, scratch);
            if (l.idx < 0) { status = LPGPU_OPTIMAL; continue; }
            D.r = l.idx;
            const Cand e = dual_enter(P, D.r, scratch);
            if (e.idx < 0) { status = LPGPU_INFEASIBLE; continue; }
            D.k = e.idx;
            D.have = true;
            continue;
        }
        if (phase == PH_DRIVEOUT) {
            if (driveout_find(P, scratch, &D.r, &D.k)) D.have = true;
            else phase = PH_MAIN;
            continue;
        }
        Cand e;
        if (P.method == LPGPU_BIGM) e = price_bigm(P, scratch);
        else if (phase == PH_ONE) e = price_row(P, P.m + 1, P.n, scratch);
        else e = price_row(P, P.m, P.n - (P.method == LPGPU_TWOPHASE ? P.n_art : 0), scratch);
        if (e.idx < 0) {
            if (phase == PH_ONE) {
                if (art_basic_positive(P)) status = LPGPU_INFEASIBLE;
                else { phase = PH_DRIVEOUT; p1 = iters; }
                continue;
            }
            if (P.method == LPGPU_BIGM && art_basic_positive(P)) status = LPGPU_INFEASIBLE;
            else status = LPGPU_OPTIMAL;
            continue;
        }
        D.k = e.idx;
        Cand l = ratio_test(P, D.k, scratch);
        if (l.idx < 0) { status = LPGPU_UNBOUNDED; continue; }
        D.r = l.idx;
        D.have = true;
    }
    if (D.have && iters >= P.max_iter) { status = LPGPU_ITER_LIMIT; D.have = false; }
    D.status = status; D.phase = phase; D.p1 = p1;
    return D;
}

}  // namespace lpgpu
