// dense_update.cuh -- the HBM-bound rank-1 elimination pass over one large tableau.
//
//   a_ij <- fma(-a_ik, a_rj', a_ij)   for every row i != r (constraint and objective rows,
//   a_rj <- a_rj'                      b column included), a_rj' = a_rj / a_rk from prow[]
//
// This is the loop the reference never wrote (Source/simplex.c:57-63); the layout it runs
// over is SimplexMatrix's dense row-major b|A (Source/matrix.c:42-72).  0.125 flop/B: the
// bound is HBM bandwidth.  Algorithmic bytes per launch: 16 * (m+R) * (n+1).
//
// Work decomposition: tasks = (column tile of 512 doubles) x (row block of TILE_ROWS rows),
// column tile fastest, handed round-robin to a persistent grid sized to a whole number of
// resident CTAs per SM.  Each thread owns two adjacent columns (one 128-bit access per row)
// and keeps UNROLL rows in flight.  Rows start 128-byte aligned (ld % 16 == 0).
#pragma once

#include "common.cuh"

namespace lpgpu {

constexpr int UPD_THREADS = 256;
constexpr int UPD_TILE_COLS = UPD_THREADS * 2;   // 512 doubles = 4 KB per row per CTA
constexpr int UPD_TILE_ROWS = 32;
constexpr int UPD_UNROLL = 8;

struct UpdateParams {
    double *T;
    size_t ld;
    int rows;               // m + R
    const double *prow;     // ld doubles
    double *colbuf;         // 2 x pcn: half st->cur = pivot column (read), other half = next entering
                            // column as updated by this pass (written; lookahead of the select kernel)
    size_t pcn;             // doubles per half, a multiple of UPD_TILE_ROWS
    double *bnext;          // b column as updated by this pass (written)
    const PivotState *st;
    int alternate;          // 1: walk the tasks backwards on odd pivots (L2 reuse of the tail)
    double *cnext;          // where the next entering column goes; nullptr = colbuf half st->cur ^ 1
};

__device__ __forceinline__ long upd_task(long t, long ntasks, bool reverse) { return reverse ? ntasks - 1 - t : t; }

// ---- variant 0: pivot-row pair in registers straight from L2, pivot column by broadcast loads
__global__ void __launch_bounds__(UPD_THREADS, 3) dense_update_direct(UpdateParams U) {
    if (!U.st->do_pivot) return;
    const int r = U.st->r;
    const int cur = U.st->cur, kn = U.st->k_next;
    const double *pcol = U.colbuf + (size_t)cur * U.pcn;
    double *cnext = U.cnext ? U.cnext : U.colbuf + (size_t)(cur ^ 1) * U.pcn;
    const bool reverse = U.alternate && (U.st->iters & 1);
    const int nct = (int)((U.ld + UPD_TILE_COLS - 1) / UPD_TILE_COLS);
    const int nrb = (U.rows + UPD_TILE_ROWS - 1) / UPD_TILE_ROWS;
    const long ntasks = (long)nct * nrb;
    for (long tt = blockIdx.x; tt < ntasks; tt += gridDim.x) {
        const long t = upd_task(tt, ntasks, reverse);
        const int c = (int)(t % nct), rb = (int)(t / nct);
        const size_t j = (size_t)c * UPD_TILE_COLS + (size_t)threadIdx.x * 2;
        if (j >= U.ld) continue;
        const bool wb = (j == 0);                                    // this thread owns the b column
        const bool wk = (kn > 0) && (j == ((size_t)kn & ~(size_t)1));  // ... or the next entering column
        const double2 pr = *reinterpret_cast<const double2 *>(U.prow + j);
        const int i0 = rb * UPD_TILE_ROWS;
        const int i1 = min(i0 + UPD_TILE_ROWS, U.rows);
        double *base = U.T + (size_t)i0 * U.ld + j;
        for (int i = i0; i < i1; i += UPD_UNROLL, base += (size_t)UPD_UNROLL * U.ld) {
            double2 v[UPD_UNROLL];
            double pc[UPD_UNROLL];
#pragma unroll
            for (int u = 0; u < UPD_UNROLL; ++u)
                if (i + u < i1) {
                    v[u] = ld_stream2(base + (size_t)u * U.ld);
                    pc[u] = __ldg(pcol + i + u);
                }
#pragma unroll
            for (int u = 0; u < UPD_UNROLL; ++u)
                if (i + u < i1) {
                    double2 o;
                    o.x = __fma_rn(-pc[u], pr.x, v[u].x);
                    o.y = __fma_rn(-pc[u], pr.y, v[u].y);
                    if (i + u == r) o = pr;
                    st_stream2(base + (size_t)u * U.ld, o);
                    if (wb) U.bnext[i + u] = o.x;
                    if (wk) cnext[i + u] = (kn & 1) ? o.y : o.x;
                }
        }
    }
}

// ---- variant 1: pivot-row tile and pivot-column tile staged in shared memory by the TMA unit
// (cp.async.bulk + mbarrier, SASS UBLKCP), double buffered one task ahead.
__global__ void __launch_bounds__(UPD_THREADS, 3) dense_update_tma_staged(UpdateParams U) {
    __shared__ __align__(128) double s_prow[2][UPD_TILE_COLS];
    __shared__ __align__(128) double s_pcol[2][UPD_TILE_ROWS];
    __shared__ __align__(8) uint64_t bar[2];

    if (!U.st->do_pivot) return;
    const int r = U.st->r;
    const int cur = U.st->cur, kn = U.st->k_next;
    const double *pcol = U.colbuf + (size_t)cur * U.pcn;
    double *cnext = U.cnext ? U.cnext : U.colbuf + (size_t)(cur ^ 1) * U.pcn;
    const bool reverse = U.alternate && (U.st->iters & 1);
    const int nct = (int)((U.ld + UPD_TILE_COLS - 1) / UPD_TILE_COLS);
    const int nrb = (U.rows + UPD_TILE_ROWS - 1) / UPD_TILE_ROWS;
    const long ntasks = (long)nct * nrb;

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
    }
    __syncthreads();

    auto issue = [&](long tt, int s) {
        const long t = upd_task(tt, ntasks, reverse);
        const int c = (int)(t % nct), rb = (int)(t / nct);
        const size_t j0 = (size_t)c * UPD_TILE_COLS;
        const uint32_t pbytes = (uint32_t)(min((size_t)UPD_TILE_COLS, U.ld - j0) * sizeof(double));
        const uint32_t cbytes = (uint32_t)(UPD_TILE_ROWS * sizeof(double));
        mbar_arrive_expect_tx(&bar[s], pbytes + cbytes);
        tma_load_1d(s_prow[s], U.prow + j0, pbytes, &bar[s]);
        tma_load_1d(s_pcol[s], pcol + (size_t)rb * UPD_TILE_ROWS, cbytes, &bar[s]);
    };

    long tt = blockIdx.x;
    if (tt < ntasks && threadIdx.x == 0) issue(tt, 0);
    for (int it = 0; tt < ntasks; tt += gridDim.x, ++it) {
        const int s = it & 1;
        // stage s^1 was released by the __syncthreads that ended the previous task
        if (tt + gridDim.x < ntasks && threadIdx.x == 0) issue(tt + gridDim.x, s ^ 1);

        const long t = upd_task(tt, ntasks, reverse);
        const int c = (int)(t % nct), rb = (int)(t / nct);
        const size_t j = (size_t)c * UPD_TILE_COLS + (size_t)threadIdx.x * 2;
        const bool active = j < U.ld;
        const bool wb = (j == 0);
        const bool wk = (kn > 0) && (j == ((size_t)kn & ~(size_t)1));
        const int i0 = rb * UPD_TILE_ROWS;
        const int i1 = min(i0 + UPD_TILE_ROWS, U.rows);
        double *base = U.T + (size_t)i0 * U.ld + j;

        // first batch of tableau loads goes out before we block on the staged tiles
        double2 v[UPD_UNROLL];
#pragma unroll
        for (int u = 0; u < UPD_UNROLL; ++u)
            if (active && i0 + u < i1) v[u] = ld_stream2(base + (size_t)u * U.ld);

        mbar_wait(&bar[s], (uint32_t)((it >> 1) & 1));
        if (active) {
            const double2 pr = *reinterpret_cast<const double2 *>(&s_prow[s][threadIdx.x * 2]);
            const double *pc = s_pcol[s];
            for (int i = i0; i < i1; i += UPD_UNROLL, base += (size_t)UPD_UNROLL * U.ld) {
                if (i != i0) {
#pragma unroll
                    for (int u = 0; u < UPD_UNROLL; ++u)
                        if (i + u < i1) v[u] = ld_stream2(base + (size_t)u * U.ld);
                }
#pragma unroll
                for (int u = 0; u < UPD_UNROLL; ++u)
                    if (i + u < i1) {
                        const double f = -pc[i + u - i0];
                        double2 o;
                        o.x = __fma_rn(f, pr.x, v[u].x);
                        o.y = __fma_rn(f, pr.y, v[u].y);
                        if (i + u == r) o = pr;
                        st_stream2(base + (size_t)u * U.ld, o);
                        if (wb) U.bnext[i + u] = o.x;
                        if (wk) cnext[i + u] = (kn & 1) ? o.y : o.x;
                    }
            }
        }
        __syncthreads();     // everyone is done reading stage s before it is refilled
    }
}

// ---- synthetic tableau written on the device (bench "value" leg; same bits as the host)
__global__ void dense_synth_kernel(double *T, size_t ld, int m, int ns, int *basis, uint64_t seed, uint64_t lp) {
    const int n = ns + m;
    const size_t total = (size_t)(m + 1) * ld;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / ld), j = (int)(e % ld);
        T[e] = (j <= n) ? synth_elem(seed, lp, m, ns, i, j) : 0.0;
        if (j == 0 && i < m) basis[i] = 1 + ns + i;
    }
}

}  // namespace lpgpu
