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
    int *task_ctr;          // 2 counters for the work-stealing variant, used alternately by launch parity
    unsigned seq;           // launch sequence number (host side)
    void *const *peer_x;    // column-sharded, peer-memory transport: the next entering column is stored into
    int n_peers;            // every rank's exchange region at byte offset peer_off instead of cnext
    size_t peer_off;
};

// where the owner of the next entering column leaves its freshly computed value for row i
__device__ __forceinline__ void put_next_col(const UpdateParams &U, double *cnext, int i, double val) {
    if (U.n_peers > 0) {
        for (int g = 0; g < U.n_peers; ++g)
            reinterpret_cast<double *>(reinterpret_cast<unsigned char *>(U.peer_x[g]) + U.peer_off)[i] = val;
        // no fence here (a system fence under full streaming-store load costs tens of microseconds):
        // the "column complete" flag is posted by a kernel that follows this one in stream order, and
        // kernel completion is a system-scope release of everything this kernel wrote
    } else {
        cnext[i] = val;
    }
}

constexpr int UPD_SPLIT = UPD_TILE_ROWS / UPD_UNROLL;   // row slices a leftover task is cut into

// The task list as a persistent grid walks it: `full` whole tasks (a multiple of the grid size, so
// every CTA gets the same number), then the leftover tasks of the last, partial round cut into
// UPD_SPLIT row slices each, so that the tail of a launch is a slice of UPD_UNROLL rows per CTA
// instead of a few CTAs finishing a whole task alone at a fraction of the HBM bandwidth.
struct UpdSchedule {
    int nct;        // column tiles
    long full;      // whole tasks
    long nvt;       // virtual tasks = full + UPD_SPLIT * leftover
};

__device__ __forceinline__ UpdSchedule upd_schedule(const UpdateParams &U) {
    UpdSchedule S;
    S.nct = (int)((U.ld + UPD_TILE_COLS - 1) / UPD_TILE_COLS);
    const long ntasks = (long)S.nct * ((U.rows + UPD_TILE_ROWS - 1) / UPD_TILE_ROWS);
    S.full = ntasks / gridDim.x * gridDim.x;
    S.nvt = S.full + (ntasks - S.full) * UPD_SPLIT;
    return S;
}

// virtual task -> column tile c, rows [i0, i1) (possibly empty), rows staged per pivot-column tile
__device__ __forceinline__ void upd_decode(const UpdSchedule &S, long tt, bool reverse, int rows, int &c, int &i0, int &i1) {
    const long vt = reverse ? S.nvt - 1 - tt : tt;
    long t;
    int len;
    if (vt < S.full) { t = vt; i0 = (int)(t / S.nct) * UPD_TILE_ROWS; len = UPD_TILE_ROWS; }
    else {
        const long u = vt - S.full;
        t = S.full + u / UPD_SPLIT;
        i0 = (int)(t / S.nct) * UPD_TILE_ROWS + (int)(u % UPD_SPLIT) * UPD_UNROLL;
        len = UPD_UNROLL;
    }
    c = (int)(t % S.nct);
    i1 = min(i0 + len, rows);
}

// ---- variant 2: pivot-row pair in registers straight from L2, pivot column by broadcast loads
__global__ void __launch_bounds__(UPD_THREADS, 3) dense_update_direct(UpdateParams U) {
    pdl_wait();
    pdl_launch_dependents();
    if (!U.st->do_pivot) return;
    const int r = U.st->r;
    const int cur = U.st->cur, kn = U.st->k_next;
    const double *pcol = U.colbuf + (size_t)cur * U.pcn;
    double *cnext = U.cnext ? U.cnext : U.colbuf + (size_t)(cur ^ 1) * U.pcn;
    const bool reverse = U.alternate && (U.st->iters & 1);
    const UpdSchedule S = upd_schedule(U);
    for (long tt = blockIdx.x; tt < S.nvt; tt += gridDim.x) {
        int c, i0, i1;
        upd_decode(S, tt, reverse, U.rows, c, i0, i1);
        const size_t j = (size_t)c * UPD_TILE_COLS + (size_t)threadIdx.x * 2;
        if (j >= U.ld) continue;
        const bool wb = (j == 0);                                    // this thread owns the b column
        const bool wk = (kn > 0) && (j == ((size_t)kn & ~(size_t)1));  // ... or the next entering column
        const double2 pr = *reinterpret_cast<const double2 *>(U.prow + j);
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
                    if (wk) put_next_col(U, cnext, i + u, (kn & 1) ? o.y : o.x);
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

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
    }
    pdl_wait();                 // everything above overlaps the tail of the select kernel
    pdl_launch_dependents();
    if (!U.st->do_pivot) return;
    const int r = U.st->r;
    const int cur = U.st->cur, kn = U.st->k_next;
    const double *pcol = U.colbuf + (size_t)cur * U.pcn;
    double *cnext = U.cnext ? U.cnext : U.colbuf + (size_t)(cur ^ 1) * U.pcn;
    const bool reverse = U.alternate && (U.st->iters & 1);
    const UpdSchedule S = upd_schedule(U);
    const long ntasks = S.nvt;
    __syncthreads();

    auto issue = [&](long tt, int s) {
        int c, i0, i1;
        upd_decode(S, tt, reverse, U.rows, c, i0, i1);
        const size_t j0 = (size_t)c * UPD_TILE_COLS;
        const uint32_t pbytes = (uint32_t)(min((size_t)UPD_TILE_COLS, U.ld - j0) * sizeof(double));
        // whole pivot-column tile from i0 on: the buffer is padded past the last row (pcn)
        const uint32_t cbytes = (uint32_t)(UPD_TILE_ROWS * sizeof(double));
        mbar_arrive_expect_tx(&bar[s], pbytes + cbytes);
        tma_load_1d(s_prow[s], U.prow + j0, pbytes, &bar[s]);
        tma_load_1d(s_pcol[s], pcol + i0, cbytes, &bar[s]);
    };

    long tt = blockIdx.x;
    if (tt < ntasks && threadIdx.x == 0) issue(tt, 0);
    for (int it = 0; tt < ntasks; tt += gridDim.x, ++it) {
        const int s = it & 1;
        // stage s^1 was released by the __syncthreads that ended the previous task
        if (tt + gridDim.x < ntasks && threadIdx.x == 0) issue(tt + gridDim.x, s ^ 1);

        int c, i0, i1;
        upd_decode(S, tt, reverse, U.rows, c, i0, i1);
        const size_t j = (size_t)c * UPD_TILE_COLS + (size_t)threadIdx.x * 2;
        const bool active = j < U.ld;
        const bool wb = (j == 0);
        const bool wk = (kn > 0) && (j == ((size_t)kn & ~(size_t)1));
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
                        if (wk) put_next_col(U, cnext, i + u, (kn & 1) ? o.y : o.x);
                    }
            }
        }
        __syncthreads();     // everyone is done reading stage s before it is refilled
    }
}

// ---- variant 0 (default): as variant 1, but tasks are claimed from a global counter instead of a fixed stride.
// SMs do not see the same HBM bandwidth (two dies, near / far L2 slices), so with a static schedule the
// fast ones go idle while the slow ones finish (ncu: SMs active 90 % of the launch); here a CTA that
// finishes early simply claims more tasks.  The claim for the task after next is issued one task
// ahead, so its latency and the TMA staging of that task's tiles overlap the current task.
// Counter protocol: launch `seq` claims from task_ctr[seq & 1] and zeroes task_ctr[(seq & 1) ^ 1],
// which only the previous (completed) and the next (not yet started) launch touch.
__global__ void __launch_bounds__(UPD_THREADS, 3) dense_update_dynamic(UpdateParams U) {
    __shared__ __align__(128) double s_prow[2][UPD_TILE_COLS];
    __shared__ __align__(128) double s_pcol[2][UPD_TILE_ROWS];
    __shared__ __align__(8) uint64_t bar[2];
    __shared__ long s_task[2];

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
    }
    pdl_wait();
    pdl_launch_dependents();
    int *ctr = U.task_ctr + (U.seq & 1u);
    if (blockIdx.x == 0 && threadIdx.x == 0) U.task_ctr[(U.seq & 1u) ^ 1u] = 0;
    if (!U.st->do_pivot) return;
    const int r = U.st->r;
    const int cur = U.st->cur, kn = U.st->k_next;
    const double *pcol = U.colbuf + (size_t)cur * U.pcn;
    double *cnext = U.cnext ? U.cnext : U.colbuf + (size_t)(cur ^ 1) * U.pcn;
    const bool reverse = U.alternate && (U.st->iters & 1);
    const UpdSchedule S = upd_schedule(U);
    const long ntasks = S.nvt;

    auto issue = [&](long tt, int s) {
        int c, i0, i1;
        upd_decode(S, tt, reverse, U.rows, c, i0, i1);
        const size_t j0 = (size_t)c * UPD_TILE_COLS;
        const uint32_t pbytes = (uint32_t)(min((size_t)UPD_TILE_COLS, U.ld - j0) * sizeof(double));
        const uint32_t cbytes = (uint32_t)(UPD_TILE_ROWS * sizeof(double));
        mbar_arrive_expect_tx(&bar[s], pbytes + cbytes);
        tma_load_1d(s_prow[s], U.prow + j0, pbytes, &bar[s]);
        tma_load_1d(s_pcol[s], pcol + i0, cbytes, &bar[s]);
    };

    if (threadIdx.x == 0) {
        const long t0 = (long)atomicAdd(ctr, 1);
        s_task[0] = t0;
        if (t0 < ntasks) issue(t0, 0);
    }
    __syncthreads();
    for (int it = 0;; ++it) {
        const int s = it & 1;
        const long tt = s_task[s];
        if (tt >= ntasks) break;
        // claim the task after this one now; its index is published and staged further down
        int claimed = 0;
        if (threadIdx.x == 0) claimed = atomicAdd(ctr, 1);

        int c, i0, i1;
        upd_decode(S, tt, reverse, U.rows, c, i0, i1);
        const size_t j = (size_t)c * UPD_TILE_COLS + (size_t)threadIdx.x * 2;
        const bool active = j < U.ld;
        const bool wb = (j == 0);
        const bool wk = (kn > 0) && (j == ((size_t)kn & ~(size_t)1));
        double *base = U.T + (size_t)i0 * U.ld + j;

        double2 v[UPD_UNROLL];
#pragma unroll
        for (int u = 0; u < UPD_UNROLL; ++u)
            if (active && i0 + u < i1) v[u] = ld_stream2(base + (size_t)u * U.ld);

        if (threadIdx.x == 0) {      // stage s^1 was released by the __syncthreads that ended the previous task
            s_task[s ^ 1] = claimed;
            if (claimed < ntasks) issue(claimed, s ^ 1);
        }
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
                        if (wk) put_next_col(U, cnext, i + u, (kn & 1) ? o.y : o.x);
                    }
            }
        }
        __syncthreads();     // stage s may be refilled, s_task[s^1] is visible
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
