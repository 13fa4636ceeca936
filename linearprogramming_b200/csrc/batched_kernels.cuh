// batched_kernels.cuh -- thousands of independent small LPs, one LP per CTA, the whole pivot
// loop resident in that CTA's shared memory.
//
// Each CTA pulls its tableau (m+R) x (n+1) from HBM with ONE bulk asynchronous copy through
// the TMA unit (cp.async.bulk + mbarrier), runs pricing / ratio test / scaling / rank-1
// elimination out of shared memory until a terminal status, and writes back status, iteration
// count, objective, x and the basis.  HBM sees each LP once on the way in and only the
// results on the way out.  The decision code is the same device function the large path uses
// (dense_select.cuh: decide), so both paths follow one rule by construction.
//
// Reference anchor: this is the loop missing from newSimplex (Source/simplex.c:40-67) applied
// to many SimplexMatrix-shaped tableaux (Source/matrix.h:13-21) at once.
#pragma once

#include "common.cuh"
#include "dense_select.cuh"

namespace lpgpu {

struct BatchedParams {
    const double *T_all;   // n_lp tableaux back to back, ld = n+1 (+16 B tail padding)
    const int *basis_all;  // n_lp * m
    long n_lp;
    int m, n, R;
    double eps;
    int max_iter, method, n_art, bland;
    int devex;             // 1: Devex reference weights (kept in shared memory; oracle pricing == 2)
    int log_cap;
    int *status, *iters;   // n_lp
    double *z;             // n_lp (may be null)
    double *x;             // n_lp * n (may be null)
    int *basis_out;        // n_lp * m (may be null)
    int *log;              // n_lp * 2*log_cap (may be null)
    double *T_final;       // n_lp tableaux (may be null)
};

constexpr int BATCH_THREADS = 128;

__host__ __device__ inline size_t batched_smem_bytes(int m, int n, int R) {
    const size_t tsz = (size_t)(m + R) * (size_t)(n + 1);
    size_t b = 16;                                   // mbarrier
    b += (tsz + 2) * sizeof(double);                 // tableau (+ lead/tail doubles of the 16 B window)
    b = (b + 15) / 16 * 16;
    b += ((size_t)n + 2) * sizeof(double);           // scaled pivot row
    b += ((size_t)m + R + 1) * sizeof(double);       // pivot column
    b += ((size_t)n + 2) * sizeof(double);           // Devex reference weights
    b += (size_t)(BATCH_THREADS / 32) * sizeof(Cand);
    b += (size_t)m * sizeof(int);                    // basis
    return (b + 15) / 16 * 16;
}

__global__ void __launch_bounds__(BATCH_THREADS) batched_simplex_kernel(BatchedParams B) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int m = B.m, n = B.n, R = B.R, rows = m + R, ld1 = n + 1;
    const size_t tsz = (size_t)rows * (size_t)ld1;
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw);
    double *tab_raw = reinterpret_cast<double *>(smem_raw + 16);
    size_t off = 16 + (tsz + 2) * sizeof(double);
    off = (off + 15) / 16 * 16;
    double *prow = reinterpret_cast<double *>(smem_raw + off);
    off += ((size_t)n + 2) * sizeof(double);
    double *pcol = reinterpret_cast<double *>(smem_raw + off);
    off += ((size_t)rows + 1) * sizeof(double);
    double *wts = reinterpret_cast<double *>(smem_raw + off);
    off += ((size_t)n + 2) * sizeof(double);
    Cand *scratch = reinterpret_cast<Cand *>(smem_raw + off);
    off += (size_t)(BATCH_THREADS / 32) * sizeof(Cand);
    int *basis = reinterpret_cast<int *>(smem_raw + off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NWARP = BATCH_THREADS / 32;

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    __syncthreads();

    uint32_t parity = 0;
    for (long p = blockIdx.x; p < B.n_lp; p += gridDim.x) {
        // ---- load: one TMA bulk copy of the 16-byte aligned window that contains the tableau
        const size_t gbyte = (size_t)p * tsz * sizeof(double);
        const size_t gal = gbyte & ~(size_t)15;
        const int lead = (int)((gbyte - gal) / sizeof(double));                 // 0 or 1
        const uint32_t nbytes = (uint32_t)(((gbyte - gal) + tsz * sizeof(double) + 15) / 16 * 16);
        if (tid == 0) {
            mbar_arrive_expect_tx(bar, nbytes);
            tma_load_1d(tab_raw, reinterpret_cast<const unsigned char *>(B.T_all) + gal, nbytes, bar);
        }
        for (int i = tid; i < m; i += BATCH_THREADS) basis[i] = B.basis_all[(size_t)p * m + i];
        if (B.devex) for (int j = tid; j <= n; j += BATCH_THREADS) wts[j] = 1.0;
        mbar_wait(bar, parity);
        parity ^= 1;
        __syncthreads();
        double *S = tab_raw + lead;

        DenseParams P;
        P.T = S; P.ld = (size_t)ld1; P.m = m; P.n = n; P.R = R; P.basis = basis;
        P.prow = prow; P.pcol = pcol; P.pcn = 0; P.bnext = nullptr; P.st = nullptr; P.log = nullptr; P.log_cap = 0;
        P.eps = B.eps; P.max_iter = B.max_iter; P.method = B.method; P.n_art = B.n_art; P.bland = B.bland;
        P.wts = B.devex ? wts : nullptr;

        int status = LPGPU_RUNNING, phase = (B.method == LPGPU_TWOPHASE) ? PH_ONE : PH_MAIN, iters = 0, p1 = 0;
        int *plog = B.log ? B.log + (size_t)p * 2 * B.log_cap : nullptr;
        while (true) {
            Decision D = decide(P, scratch, status, phase, iters, p1);
            status = D.status; phase = D.phase; p1 = D.p1;
            if (!D.have) break;
            const int r = D.r, k = D.k;
            // ---- pivot-row scaling and pivot-column snapshot
            const double piv = S[(size_t)r * ld1 + k];
            for (int i = tid; i < rows; i += BATCH_THREADS) pcol[i] = S[(size_t)i * ld1 + k];
            for (int j = tid; j <= n; j += BATCH_THREADS) prow[j] = div_skip_zero(S[(size_t)r * ld1 + j], piv);
            if (tid == 0) {
                basis[r] = k;
                if (plog && iters < B.log_cap) { plog[2 * iters] = k; plog[2 * iters + 1] = r; }
            }
            __syncthreads();
            if (B.devex) {       // reference weights follow the scaled pivot row (devex_weight: the oracle's devex_update)
                const double wk = wts[k];
                __syncthreads();
                for (int j = 1 + tid; j <= n; j += BATCH_THREADS) wts[j] = devex_weight(wts[j], prow[j], wk, j == k);
            }
            // ---- rank-1 elimination: a warp per row, lanes across columns (conflict-free)
            for (int i = warp; i < rows; i += NWARP) {
                double *row = S + (size_t)i * ld1;
                if (i == r) {
                    for (int j = lane; j <= n; j += 32) row[j] = prow[j];
                } else {
                    const double f = -pcol[i];
                    for (int j = lane; j <= n; j += 32) row[j] = __fma_rn(f, prow[j], row[j]);
                }
            }
            ++iters;
            __syncthreads();
        }

        // ---- results
        if (tid == 0) {
            B.status[p] = status;
            B.iters[p] = iters;
            if (B.z) B.z[p] = -S[(size_t)m * ld1];
        }
        if (B.x) {
            double *xo = B.x + (size_t)p * n;
            for (int j = tid; j < n; j += BATCH_THREADS) xo[j] = 0.0;
            __syncthreads();
            for (int i = tid; i < m; i += BATCH_THREADS) {
                const int j = basis[i];
                if (j >= 1 && j <= n) xo[j - 1] = S[(size_t)i * ld1];
            }
        }
        if (B.basis_out)
            for (int i = tid; i < m; i += BATCH_THREADS) B.basis_out[(size_t)p * m + i] = basis[i];
        if (B.T_final) {
            double *to = B.T_final + (size_t)p * tsz;
            for (size_t e = tid; e < tsz; e += BATCH_THREADS) to[e] = S[e];
        }
        fence_proxy_async();  // generic-proxy writes above are ordered before the next TMA write
        __syncthreads();      // shared memory is reused by the next LP of this CTA
    }
}

// synthetic feasible LPs written on the device (same bits as the host generator)
__global__ void batched_synth_kernel(double *T_all, int *basis_all, long n_lp, int m, int ns, uint64_t seed, uint64_t lp0) {
    const int n = ns + m, ld1 = n + 1;
    const size_t tsz = (size_t)(m + 1) * ld1;
    const size_t total = (size_t)n_lp * tsz;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t p = e / tsz, w = e % tsz;
        const int i = (int)(w / ld1), j = (int)(w % ld1);
        T_all[e] = synth_elem(seed, lp0 + p, m, ns, i, j);
        if (j == 0 && i < m) basis_all[p * m + i] = 1 + ns + i;
    }
}

}  // namespace lpgpu
