// batched_regx.cuh -- batched path, register-resident kernels for everything that is not the plain method:
// two objective rows (Big-M, two-phase incl. the drive-out), the dual simplex, and Bland's rule.
//
// Same layout as batched_reg_kernel (batched_reg.cuh): one LP per CTA of WARPS warps, the tableau in REGISTERS for the
// whole pivot loop (warp w, lane l: a[w*RT + i][1 + l + 32c]; every warp a replica of the objective row(s); lane i of
// warp w the b / basis entry of row w*RT + i), shared memory only for the entering column, the pivot row that crosses
// warps and the staging window the TMA unit prefetches the CTA's next LP into.  What differs is the decision:
//   MODE 1 (R = 2): lexicographic (snap(sigmaM), sigma) pricing for Big-M; phase one on row m+1, infeasibility test,
//          drive-out pivots, main phase without the artificial columns for the two-phase method;
//   MODE 2 (dual): leaving row = most negative b, entering column = min sigma_j / a_rj over a_rj < -eps;
//   BL: Bland's rule -- the lowest eligible column enters.
// The rule is DESIGN.md section 3, the operations are those of decide() (dense_select.cuh) on the same operands in the
// same order, so these kernels leave the bits of the shared-memory kernel and of the oracle (parity tests: force_smem
// as the cross-check).  Round 1 sent all of these to the shared-memory kernel, 5-6x slower.
// Reference gap: both cases of Source/simplex.c:58-62 and the menu stub Source/router.c:32-34.
#pragma once

#include "batched_reg.cuh"
#include "block_common.cuh"

namespace lpgpu {

// scratch of the cross-warp reductions: two sets, used alternately (a set is reused only after another barrier)
template <int WARPS>
struct RegxRed {
    unsigned long long key[2][WARPS];
    int idx[2][WARPS];
    double pay[2][2 * WARPS];
};

// CTA-wide arg-max / arg-min on an order-preserving key, ties to the lowest idx; the winner's two payload doubles come
// back in every thread.  WARPS == 1: warp reductions only, no barrier.
template <int WARPS>
__device__ __forceinline__ int regx_block_arg(unsigned long long key, bool has, int idx, bool want_max, double &p0, double &p1,
                                              RegxRed<WARPS> &red, int &nred) {
    constexpr unsigned FULL = 0xffffffffu;
    const int w = warp_arg_reduce(key, has, idx, want_max);
    const unsigned who = __ballot_sync(FULL, has && idx == w);
    const int sl = who ? __ffs(who) - 1 : 0;
    p0 = __shfl_sync(FULL, p0, sl);
    p1 = __shfl_sync(FULL, p1, sl);
    if constexpr (WARPS == 1) return w;
    else {
        key = __shfl_sync(FULL, key, sl);
        const int set = nred & 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        ++nred;
        if (lane == 0) { red.key[set][warp] = want_max ? key : ~key; red.idx[set][warp] = w; red.pay[set][2 * warp] = p0; red.pay[set][2 * warp + 1] = p1; }
        __syncthreads();
        int best = -1, bw = 0;
        unsigned long long bk = 0;
#pragma unroll
        for (int q = 0; q < WARPS; ++q) {
            const int i2 = red.idx[set][q];
            const unsigned long long k2 = red.key[set][q];
            if (i2 >= 0 && (best < 0 || k2 > bk || (k2 == bk && i2 < best))) { best = i2; bk = k2; bw = q; }
        }
        p0 = red.pay[set][2 * bw];
        p1 = red.pay[set][2 * bw + 1];
        return best;
    }
}

// pricing over a lane's CT entries and across the warp; mode: PRICE_COST / PRICE_BIGM / PRICE_PHASE1 (block_common.cuh);
// columns above jmax are not priced.  Returns the entering column (1-based) or -1.
template <int CT, bool BL>
__device__ __forceinline__ int regx_price(const double (&obj)[CT], const double (&objm)[CT], int mode, int jmax, double eps, int lane) {
    constexpr unsigned FULL = 0xffffffffu;
    unsigned long long kp = 0, kq = 0;
    int k = -1;
#pragma unroll
    for (int c = 0; c < CT; ++c) {
        const int j = 1 + lane + 32 * c;
        if (j > jmax) continue;
        unsigned long long tp = 0, tq;
        bool ok;
        if (mode == PRICE_BIGM) {
            double p = objm[c];
            if (fabs(p) <= eps) p = 0.0;
            ok = p > 0.0 || (p == 0.0 && obj[c] > eps);
            tp = BL ? 0ull : key_ordered(p);
            tq = BL ? 1ull : key_ordered(obj[c]);
        } else {
            const double q = (mode == PRICE_PHASE1) ? objm[c] : obj[c];
            ok = q > eps;
            tq = BL ? 1ull : key_ordered(q);
        }
        if (ok && (k < 0 || tp > kp || (tp == kp && tq > kq))) { kp = tp; kq = tq; k = j; }      // ascending j: ties stay on the lowest
    }
    // two-key warp arg-max, ties to the lowest column
    bool in = k >= 0;
    if (__ballot_sync(FULL, in) == 0u) return -1;
    if (mode == PRICE_BIGM && !BL) {
        const unsigned h = __reduce_max_sync(FULL, in ? (unsigned)(kp >> 32) : 0u);
        in = in && (unsigned)(kp >> 32) == h;
        const unsigned l = __reduce_max_sync(FULL, in ? (unsigned)kp : 0u);
        in = in && (unsigned)kp == l;
    }
    const unsigned h = __reduce_max_sync(FULL, in ? (unsigned)(kq >> 32) : 0u);
    in = in && (unsigned)(kq >> 32) == h;
    const unsigned l = __reduce_max_sync(FULL, in ? (unsigned)kq : 0u);
    in = in && (unsigned)kq == l;
    return (int)__reduce_min_sync(FULL, in ? (unsigned)k : 0x7fffffffu);
}

template <int WARPS, int RT, int CT, int MODE, bool BL>
__global__ void __launch_bounds__(WARPS * 32, (WARPS == 1 ? 10 : 3)) batched_regx_kernel(BatchedParams B) {
    using Cfg = RegCfg<WARPS, RT, CT>;
    constexpr bool R2 = (MODE == 1), DUALM = (MODE == 2);
    constexpr int RR = R2 ? 2 : 1;
    constexpr unsigned FULL = 0xffffffffu;
    static_assert(RT % 2 == 0 && RT <= 32, "rows per warp: even (128-bit column reads), one b per lane");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int m = B.m, n = B.n, ld1 = n + 1;
    const size_t tsz = (size_t)(m + RR) * (size_t)ld1;
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw);
    double *stage = reinterpret_cast<double *>(smem_raw + 16);
    __shared__ __align__(16) double pcs[Cfg::MAX_M];                // entering column
    __shared__ __align__(16) double prs[Cfg::MAX_N + 2];            // a row that crosses warps (unscaled or scaled), or the x scatter
    __shared__ RegxRed<WARPS> red;
    static_assert(Cfg::MAX_N + 2 >= Cfg::MAX_M, "x scatter fits the row buffer");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row0 = warp * RT;
    const int myrow = row0 + lane;
    const bool has_row = (lane < RT) && (myrow < m);
    const double eps = B.eps;
    const int nreal = n - B.n_art;
    const bool bigm = R2 && B.method == LPGPU_BIGM;

    auto block_sync = [&]() {
        if constexpr (WARPS == 1) __syncwarp();
        else __syncthreads();
    };
    auto issue_load = [&](long p) {
        const size_t gbyte = (size_t)p * tsz * sizeof(double);
        const size_t gal = gbyte & ~(size_t)15;
        const uint32_t nbytes = (uint32_t)(((gbyte - gal) + tsz * sizeof(double) + 15) / 16 * 16);
        mbar_arrive_expect_tx(bar, nbytes);
        tma_load_1d(stage, reinterpret_cast<const unsigned char *>(B.T_all) + gal, nbytes, bar);
    };

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
        if ((long)blockIdx.x < B.n_lp) issue_load(blockIdx.x);
    }
    block_sync();

    uint32_t parity = 0;
    int nred = 0;
    for (long p = blockIdx.x; p < B.n_lp; p += gridDim.x) {
        mbar_wait(bar, parity);
        parity ^= 1;
        const double *S = stage + (int)((((size_t)p * tsz * sizeof(double)) & 15) / sizeof(double));
        double v[RT][CT], obj[CT], objm[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            const int j = 1 + lane + 32 * c;
            const bool jc = j <= n;
#pragma unroll
            for (int i = 0; i < RT; ++i) v[i][c] = (jc && row0 + i < m) ? S[(size_t)(row0 + i) * ld1 + j] : 0.0;
            obj[c] = jc ? S[(size_t)m * ld1 + j] : 0.0;
            objm[c] = (R2 && jc) ? S[(size_t)(m + 1) * ld1 + j] : 0.0;
        }
        double bi = has_row ? S[(size_t)myrow * ld1] : 0.0;
        double zc = S[(size_t)m * ld1], zcm = R2 ? S[(size_t)(m + 1) * ld1] : 0.0;
        int bas = has_row ? B.basis_all[(size_t)p * m + myrow] : 0;
        fence_proxy_async();
        block_sync();
        if (tid == 0 && p + gridDim.x < B.n_lp) issue_load(p + gridDim.x);

        int status = LPGPU_RUNNING, iters = 0;
        int phase = (R2 && B.method == LPGPU_TWOPHASE) ? PH_ONE : PH_MAIN;
        int *plog = B.log ? B.log + (size_t)p * 2 * B.log_cap : nullptr;
        // the unscaled row `row` for every warp: its owner writes it to prs, one barrier (WARPS == 1: registers only)
        auto fetch_row = [&](int row, double (&rv)[CT]) {
            const int rw_ = row / RT;
            if constexpr (WARPS == 1) reg_row_get<0, RT, RT, CT>(v, row, rv);
            else {
                __syncthreads();                        // earlier readers of prs are done
                if (warp == rw_) {
                    double t[CT];
                    reg_row_get<0, RT, RT, CT>(v, row - rw_ * RT, t);
#pragma unroll
                    for (int c = 0; c < CT; ++c) prs[lane + 32 * c] = t[c];
                }
                __syncthreads();
#pragma unroll
                for (int c = 0; c < CT; ++c) rv[c] = prs[lane + 32 * c];
            }
        };
        while (true) {
            const int pmode = bigm ? PRICE_BIGM : (phase == PH_ONE ? PRICE_PHASE1 : PRICE_COST);
            const int jmax = (R2 && !bigm && phase != PH_ONE) ? nreal : n;       // two-phase main phase: no artificial columns
            int k = -1, r = -1;
            bool forced = false;                       // the row is given (dual, drive-out): no ratio test
            double rraw[CT];                           // forced: the unscaled row r
            double d0 = 0.0, d1 = 0.0;
            if constexpr (DUALM) {
                // ---- leaving row: most negative b_i < -eps, ties to the lowest row
                r = regx_block_arg<WARPS>(key_ordered(bi), has_row && bi < -eps, myrow, false, d0, d1, red, nred);
                if (r < 0) { status = LPGPU_OPTIMAL; break; }
                fetch_row(r, rraw);
                // ---- entering column: min sigma_j / a_rj over a_rj < -eps, ties to the lowest column (every warp: replicas)
                unsigned long long bk = 0;
                int kk = -1;
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                    const int j = 1 + lane + 32 * c;
                    if (j <= n && rraw[c] < -eps) {
                        const unsigned long long t = ~key_ordered(obj[c] / rraw[c]);       // arg-min = arg-max of the complement
                        if (kk < 0 || t > bk) { bk = t; kk = j; }
                    }
                }
                k = warp_arg_reduce(bk, kk >= 0, kk, true);
                if (k < 0) { status = LPGPU_INFEASIBLE; break; }
                forced = true;
            } else if (phase == PH_DRIVEOUT) {
                // ---- drive-out: lowest row whose basic variable is artificial and that has a non-artificial entry |a| > eps
                int start = 0;
                while (true) {
                    r = regx_block_arg<WARPS>(1ull, has_row && bas > nreal && myrow >= start, myrow, true, d0, d1, red, nred);
                    if (r < 0) break;
                    fetch_row(r, rraw);
                    int kk = -1;
#pragma unroll
                    for (int c = CT - 1; c >= 0; --c) {
                        const int j = 1 + lane + 32 * c;
                        if (j <= nreal && fabs(rraw[c]) > eps) kk = j;
                    }
                    k = warp_arg_reduce(1ull, kk >= 0, kk, true);           // equal keys: the lowest column
                    if (k > 0) break;
                    start = r + 1;
                }
                if (r < 0) { phase = PH_MAIN; continue; }
                forced = true;
            } else {
                k = regx_price<CT, BL>(obj, objm, pmode, jmax, eps, lane);
                if (k < 0) {
                    // nothing prices in: optimal -- unless Big-M or phase one left an artificial basic at a positive level, or
                    // phase one is over (drive-out next)
                    if (bigm || phase == PH_ONE) {
                        const int bad = regx_block_arg<WARPS>(1ull, has_row && bas > nreal && bi > eps, myrow, true, d0, d1, red, nred);
                        if (bad >= 0) { status = LPGPU_INFEASIBLE; break; }
                    }
                    if (phase == PH_ONE) { phase = PH_DRIVEOUT; continue; }
                    status = LPGPU_OPTIMAL;
                    break;
                }
            }
            // ---- entering column: its owner lane spills it; sigma_k (and sigmaM_k) by shuffle
            const int kc = (k - 1) >> 5, kl = (k - 1) & 31;
            double sig = 0.0, sigm = 0.0;
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (kc == c) {                                   // warp-uniform
                    sig = obj[c];
                    sigm = objm[c];
                    if (lane == kl) {
#pragma unroll
                        for (int i = 0; i < RT; ++i) pcs[row0 + i] = v[i][c];
                    }
                }
            sig = __shfl_sync(FULL, sig, kl);
            sigm = __shfl_sync(FULL, sigm, kl);
            __syncwarp();
            const double mine = has_row ? pcs[myrow] : 0.0;     // a_ik of this lane's row
            double piv, brp;
            if (!forced) {
                // ---- ratio test: min b_i / a_ik over a_ik > eps, ties to the lowest row; the winner carries a_rk along
                const bool elig = mine > eps;
                const double theta = ratio_quotient(bi, mine, elig);
                d0 = theta; d1 = mine;
                r = regx_block_arg<WARPS>(key_ordered(theta), elig, myrow, false, d0, d1, red, nred);
                if (r < 0) { status = LPGPU_UNBOUNDED; break; }
                brp = d0; piv = d1;
            } else {
                d0 = bi; d1 = mine;
                regx_block_arg<WARPS>(1ull, has_row && myrow == r, myrow, true, d0, d1, red, nred);      // (b_r, a_rk) to everyone
                piv = d1;
                brp = div_skip_zero(d0, piv);
            }
            if (iters >= B.max_iter) { status = LPGPU_ITER_LIMIT; break; }
            const int rw = r / RT, ri = r - rw * RT;
            // ---- pivot-row scaling a_rj / a_rk (one IEEE division each), then to every warp
            double prow[CT];
            if constexpr (WARPS == 1) {
                if (!forced) reg_row_get<0, RT, RT, CT>(v, ri, rraw);
#pragma unroll
                for (int c = 0; c < CT; ++c) prow[c] = div_skip_zero(rraw[c], piv);
            } else {
                if (!forced) fetch_row(r, rraw);
#pragma unroll
                for (int c = 0; c < CT; ++c) prow[c] = div_skip_zero(rraw[c], piv);
            }
            // ---- objective rows, b, z; then the rank-1 elimination out of registers
#pragma unroll
            for (int c = 0; c < CT; ++c) {
                obj[c] = __fma_rn(-sig, prow[c], obj[c]);
                if (R2) objm[c] = __fma_rn(-sigm, prow[c], objm[c]);
            }
            constexpr int EB = (RT % 8 == 0) ? 8 : 2;
#pragma unroll
            for (int i0 = 0; i0 < RT; i0 += EB) {
                double2 f2[EB / 2];
#pragma unroll
                for (int u = 0; u < EB / 2; ++u) f2[u] = *reinterpret_cast<const double2 *>(pcs + row0 + i0 + 2 * u);
#pragma unroll
                for (int u = 0; u < EB / 2; ++u)
#pragma unroll
                    for (int c = 0; c < CT; ++c) {
                        v[i0 + 2 * u][c] = __fma_rn(-f2[u].x, prow[c], v[i0 + 2 * u][c]);
                        v[i0 + 2 * u + 1][c] = __fma_rn(-f2[u].y, prow[c], v[i0 + 2 * u + 1][c]);
                    }
            }
            bi = __fma_rn(-mine, brp, bi);
            zc = __fma_rn(-sig, brp, zc);
            if (R2) zcm = __fma_rn(-sigm, brp, zcm);
            if (warp == rw) {
                reg_row_set<0, RT, RT, CT>(v, ri, prow);
                if (lane == ri) { bi = brp; bas = k; }
            }
            if (tid == 0 && plog && iters < B.log_cap) { plog[2 * iters] = k; plog[2 * iters + 1] = r; }
            ++iters;
            __syncwarp();          // pcs (per warp) is rewritten by the next pivot
        }

        // ---- results
        if (tid == 0) {
            B.status[p] = status;
            B.iters[p] = iters;
            if (B.z) B.z[p] = -zc;
        }
        if (B.x) {
            double *xs = prs;
            block_sync();
            for (int j = tid; j < n; j += Cfg::THREADS) xs[j] = 0.0;
            block_sync();
            if (has_row && bas >= 1 && bas <= n) xs[bas - 1] = bi;
            block_sync();
            double *xo = B.x + (size_t)p * n;
            for (int j = tid; j < n; j += Cfg::THREADS) xo[j] = xs[j];
        }
        if (B.basis_out && has_row) B.basis_out[(size_t)p * m + myrow] = bas;
        if (B.T_final) {
            double *to = B.T_final + (size_t)p * tsz;
#pragma unroll
            for (int c = 0; c < CT; ++c) {
                const int j = 1 + lane + 32 * c;
                if (j <= n) {
#pragma unroll
                    for (int i = 0; i < RT; ++i)
                        if (row0 + i < m) to[(size_t)(row0 + i) * ld1 + j] = v[i][c];
                    if (warp == 0) {
                        to[(size_t)m * ld1 + j] = obj[c];
                        if (R2) to[(size_t)(m + 1) * ld1 + j] = objm[c];
                    }
                }
            }
            if (has_row) to[(size_t)myrow * ld1] = bi;
            if (tid == 0) { to[(size_t)m * ld1] = zc; if (R2) to[(size_t)(m + 1) * ld1] = zcm; }
        }
        block_sync();          // prs / pcs are reused by the next LP of this CTA
    }
}

}  // namespace lpgpu
