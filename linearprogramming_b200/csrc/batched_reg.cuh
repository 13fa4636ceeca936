// batched_reg.cuh -- batched path, register-resident variant (plain method, R = 1).
//
// One LP per CTA of WARPS warps.  The tableau lives in REGISTERS for the whole pivot loop:
//   warp w, lane l holds  a[w*RT + i][1 + l + 32c]   for i < RT, c < CT      (v[RT][CT])
//   every warp holds a replica of the objective row   sigma[1 + l + 32c]      (obj[CT])
//   lane i < RT of warp w holds b[w*RT + i] and basis[w*RT + i]; every thread holds the z cell
// so a pivot is RT*CT independent DFMAs per thread with no tableau traffic at all.  Shared memory
// carries only the entering column (written by the one lane per warp that owns it, read back as
// broadcasts), the scaled pivot row when WARPS > 1, and the staging buffer into which the TMA unit
// prefetches the NEXT LP of this CTA (one cp.async.bulk per LP) while the current one is pivoting.
// With WARPS == 1 the loop has no block barrier at all: arg-max / arg-min are warp shuffles.
//
// Same rule and same operations as the shared-memory kernel and the oracle:
// a_rj' = a_rj / a_rk, a_ij = fma(-a_ik, a_rj', a_ij); ties to the lowest index.  This is the loop
// missing from newSimplex (Source/simplex.c:40-67) over SimplexMatrix-shaped tableaux
// (Source/matrix.h:13-21).
#pragma once

#include "batched_kernels.cuh"

namespace lpgpu {

template <int WARPS, int RT, int CT>
struct RegCfg {
    static constexpr int THREADS = WARPS * 32;
    static constexpr int MAX_M = WARPS * RT;
    static constexpr int MAX_N = 32 * CT;
    // scratch used inside the pivot loop (entering column; when WARPS > 1 also the scaled pivot row, the
    // per-warp ratio winners and each warp's unscaled candidate row); the x scatter of the result phase
    // reuses the same bytes
    static constexpr size_t LOOP_BYTES =
        (size_t)MAX_M * 8 + (WARPS > 1 ? ((size_t)MAX_N + 2) * 8 + (size_t)WARPS * 24 + (size_t)WARPS * MAX_N * 8 : 0);
    static constexpr size_t SCRATCH_BYTES = LOOP_BYTES > (size_t)MAX_N * 8 ? LOOP_BYTES : (size_t)MAX_N * 8;
    __host__ __device__ static size_t smem_bytes(int m, int n) {
        size_t b = 16;                                              // mbarrier
        b += ((size_t)(m + 1) * (size_t)(n + 1) + 2) * sizeof(double);  // staging window
        return (b + 15) / 16 * 16;                                  // dynamic part; the loop scratch is static
    }
};

// static-index access to "row ri" of a register array through a tree of warp-uniform branches
template <int LO, int HI, int RT, int CT>
__device__ __forceinline__ void reg_row_get(const double (&v)[RT][CT], int ri, double (&out)[CT]) {
    if constexpr (HI - LO == 1) {
#pragma unroll
        for (int c = 0; c < CT; ++c) out[c] = v[LO][c];
    } else {
        constexpr int MID = (LO + HI) / 2;
        if (ri < MID) reg_row_get<LO, MID, RT, CT>(v, ri, out);
        else reg_row_get<MID, HI, RT, CT>(v, ri, out);
    }
}

template <int LO, int HI, int RT, int CT>
__device__ __forceinline__ void reg_row_set(double (&v)[RT][CT], int ri, const double (&in)[CT]) {
    if constexpr (HI - LO == 1) {
#pragma unroll
        for (int c = 0; c < CT; ++c) v[LO][c] = in[c];
    } else {
        constexpr int MID = (LO + HI) / 2;
        if (ri < MID) reg_row_set<LO, MID, RT, CT>(v, ri, in);
        else reg_row_set<MID, HI, RT, CT>(v, ri, in);
    }
}

// Warp-wide arg-max / arg-min on 64-bit order-preserving keys with three redux.sync operations
// (high word, low word among the high-word winners, lowest index among the exact ties) instead of a
// five-level shuffle tree.  has = this lane holds a candidate; idx >= 0.  Returns the winning idx
// (the same in every lane) or -1 when no lane has a candidate.
__device__ __forceinline__ int warp_arg_reduce(unsigned long long key, bool has, int idx, bool want_max) {
    constexpr unsigned FULL = 0xffffffffu;
    if (!want_max) key = ~key;                                       // arg-min = arg-max of the complement
    const unsigned hi = has ? (unsigned)(key >> 32) : 0u, lo = (unsigned)key;
    const unsigned any = __ballot_sync(FULL, has);
    if (any == 0u) return -1;
    const unsigned mh = __reduce_max_sync(FULL, hi);
    const bool in1 = has && hi == mh;
    const unsigned ml = __reduce_max_sync(FULL, in1 ? lo : 0u);
    const bool in2 = in1 && lo == ml;
    return (int)__reduce_min_sync(FULL, in2 ? (unsigned)idx : 0x7fffffffu);
}

// positive doubles order like their bit patterns
__device__ __forceinline__ unsigned long long key_positive(double q) { return (unsigned long long)__double_as_longlong(q); }

// any finite double: flip the sign bit of non-negatives, all bits of negatives; -0 is folded into +0
__device__ __forceinline__ unsigned long long key_ordered(double x) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x + 0.0);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

// pivot-row scaling of one lane's CT entries: a_rj / a_rk, each the correctly rounded IEEE quotient
// through ONE reciprocal shared by the row (common.cuh: div_by_pivot); the plain divide sequence only
// for out-of-range operands.  (b_r / a_rk is not recomputed: it is the winning ratio of the ratio test.)
template <int CT>
__device__ __forceinline__ void scale_row(const double (&rv)[CT], double piv, double (&out)[CT]) {
    const PivotRecip R = pivot_recip(piv);
    bool ok = div_in_range(piv);
#pragma unroll
    for (int c = 0; c < CT; ++c) {
        out[c] = div_by_pivot(rv[c], R);
        ok = ok && div_in_range(rv[c]);
    }
    if (!__all_sync(0xffffffffu, ok)) {
#pragma unroll
        for (int c = 0; c < CT; ++c) out[c] = div_skip_zero(rv[c], piv);
    }
}

// theta = b / a for the ratio test.  Lanes without a candidate divide 1 by 1 and exact zeros are passed
// through (0 / a = 0 with the sign of b, a > 0), so the divide's slow path -- taken for a zero
// numerator or denominator -- is not dragged in by lanes whose quotient is never used.
__device__ __forceinline__ double ratio_quotient(double b, double a, bool elig) {
    const bool dv = elig && b != 0.0;
    double num = dv ? b : 1.0, den = dv ? a : 1.0;
    asm("" : "+d"(num), "+d"(den));      // opaque: otherwise the selects are folded away and 0 / 0 is divided after all
    const double q = num / den;
    return dv ? q : (elig ? b : 0.0);
}

// pricing: largest sigma_j > eps over this lane's CT entries, then across the warp; ties to the lowest j
template <int CT>
__device__ __forceinline__ int price_row(const double (&obj)[CT], double eps, int lane) {
    double pq = 0.0;
    int k = -1;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (obj[c] > eps && (k < 0 || obj[c] > pq)) { pq = obj[c]; k = 1 + lane + 32 * c; }
    return warp_arg_reduce(key_positive(pq), k >= 0, k, true);
}

template <int WARPS, int RT, int CT>
__global__ void __launch_bounds__(WARPS * 32, (WARPS == 1 ? 12 : WARPS == 2 ? 8 : 3)) batched_reg_kernel(BatchedParams B) {
    using Cfg = RegCfg<WARPS, RT, CT>;
    static_assert(RT % 2 == 0 && RT <= 32, "rows per warp: even (128-bit column reads), one b per lane");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int m = B.m, n = B.n, ld1 = n + 1;
    const size_t tsz = (size_t)(m + 1) * (size_t)ld1;
    // loop scratch: STATIC shared memory, so every access of the loop has a compile-time address (the
    // arithmetic of a run-time offset was being rematerialised at each of them); dynamic shared memory:
    // mbarrier | staging window
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw);
    __shared__ __align__(16) unsigned char scratch_raw[(Cfg::SCRATCH_BYTES + 15) / 16 * 16];
    double *pcs = reinterpret_cast<double *>(scratch_raw);          // entering column, MAX_M
    double *prs = pcs + Cfg::MAX_M;                                 // WARPS > 1: scaled pivot row, MAX_N (+ 2 spare)
    double *redq = prs + Cfg::MAX_N + 2;                            // WARPS > 1: per-warp ratio winners (theta, a_rk, row)
    double *redp = redq + WARPS;
    int *redi = reinterpret_cast<int *>(redp + WARPS);
    double *raws = redp + 2 * WARPS;                                // WARPS > 1: each warp's candidate row, unscaled
    double *stage = reinterpret_cast<double *>(smem_raw + 16);
    double *xs = pcs;                                               // result phase: x scatter (loop scratch is dead)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row0 = warp * RT;                    // first tableau row of this warp
    const int myrow = row0 + lane;                 // the row whose b / basis this lane keeps
    const bool has_row = (lane < RT) && (myrow < m);
    const double eps = B.eps;

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
    for (long p = blockIdx.x; p < B.n_lp; p += gridDim.x) {
        // ---- staging buffer -> registers
        mbar_wait(bar, parity);
        parity ^= 1;
        const double *S = stage + (int)((((size_t)p * tsz * sizeof(double)) & 15) / sizeof(double));
        double v[RT][CT], obj[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            const int j = 1 + lane + 32 * c;
            const bool jc = j <= n;
#pragma unroll
            for (int i = 0; i < RT; ++i) v[i][c] = (jc && row0 + i < m) ? S[(size_t)(row0 + i) * ld1 + j] : 0.0;
            obj[c] = jc ? S[(size_t)m * ld1 + j] : 0.0;
        }
        double bi = has_row ? S[(size_t)myrow * ld1] : 0.0;
        double zc = S[(size_t)m * ld1];
        int bas = has_row ? B.basis_all[(size_t)p * m + myrow] : 0;
        // everyone has read the staging buffer: hand it to the TMA unit for this CTA's next LP
        fence_proxy_async();
        block_sync();
        if (tid == 0 && p + gridDim.x < B.n_lp) issue_load(p + gridDim.x);

        int status = LPGPU_RUNNING, iters = 0;
        int *plog = B.log ? B.log + (size_t)p * 2 * B.log_cap : nullptr;
        // ---- pricing of the first pivot (every warp, replicated row); later pivots are priced inside the
        // loop, right after the objective row is updated and BEFORE the elimination's DFMAs are issued, so
        // that the reductions' latency runs under them
        int k = price_row<CT>(obj, eps, lane);
        while (true) {
            if (k < 0) { status = LPGPU_OPTIMAL; break; }
            const int kc = (k - 1) >> 5, kl = (k - 1) & 31;
            // ---- entering column: its owner lane spills it; sigma_k by shuffle
            double sig = 0.0;
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (kc == c) {                                   // warp-uniform
                    sig = obj[c];
                    if (lane == kl) {
#pragma unroll
                        for (int i = 0; i < RT; ++i) pcs[row0 + i] = v[i][c];
                    }
                }
            sig = __shfl_sync(0xffffffffu, sig, kl);
            __syncwarp();
            // ---- ratio test: min b_i / a_ik over a_ik > eps, ties to the lowest row
            const double mine = has_row ? pcs[myrow] : 0.0;
            const bool elig = mine > eps;
            const double theta = ratio_quotient(bi, mine, elig);
            int r = warp_arg_reduce(key_ordered(theta), elig, myrow, false);
            double prow[CT], brp;
            int rw, ri;
            if constexpr (WARPS == 1) {
                if (r < 0) { status = LPGPU_UNBOUNDED; break; }
                if (iters >= B.max_iter) { status = LPGPU_ITER_LIMIT; break; }
                rw = 0;
                ri = r;
                // ---- pivot-row scaling a_rj / a_rk; b_r / a_rk is the winning ratio itself
                const double piv = __shfl_sync(0xffffffffu, mine, ri);
                brp = __shfl_sync(0xffffffffu, theta, ri);
                double rv[CT];
                reg_row_get<0, RT, RT, CT>(v, ri, rv);
                scale_row<CT>(rv, piv, prow);
            } else {
                // every warp publishes its own winner (ratio, pivot element, row index and the UNSCALED row)
                // before the first barrier, so that after it the four warps can scale one 32-column chunk
                // each instead of waiting for the owner warp to divide the whole row
                const int lr = r >= 0 ? r - row0 : 0;
                const double tr = __shfl_sync(0xffffffffu, theta, lr);
                const double pv = __shfl_sync(0xffffffffu, mine, lr);
                if (r >= 0) {                                    // warp-uniform
                    double rv[CT];
                    reg_row_get<0, RT, RT, CT>(v, lr, rv);
#pragma unroll
                    for (int c = 0; c < CT; ++c) raws[warp * Cfg::MAX_N + lane + 32 * c] = rv[c];
                }
                if (lane == 0) { redq[warp] = tr; redp[warp] = pv; redi[warp] = r; }
                __syncthreads();
                // rows ascend with the warp index, so "first strictly smaller" keeps ties on the lowest row
                r = -1;
                rw = 0;
                double best = 0.0;
#pragma unroll
                for (int w = 0; w < WARPS; ++w) {
                    const int rw_ = redi[w];
                    const double tw = redq[w];
                    if (rw_ >= 0 && (r < 0 || tw < best)) { r = rw_; best = tw; rw = w; }
                }
                if (r < 0) { status = LPGPU_UNBOUNDED; break; }
                if (iters >= B.max_iter) { status = LPGPU_ITER_LIMIT; break; }
                ri = r - rw * RT;
                const double piv = redp[rw];
                brp = best;
#pragma unroll
                for (int c = 0; c < CT; ++c)
                    if (c % WARPS == warp)                       // chunk c belongs to warp c mod WARPS
                        prs[lane + 32 * c] = div_skip_zero(raws[rw * Cfg::MAX_N + lane + 32 * c], piv);
                __syncthreads();
#pragma unroll
                for (int c = 0; c < CT; ++c) prow[c] = prs[lane + 32 * c];
            }
            // ---- objective row first, then the next pivot's pricing, then the elimination
#pragma unroll
            for (int c = 0; c < CT; ++c) obj[c] = __fma_rn(-sig, prow[c], obj[c]);
            const int knext = price_row<CT>(obj, eps, lane);
            // ---- rank-1 elimination out of registers
            // entering column read back as 128-bit broadcasts, EB rows per batch so that the loads of a
            // batch are all in flight before its first DFMA needs one
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
            if (warp == rw) {
                reg_row_set<0, RT, RT, CT>(v, ri, prow);
                if (lane == ri) { bi = brp; bas = k; }
            }
            if (tid == 0 && plog && iters < B.log_cap) { plog[2 * iters] = k; plog[2 * iters + 1] = r; }
            ++iters;
            k = knext;
            // pcs (per warp) is rewritten by the next pivot; prs / redq / raws are protected by the two
            // block barriers of the next pivot, so no third one is needed
            __syncwarp();
        }

        // ---- results
        if (tid == 0) {
            B.status[p] = status;
            B.iters[p] = iters;
            if (B.z) B.z[p] = -zc;
        }
        if (B.x) {
            block_sync();      // xs aliases the loop scratch: every warp has left the pivot loop
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
                    if (warp == 0) to[(size_t)m * ld1 + j] = obj[c];
                }
            }
            if (has_row) to[(size_t)myrow * ld1] = bi;
            if (tid == 0) to[(size_t)m * ld1] = zc;
        }
        block_sync();          // xs / pcs are reused by the next LP of this CTA
    }
}

}  // namespace lpgpu
