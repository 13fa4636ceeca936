// dense_block_mma.cuh -- the pass that applies a block of pivots, with the FP64 matrix instruction (DMMA) as the FMA engine.
//
// blk_flush_sp_kernel (dense_block.cuh) runs every tableau element through the block's K fused multiply-adds with scalar
// DFMAs: three 64-bit register operands per FMA, two of them fetched from shared memory every step.  That caps the FP64
// pipe at about 45 of its 64 DFMA / clock / SM (tools/dfma_peak.cu, tools/chain_bench.cu), and at K = 32 the pass is
// as long on the FP64 pipe as on HBM -- it is bounded by both and reaches neither.
//
// mma.sync.aligned.m8n8k4.f64 (SASS DMMA.8x8x4) computes, per output element,
//     d = fma(a3, b3, fma(a2, b2, fma(a1, b1, fma(a0, b0, c))))
// -- the CHAIN of IEEE fused multiply-adds in ascending k, one rounding per step; tools/dmma_probe.cu checks exactly that,
// bit for bit, on 6.4e8 random elements per run incl. cancellation and zero-padded k (0 mismatches on the B200), and every
// parity test of the blocked path checks it again against the oracle.  With A = -F[s..s+3][rows] (8 x 4) and
// B = P[s..s+3][cols] (4 x 8) one instruction therefore IS four consecutive steps of the pivot chain on an 8 x 8 patch
// of the tableau: same operands, same order, same roundings as the rank-1 path -- not a re-associated contraction.
// A thread feeds 1 + 1 operand registers per 8 FMAs instead of 3 per FMA, so the FP64 side runs near its nominal rate and
// the pass becomes what the one-pass-per-pivot kernel already is: bounded by HBM alone.
//
// Steps are padded to a multiple of 4 with F = +0, P = +0: fma(-0, +0, c) = c for every c (signed zeros included).
// Layout: CTA tile 64 rows x 128 columns, 4 warps stacked over the rows (16 rows each); a warp's tile is two halves of
// 16 rows x 64 columns = 2 x 8 MMA patches (32 accumulator doubles per half).  While the chain runs on one half the
// loads of the other half -- of this tile or of the CTA's next tile -- are in flight (as in blk_flush_sp_kernel).
// Accumulator fragment of patch (i, j): thread (g = lane / 4, q = lane % 4) holds row 8 i + g, columns 8 j + 2 q, +1 --
// one 128-bit access; a warp-level access touches 8 rows x 64 contiguous bytes (whole 32-byte sectors).
// Operand tiles in shared memory carry 4 doubles of padding per step row: a fragment load is a 64-bit access of lane (g, q) at
// word address q * stride + 2 g, served per half warp (g = 0..3 / 4..7); with stride = 8 (mod 32) words the four step rows of a
// half warp fall into four disjoint groups of 8 banks -- conflict-free.  (8 doubles of padding, stride = 16 (mod 32) words,
// put steps q and q + 2 on the same banks: ncu showed 26 M bank conflicts in 48 M shared-memory wavefronts and the pass
// stalled on mio_throttle / short_scoreboard at 12 % issue utilisation, profiles/r2_mma_pass_pad8_raw.csv.)
#pragma once

#include "dense_block.cuh"

namespace lpgpu {

constexpr int MM_FS = FL_ROWS + 4;     // doubles per step row of the staged multipliers
constexpr int MM_PS = FL_COLS + 4;     // doubles per step row of the staged pivot-row segments
constexpr size_t MM_SMEM_BYTES = ((size_t)BLK_KMAX * MM_PS + 2 * (size_t)BLK_KMAX * MM_FS) * 8;
constexpr int MM_NI = 2, MM_NJ = 8;    // MMA patches per half: 2 (rows) x 8 (columns)

__device__ __forceinline__ void dmma_8x8x4(double2 &d, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(d.x), "+d"(d.y) : "d"(a), "d"(b));
}

// row0: first row of the warp's 16; c0: first column (inside the tableau) of the half
__device__ __forceinline__ void mm_load_half(const FlushParams &U, int row0, size_t c0, int ncols_left, int g, int q, double2 (&acc)[MM_NI][MM_NJ], uint64_t pol) {
#pragma unroll
    for (int i = 0; i < MM_NI; ++i) {
        const int row = row0 + 8 * i + g;
        const double *tp = U.Tsrc + (size_t)row * U.ld + c0 + 2 * q;
        const bool rowok = row < U.rows;
#pragma unroll
        for (int j = 0; j < MM_NJ; ++j)
            acc[i][j] = (rowok && 8 * j + 2 * q < ncols_left) ? ld_stream2_hint(tp + 8 * j, pol) : make_double2(0.0, 0.0);
    }
}

__device__ __forceinline__ void mm_store_half(const FlushParams &U, int row0, size_t c0, int ncols_left, int g, int q, double2 (&acc)[MM_NI][MM_NJ], uint64_t pol) {
#pragma unroll
    for (int i = 0; i < MM_NI; ++i) {
        const int row = row0 + 8 * i + g;
        double *tp = U.T + (size_t)row * U.ld + c0 + 2 * q;
        const bool rowok = row < U.rows;
#pragma unroll
        for (int j = 0; j < MM_NJ; ++j)
            if (rowok && 8 * j + 2 * q < ncols_left) st_stream2_hint(tp + 8 * j, acc[i][j], pol);
    }
}

// kc4 = steps rounded up to a multiple of 4 (the padding rows of sF / sP are zero)
// fw: staged multipliers, already offset to (step q, this warp's first row + g); pw: staged pivot rows at (step q, first column of the half + g)
__device__ __forceinline__ void mm_chain_half(int kc4, const double *fw, const double *pw, double2 (&acc)[MM_NI][MM_NJ]) {
    for (int s0 = 0; s0 < kc4; s0 += 4) {
        double a[MM_NI], b[MM_NJ];
#pragma unroll
        for (int i = 0; i < MM_NI; ++i) a[i] = -fw[(size_t)s0 * MM_FS + 8 * i];
#pragma unroll
        for (int j = 0; j < MM_NJ; ++j) b[j] = pw[(size_t)s0 * MM_PS + 8 * j];
#pragma unroll
        for (int i = 0; i < MM_NI; ++i)
#pragma unroll
            for (int j = 0; j < MM_NJ; ++j) dmma_8x8x4(acc[i][j], a[i], b[j]);
    }
}

__global__ void __launch_bounds__(FL_THREADS, 3) blk_flush_mma_kernel(FlushParams U) {
    extern __shared__ __align__(128) unsigned char fl_smem[];
    double *sP = reinterpret_cast<double *>(fl_smem);                                 // [BLK_KMAX][MM_PS]
    double *sF = sP + (size_t)BLK_KMAX * MM_PS;                                       // [2][BLK_KMAX][MM_FS]
    __shared__ __align__(8) uint64_t bar_p, bar_f[2];
    __shared__ long s_task[2];

    if (threadIdx.x == 0) {
        mbar_init(&bar_p, 1);
        mbar_init(&bar_f[0], 1);
        mbar_init(&bar_f[1], 1);
        fence_mbar_init();
    }
    pdl_wait();
    int *ctr = U.task_ctr + (U.seq & 1u);
    if (blockIdx.x == 0 && threadIdx.x == 0) U.task_ctr[(U.seq & 1u) ^ 1u] = 0;
    if (U.flag) {       // pipelined blocks: CTAs that land on the SMs reserved for the step cluster leave (see blk_flush_ring_kernel)
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        const int want = (int)(U.seq + 1u);
        if ((U.resv[(smid >> 5) & 7u] >> (smid & 31u)) & 1u) {
            __shared__ int s_leave;
            if (threadIdx.x == 0) {
                unsigned long long t0, now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
                bool leave = false;
                do {
                    leave = (*(volatile int *)U.flag - want) >= 0;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                } while (!leave && now - t0 < 20000ull);
                s_leave = leave ? 1 : 0;
            }
            __syncthreads();
            if (s_leave) return;
        } else if (threadIdx.x == 0) atomicMax(U.flag, want);
    }
    const int kc = *U.kc_ptr;
    if (kc <= 0 && U.Tsrc == U.T) return;
    const int kc4 = (kc + 3) & ~3;
    // padding steps (kc .. kc4-1): exact zeros, written once -- the bulk copies below only ever fill steps < kc
    for (int e = (int)threadIdx.x; e < (kc4 - kc) * MM_PS; e += FL_THREADS) sP[(size_t)kc * MM_PS + e] = 0.0;
    for (int e = (int)threadIdx.x; e < (kc4 - kc) * MM_FS; e += FL_THREADS) {
        sF[(size_t)kc * MM_FS + e] = 0.0;
        sF[(size_t)(BLK_KMAX + kc) * MM_FS + e] = 0.0;
    }
    fence_proxy_async();

    const int nct = (int)((U.ld + FL_COLS - 1) / FL_COLS);
    const int nrt = (U.rows + FL_ROWS - 1) / FL_ROWS;
    const int nsb = (nrt + U.rb - 1) / U.rb;
    const long ntasks = (long)nct * nsb;
    const int lane = threadIdx.x & 31, wr = threadIdx.x >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t FBYTES = (uint32_t)(kc * FL_ROWS * 8);
    auto issue_f = [&](int rt, int fb) {
        mbar_arrive_expect_tx(&bar_f[fb], FBYTES);
        for (int s = 0; s < kc; ++s)
            tma_load_1d(sF + ((size_t)fb * BLK_KMAX + s) * MM_FS, U.Fs + (size_t)s * U.pcn + (size_t)rt * FL_ROWS, (uint32_t)(FL_ROWS * 8), &bar_f[fb]);
    };
    auto tile_of = [&](long tt, int k) {             // k-th tile of task tt
        FlTile t;
        t.j0 = (size_t)(tt % nct) * FL_COLS;
        t.ncols = (int)min((size_t)FL_COLS, U.ld - t.j0);
        t.rt = (int)(tt / nct) * U.rb + k;
        return t;
    };

    if (threadIdx.x == 0) s_task[0] = (long)atomicAdd(ctr, 1);
    __syncthreads();
    uint32_t par_p = 0, par_f = 0;
    const uint64_t pol = l2_policy(U.l2first ? 1 : 0);
    double2 acc0[MM_NI][MM_NJ], acc1[MM_NI][MM_NJ];    // left and right half (64 columns each) of the warp's 16 rows
    bool pre = false;                                  // acc0 already holds (is being loaded with) the left half of the coming tile
    for (int it = 0;; ++it) {
        const long tt = s_task[it & 1];
        if (tt >= ntasks) break;
        FlTile t = tile_of(tt, 0);
        const int rt0 = t.rt, rt1 = min(rt0 + U.rb, nrt);
        if (threadIdx.x == 0) {
            s_task[(it & 1) ^ 1] = (long)atomicAdd(ctr, 1);            // next task; visible after the barriers below
            mbar_arrive_expect_tx(&bar_p, (uint32_t)(kc * t.ncols * 8));
            for (int s = 0; s < kc; ++s) tma_load_1d(sP + (size_t)s * MM_PS, U.Pm + (size_t)s * U.ld + t.j0, (uint32_t)(t.ncols * 8), &bar_p);
            issue_f(rt0, 0);
        }
        for (int rt = rt0; rt < rt1; ++rt) {
            const int fb = (rt - rt0) & 1;
            t.rt = rt;
            const int row0 = rt * FL_ROWS + 16 * wr;
            if (!pre) mm_load_half(U, row0, t.j0, t.ncols, g, q, acc0, pol);
            mm_load_half(U, row0, t.j0 + 64, t.ncols - 64, g, q, acc1, pol);     // in flight while the chain runs on the left half
            if (threadIdx.x == 0 && rt + 1 < rt1) issue_f(rt + 1, fb ^ 1);
            if (rt == rt0) { mbar_wait(&bar_p, par_p); par_p ^= 1; }
            mbar_wait(&bar_f[fb], (par_f >> fb) & 1u);
            par_f ^= 1u << fb;
            const double *fw = sF + ((size_t)fb * BLK_KMAX + q) * MM_FS + 16 * wr + g;
            const double *pw = sP + (size_t)q * MM_PS + g;
            mm_chain_half(kc4, fw, pw, acc0);
            mm_store_half(U, row0, t.j0, t.ncols, g, q, acc0, pol);
            // the tile after this one (same task, or the first tile of the next task once its claim is visible): its left half
            // is requested now and arrives while the chain runs on the right half
            FlTile nx = t;
            pre = true;
            if (rt + 1 < rt1) nx.rt = rt + 1;
            else {
                const long nt = s_task[(it & 1) ^ 1];
                pre = (rt > rt0) && nt < ntasks;       // the claim was published before tile rt0's closing barrier
                if (pre) nx = tile_of(nt, 0);
            }
            if (pre) mm_load_half(U, nx.rt * FL_ROWS + 16 * wr, nx.j0, nx.ncols, g, q, acc0, pol);
            mm_chain_half(kc4, fw, pw + 64, acc1);
            mm_store_half(U, row0, t.j0 + 64, t.ncols - 64, g, q, acc1, pol);
            __syncthreads();       // sF[fb] may be refilled (two tiles ahead), and at the last tile sP and s_task
        }
    }
}

}  // namespace lpgpu
