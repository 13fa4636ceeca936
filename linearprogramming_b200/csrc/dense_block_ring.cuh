// dense_block_ring.cuh -- the pass that applies a block of pivots, with the bytes in flight decoupled from the registers.
//
// blk_flush_sp_kernel and blk_flush_mma_kernel fetch their tiles with ordinary loads: a CTA's accumulators ARE its loads in
// flight, a warp that waits for them does nothing else, and both kernels end up serialising memory and arithmetic
// (ncu of the DMMA pass, profiles/r2_mma_pass_raw.csv: the DMMA pipe is busy 37 % of the time, every warp spends 85 % of
// its life waiting for its tile, 3.4 TB/s).  Here ONE persistent CTA per SM is split into two producer warps and
// RG_NW compute warps:
//   * the OPERAND producer claims tasks (column tile x row super-block, the same global counter as the other pass kernels)
//     and fetches, with TMA bulk copies, the kc pivot-row segments of the column tile (double buffered across tasks) and the
//     multipliers of each 64-row tile (RG_NF slots, kc x 512 B each);
//   * the DATA producer keeps a ring of RG_ND slabs of 16 rows x 128 columns full: one 1 KB bulk copy per row, issued by
//     16 lanes at once (one lane issuing them in turn capped the pass at 4.8 TB/s).  Nothing the producers move ever touches
//     a register, and the ring runs ahead of the arithmetic by 5 x 16 KB per SM -- more than the bandwidth-latency product of
//     the SM's share of HBM;
//   * a compute warp takes the next slab that is complete (mbarrier), pulls it into registers as DMMA accumulator
//     fragments (conflict-free LDS.128: rows are padded to 1088 B), hands the data slot back at once, runs the chain with
//     mma.sync.m8n8k4.f64 -- four consecutive steps of the pivot chain per instruction, in pivot order, one rounding per
//     step, exactly the fused multiply-adds of the rank-1 path (dense_block_mma.cuh has the argument, tools/dmma_probe.cu the
//     bit-for-bit check) -- and streams the result to the tableau with 128-bit stores that cover whole 32-byte sectors.
// At 61 .. 63 FMA / clock / SM (measured for this operand mix, tools/dmma_probe.cu) the chain of a slab takes 1 024 clocks
// of the SM, its 32 KB of HBM traffic 1 456 clocks at the measured copy peak: the pass is bounded by HBM alone, also when
// 16 or 32 SMs are left to the step kernels.  Same results as every other pass kernel, bit for bit (tests/test_gpu_parity.py).
// Reference: the loop missing from newSimplex (Source/simplex.c:40-67); layout Source/matrix.c:42-72.
#pragma once

#include "dense_block.cuh"
#include "dense_block_mma.cuh"

namespace lpgpu {

constexpr int RG_NW = 8;                          // compute warps
constexpr int RG_THREADS = (RG_NW + 2) * 32;      // + two producer warps: warp RG_NW moves the tableau, warp RG_NW + 1 the operands
constexpr int RG_SR = 16;                         // rows per slab (two DMMA row patches)
constexpr int RG_TS = FL_ROWS / RG_SR;            // slabs per row tile (the unit the multipliers are fetched in): 4
constexpr int RG_NJ = FL_COLS / 8;                // DMMA column patches per slab: 16
constexpr int RG_DS = FL_COLS + 8;                // doubles per slab row in shared memory: 1088 B = 64 (mod 128): the 8 lanes of an
                                                  // LDS.128 wavefront (2 rows x 64 B) fall on distinct banks
constexpr int RG_FS = FL_ROWS + 4;                // doubles per step row of a row tile's multipliers (stride = 8 words mod 32, see dense_block_mma.cuh)
constexpr int RG_PS = FL_COLS + 4;                // doubles per step row of the pivot-row segments
constexpr int RG_ND = 5;                          // data slots in flight
constexpr int RG_NF = 4;                          // multiplier slots (row tiles): the RG_NW slabs at work and the RG_ND in flight span at most 4 tiles
constexpr size_t RG_P_BYTES = (size_t)BLK_KMAX * RG_PS * 8;
constexpr size_t RG_D_BYTES = (size_t)RG_SR * RG_DS * 8;
constexpr size_t RG_F_BYTES = (size_t)BLK_KMAX * RG_FS * 8;
constexpr size_t RG_SMEM_BYTES = 2 * RG_P_BYTES + RG_ND * RG_D_BYTES + RG_NF * RG_F_BYTES;
static_assert(RG_SMEM_BYTES <= 227 * 1024 - 1024, "ring pass: shared memory");
static_assert((RG_NW + RG_ND + 2 * (RG_TS - 1)) / RG_TS <= RG_NF, "ring pass: the slabs at work and in flight (consecutive sequence numbers) must not span more row tiles than there are multiplier slots");

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_arrive_n(uint64_t *bar, uint32_t n) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n) : "memory");
}

__device__ __forceinline__ void tma_load_1d_hint(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
                 : "memory");
}

// a task of the pass: column tile x row super-block, cut into slabs of RG_SR rows
struct RgTask {
    size_t j0;
    int ncols, row0, nslab;
};

__device__ __forceinline__ RgTask rg_task(const FlushParams &U, long tt, int nct) {
    RgTask t;
    t.j0 = (size_t)(tt % nct) * FL_COLS;
    t.ncols = (int)min((size_t)FL_COLS, U.ld - t.j0);                  // a multiple of 16 (ld % 16 == 0)
    t.row0 = (int)(tt / nct) * U.rb * FL_ROWS;
    const int row1 = min(t.row0 + U.rb * FL_ROWS, U.rows);
    t.nslab = (row1 - t.row0 + RG_SR - 1) / RG_SR;
    return t;
}

__global__ void __launch_bounds__(RG_THREADS, 1) blk_flush_ring_kernel(FlushParams U) {
    extern __shared__ __align__(128) unsigned char fl_smem[];
    double *sP = reinterpret_cast<double *>(fl_smem);                                             // [2][BLK_KMAX][RG_PS]
    double *sD = reinterpret_cast<double *>(fl_smem + 2 * RG_P_BYTES);                           // [RG_ND][RG_SR][RG_DS]
    double *sF = reinterpret_cast<double *>(fl_smem + 2 * RG_P_BYTES + RG_ND * RG_D_BYTES);      // [RG_NF][BLK_KMAX][RG_FS]
    __shared__ __align__(8) uint64_t full_p[2], empty_p[2], full_d[RG_ND], empty_d[RG_ND], full_f[RG_NF], empty_f[RG_NF];
    __shared__ long s_task[2];
    // Slabs / row tiles the producers have issued so far.  A slot is consumed by a different warp every time round the ring,
    // and an mbarrier only tells "the phase of this parity is over" apart from "it is not" for a waiter that is at most one
    // phase away: a warp that finished early and waits for slab seq while slab seq - RG_ND has not even landed would see the
    // parity of the phase before and walk through.  So a warp first waits until its slab (tile) has been ISSUED -- which
    // implies that the slot's previous contents were consumed, i.e. had landed -- and only then for the bytes.
    __shared__ volatile unsigned s_issued_d, s_issued_f;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        s_issued_d = 0; s_issued_f = 0;
        for (int b = 0; b < 2; ++b) { mbar_init(&full_p[b], 1); mbar_init(&empty_p[b], RG_NW + 1); }     // compute warps + the data producer
        for (int b = 0; b < RG_ND; ++b) { mbar_init(&full_d[b], 1); mbar_init(&empty_d[b], 1); }
        for (int b = 0; b < RG_NF; ++b) { mbar_init(&full_f[b], 1); mbar_init(&empty_f[b], RG_TS); }
        fence_mbar_init();
    }
    pdl_wait();
    int *ctr = U.task_ctr + (U.seq & 1u);
    if (blockIdx.x == 0 && threadIdx.x == 0) U.task_ctr[(U.seq & 1u) ^ 1u] = 0;
    if (U.flag) {
        // Pipelined blocks: the SMs in U.resv are left to the step kernels of the next block.  A CTA that lands on one
        // leaves as soon as it knows that some CTA of this launch works elsewhere (any one worker finishes the pass:
        // tasks are claimed from a counter); if none shows up within 20 us it works after all -- placement only ever
        // costs time, never the result.
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
    // padding steps (kc .. kc4-1) of every operand buffer: exact zeros, written once -- the bulk copies only ever fill steps < kc
    for (int e = (int)threadIdx.x; e < (kc4 - kc) * RG_PS; e += RG_THREADS) {
        sP[(size_t)kc * RG_PS + e] = 0.0;
        sP[(size_t)(BLK_KMAX + kc) * RG_PS + e] = 0.0;
    }
    for (int e = (int)threadIdx.x; e < (kc4 - kc) * RG_FS; e += RG_THREADS)
        for (int b = 0; b < RG_NF; ++b) sF[((size_t)b * BLK_KMAX + kc) * RG_FS + e] = 0.0;
    __syncthreads();            // barriers initialised, padding written: the roles part here and meet only through mbarriers

    const int nct = (int)((U.ld + FL_COLS - 1) / FL_COLS);
    const int nrt = (U.rows + FL_ROWS - 1) / FL_ROWS;
    const long ntasks = (long)nct * ((nrt + U.rb - 1) / U.rb);
    const uint64_t pol = l2_policy(U.l2first ? 1 : 0);
    const uint64_t pol_keep = l2_policy(U.l2first ? 2 : 0);      // the slices: the step kernels re-read them all the time (evict last)
    const unsigned ND = (unsigned)min(max(U.ring_nd, 2), RG_ND);     // data slots in use
    auto wait_issued = [](volatile unsigned *ctr_issued, unsigned seq) {
        unsigned spin = 0;
        while ((int)(*ctr_issued - seq) <= 0)
            if (++spin > (1u << 28)) __trap();
    };

    if (warp == RG_NW + 1) {
        // ================================================================ operand producer: claims the tasks; pivot-row segments per
        // task, multipliers per row tile.  One lane issues (a bulk copy is one instruction per warp anyway).
        if (lane != 0) return;
        unsigned tseq = 0;                                  // row tiles issued so far by this CTA
        for (unsigned it = 0;; ++it) {
            const int pb = (int)(it & 1u);
            const long tt = (long)atomicAdd(ctr, 1);
            if (it >= 2) mbar_wait(&empty_p[pb], ((it >> 1) - 1u) & 1u);       // every consumer of the task two back is done with it
            s_task[pb] = tt;
            if (tt >= ntasks) { mbar_arrive(&full_p[pb]); break; }              // the sentinel: nothing more to come
            const RgTask t = rg_task(U, tt, nct);
            const uint32_t rowbytes = (uint32_t)(t.ncols * 8);
            mbar_arrive_expect_tx(&full_p[pb], (uint32_t)kc * rowbytes);
            for (int s = 0; s < kc; ++s)
                tma_load_1d_hint(sP + ((size_t)pb * BLK_KMAX + s) * RG_PS, U.Pm + (size_t)s * U.ld + t.j0, rowbytes, &full_p[pb], pol_keep);
            const int ntile = (t.nslab + RG_TS - 1) / RG_TS;
            for (int ti = 0; ti < ntile; ++ti, ++tseq) {
                const unsigned fs = tseq % RG_NF;
                if (tseq >= (unsigned)RG_NF) mbar_wait(&empty_f[fs], (tseq / RG_NF - 1u) & 1u);
                mbar_arrive_expect_tx(&full_f[fs], (uint32_t)(kc * FL_ROWS * 8));
                __threadfence_block();
                s_issued_f = tseq + 1u;
                const double *src = U.Fs + (size_t)(t.row0 + ti * FL_ROWS);
                for (int s = 0; s < kc; ++s)
                    tma_load_1d_hint(sF + ((size_t)fs * BLK_KMAX + s) * RG_FS, src + (size_t)s * U.pcn, (uint32_t)(FL_ROWS * 8), &full_f[fs], pol_keep);
            }
        }
        return;
    }
    if (warp == RG_NW) {
        // ================================================================ data producer: keeps the ring of slabs full
        unsigned seq = 0;                                   // slabs issued so far by this CTA
        for (unsigned it = 0;; ++it) {
            const int pb = (int)(it & 1u);
            mbar_wait(&full_p[pb], (it >> 1) & 1u);          // the task has been claimed (its number is in s_task)
            const long tt = s_task[pb];
            if (tt >= ntasks) break;
            const RgTask t = rg_task(U, tt, nct);
            const uint32_t rowbytes = (uint32_t)(t.ncols * 8);
            for (int q = 0; q < t.nslab; ++q, ++seq) {
                const unsigned ds = seq % ND;
                if (seq >= ND) mbar_wait(&empty_d[ds], (seq / ND - 1u) & 1u);
                const int row0 = t.row0 + q * RG_SR;
                const int nr = min(RG_SR, U.rows - row0);
                if (lane == 0) {
                    mbar_arrive_expect_tx(&full_d[ds], (uint32_t)nr * rowbytes);
                    __threadfence_block();
                    s_issued_d = seq + 1u;
                }
                __syncwarp();
                if (lane < nr)
                    tma_load_1d_hint(sD + ((size_t)ds * RG_SR + lane) * RG_DS, U.Tsrc + (size_t)(row0 + lane) * U.ld + t.j0, rowbytes, &full_d[ds], pol);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_p[pb]);
        }
        return;
    }

    // ==================================================================== compute warps
    const int g = lane >> 2, q4 = lane & 3;
    unsigned base = 0, tbase = 0;                           // slabs / row tiles of the tasks before this one
    for (unsigned it = 0;; ++it) {
        const int pb = (int)(it & 1u);
        mbar_wait(&full_p[pb], (it >> 1) & 1u);
        const long tt = s_task[pb];
        if (tt >= ntasks) break;
        const RgTask t = rg_task(U, tt, nct);
        const double *pw = sP + ((size_t)pb * BLK_KMAX + q4) * RG_PS + g;
        // the slabs of this task whose sequence number falls to this warp
        int q = (int)((warp + RG_NW - (int)(base % RG_NW)) % RG_NW);
        for (; q < t.nslab; q += RG_NW) {
            const unsigned seq = base + (unsigned)q, ds = seq % ND;
            const unsigned tseq = tbase + (unsigned)(q / RG_TS), fs = tseq % RG_NF;
            const int row0 = t.row0 + q * RG_SR;
            wait_issued(&s_issued_d, seq);
            mbar_wait(&full_d[ds], (seq / ND) & 1u);
            double2 acc[2][RG_NJ];
            const double *dp = sD + ((size_t)ds * RG_SR + g) * RG_DS + 2 * q4;
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < RG_NJ; ++j) acc[i][j] = *reinterpret_cast<const double2 *>(dp + (size_t)(8 * i) * RG_DS + 8 * j);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_d[ds]);          // the data slot can be refilled while the chain runs
            wait_issued(&s_issued_f, tseq);
            mbar_wait(&full_f[fs], (tseq / RG_NF) & 1u);
            const double *fw = sF + ((size_t)fs * BLK_KMAX + q4) * RG_FS + (q % RG_TS) * RG_SR + g;
#pragma unroll 2
            for (int s0 = 0; s0 < kc4; s0 += 4) {
                double a[2], b[RG_NJ];
                a[0] = -fw[(size_t)s0 * RG_FS];
                a[1] = -fw[(size_t)s0 * RG_FS + 8];
#pragma unroll
                for (int j = 0; j < RG_NJ; ++j) b[j] = pw[(size_t)s0 * RG_PS + 8 * j];
#pragma unroll
                for (int j = 0; j < RG_NJ; ++j) {
                    dmma_8x8x4(acc[0][j], a[0], b[j]);
                    dmma_8x8x4(acc[1][j], a[1], b[j]);
                }
            }
            __syncwarp();
            // one release per slab; the last slab of a task also stands in for the slabs its (partial) row tile does not have
            if (lane == 0) mbar_arrive_n(&empty_f[fs], (q == t.nslab - 1) ? (uint32_t)(RG_TS - q % RG_TS) : 1u);
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int row = row0 + 8 * i + g;
                double *tp = U.T + (size_t)row * U.ld + t.j0 + 2 * q4;
                if (row < U.rows) {
#pragma unroll
                    for (int j = 0; j < RG_NJ; ++j)
                        if (8 * j + 2 * q4 < t.ncols) st_stream2_hint(tp + 8 * j, acc[i][j], pol);
                }
            }
        }
        // a row tile whose slabs all went to other warps still counts: every slot is released RG_TS times, once per slab
        base += (unsigned)t.nslab;
        tbase += (unsigned)((t.nslab + RG_TS - 1) / RG_TS);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_p[pb]);
    }
}

}  // namespace lpgpu
