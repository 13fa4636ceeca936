// common.cuh -- shared device helpers for the sm_100a simplex kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lpgpu.h"

namespace lpgpu {

// ------------------------------------------------------------------ device state
// phases of the pivot state machine
enum { PH_MAIN = 0, PH_ONE = 1, PH_DRIVEOUT = 2 };

struct PivotState {
    int status;        // LPGPU_* status
    int iters;         // pivots done
    int phase;         // PH_*
    int phase1_iters;
    int r;             // leaving row of the pivot to apply
    int k;             // entering column (1-based)
    int do_pivot;      // 1: the elimination kernel that follows has work
    int budget;        // pivots still allowed in this run call
    int k_next;        // lookahead: entering column of the NEXT pivot (0 = unknown -> full decision)
    int cur;           // which half of colbuf holds the pivot column of the pivot to apply
    int blk_t;         // blocked path: pivots decided since the tableau was last brought up to date
    int blk_seq;       // blocked path: sequence number of the current block (stamps pivot rows)
};

// ------------------------------------------------------------------ candidates
// Both pricing (max reduced cost) and the ratio test (min b/a) are "maximise (p, q), ties
// to the lowest index" reductions: the ratio test feeds q = -(b/a), an exact negation.
struct Cand {
    double p, q;
    int idx;           // -1 = none
};

__device__ __forceinline__ bool cand_better(const Cand &a, const Cand &b) {
    if (a.idx < 0) return false;
    if (b.idx < 0) return true;
    if (a.p != b.p) return a.p > b.p;
    if (a.q != b.q) return a.q > b.q;
    return a.idx < b.idx;
}

__device__ __forceinline__ Cand cand_none() {
    Cand c;
    c.p = 0.0; c.q = 0.0; c.idx = -1;
    return c;
}

__device__ __forceinline__ Cand cand_shfl_xor(const Cand &c, int mask) {
    Cand o;
    o.p = __shfl_xor_sync(0xffffffffu, c.p, mask);
    o.q = __shfl_xor_sync(0xffffffffu, c.q, mask);
    o.idx = __shfl_xor_sync(0xffffffffu, c.idx, mask);
    return o;
}

__device__ __forceinline__ Cand warp_best(Cand c) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        Cand o = cand_shfl_xor(c, s);
        if (cand_better(o, c)) c = o;
    }
    return c;
}

// Block-wide best candidate: warp shuffles, then a tree over the per-warp winners in
// shared memory.  scratch must hold blockDim.x/32 entries.  Result valid in all threads.
__device__ __forceinline__ Cand block_best(Cand c, Cand *scratch) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
    c = warp_best(c);
    __syncthreads();                       // scratch may still be read from a previous call
    if (lane == 0) scratch[warp] = c;
    __syncthreads();
    Cand v = (lane < nwarp) ? scratch[lane] : cand_none();
    v = warp_best(v);
    return v;
}

// ------------------------------------------------------------------ pivot-row scaling
// a / piv with one IEEE division; exact zeros (most of a tableau row early on) skip the divide
// sequence, whose slow path they would otherwise drag the whole warp through
__device__ __forceinline__ double div_skip_zero(double a, double piv) {
    return (a == 0.0) ? ((piv > 0.0) ? a : -a) : a / piv;
}

// Several quotients by ONE pivot without one divide sequence each: with y = RN(1/piv) (__drcp_rn) two
// Markstein correction steps  q0 = RN(a*y), r = fma(-q, piv, a), q = fma(r, y, q)  reproduce the
// correctly rounded IEEE quotient a / piv bit for bit -- checked against a / piv on 4.3e10 operand
// pairs incl. adversarial significands (tools/div_selftest.cu: 0 mismatches, even after one step).
// The steps are branch-free, so the quotients of a pivot row overlap instead of queueing behind
// each other's slow-path checks.  Operands outside [1e-90, 1e90] (never seen in a tableau, but legal
// input) are reported through in_range so the caller falls back to the plain division.
struct PivotRecip {
    double piv, y;
};

__device__ __forceinline__ PivotRecip pivot_recip(double piv) {
    PivotRecip R;
    R.piv = piv;
    R.y = __drcp_rn(piv);
    return R;
}

__device__ __forceinline__ bool div_in_range(double a) {
    const double t = fabs(a);
    return a == 0.0 || (t > 1e-90 && t < 1e90);
}

__device__ __forceinline__ double div_by_pivot(double a, const PivotRecip &R) {
    double q = __dmul_rn(a, R.y);
    q = __fma_rn(__fma_rn(-q, R.piv, a), R.y, q);
    q = __fma_rn(__fma_rn(-q, R.piv, a), R.y, q);
    return (a == 0.0) ? ((R.piv > 0.0) ? a : -a) : q;          // exact signed zero of the IEEE quotient
}

// ------------------------------------------------------------------ synthetic data
// Counter-based generator shared bit-for-bit with the host (and restated in the oracle).
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__host__ __device__ __forceinline__ double uniform64(uint64_t seed, uint64_t lp, uint64_t elem,
                                                     double lo, double hi) {
    uint64_t h = mix64(mix64(mix64(seed) ^ lp) ^ elem);
    double u = (double)(h >> 11) * 0x1.0p-53;
#ifdef __CUDA_ARCH__
    return __fma_rn(u, hi - lo, lo);
#else
    return __builtin_fma(u, hi - lo, lo);
#endif
}

// Element (i, j) of the synthetic feasible tableau (m+1) x (n+1), n = ns + m.
__host__ __device__ __forceinline__ double synth_elem(uint64_t seed, uint64_t lp, int m, int ns, int i, int j) {
    const uint64_t nA = (uint64_t)m * (uint64_t)ns;
    if (i < m) {
        if (j == 0) return uniform64(seed, lp, nA + (uint64_t)i, 1.0, 2.0) * (double)ns;
        if (j <= ns) return uniform64(seed, lp, (uint64_t)i * (uint64_t)ns + (uint64_t)(j - 1), 0.1, 1.1);
        return (j - 1 - ns == i) ? 1.0 : 0.0;
    }
    if (j >= 1 && j <= ns) return uniform64(seed, lp, nA + (uint64_t)m + (uint64_t)(j - 1), 0.1, 1.1);
    return 0.0;
}

// ------------------------------------------------------------------ TMA / mbarrier PTX
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

// make mbarrier.init visible to the async proxy (TMA unit)
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}

// Bounded: a protocol error (a copy that never lands) traps after about 2^28 polls -- seconds -- instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    for (uint32_t spin = 0; spin < (1u << 28); ++spin) {
        uint32_t ok;
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (ok) return;
    }
    __trap();
}

// 1-D bulk asynchronous copy global -> shared through the TMA unit (SASS: UBLKCP).
// dst, src 16-byte aligned; bytes a multiple of 16.  Completion is counted on `bar`.
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// 1-D bulk asynchronous copy shared -> global.
__device__ __forceinline__ void tma_store_1d(void *dst, const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ------------------------------------------------------------------ cluster entry barrier (split)
// Distributed shared memory of another CTA may only be touched once that CTA has started.  Every cluster kernel
// arrives at kernel entry and waits right before its first possible remote access: by then all CTAs of the
// cluster have long been scheduled, so the wait costs nothing on the critical path, and the rule holds.
__device__ __forceinline__ void cluster_entry_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_entry_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ programmatic dependent launch
// Kernels launched with cudaLaunchAttributeProgrammaticStreamSerialization may become resident while
// the kernel before them in the stream is still draining; pdl_wait() blocks until that kernel has
// completed and its writes are visible (a no-op in a normal launch), pdl_launch_dependents() lets the
// kernel AFTER this one start its own prologue early.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ------------------------------------------------------------------ streaming 128-bit access
// The tableau is read once and written once per pivot: keep it out of L1.
__device__ __forceinline__ double2 ld_stream2(const double *p) {
    double2 v;
    asm volatile("ld.global.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream2(double *p, double2 v) {
    asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

// ------------------------------------------------------------------ L2 eviction priorities
// kind: 0 = normal, 1 = evict first (data streamed once), 2 = evict last (small working set re-read all the time)
__device__ __forceinline__ uint64_t l2_policy(int kind) {
    uint64_t pol;
    if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ double2 ld_stream2_hint(const double *p, uint64_t pol) {
    double2 v;
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void st_stream2_hint(double *p, double2 v, uint64_t pol) {
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(v.x), "d"(v.y), "l"(pol) : "memory");
}
__device__ __forceinline__ double ld_hint(const double *p, uint64_t pol) {
    double v;
    asm volatile("ld.global.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void st_hint(double *p, double v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;" ::"l"(p), "d"(v), "l"(pol) : "memory");
}

}  // namespace lpgpu
