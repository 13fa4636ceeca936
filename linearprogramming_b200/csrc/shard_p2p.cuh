// shard_p2p.cuh -- peer-memory transport of the column-sharded path: the per-pivot exchanges are plain
// stores into the other GPUs' memory over NVLink / NVSwitch, issued from inside the kernels that
// produce the data, plus a sequence-numbered flag per message; no collective library call, no extra
// launch on the critical path except a one-warp resolve and a one-warp publish.
//
// Every rank owns one "exchange region" (a single cudaMalloc, exported with cudaIpcGetMemHandle and
// mapped by every peer with cudaIpcOpenMemHandle):
//
//   pairs[2][64]      32-byte candidates (p, q, global column, 0), slot = [iteration parity][sender]
//   pair_flag[2][64]  u32, = iteration + 1 once the sender's candidate of that iteration has landed
//   col_flag[2]       u32, = iteration + 1 once the entering column of that iteration has landed
//   cols[2][pcn]      the entering column, written by its owner during the elimination pass
//
// Iterations are counted identically on every rank (the host loops are replicas), double buffering
// by parity is enough because a rank can run at most one iteration ahead of its slowest peer: it
// cannot leave resolve(i) before every peer has posted candidate(i).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lpgpu {

constexpr int XMAXW = 64;                       // ranks
constexpr int LPGPU_EXCHANGE_TIMEOUT = 5;       // internal status: a peer never posted (see shard.cu)

struct XLayout {
    __host__ __device__ static size_t pairs(int par) { return (size_t)par * XMAXW * 32; }
    __host__ __device__ static size_t pair_flag(int par) { return 4096 + (size_t)par * XMAXW * 4; }
    __host__ __device__ static size_t col_flag(int par) { return 4608 + (size_t)par * 128; }
    __host__ __device__ static size_t cols(int par, size_t pcn) { return 5120 + (size_t)par * pcn * 8; }
    // flag-in-payload ("LL") exchange of the blocked path: every 64-bit word carries 32 bits of payload and the 32-bit sequence
    // number of its message, so a message needs no fence and no separate flag (see ll_put / ll_wait below)
    __host__ __device__ static size_t ll_base(size_t pcn) { return (5120 + 2 * pcn * 8 + 255) / 256 * 256; }
    __host__ __device__ static size_t ll_pairs(int par, size_t pcn) { return ll_base(pcn) + (size_t)par * XMAXW * 64; }     // [sender][8 words]
    __host__ __device__ static size_t ll_cols(int par, size_t pcn) { return ll_base(pcn) + 2 * XMAXW * 64 + (size_t)par * pcn * 16; }   // [row][2 words]
    __host__ __device__ static size_t bytes(size_t pcn) { return ll_base(pcn) + 2 * XMAXW * 64 + 2 * pcn * 16; }
};

template <typename T>
__device__ __forceinline__ T *xptr(void *region, size_t off) {
    return reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(region) + off);
}

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// ---- flag-in-payload words.  A release flag after the data makes the sender wait until its data stores have been acknowledged
// by the peer (a fence per message: 5 us for a candidate, 12 us for a column at N = 8, profiles/r2_shard_step_phases_n8.txt).
// Here each aligned 64-bit store is its own message: (sequence number << 32) | 32 payload bits -- single-copy atomic, so the
// reader that sees the expected sequence number in a word has that word's payload, whatever the order the words arrive in.
__device__ __forceinline__ void ll_put(unsigned long long *p, unsigned data, unsigned seq) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(((unsigned long long)seq << 32) | data) : "memory");
}
__device__ __forceinline__ void ll_put_f64(unsigned long long *p2, double v, unsigned seq) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    ll_put(p2, (unsigned)b, seq);
    ll_put(p2 + 1, (unsigned)(b >> 32), seq);
}
// waits until the word carries `seq`; a peer that died makes the kernel trap after about 30 s instead of hanging the GPU
__device__ __forceinline__ unsigned ll_wait(const unsigned long long *p, unsigned seq) {
    unsigned long long t0 = 0;
    for (unsigned spin = 0;; ++spin) {
        unsigned long long w;
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
        if ((unsigned)(w >> 32) == seq) return (unsigned)w;
        if ((spin & 1023u) == 1023u) {
            const unsigned long long now = global_ns();
            if (!t0) t0 = now;
            else if (now - t0 > 30ull * 1000 * 1000 * 1000) __trap();
        }
    }
}
__device__ __forceinline__ double ll_wait_f64(const unsigned long long *p2, unsigned seq) {
    const unsigned lo = ll_wait(p2, seq), hi = ll_wait(p2 + 1, seq);
    return __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
}

// Spin until *flag >= want (sequence numbers only grow).  Gives up after `timeout_ns` so that a peer
// that died cannot hang this GPU; returns false in that case.
__device__ __forceinline__ bool spin_until(const unsigned *flag, unsigned want, unsigned long long timeout_ns) {
    if ((int)(ld_acquire_sys(flag) - want) >= 0) return true;
    const unsigned long long t0 = global_ns();
    while ((int)(ld_acquire_sys(flag) - want) < 0) {
        __nanosleep(64);
        if (global_ns() - t0 > timeout_ns) return false;
    }
    return true;
}

}  // namespace lpgpu
