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
    __host__ __device__ static size_t bytes(size_t pcn) { return 5120 + 2 * pcn * 8; }
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
