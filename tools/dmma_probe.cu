// dmma_probe.cu -- does mma.sync.m8n8k4.f64 on sm_100a perform, per output element, the CHAIN of IEEE fused multiply-adds
//   d = fma(a3, b3, fma(a2, b2, fma(a1, b1, fma(a0, b0, c))))          (k ascending, one rounding per step)?
// If it does, a block of K pivots can be applied to a tableau tile by K/4 such instructions and leave exactly the bits
// of the K rank-1 eliminations done one after the other -- with 1 A and 1 B operand register per
// 8 FMAs instead of the 3 x 64-bit operands per DFMA that cap the FP64 pipe at about 45 of 64 DFMA / clock / SM.
// Part 1 compares bits on random operands (several exponent spreads, cancellation cases); part 2 measures the rate.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/dmma_probe tools/dmma_probe.cu && tools/dmma_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// random double: full random mantissa, sign, exponent uniform in [-spread, spread]
__device__ double rnd(uint64_t key, int spread) {
    const uint64_t h = mix64(key);
    const double mant = 1.0 + (double)(h >> 12) * 0x1.0p-52;
    const int e = spread ? (int)(mix64(h) % (uint64_t)(2 * spread + 1)) - spread : 0;
    const double v = ldexp(mant, e);
    return (h & 1) ? -v : v;
}

// counts[0] = tiles, [1] = elements that differ from the ascending chain, [2] = from the descending chain,
// [3] = elements where ascending and descending chains differ from each other (the test discriminates there)
__global__ void exact_kernel(uint64_t seed, int spread, int mode, unsigned long long *counts) {
    const int lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
    const uint64_t tile = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    __shared__ double sA[8][8][4], sB[8][4][8];            // per warp of the block (8 warps)
    const int w = threadIdx.x >> 5;
    // A[row g][k q], B[k q][col g]
    double a = rnd(seed ^ (tile * 1315423911ull + 17 * lane + 1), spread);
    double b = rnd(seed ^ (tile * 2654435761ull + 31 * lane + 7), spread);
    double c0 = rnd(seed ^ (tile * 40503ull + 2 * lane + 3), spread), c1 = rnd(seed ^ (tile * 40503ull + 2 * lane + 4), spread);
    if (mode == 2 && q >= 2) { a = 0.0; }                  // zero padding of the k dimension (a = +0 -> -0 below)
    a = -a;                                                // the pass feeds A = -F
    sA[w][g][q] = a;
    sB[w][q][g] = b;
    __syncwarp();
    if (mode == 1) {            // cancellation: c = -(a0 b0) (1 + tiny), so that the later products meet a small accumulator
        c0 = -(sA[w][g][0] * sB[w][0][2 * q]) * (1.0 + 0x1.0p-30 * (double)(lane + 1));
        c1 = -(sA[w][g][0] * sB[w][0][2 * q + 1]) * (1.0 - 0x1.0p-33 * (double)(lane + 1));
    }
    double d0 = c0, d1 = c1;
    dmma(d0, d1, a, b);
    // reference chains for this thread's two outputs: row g, cols 2q, 2q+1
    double asc[2] = {c0, c1}, desc[2] = {c0, c1};
    for (int e = 0; e < 2; ++e) {
        const int col = 2 * q + e;
        for (int k = 0; k < 4; ++k) asc[e] = __fma_rn(sA[w][g][k], sB[w][k][col], asc[e]);
        for (int k = 3; k >= 0; --k) desc[e] = __fma_rn(sA[w][g][k], sB[w][k][col], desc[e]);
    }
    unsigned long long bad_asc = 0, bad_desc = 0, differ = 0;
    const double d[2] = {d0, d1};
    for (int e = 0; e < 2; ++e) {
        bad_asc += __double_as_longlong(d[e]) != __double_as_longlong(asc[e]);
        bad_desc += __double_as_longlong(d[e]) != __double_as_longlong(desc[e]);
        differ += __double_as_longlong(asc[e]) != __double_as_longlong(desc[e]);
    }
    if (lane == 0) atomicAdd(&counts[0], 1ull);
    if (bad_asc) atomicAdd(&counts[1], bad_asc);
    if (bad_desc) atomicAdd(&counts[2], bad_desc);
    if (differ) atomicAdd(&counts[3], differ);
}

// rate: every thread owns NT 8x8 output tiles (2 accumulators each); per iteration one DMMA per tile.
// SM = 0: operands stay in registers; 1: A (RA frags) and B (RB frags) re-read from shared memory every iteration,
// as the pass would (warp tile = 8*RA rows x 8*RB columns, NT = RA*RB).
template <int RA, int RB, int SM>
__global__ void __launch_bounds__(128) rate_kernel(int iters, double *out) {
    constexpr int NT = RA * RB;
    __shared__ double sF[16][72], sP[16][136];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, g = lane >> 2, q = lane & 3;
    for (int i = threadIdx.x; i < 16 * 72; i += blockDim.x) sF[0][i] = 1e-9 * (i % 97);
    for (int i = threadIdx.x; i < 16 * 136; i += blockDim.x) sP[0][i] = 1e-3 * (i % 101);
    __syncthreads();
    double acc[NT][2];
#pragma unroll
    for (int t = 0; t < NT; ++t) { acc[t][0] = 1.0 + t; acc[t][1] = 2.0 + t + lane; }
    double a[RA], b[RB];
#pragma unroll
    for (int i = 0; i < RA; ++i) a[i] = 1e-9 * (lane + i);
#pragma unroll
    for (int j = 0; j < RB; ++j) b[j] = 1e-3 * (lane + 3 * j);
    for (int it = 0; it < iters; ++it) {
        const int s0 = (it & 3) * 4;
        if (SM) {
#pragma unroll
            for (int i = 0; i < RA; ++i) a[i] = -sF[s0 + q][(8 * i + g + 16 * w) & 63];
#pragma unroll
            for (int j = 0; j < RB; ++j) b[j] = sP[s0 + q][8 * j + g];
        }
#pragma unroll
        for (int i = 0; i < RA; ++i)
#pragma unroll
            for (int j = 0; j < RB; ++j) dmma(acc[i * RB + j][0], acc[i * RB + j][1], a[i], b[j]);
    }
    double s = 0.0;
#pragma unroll
    for (int t = 0; t < NT; ++t) s += acc[t][0] + acc[t][1];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// the DFMA form of the same work (16 x 4 thread tile, operands from shared memory) for comparison in the same run
__global__ void __launch_bounds__(128) dfma_rate_kernel(int iters, double *out) {
    __shared__ double sF[16][64], sP[16][128];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 16 * 64; i += blockDim.x) sF[0][i] = 1e-9 * (i % 97);
    for (int i = threadIdx.x; i < 16 * 128; i += blockDim.x) sP[0][i] = 1e-3 * (i % 101);
    __syncthreads();
    double acc[16][4];
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 1.0 + r + c + lane;
    for (int it = 0; it < iters; ++it) {
        const int s = it & 15;
        const double2 p0 = *reinterpret_cast<const double2 *>(&sP[s][2 * lane]), p1 = *reinterpret_cast<const double2 *>(&sP[s][64 + 2 * lane]);
#pragma unroll
        for (int r = 0; r < 16; r += 2) {
            const double2 f = *reinterpret_cast<const double2 *>(&sF[s][(16 * w + r) & 63]);
            acc[r][0] = __fma_rn(-f.x, p0.x, acc[r][0]); acc[r][1] = __fma_rn(-f.x, p0.y, acc[r][1]);
            acc[r][2] = __fma_rn(-f.x, p1.x, acc[r][2]); acc[r][3] = __fma_rn(-f.x, p1.y, acc[r][3]);
            acc[r + 1][0] = __fma_rn(-f.y, p0.x, acc[r + 1][0]); acc[r + 1][1] = __fma_rn(-f.y, p0.y, acc[r + 1][1]);
            acc[r + 1][2] = __fma_rn(-f.y, p1.x, acc[r + 1][2]); acc[r + 1][3] = __fma_rn(-f.y, p1.y, acc[r + 1][3]);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) s += acc[r][c];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    printf("device: %s, %d SMs\n", prop.name, prop.multiProcessorCount);
    unsigned long long *counts;
    cudaMalloc(&counts, 4 * sizeof(unsigned long long));
    unsigned long long total_bad = 0;
    for (int mode = 0; mode < 3; ++mode)
        for (int spread : {0, 4, 30, 200}) {
            cudaMemset(counts, 0, 4 * sizeof(unsigned long long));
            exact_kernel<<<32768, 256>>>(0xC0FFEEull + 977 * spread + mode, spread, mode, counts);
            unsigned long long h[4];
            cudaMemcpy(h, counts, sizeof h, cudaMemcpyDeviceToHost);
            printf("exact mode %d spread %3d: %llu tiles, %llu elements; != ascending chain %llu, != descending chain %llu, (asc != desc on %llu)\n",
                   mode, spread, h[0], h[0] * 64, h[1], h[2], h[3]);
            total_bad += h[1];
        }
    printf("DMMA == ascending chain of IEEE FMAs: %s\n", total_bad == 0 ? "YES (0 mismatches)" : "NO");
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }

    double *out;
    cudaMalloc(&out, (size_t)148 * 16 * 128 * sizeof(double));
    const int iters = 20000;
    for (int cps : {1, 2, 3, 4}) {       // CTAs of 4 warps per SM
        const int grid = prop.multiProcessorCount * cps;
        auto report = [&](const char *name, float ms, double fma_per_thread_iter) {
            const double flops = 2.0 * fma_per_thread_iter * iters * 128.0 * grid;
            printf("%-34s %2d warps/SM: %7.3f ms  %6.2f TFLOP/s  %5.1f FMA/clk/SM @1.965GHz\n", name, 4 * cps, ms, flops / ms / 1e9,
                   flops / 2.0 / (ms * 1e-3) / prop.multiProcessorCount / 1.965e9);
        };
        report("dmma 4x8 tiles (32x64) regs", time_ms([&] { rate_kernel<4, 8, 0><<<grid, 128>>>(iters, out); }), 32 * 8);
        report("dmma 4x8 tiles (32x64) smem", time_ms([&] { rate_kernel<4, 8, 1><<<grid, 128>>>(iters, out); }), 32 * 8);
        report("dmma 8x4 tiles (64x32) smem", time_ms([&] { rate_kernel<8, 4, 1><<<grid, 128>>>(iters, out); }), 32 * 8);
        report("dmma 2x16 tiles (16x128) smem", time_ms([&] { rate_kernel<2, 16, 1><<<grid, 128>>>(iters, out); }), 32 * 8);
        report("dmma 2x8 tiles (16x64) smem", time_ms([&] { rate_kernel<2, 8, 1><<<grid, 128>>>(iters, out); }), 16 * 8);
        report("dfma 16x4 thread tile smem", time_ms([&] { dfma_rate_kernel<<<grid, 128>>>(4 * iters, out); }), 64 * 4);
    }
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
