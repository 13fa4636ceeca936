// dfma_peak.cu -- measured FP64 FMA throughput of the device (register-only dependent-free chains), as a function of the
// number of resident warps per SM.  The pass of the blocked path (blk_flush_*_kernel) runs 2 K flops per element; this is
// the ceiling its FP64 share is compared with.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/dfma_peak tools/dfma_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ACC>
__global__ void dfma_kernel(double *out, double f, double p, int iters) {
    double a[ACC];
#pragma unroll
    for (int i = 0; i < ACC; ++i) a[i] = (double)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ACC; ++i) a[i] = __fma_rn(-f, p, a[i]);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < ACC; ++i) s += a[i];
    if (s == 123.456) out[0] = s;
}

// outer-product shape of the pass: R rows x 4 columns per thread, a[r][c] = fma(-f[r], p[c], a[r][c]); operands from
// registers (SMEM = 0) or fetched from shared memory every step as the pass does (SMEM = 1: f broadcast, p per lane)
template <int R, int SMEM>
__global__ void outer_kernel(double *out, const double *src, int iters) {
    __shared__ __align__(16) double sf[32 * 64], sp[32 * 128];
    for (int i = threadIdx.x; i < 32 * 64; i += blockDim.x) sf[i] = src[i % 97] * 1e-9;
    for (int i = threadIdx.x; i < 32 * 128; i += blockDim.x) sp[i] = src[i % 89] * 1e-9;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double a[R][4];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) a[r][c] = (double)(threadIdx.x + r * 4 + c);
    double f[R], p[4];
#pragma unroll
    for (int r = 0; r < R; ++r) f[r] = sf[(warp * R + r) % 64];
#pragma unroll
    for (int c = 0; c < 4; ++c) p[c] = sp[2 * lane + (c & 1) + 64 * (c >> 1)];
    for (int it = 0; it < iters; ++it) {
        const int s = it & 31;
        if (SMEM) {
#pragma unroll
            for (int c = 0; c < 4; c += 2) {
                const double2 v = *reinterpret_cast<const double2 *>(sp + s * 128 + 2 * lane + 64 * (c >> 1));
                p[c] = v.x; p[c + 1] = v.y;
            }
#pragma unroll
            for (int r = 0; r < R; r += 2) {
                const double2 v = *reinterpret_cast<const double2 *>(sf + s * 64 + ((warp * R) % 64) + r);
                f[r] = v.x; f[r + 1] = v.y;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) a[r][c] = __fma_rn(-f[r], p[c], a[r][c]);
    }
    double sum = 0.0;
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) sum += a[r][c];
    if (sum == 123.456) out[0] = sum;
}

template <int R, int SMEM>
static void run_outer(int sms, int threads, int ctas_per_sm, int iters) {
    double *out, *src;
    cudaMalloc(&out, 8);
    cudaMalloc(&src, 128 * 8);
    double h[128];
    for (int i = 0; i < 128; ++i) h[i] = 1.0 + i * 0.001;
    cudaMemcpy(src, h, sizeof h, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    outer_kernel<R, SMEM><<<sms * ctas_per_sm, threads>>>(out, src, 10);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    outer_kernel<R, SMEM><<<sms * ctas_per_sm, threads>>>(out, src, iters);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * R * 4 * (double)iters * threads * sms * ctas_per_sm;
    printf("outer %2d x 4, operands from %s, threads %4d x %d CTAs/SM (%2d warps/SM): %8.3f ms  %7.2f TFLOP/s  %6.2f DFMA/clk/SM (%s)\n", R,
           SMEM ? "shared memory" : "registers    ", threads, ctas_per_sm, threads / 32 * ctas_per_sm, ms, flops / ms / 1e9,
           flops / 2 / (ms * 1e-3) / sms / 1.965e9, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(src);
}

template <int ACC>
static void run(int sms, int threads, int ctas_per_sm, int iters) {
    double *out;
    cudaMalloc(&out, 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    dfma_kernel<ACC><<<sms * ctas_per_sm, threads>>>(out, 1.0000001, 0.9999999, 10);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    dfma_kernel<ACC><<<sms * ctas_per_sm, threads>>>(out, 1.0000001, 0.9999999, iters);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * ACC * (double)iters * threads * sms * ctas_per_sm;
    printf("acc %2d  threads %4d x %d CTAs/SM (%2d warps/SM): %8.3f ms  %7.2f TFLOP/s  %6.2f DFMA/clk/SM at 1.965 GHz\n", ACC, threads, ctas_per_sm,
           threads / 32 * ctas_per_sm, ms, flops / ms / 1e9, flops / 2 / (ms * 1e-3) / sms / 1.965e9);
    cudaFree(out);
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    printf("%s, %d SMs\n", prop.name, sms);
    const int iters = 20000;
    run<16>(sms, 128, 1, iters);
    run<16>(sms, 128, 2, iters);
    run<16>(sms, 128, 3, iters);
    run<16>(sms, 128, 4, iters);
    run<16>(sms, 256, 4, iters);
    run<16>(sms, 1024, 2, iters);
    run<64>(sms, 128, 1, iters);
    run<64>(sms, 128, 3, iters);
    run<32>(sms, 256, 2, iters);
    run_outer<16, 0>(sms, 128, 3, iters);
    run_outer<16, 1>(sms, 128, 3, iters);
    run_outer<16, 1>(sms, 128, 1, iters);
    run_outer<8, 0>(sms, 256, 2, iters);
    run_outer<8, 1>(sms, 256, 2, iters);
    run_outer<8, 1>(sms, 128, 4, iters);
    run_outer<8, 1>(sms, 128, 2, iters);
    run_outer<4, 1>(sms, 512, 1, iters);
    return 0;
}
