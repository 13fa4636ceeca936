// chain_bench.cu -- the pass's inner loop (fl_chain_rows and candidate rewrites) alone: operands resident in shared memory,
// accumulators in registers, no global traffic.  Reports DFMA / clock / SM so that loop structures can be compared with
// the 64 the pipe can issue and the 43.5 of the plain outer-product loop (tools/dfma_peak.cu).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false --expt-relaxed-constexpr -Xptxas -v -o tools/chain_bench tools/chain_bench.cu
#include <cstdio>
#include "../linearprogramming_b200/csrc/dense_block.cuh"

using namespace lpgpu;

// candidate 1: one step at a time, the multipliers of the NEXT row group requested before the DFMAs of this one
template <int NR>
__device__ __forceinline__ void chain_v1(int kc, const double *fw, const double *pw, double2 (&acc)[NR][2]) {
    for (int s = 0; s < kc; ++s) {
        double2 p[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) p[q] = *reinterpret_cast<const double2 *>(pw + (size_t)s * FL_COLS + 64 * q);
        double2 f[2], fn[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) f[u] = *reinterpret_cast<const double2 *>(fw + (size_t)s * FL_ROWS + 2 * u);
#pragma unroll
        for (int rg = 0; rg < NR; rg += 4) {
            if (rg + 4 < NR) {
#pragma unroll
                for (int u = 0; u < 2; ++u) fn[u] = *reinterpret_cast<const double2 *>(fw + (size_t)s * FL_ROWS + rg + 4 + 2 * u);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    acc[rg + 2 * u][q].x = __fma_rn(-f[u].x, p[q].x, acc[rg + 2 * u][q].x);
                    acc[rg + 2 * u][q].y = __fma_rn(-f[u].x, p[q].y, acc[rg + 2 * u][q].y);
                    acc[rg + 2 * u + 1][q].x = __fma_rn(-f[u].y, p[q].x, acc[rg + 2 * u + 1][q].x);
                    acc[rg + 2 * u + 1][q].y = __fma_rn(-f[u].y, p[q].y, acc[rg + 2 * u + 1][q].y);
                }
            asm volatile("" ::: "memory");
#pragma unroll
            for (int u = 0; u < 2; ++u) f[u] = fn[u];
        }
    }
}

// candidate 2: one step at a time, all 16 multipliers of the step requested up front (8 broadcast LDS.128), no fences
template <int NR>
__device__ __forceinline__ void chain_v2(int kc, const double *fw, const double *pw, double2 (&acc)[NR][2]) {
#pragma unroll 1
    for (int s = 0; s < kc; ++s) {
        double2 p[2], f[NR / 2];
#pragma unroll
        for (int q = 0; q < 2; ++q) p[q] = *reinterpret_cast<const double2 *>(pw + (size_t)s * FL_COLS + 64 * q);
#pragma unroll
        for (int u = 0; u < NR / 2; ++u) f[u] = *reinterpret_cast<const double2 *>(fw + (size_t)s * FL_ROWS + 2 * u);
#pragma unroll
        for (int u = 0; u < NR / 2; ++u)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                acc[2 * u][q].x = __fma_rn(-f[u].x, p[q].x, acc[2 * u][q].x);
                acc[2 * u][q].y = __fma_rn(-f[u].x, p[q].y, acc[2 * u][q].y);
                acc[2 * u + 1][q].x = __fma_rn(-f[u].y, p[q].x, acc[2 * u + 1][q].x);
                acc[2 * u + 1][q].y = __fma_rn(-f[u].y, p[q].y, acc[2 * u + 1][q].y);
            }
    }
}

// candidate 3: as the current loop (step pairs, row groups of 4) but column-major inside a group: p fixed, rows vary
template <int NR>
__device__ __forceinline__ void chain_v3(int kc, const double *fw, const double *pw, double2 (&acc)[NR][2]) {
    for (int s = 0; s + 1 < kc; s += 2) {
        double2 p[2][2];
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int q = 0; q < 2; ++q) p[e][q] = *reinterpret_cast<const double2 *>(pw + (size_t)(s + e) * FL_COLS + 64 * q);
#pragma unroll
        for (int rg = 0; rg < NR; rg += 4) {
            double2 f[2][2];
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int u = 0; u < 2; ++u) f[e][u] = *reinterpret_cast<const double2 *>(fw + (size_t)(s + e) * FL_ROWS + rg + 2 * u);
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        acc[rg + 2 * u][q].x = __fma_rn(-f[e][u].x, p[e][q].x, acc[rg + 2 * u][q].x);
                        acc[rg + 2 * u + 1][q].x = __fma_rn(-f[e][u].y, p[e][q].x, acc[rg + 2 * u + 1][q].x);
                    }
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        acc[rg + 2 * u][q].y = __fma_rn(-f[e][u].x, p[e][q].y, acc[rg + 2 * u][q].y);
                        acc[rg + 2 * u + 1][q].y = __fma_rn(-f[e][u].y, p[e][q].y, acc[rg + 2 * u + 1][q].y);
                    }
                }
            asm volatile("" ::: "memory");
        }
    }
}

template <int V, int NR>
__global__ void __launch_bounds__(128, 3) chain_kernel(double *out, const double *src, int reps) {
    extern __shared__ __align__(128) unsigned char smem[];
    double *sP = reinterpret_cast<double *>(smem);
    double *sF = sP + BLK_KMAX * FL_COLS;
    for (int i = threadIdx.x; i < BLK_KMAX * FL_COLS; i += blockDim.x) sP[i] = src[i % 89] * 1e-9;
    for (int i = threadIdx.x; i < 2 * BLK_KMAX * FL_ROWS; i += blockDim.x) sF[i] = src[i % 97] * 1e-9;
    __syncthreads();
    const int lane = threadIdx.x & 31, wr = threadIdx.x >> 5;
    double2 acc[NR][2];
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int q = 0; q < 2; ++q) acc[r][q] = make_double2(threadIdx.x + r, q + 0.5);
    const double *fw = sF + (FL_TR * wr) % FL_ROWS, *pw = sP + 2 * lane;
    for (int it = 0; it < reps; ++it) {
        if (V == 0) fl_chain_rows<NR>(32, fw, pw, acc);
        else if (V == 1) chain_v1<NR>(32, fw, pw, acc);
        else if (V == 2) chain_v2<NR>(32, fw, pw, acc);
        else chain_v3<NR>(32, fw, pw, acc);
    }
    double sum = 0.0;
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int q = 0; q < 2; ++q) sum += acc[r][q].x + acc[r][q].y;
    if (sum == 123.456) out[0] = sum;
}

template <int V, int NR>
static void run(const char *name, int sms, int ctas, int reps) {
    double *out, *src;
    cudaMalloc(&out, 8); cudaMalloc(&src, 128 * 8);
    double h[128];
    for (int i = 0; i < 128; ++i) h[i] = 1.0 + i * 0.001;
    cudaMemcpy(src, h, sizeof h, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(chain_kernel<V, NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FL_SMEM_BYTES);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    chain_kernel<V, NR><<<sms * ctas, 128, FL_SMEM_BYTES>>>(out, src, 2);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    chain_kernel<V, NR><<<sms * ctas, 128, FL_SMEM_BYTES>>>(out, src, reps);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double dfma = (double)NR * 4 * 32 * reps * 128.0 * sms * ctas;
    printf("%-44s rows/thread %2d, %d CTAs/SM: %8.3f ms %7.2f TFLOP/s %6.2f DFMA/clk/SM (%s)\n", name, NR, ctas, ms, 2 * dfma / ms / 1e9,
           dfma / (ms * 1e-3) / sms / 1.965e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount, reps = 400;
    run<0, 16>("fl_chain_rows (current: step pairs, fences)", sms, 3, reps);
    run<1, 16>("v1: single step, next group's f prefetched", sms, 3, reps);
    run<2, 16>("v2: single step, all f up front, no fences", sms, 3, reps);
    run<3, 16>("v3: step pairs, column-major in a group", sms, 3, reps);
    run<0, 8>("fl_chain_rows", sms, 3, reps);
    run<2, 8>("v2", sms, 3, reps);
    run<0, 16>("fl_chain_rows, 1 CTA/SM", sms, 1, reps);
    run<2, 16>("v2, 1 CTA/SM", sms, 1, reps);
    return 0;
}
