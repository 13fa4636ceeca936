// div_selftest.cu -- is  a / b  (IEEE, correctly rounded) reproduced bit for bit by two Markstein
// correction steps on the correctly rounded reciprocal y = RN(1/b)?
//     q0 = RN(a*y); r0 = fma(-q0, b, a); q1 = fma(r0, y, q0); r1 = fma(-q1, b, a); q2 = fma(r1, y, q1)
// Counter-based random operands over several exponent windows plus adversarial significands
// (all ones, one above a power of two, ...).  Prints the number of mismatches per class.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/div_selftest tools/div_selftest.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__host__ __device__ inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ inline double div2(double a, double b, double y) {
    const double q0 = __dmul_rn(a, y);
    const double r0 = __fma_rn(-q0, b, a);
    const double q1 = __fma_rn(r0, y, q0);
    const double r1 = __fma_rn(-q1, b, a);
    return __fma_rn(r1, y, q1);
}

__device__ inline double div1(double a, double b, double y) {
    const double q0 = __dmul_rn(a, y);
    const double r0 = __fma_rn(-q0, b, a);
    return __fma_rn(r0, y, q0);
}

// cls 0: a, b with random significands, exponents within +-40; 1: +-300; 2: b significand all ones /
// near all ones; 3: b = 1 + few ulps; 4: a significand patterned
__device__ inline void make(uint64_t idx, int cls, double &a, double &b) {
    const uint64_t h1 = mix64(idx * 2 + 1), h2 = mix64(idx * 2 + 2), h3 = mix64(idx ^ 0xABCDEF);
    uint64_t ma = h1 & 0xFFFFFFFFFFFFFull, mb = h2 & 0xFFFFFFFFFFFFFull;
    int ea = 1023, eb = 1023;
    const int span = (cls == 1) ? 300 : 40;
    ea += (int)(h3 % (2 * span + 1)) - span;
    eb += (int)((h3 >> 20) % (2 * span + 1)) - span;
    if (cls == 2) mb = 0xFFFFFFFFFFFFFull - (h2 & 0xFF);
    if (cls == 3) mb = (h2 & 0xFF);
    if (cls == 4) ma = (h1 & 1) ? 0xFFFFFFFFFFFFFull - (h1 >> 40 & 0xFF) : (h1 >> 40 & 0xFF);
    const uint64_t sa = (h3 >> 62) & 1;
    a = __longlong_as_double((long long)((sa << 63) | ((uint64_t)ea << 52) | ma));
    b = __longlong_as_double((long long)(((uint64_t)eb << 52) | mb));        // pivot elements are positive
}

__global__ void check(uint64_t base, uint64_t n, int cls, unsigned long long *bad2, unsigned long long *bad1) {
    unsigned long long m2 = 0, m1 = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        double a, b;
        make(base + i, cls, a, b);
        const double ref = a / b;
        const double y = __drcp_rn(b);
        m2 += (__double_as_longlong(div2(a, b, y)) != __double_as_longlong(ref));
        m1 += (__double_as_longlong(div1(a, b, y)) != __double_as_longlong(ref));
    }
    if (m2) atomicAdd(bad2, m2);
    if (m1) atomicAdd(bad1, m1);
}

int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], nullptr, 10) : (1ull << 32);
    unsigned long long *d, h[2];
    cudaMalloc(&d, 16);
    for (int cls = 0; cls < 5; ++cls) {
        cudaMemset(d, 0, 16);
        check<<<148 * 8, 256>>>(0x1000000000ull * cls, n, cls, d, d + 1);
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("class %d: %llu samples, mismatches two-step %llu, one-step %llu\n", cls, (unsigned long long)n, h[0], h[1]);
    }
    return cudaDeviceSynchronize() != cudaSuccess;
}
