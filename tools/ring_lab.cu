// ring_lab.cu -- the pass kernels side by side on a synthetic tableau, outside the library: bits against blk_flush_sp_kernel,
// time per launch.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -lineinfo -I linearprogramming_b200/csrc
//                         -o tools/ring_lab tools/ring_lab.cu && tools/ring_lab [rows] [ld] [kc] [reps]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "dense_block_ring.cuh"

using namespace lpgpu;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__global__ void fill(double *p, size_t n, uint64_t seed, double lo, double hi) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = uniform64(seed, 1, i, lo, hi);
}

int main(int argc, char **argv) {
    const int rows = argc > 1 ? atoi(argv[1]) : 8193;
    const size_t ld = argc > 2 ? (size_t)atol(argv[2]) : 16400;
    const int kc = argc > 3 ? atoi(argv[3]) : 32;
    const int reps = argc > 4 ? atoi(argv[4]) : 10;
    const size_t pcn = ((size_t)rows + 31) / 32 * 32 + 32;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    double *T0, *Ta, *Tb, *Fs, *Pm;
    int *kcp, *ctr;
    const size_t tn = (size_t)rows * ld;
    CK(cudaMalloc(&T0, tn * 8)); CK(cudaMalloc(&Ta, tn * 8)); CK(cudaMalloc(&Tb, tn * 8));
    CK(cudaMalloc(&Fs, pcn * BLK_KMAX * 8)); CK(cudaMalloc(&Pm, ld * BLK_KMAX * 8));
    CK(cudaMalloc(&kcp, 8)); CK(cudaMalloc(&ctr, 8));
    CK(cudaMemset(ctr, 0, 8));
    CK(cudaMemcpy(kcp, &kc, 4, cudaMemcpyHostToDevice));
    fill<<<1024, 256>>>(T0, tn, 1, -2.0, 2.0);
    fill<<<256, 256>>>(Fs, pcn * BLK_KMAX, 2, -1.0, 1.0);
    fill<<<256, 256>>>(Pm, ld * BLK_KMAX, 3, -1.0, 1.0);
    CK(cudaDeviceSynchronize());
    CK(cudaFuncSetAttribute(blk_flush_sp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FL_SMEM_BYTES));
    CK(cudaFuncSetAttribute(blk_flush_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MM_SMEM_BYTES));
    CK(cudaFuncSetAttribute(blk_flush_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RG_SMEM_BYTES));
    FlushParams U;
    memset(&U, 0, sizeof U);
    U.Tsrc = T0; U.kc_ptr = kcp; U.seq_ptr = kcp; U.flag = nullptr; U.l2first = 1; U.ld = ld; U.rows = rows; U.Fs = Fs; U.pcn = pcn; U.Pm = Pm;
    U.task_ctr = ctr; U.seq = 0; U.pf = 1; U.ring_nd = getenv("LPGPU_RING_ND") ? atoi(getenv("LPGPU_RING_ND")) : RG_ND;
    const long nct = (long)((ld + FL_COLS - 1) / FL_COLS), nrt = (rows + FL_ROWS - 1) / FL_ROWS;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto run = [&](int which, double *out, float *ms) -> int {
        U.T = out;
        U.rb = which == 2 ? 2 * FL_RB : FL_RB;
        const long tasks = nct * ((nrt + U.rb - 1) / U.rb);
        CK(cudaEventRecord(e0));
        for (int r = 0; r < reps; ++r) {
            const int grid = (int)std::min<long>(tasks, (long)sms * (which == 2 ? 1 : 3));
            if (which == 0) blk_flush_sp_kernel<<<grid, FL_THREADS, FL_SMEM_BYTES>>>(U);
            else if (which == 1) blk_flush_mma_kernel<<<grid, FL_THREADS, MM_SMEM_BYTES>>>(U);
            else blk_flush_ring_kernel<<<grid, RG_THREADS, RG_SMEM_BYTES>>>(U);
            U.seq++;
        }
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        CK(cudaEventElapsedTime(ms, e0, e1));
        *ms /= reps;
        return 0;
    };
    const char *names[3] = {"dfma_sp", "mma", "ring"};
    std::vector<double> ha(tn), hb(tn);
    float ms = 0.f;
    if (run(0, Ta, &ms)) return 1;
    printf("%-8s %8.4f ms  %7.1f GB/s\n", names[0], ms, 16.0 * tn / ms / 1e6);
    CK(cudaMemcpy(ha.data(), Ta, tn * 8, cudaMemcpyDeviceToHost));
    for (int which = 1; which <= 2; ++which) {
        CK(cudaMemset(Tb, 0xff, tn * 8));
        if (run(which, Tb, &ms)) return 1;
        CK(cudaMemcpy(hb.data(), Tb, tn * 8, cudaMemcpyDeviceToHost));
        size_t bad = 0, first = 0;
        for (size_t i = 0; i < tn; ++i)
            if (memcmp(&ha[i], &hb[i], 8)) { if (!bad) first = i; ++bad; }
        printf("%-8s %8.4f ms  %7.1f GB/s  mismatches vs dfma_sp: %zu", names[which], ms, 16.0 * tn / ms / 1e6, bad);
        if (bad) printf("  (first at row %zu col %zu: %.17g vs %.17g)", first / ld, first % ld, hb[first], ha[first]);
        printf("\n");
    }
    return 0;
}
