// noalloc.cu -- lpgpu_pivot_noalloc: one forced simplex pivot on a caller-owned device tableau, allocating nothing.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the allocation-free variant of
// lpgpu_dense_pivot (dense.cu): that entry point stages the pivot column and the scaled pivot row in buffers that
// lpgpu_dense_create obtained with cudaMalloc; here they live in ONE file-scope device array of a fixed, modest size
// (LPGPU_NOALLOC_MAX_ROWS + LPGPU_NOALLOC_MAX_COLS doubles = 1.25 MB, held by every GPU that loads the library until
// the process exits).  The caller passes the tableau (input and output, device memory) and the pivot position.
//
//   stage kernel:      col[i]  = a_ik  for every row i (constraint and objective rows)
//                      prow[j] = a_rj / a_rk  (exact signed zero for a_rj == 0, as div_skip_zero in common.cuh)
//   eliminate kernel:  a_ij <- fma(-col[i], prow[j], a_ij)  for i != r;   a_rj <- prow[j]
// -- the same operations on the same operands as the forced-pivot path of dense_select_kernel + dense_update_* (the
// rank-1 path), so the result is bit-identical to lpgpu_dense_pivot's (tests/test_zz_no_alloc.py).
//
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current (scratch_of below).
// Every call on a device shares the array: calls that overlap on the GPU (different streams) race.  Nothing in it has
// to start at zero -- both regions are rewritten in full for the requested shape before they are read.
#include <mutex>

#include "common.cuh"
#include "ctx.h"

namespace {

constexpr size_t NOALLOC_DOUBLES = (size_t)LPGPU_NOALLOC_MAX_ROWS + (size_t)LPGPU_NOALLOC_MAX_COLS;
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];      // [0, MAX_ROWS): pivot column, then the scaled pivot row

constexpr int NA_MAX_DEVICES = 64;
double *g_scratch_ptr[NA_MAX_DEVICES];       // device address of the array, per device (nullptr = not looked up yet)
std::mutex g_scratch_mu;

// device address of the scratch array on `device` (which must be current); no allocation
cudaError_t scratch_of(int device, double **out) {
    if (device < 0 || device >= NA_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[device]) {
        double *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

constexpr int NA_THREADS = 256;

__global__ void __launch_bounds__(NA_THREADS) noalloc_stage_kernel(const double *T, size_t ld, int rows, int ncols, int r, int k, double *col, double *prow) {
    const double piv = T[(size_t)r * ld + (size_t)k];
    const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = gtid; i < (size_t)rows; i += gstride) col[i] = T[i * ld + (size_t)k];
    const double *xr = T + (size_t)r * ld;
    for (size_t j = gtid; j < (size_t)ncols; j += gstride) prow[j] = lpgpu::div_skip_zero(xr[j], piv);
}

// CTA = 8 rows x 512 columns per trip; a thread owns two adjacent columns where the row start allows a 128-bit access
__global__ void __launch_bounds__(NA_THREADS) noalloc_eliminate_kernel(double *T, size_t ld, int rows, int ncols, int r, const double *col, const double *prow) {
    constexpr int RB = 8;
    const int nct = (ncols + 2 * NA_THREADS - 1) / (2 * NA_THREADS);
    const long ntasks = (long)nct * ((rows + RB - 1) / RB);
    const bool vec = (ld % 2 == 0) && ((reinterpret_cast<uintptr_t>(T) & 15u) == 0);
    for (long tt = blockIdx.x; tt < ntasks; tt += gridDim.x) {
        const int j = (int)(tt % nct) * 2 * NA_THREADS + 2 * (int)threadIdx.x;
        const int i0 = (int)(tt / nct) * RB;
        if (j >= ncols) continue;
        const bool two = j + 1 < ncols;
        const double p0 = prow[j], p1 = two ? prow[j + 1] : 0.0;
#pragma unroll
        for (int u = 0; u < RB; ++u) {
            const int i = i0 + u;
            if (i >= rows) break;
            double *a = T + (size_t)i * ld + (size_t)j;
            if (i == r) {
                a[0] = p0;
                if (two) a[1] = p1;
            } else {
                const double f = -col[i];
                if (two && vec) {
                    double2 v = *reinterpret_cast<double2 *>(a);
                    v.x = __fma_rn(f, p0, v.x);
                    v.y = __fma_rn(f, p1, v.y);
                    *reinterpret_cast<double2 *>(a) = v;
                } else {
                    a[0] = __fma_rn(f, p0, a[0]);
                    if (two) a[1] = __fma_rn(f, p1, a[1]);
                }
            }
        }
    }
}

}  // namespace

extern "C" int lpgpu_pivot_noalloc(lpgpu_ctx *ctx, double *T_dev, size_t ld, int rows, int ncols, int r, int k) {
    using lpgpu::fail;
    if (!ctx || !T_dev) return LPGPU_EINVAL;
    if (rows < 1 || ncols < 1 || ld < (size_t)ncols) return fail(ctx, LPGPU_EINVAL, "pivot_noalloc: bad shape rows=%d ncols=%d ld=%zu", rows, ncols, ld);
    if (rows > LPGPU_NOALLOC_MAX_ROWS || ncols > LPGPU_NOALLOC_MAX_COLS)
        return fail(ctx, LPGPU_EINVAL, "pivot_noalloc: %d rows x %d columns exceed the fixed scratch array (at most %d rows incl. objective rows, %d columns incl. b); "
                                       "use lpgpu_dense_pivot for larger tableaux", rows, ncols, LPGPU_NOALLOC_MAX_ROWS, LPGPU_NOALLOC_MAX_COLS);
    if (r < 0 || r >= rows || k < 1 || k >= ncols) return fail(ctx, LPGPU_EINVAL, "pivot_noalloc: (r=%d,k=%d) out of range", r, k);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    double *scratch = nullptr;
    LP_CUDA(ctx, scratch_of(ctx->device, &scratch));
    double *col = scratch, *prow = scratch + LPGPU_NOALLOC_MAX_ROWS;
    const int want = (int)(((size_t)rows + ncols + NA_THREADS - 1) / NA_THREADS);
    const int gs = want < 1 ? 1 : (want > ctx->sm_count * 4 ? ctx->sm_count * 4 : want);
    noalloc_stage_kernel<<<gs, NA_THREADS, 0, ctx->stream>>>(T_dev, ld, rows, ncols, r, k, col, prow);
    LP_LAUNCH_CHECK(ctx);
    const long nct = (ncols + 2 * NA_THREADS - 1) / (2 * NA_THREADS), ntasks = nct * ((rows + 7) / 8);
    const long cap = (long)ctx->sm_count * 8;
    const int ge = (int)(ntasks < cap ? ntasks : cap);
    noalloc_eliminate_kernel<<<ge, NA_THREADS, 0, ctx->stream>>>(T_dev, ld, rows, ncols, r, col, prow);
    LP_LAUNCH_CHECK(ctx);
    return LPGPU_OK;
}
