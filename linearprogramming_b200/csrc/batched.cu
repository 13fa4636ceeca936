// batched.cu -- C-ABI entry points of the batched path (include/lpgpu.h): many independent
// small LPs, one per CTA, pivot loop resident in shared memory (batched_kernels.cuh).
#include <algorithm>
#include <new>

#include "batched_kernels.cuh"
#include "batched_reg.cuh"
#include "batched_regx.cuh"
#include "ctx.h"

using namespace lpgpu;

struct lpgpu_batched {
    lpgpu_ctx *ctx;
    long n_lp;
    int m, n, R;
    size_t tsz;           // doubles per tableau
    double *T_all;
    int *basis_all;
    int *status, *iters, *basis_out, *log;
    double *z, *x, *T_final;
    int log_cap;          // active limit = opts->log_cap of the last run (stride of the log, download length)
    int log_alloc;        // what the device buffer has room for
    bool want_final;
    cudaEvent_t e0, e1;
};

namespace {

using RegSmall = RegCfg<1, 32, 2>;     // m <= 32, n <= 64: one warp per LP, no block barrier in the loop
using RegWide = RegCfg<4, 16, 4>;      // m <= 64, n <= 128: four warps per LP

bool batched_opts_ok(const lpgpu_opts *o, int n, int R) {
    if (!o) return false;
    if (o->method != LPGPU_PLAIN && o->method != LPGPU_BIGM && o->method != LPGPU_TWOPHASE && o->method != LPGPU_DUAL) return false;
    if ((o->method == LPGPU_PLAIN || o->method == LPGPU_DUAL) != (R == 1)) return false;
    if (o->n_art < 0 || o->n_art > n || o->max_iter < 0 || o->log_cap < 0 || !(o->eps >= 0.0)) return false;
    if ((o->variant & (LPGPU_VARIANT_BLAND | LPGPU_VARIANT_DEVEX)) && o->method == LPGPU_DUAL) return false;      // pricing rules: primal methods (include/lpgpu.h)
    if ((o->variant & LPGPU_VARIANT_BLAND) && (o->variant & LPGPU_VARIANT_DEVEX)) return false;
    return true;
}

}  // namespace

extern "C" {

int lpgpu_batched_create(lpgpu_ctx *ctx, long n_lp, int m, int n, int R, lpgpu_batched **out) {
    if (!ctx || !out) return LPGPU_EINVAL;
    if (n_lp < 0 || m < 1 || n < 1 || (R != 1 && R != 2))
        return fail(ctx, LPGPU_EINVAL, "batched_create: bad shape n_lp=%ld m=%d n=%d R=%d", n_lp, m, n, R);
    const size_t smem = batched_smem_bytes(m, n, R);
    if (smem > 227 * 1024)
        return fail(ctx, LPGPU_EINVAL, "batched_create: tableau needs %zu B of shared memory (> 227 KB); use the dense path", smem);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    lpgpu_batched *b = new (std::nothrow) lpgpu_batched();
    if (!b) return fail(ctx, LPGPU_ENOMEM, "batched_create: host allocation failed");
    b->ctx = ctx; b->n_lp = n_lp; b->m = m; b->n = n; b->R = R;
    b->tsz = (size_t)(m + R) * (size_t)(n + 1);
    b->T_all = nullptr; b->basis_all = nullptr; b->status = b->iters = b->basis_out = b->log = nullptr;
    b->z = b->x = b->T_final = nullptr; b->log_cap = 0; b->log_alloc = 0; b->want_final = false; b->e0 = b->e1 = nullptr;
    const size_t nl = (size_t)std::max(1L, n_lp);
    cudaError_t e = cudaSuccess;
    if (e == cudaSuccess) e = cudaMalloc(&b->T_all, nl * b->tsz * sizeof(double) + 16);
    if (e == cudaSuccess) e = cudaMalloc(&b->basis_all, nl * m * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&b->status, nl * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&b->iters, nl * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&b->basis_out, nl * m * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&b->z, nl * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&b->x, nl * n * sizeof(double));
    if (e == cudaSuccess) e = cudaEventCreate(&b->e0);
    if (e == cudaSuccess) e = cudaEventCreate(&b->e1);
    if (e == cudaSuccess) e = cudaMemsetAsync(reinterpret_cast<unsigned char *>(b->T_all) + nl * b->tsz * sizeof(double), 0, 16, ctx->stream);
    // dynamic shared memory up to the 227 KB per-CTA limit minus what the kernel declares statically
    auto allow_max_dynamic = [&](const void *kernel) {
        cudaFuncAttributes fa;
        if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, kernel);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - fa.sharedSizeBytes));
    };
    allow_max_dynamic(reinterpret_cast<const void *>(batched_simplex_kernel));
    allow_max_dynamic(reinterpret_cast<const void *>(batched_reg_kernel<1, 32, 2>));
    allow_max_dynamic(reinterpret_cast<const void *>(batched_reg_kernel<4, 16, 4>));
    if (e != cudaSuccess) {
        lpgpu_batched_destroy(b);
        return fail(ctx, e == cudaErrorMemoryAllocation ? LPGPU_ENOMEM : LPGPU_ECUDA, "batched_create: %s", cudaGetErrorString(e));
    }
    *out = b;
    return LPGPU_OK;
}

void lpgpu_batched_destroy(lpgpu_batched *b) {
    if (!b) return;
    cudaSetDevice(b->ctx->device);
    cudaStreamSynchronize(b->ctx->stream);
    cudaFree(b->T_all); cudaFree(b->basis_all); cudaFree(b->status); cudaFree(b->iters);
    cudaFree(b->basis_out); cudaFree(b->log); cudaFree(b->z); cudaFree(b->x); cudaFree(b->T_final);
    if (b->e0) cudaEventDestroy(b->e0);
    if (b->e1) cudaEventDestroy(b->e1);
    delete b;
}

int lpgpu_batched_upload(lpgpu_batched *b, const double *T_all, const int *basis_all) {
    if (!b || !T_all || !basis_all) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = b->ctx;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    if (b->n_lp > 0) {
        LP_CUDA(ctx, cudaMemcpyAsync(b->T_all, T_all, (size_t)b->n_lp * b->tsz * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        LP_CUDA(ctx, cudaMemcpyAsync(b->basis_all, basis_all, (size_t)b->n_lp * b->m * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    }
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

int lpgpu_batched_synth(lpgpu_batched *b, uint64_t seed, uint64_t lp0, int ns) {
    if (!b) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = b->ctx;
    if (b->R != 1 || ns < 1 || ns + b->m != b->n) return fail(ctx, LPGPU_EINVAL, "batched_synth: needs R=1 and n = ns + m");
    if (b->n_lp == 0) return LPGPU_OK;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    batched_synth_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(b->T_all, b->basis_all, b->n_lp, b->m, ns, seed, lp0);
    LP_LAUNCH_CHECK(ctx);
    return LPGPU_OK;
}

int lpgpu_batched_run(lpgpu_batched *b, const lpgpu_opts *opts, float *kernel_ms) {
    if (!b) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = b->ctx;
    if (!batched_opts_ok(opts, b->n, b->R)) return fail(ctx, LPGPU_EINVAL, "batched_run: bad options (method/R/n_art)");
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    if (kernel_ms) *kernel_ms = 0.f;
    if (b->n_lp == 0) return LPGPU_OK;
    const size_t nl = (size_t)b->n_lp;
    if (opts->log_cap > b->log_alloc) {
        LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(b->log);
        b->log = nullptr;
        b->log_alloc = 0;
        LP_CUDA(ctx, cudaMalloc(&b->log, nl * 2 * (size_t)opts->log_cap * sizeof(int)));
        b->log_alloc = opts->log_cap;
    }
    b->log_cap = opts->log_cap;       // stride and download length: the caller's pivot_log holds n_lp * 2 * opts->log_cap ints
    if (opts->log_cap > 0) LP_CUDA(ctx, cudaMemsetAsync(b->log, 0xff, nl * 2 * (size_t)b->log_cap * sizeof(int), ctx->stream));
    const bool want_final = (opts->variant & LPGPU_VARIANT_BATCH_KEEP_FINAL) != 0;   // parity tests: keep the final tableaux
    if (want_final && !b->T_final) LP_CUDA(ctx, cudaMalloc(&b->T_final, nl * b->tsz * sizeof(double)));
    b->want_final = want_final;

    BatchedParams B;
    B.T_all = b->T_all; B.basis_all = b->basis_all; B.n_lp = b->n_lp; B.m = b->m; B.n = b->n; B.R = b->R;
    B.eps = opts->eps; B.max_iter = opts->max_iter; B.method = opts->method; B.n_art = opts->n_art;
    B.bland = (opts->variant & LPGPU_VARIANT_BLAND) ? 1 : 0;
    B.devex = (opts->variant & LPGPU_VARIANT_DEVEX) ? 1 : 0;       // Devex: the shared-memory kernel keeps the weights next to the tableau
    B.log_cap = b->log_cap; B.status = b->status; B.iters = b->iters; B.z = b->z; B.x = b->x;
    B.basis_out = b->basis_out; B.log = opts->log_cap > 0 ? b->log : nullptr;
    B.T_final = want_final ? b->T_final : nullptr;

    // Kernel choice.  Shapes that fit a register tile run register-resident: the plain method through batched_reg_kernel,
    // Big-M / two-phase / dual / Bland through batched_regx_kernel (batched_regx.cuh); other shapes and
    // LPGPU_VARIANT_BATCH_SMEM: the shared-memory kernel.  Grid = a whole number of resident CTAs per SM; each CTA walks LPs
    // p = blockIdx, blockIdx + grid, ...
    const bool bland = (opts->variant & LPGPU_VARIANT_BLAND) != 0;
    const bool force_smem = (opts->variant & (LPGPU_VARIANT_BATCH_SMEM | LPGPU_VARIANT_DEVEX)) != 0;
    const bool plain = opts->method == LPGPU_PLAIN && b->R == 1 && !bland;
    const int shape = force_smem ? 0 : (b->m <= RegSmall::MAX_M && b->n <= RegSmall::MAX_N) ? 1 : (b->m <= RegWide::MAX_M && b->n <= RegWide::MAX_N) ? 2 : 0;
    typedef void (*bkernel_t)(BatchedParams);
    bkernel_t kern = batched_simplex_kernel;
    size_t smem = batched_smem_bytes(b->m, b->n, b->R);
    int threads = BATCH_THREADS;
    if (shape && plain) {
        kern = shape == 1 ? (bkernel_t)batched_reg_kernel<1, 32, 2> : (bkernel_t)batched_reg_kernel<4, 16, 4>;
    } else if (shape) {
        const int mode = opts->method == LPGPU_DUAL ? 2 : (b->R == 2 ? 1 : 0);
        if (shape == 1)
            kern = mode == 2 ? (bkernel_t)batched_regx_kernel<1, 32, 2, 2, false>
                 : mode == 1 ? (bland ? (bkernel_t)batched_regx_kernel<1, 32, 2, 1, true> : (bkernel_t)batched_regx_kernel<1, 32, 2, 1, false>)
                             : (bkernel_t)batched_regx_kernel<1, 32, 2, 0, true>;
        else
            kern = mode == 2 ? (bkernel_t)batched_regx_kernel<4, 16, 4, 2, false>
                 : mode == 1 ? (bland ? (bkernel_t)batched_regx_kernel<4, 16, 4, 1, true> : (bkernel_t)batched_regx_kernel<4, 16, 4, 1, false>)
                             : (bkernel_t)batched_regx_kernel<4, 16, 4, 0, true>;
    }
    if (shape) {
        smem = shape == 1 ? RegSmall::smem_bytes(b->m + b->R - 1, b->n) : RegWide::smem_bytes(b->m + b->R - 1, b->n);     // staging window of (m+R) rows
        threads = shape == 1 ? RegSmall::THREADS : RegWide::THREADS;
    }
    int occ = 0;
    LP_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem));
    if (occ < 1) return fail(ctx, LPGPU_EINVAL, "batched_run: kernel does not fit (smem %zu B)", smem);
    const long grid = std::min(b->n_lp, (long)ctx->sm_count * occ);
    LP_CUDA(ctx, cudaEventRecord(b->e0, ctx->stream));
    kern<<<(unsigned)grid, threads, smem, ctx->stream>>>(B);
    LP_LAUNCH_CHECK(ctx);
    LP_CUDA(ctx, cudaEventRecord(b->e1, ctx->stream));
    if (kernel_ms) {
        LP_CUDA(ctx, cudaEventSynchronize(b->e1));
        LP_CUDA(ctx, cudaEventElapsedTime(kernel_ms, b->e0, b->e1));
    }
    return LPGPU_OK;
}

int lpgpu_batched_download(lpgpu_batched *b, int *status, int *iters, double *z, double *x, int *basis_out,
                           int *pivot_log, double *T_final) {
    if (!b) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = b->ctx;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t nl = (size_t)b->n_lp;
    if (nl > 0) {
        if (status) LP_CUDA(ctx, cudaMemcpyAsync(status, b->status, nl * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        if (iters) LP_CUDA(ctx, cudaMemcpyAsync(iters, b->iters, nl * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        if (z) LP_CUDA(ctx, cudaMemcpyAsync(z, b->z, nl * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        if (x) LP_CUDA(ctx, cudaMemcpyAsync(x, b->x, nl * b->n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        if (basis_out) LP_CUDA(ctx, cudaMemcpyAsync(basis_out, b->basis_out, nl * b->m * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        if (pivot_log && b->log_cap > 0)
            LP_CUDA(ctx, cudaMemcpyAsync(pivot_log, b->log, nl * 2 * (size_t)b->log_cap * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        if (T_final) {
            if (!b->want_final || !b->T_final) return fail(ctx, LPGPU_EINVAL, "batched_download: final tableaux were not kept (LPGPU_VARIANT_BATCH_KEEP_FINAL)");
            LP_CUDA(ctx, cudaMemcpyAsync(T_final, b->T_final, nl * b->tsz * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        }
    }
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

int lpgpu_solve_batched(lpgpu_ctx *ctx, long n_lp, int m, int n, int R, const double *T_all, const int *basis_all,
                        const lpgpu_opts *opts, int *status, int *iters, double *z, double *x, int *basis_out,
                        int *pivot_log) {
    if (!ctx || !opts || !status) return LPGPU_EINVAL;
    if (n_lp > 0 && (!T_all || !basis_all)) return LPGPU_EINVAL;
    if (!batched_opts_ok(opts, n, R)) return fail(ctx, LPGPU_EINVAL, "solve_batched: bad options (method/R/n_art)");
    if (n_lp == 0) return LPGPU_OK;
    lpgpu_batched *b = nullptr;
    int rc = lpgpu_batched_create(ctx, n_lp, m, n, R, &b);
    if (rc != LPGPU_OK) return rc;
    rc = lpgpu_batched_upload(b, T_all, basis_all);
    if (rc == LPGPU_OK) rc = lpgpu_batched_run(b, opts, nullptr);
    if (rc == LPGPU_OK) rc = lpgpu_batched_download(b, status, iters, z, x, basis_out, pivot_log, nullptr);
    lpgpu_batched_destroy(b);
    return rc;
}

}  // extern "C"
