// api.cu -- context management and host-side helpers of liblpgpu.so (include/lpgpu.h).
#include <new>
#include <string.h>

#include "common.cuh"
#include "ctx.h"

using namespace lpgpu;

extern "C" {

void lpgpu_default_opts(lpgpu_opts *o) {
    if (!o) return;
    memset(o, 0, sizeof *o);
    o->eps = 1e-9;
    o->max_iter = 0x7fffffff;
    o->method = LPGPU_PLAIN;
    o->check_every = 64;
}

size_t lpgpu_abi_sizeof(int which) { return which == 0 ? sizeof(lpgpu_opts) : which == 1 ? sizeof(lpgpu_result) : 0; }

int lpgpu_create(lpgpu_ctx **out, int device, void *stream) {
    if (!out) return LPGPU_EINVAL;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1 || device < 0 || device >= count) {
        cudaGetLastError();
        return LPGPU_ENODEV;       // no CPU fallback: the caller must fail loudly
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return LPGPU_ENODEV;
    if (prop.major != 10) return LPGPU_ENODEV;      // built for sm_100a only
    if (cudaSetDevice(device) != cudaSuccess) return LPGPU_ENODEV;
    lpgpu_ctx *ctx = new (std::nothrow) lpgpu_ctx();
    if (!ctx) return LPGPU_ENOMEM;
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->cc_major = prop.major;
    ctx->cc_minor = prop.minor;
    ctx->hbm_bytes = prop.totalGlobalMem;
    ctx->launches = 0;
    ctx->solve_cache = nullptr;
    ctx->green_pass = nullptr; ctx->green_sms = 0; ctx->green_state = 0;
    ctx->err[0] = '\0';
    if (stream) {
        ctx->stream = reinterpret_cast<cudaStream_t>(stream);
        ctx->own_stream = false;
    } else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete ctx;
            return LPGPU_ECUDA;
        }
        ctx->own_stream = true;
    }
    *out = ctx;
    return LPGPU_OK;
}

void lpgpu_destroy(lpgpu_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->solve_cache) lpgpu_dense_destroy(ctx->solve_cache);
    if (ctx->green_pass) {          // the green context of the pipelined pass (dense.cu); its streams went with the tableau objects
        typedef int (*fn_gdestroy)(void *);
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuGreenCtxDestroy", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && f)
            reinterpret_cast<fn_gdestroy>(f)(ctx->green_pass);
        cudaGetLastError();
    }
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *lpgpu_last_error(const lpgpu_ctx *ctx) { return ctx ? ctx->err : "no context"; }

int lpgpu_device_info(const lpgpu_ctx *ctx, int *sm_count, size_t *hbm_bytes, int *cc_major, int *cc_minor) {
    if (!ctx) return LPGPU_EINVAL;
    if (sm_count) *sm_count = ctx->sm_count;
    if (hbm_bytes) *hbm_bytes = ctx->hbm_bytes;
    if (cc_major) *cc_major = ctx->cc_major;
    if (cc_minor) *cc_minor = ctx->cc_minor;
    return LPGPU_OK;
}

long lpgpu_launch_count(const lpgpu_ctx *ctx) { return ctx ? ctx->launches : 0; }

int lpgpu_sync(lpgpu_ctx *ctx) {
    if (!ctx) return LPGPU_EINVAL;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

void *lpgpu_host_alloc(lpgpu_ctx *ctx, size_t bytes) {
    if (!ctx) return nullptr;
    void *p = nullptr;
    cudaSetDevice(ctx->device);
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        fail(ctx, LPGPU_ENOMEM, "host_alloc: %zu bytes of pinned memory refused", bytes);
        return nullptr;
    }
    return p;
}

void lpgpu_host_free(lpgpu_ctx *ctx, void *p) {
    if (p) cudaFreeHost(p);
}

// ------------------------------------------------------------------ host helpers (no GPU)

int lpgpu_count_columns(int m, int ns, const double *b, const int *rel, int *n_art) {
    if (m < 1 || ns < 1 || !b || !rel) return LPGPU_EINVAL;
    int slack = 0, art = 0;
    for (int i = 0; i < m; ++i) {
        int code = rel[i];
        if (code != -2 && code != 2 && code != 3) return LPGPU_EINVAL;
        if (code != 3 && b[i] < 0) code = -code;      // row will be negated
        slack += (code != 3);
        art += (code != -2);
    }
    if (n_art) *n_art = art;
    return ns + slack + art;
}

int lpgpu_build_tableau(int m, int ns, const double *A, size_t lda, const double *b, const double *c, const int *rel,
                        int R, double *T, size_t ld, int *basis, int *n_art) {
    if (!A || !b || !c || !rel || !T || !basis || (R != 1 && R != 2) || lda < (size_t)ns) return LPGPU_EINVAL;
    int art = 0;
    const int n = lpgpu_count_columns(m, ns, b, rel, &art);
    if (n < 0) return n;
    if (ld < (size_t)n + 1) return LPGPU_EINVAL;
    if (R == 1 && art != 0) return LPGPU_EINVAL;     // plain method needs a slack basis
    for (int i = 0; i < m + R; ++i) memset(T + (size_t)i * ld, 0, sizeof(double) * ((size_t)n + 1));
    const int first_art = n - art + 1;
    int sl = ns + 1, ar = first_art;
    for (int i = 0; i < m; ++i) {
        double *row = T + (size_t)i * ld;
        const bool flip = b[i] < 0;                   // b_i >= 0 by negating the row
        const double sgn = flip ? -1.0 : 1.0;
        int code = rel[i];
        if (flip && code != 3) code = -code;
        row[0] = sgn * b[i];
        for (int j = 0; j < ns; ++j) row[1 + j] = sgn * A[(size_t)i * lda + j];
        if (code == -2) {
            row[sl] = 1.0;
            basis[i] = sl++;
        } else {
            if (code == 2) row[sl++] = -1.0;
            row[ar] = 1.0;
            basis[i] = ar++;
        }
    }
    double *oc = T + (size_t)m * ld;
    for (int j = 0; j < ns; ++j) oc[1 + j] = c[j];
    if (R == 2) {
        double *om = T + (size_t)(m + 1) * ld;
        for (int j = first_art; j <= n; ++j) om[j] = -1.0;
        // price the artificial basis out of the M-row: add each artificial-basic row, ascending
        for (int i = 0; i < m; ++i) {
            if (basis[i] < first_art) continue;
            const double *row = T + (size_t)i * ld;
            for (int j = 0; j <= n; ++j) om[j] += row[j];
        }
    }
    if (n_art) *n_art = art;
    return n;
}

void lpgpu_synth_host(uint64_t seed, uint64_t lp_index, int m, int ns, double *T, size_t ld, int *basis) {
    const int n = ns + m;
    for (int i = 0; i <= m; ++i)
        for (int j = 0; j <= n; ++j) T[(size_t)i * ld + j] = synth_elem(seed, lp_index, m, ns, i, j);
    for (int i = 0; i < m; ++i) basis[i] = 1 + ns + i;
}

}  // extern "C"
