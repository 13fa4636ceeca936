// ctx.h -- host-side context shared by the translation units of liblpgpu.so.
#pragma once

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "../../include/lpgpu.h"

struct lpgpu_dense;

struct lpgpu_ctx {
    int device;
    cudaStream_t stream;
    bool own_stream;
    int sm_count;
    int cc_major, cc_minor;
    size_t hbm_bytes;
    long launches;
    void *green_pass;           // green context (driver API) over the SMs the step kernels do NOT get: the pipelined pass runs in it
    int green_sms;              // SMs in it
    int green_state;            // 0 = not tried yet, 1 = ready, -1 = unavailable
    lpgpu_dense *solve_cache;   // tableau object kept between lpgpu_solve_dense calls of one shape (1 GB cudaMalloc / cudaFree per call otherwise)
    char err[512];
};

namespace lpgpu {

inline int fail(lpgpu_ctx *ctx, int code, const char *fmt, ...) {
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof ctx->err, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace lpgpu

// CUDA call -> LPGPU_ECUDA with file:line in the context's error string
#define LP_CUDA(ctx, call)                                                                         \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return lpgpu::fail((ctx), e__ == cudaErrorMemoryAllocation ? LPGPU_ENOMEM : LPGPU_ECUDA, \
                               "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
    } while (0)

#define LP_LAUNCH_CHECK(ctx)                                                                       \
    do {                                                                                           \
        (ctx)->launches++;                                                                         \
        LP_CUDA((ctx), cudaGetLastError());                                                        \
    } while (0)
