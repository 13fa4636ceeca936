/*
 * lpgpu.h -- C ABI of the B200 (sm_100a) dense-tableau simplex engine.
 *
 * This is the drop-in boundary.  The reference (SomeBottle/LinearProgramming) has no
 * plugin / FFI layer: its boundary is the plain C link-level API of Source/simplex.h:9-13
 * (LPStandardize, LPAlign, newSimplex) over the SimplexMatrix storage of
 * Source/matrix.h:13-25.  newSimplex builds that matrix and frees it without solving
 * (Source/simplex.c:40,67; empty Big-M / two-phase cases at :57-63, empty AddBigM at
 * :79-82).  The functions below are what a C host calls in that gap; the host-side
 * replacement of newSimplex that does so lives in linearprogramming_b200/host/ and
 * INTEGRATION.md shows the binding.
 *
 * Plain pointers and sizes only; no C++/torch types.  Every function returns
 * LPGPU_OK (0) or a negative LPGPU_E* code and never aborts or throws -- the reference's
 * convention is "valid flags + message, never abort" (Source/public.h:65,
 * Source/numOprts.h:29); lpgpu_last_error() carries the message.  There is no CPU
 * fallback: without a CUDA device lpgpu_create fails with LPGPU_ENODEV.
 *
 * Tableau layout (that of SimplexMatrix: cMatrix[i][0] = b_i, cMatrix[i][j+1] = a_ij,
 * Source/matrix.c:42-72), row-major, leading dimension ld >= n+1 doubles:
 *     rows 0..m-1     [ b_i | a_i1 .. a_in ]
 *     row  m          [ -z  | sigma_1 .. sigma_n ]       reduced costs c_j - z_j
 *     row  m+1 (R=2)  [ -zM | sigmaM_1 .. sigmaM_n ]     Big-M row / phase-1 row
 * The LP is "max" with b >= 0 (Source/simplex.c:99-106,178-186) and the tableau must be
 * canonical: basis[i] (1-based column) is the unit column e_i with zero objective entries.
 * Artificial columns are the LAST n_art columns.  The pivot rule is stated in DESIGN.md
 * section 3 (the reference defines none).
 */
#ifndef LPGPU_H
#define LPGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* return codes */
#define LPGPU_OK        0
#define LPGPU_EINVAL   (-1)   /* bad argument                                   */
#define LPGPU_ENODEV   (-2)   /* no usable CUDA device / wrong architecture     */
#define LPGPU_ECUDA    (-3)   /* CUDA runtime error (see lpgpu_last_error)      */
#define LPGPU_ENOMEM   (-4)   /* device or host allocation failed               */
#define LPGPU_ENCCL    (-5)   /* NCCL error (column-sharded path)               */

/* solve status */
#define LPGPU_OPTIMAL     0
#define LPGPU_UNBOUNDED   1
#define LPGPU_INFEASIBLE  2
#define LPGPU_ITER_LIMIT  3
#define LPGPU_RUNNING     4   /* lpgpu_dense_run stopped on its pivot budget    */

/* method: how the objective rows are priced */
#define LPGPU_PLAIN     0     /* R = 1                                           */
#define LPGPU_BIGM      1     /* R = 2, (sigmaM, sigma) lexicographic            */
#define LPGPU_TWOPHASE  2     /* R = 2, phase 1 on row m+1, then row m           */
#define LPGPU_DUAL      3     /* R = 1, dual simplex: the tableau must be dual feasible (every sigma_j
                               * <= 0), b may be negative; leaving row = most negative b_i, entering
                               * column = min sigma_j / a_rj over a_rj < 0      */

/* ---- lpgpu_opts.variant: every public bit has a name (0 = the default path) -----------------------------
 * Pricing rule (primal methods, dense and batched paths, column-sharded path):
 *   BLAND   the LOWEST eligible column enters (default: the largest reduced cost, ties to the lowest column).  With
 *           the lowest-row tie-break of the ratio test it cannot cycle (Beale's example runs into the iteration
 *           limit under the default rule).  LPGPU_EINVAL for the dual simplex.  SURVEY.md 8 (f4): the reference has
 *           no pricing at all (Source/simplex.c:57-63).
 *   DEVEX   reference-framework pricing: sigma_j^2 / w_j is maximised over the eligible columns, the weights w_j (all 1
 *           after lpgpu_dense_reset) follow Devex's update from every pivot's scaled row.  Dense
 *           path (rank-1, blocked, pipelined) and the batched shared-memory kernel; primal methods; LPGPU_EINVAL for the
 *           dual simplex, together with BLAND, and on the column-sharded path.
 * Dense path, which kernels run (all bit-identical):                                        */
#define LPGPU_VARIANT_UPD_STATIC    0x1        /* rank-1 path: TMA-staged elimination, static task stride (default: work stealing) */
#define LPGPU_VARIANT_UPD_DIRECT    0x2        /* rank-1 path: direct-load elimination kernel                                      */
#define LPGPU_VARIANT_UPD_MASK      0xf
#define LPGPU_VARIANT_DFMA_PASS     0x80       /* blocked path: the pass with scalar DFMAs, software-pipelined in half tiles (default: the ring pass -- TMA ring of slabs + DMMA chain, dense_block_ring.cuh) */
#define LPGPU_VARIANT_DMMA_PASS     0x100      /* blocked path: DMMA chain on tiles fetched with ordinary loads (the ring pass's predecessor) */
#define LPGPU_VARIANT_NO_ALTERNATE  0x200      /* rank-1 path: do not alternate the sweep direction between pivots                 */
#define LPGPU_VARIANT_NO_PDL        0x400      /* plain stream-ordered launches (default: programmatic dependent launch)           */
#define LPGPU_VARIANT_RANK1         0x800      /* one pass over the tableau per pivot (default: one pass per block of pivots)      */
#define LPGPU_VARIANT_PIPELINED     0x1000     /* blocked path: always pipeline the pass of a block with the next block's decisions */
#define LPGPU_VARIANT_INPLACE       0x2000     /* blocked path: never pipeline (steps, pass, steps, ... on one stream, in place)    */
                                               /* neither bit: pipelined from 0.8 GB of tableau on                                 */
#define LPGPU_VARIANT_STEP_CL8      0x8000     /* blocked path: step kernel as a portable 8-CTA cluster (default: 16 CTAs)         */
#define LPGPU_VARIANT_STEP_PERSISTENT 0x40    /* pipelined blocks: one persistent step-kernel launch per block (default: one cluster launch per pivot, measured faster) */
#define LPGPU_VARIANT_BLOCK_K(k)    (((k) & 0x3f) << 16)   /* blocked path: pivots per block, 1..32 (0 = 32 in place, 28 pipelined) */
#define LPGPU_VARIANT_BLOCK_K_MASK  0x3f0000
#define LPGPU_VARIANT_BLAND         0x400000
#define LPGPU_VARIANT_DEVEX         0x800000
/* Batched path only (distinct from the dense bits, so one opts struct can be handed to both paths):                */
#define LPGPU_VARIANT_BATCH_KEEP_FINAL 0x1000000   /* keep the final tableaux on the device (lpgpu_batched_download T_final)  */
#define LPGPU_VARIANT_BATCH_SMEM       0x2000000   /* shared-memory kernel even where a register-resident one applies        */

typedef struct lpgpu_ctx lpgpu_ctx;         /* one per process and GPU           */
typedef struct lpgpu_dense lpgpu_dense;     /* one large device-resident tableau */
typedef struct lpgpu_batched lpgpu_batched; /* many small tableaux               */

typedef struct {
    double eps;         /* pricing / ratio / snap tolerance (1e-9)                       */
    int max_iter;       /* total pivot cap -> LPGPU_ITER_LIMIT                           */
    int method;         /* LPGPU_PLAIN | LPGPU_BIGM | LPGPU_TWOPHASE | LPGPU_DUAL        */
    int n_art;          /* number of trailing artificial columns                         */
    int log_cap;        /* pivots to record in the pivot log (0 = none)                  */
    int check_every;    /* dense: pivots enqueued between host status reads (0 = 64)     */
    int variant;        /* OR of LPGPU_VARIANT_* bits (0 = default)                      */
    int time_kernels;   /* dense: bracket every elimination launch with CUDA events      */
    int results_only;   /* lpgpu_solve_dense: do not copy the final tableau back to T    */
} lpgpu_opts;

typedef struct {
    int status;
    int iters;          /* pivots so far (phase 1 + drive-out + main)                    */
    int phase1_iters;
    int phase;          /* 0 main, 1 phase one, 2 drive-out                              */
    double z;           /* -T[m][0]                                                      */
    double zM;          /* -T[m+1][0] (R = 2) else 0                                     */
    int n_logged;       /* min(iters, log_cap)                                           */
    int update_launches;  /* elimination kernels that did work during this call          */
    float update_ms;      /* their summed device time (time_kernels = 1) else 0          */
    float select_ms;      /* summed device time of the pricing/ratio kernels, likewise   */
} lpgpu_result;

void lpgpu_default_opts(lpgpu_opts *o);
/* sizeof(lpgpu_opts) for which = 0, sizeof(lpgpu_result) for which = 1 (binding self-check) */
size_t lpgpu_abi_sizeof(int which);

/* ---- context --------------------------------------------------------------------- */
/* stream: a cudaStream_t to launch on (e.g. torch's current stream), or NULL to make
 * the context create its own.  Fails with LPGPU_ENODEV unless the device is sm_100. */
int lpgpu_create(lpgpu_ctx **ctx, int device, void *stream);
void lpgpu_destroy(lpgpu_ctx *ctx);
const char *lpgpu_last_error(const lpgpu_ctx *ctx);
int lpgpu_device_info(const lpgpu_ctx *ctx, int *sm_count, size_t *hbm_bytes, int *cc_major, int *cc_minor);
long lpgpu_launch_count(const lpgpu_ctx *ctx);   /* kernels launched through this ctx   */
int lpgpu_sync(lpgpu_ctx *ctx);
/* pinned host memory for the e2e path (cudaHostAlloc / cudaFreeHost) */
void *lpgpu_host_alloc(lpgpu_ctx *ctx, size_t bytes);
void lpgpu_host_free(lpgpu_ctx *ctx, void *p);

/* ---- one large tableau: what newSimplex calls after CreateSMatrix ----------------- */
/* One-shot: upload, pivot to a terminal status (or max_iter), download.  T and basis are
 * updated in place (opts->results_only = 1 leaves T untouched and returns basis, x, z only);
 * pivot_log (2*log_cap ints: column, row) and x (n doubles) may be NULL.
 * The device-side tableau object is kept in the context and reused by the next call of the same
 * shape (released by lpgpu_destroy or by a call with another shape).
 * Replaces the missing loop between Source/simplex.c:40 and :67. */
int lpgpu_solve_dense(lpgpu_ctx *ctx, int m, int n, int R, double *T, size_t ld, int *basis,
                      const lpgpu_opts *opts, int *pivot_log, double *x, lpgpu_result *res);

/* Device-resident form of the same path (bench "value" leg, windowed parity tests). */
int lpgpu_dense_create(lpgpu_ctx *ctx, int m, int n, int R, lpgpu_dense **out);
void lpgpu_dense_destroy(lpgpu_dense *d);
int lpgpu_dense_upload(lpgpu_dense *d, const double *T, size_t ld, const int *basis);
/* The same without blocking the host (pinned buffers: lpgpu_host_alloc): the copy runs on a stream of
 * its own, ordered after everything enqueued for this tableau before and before everything enqueued
 * for it afterwards, so the upload of the next problem into one tableau object overlaps
 * lpgpu_dense_run on another.  T and basis must stay untouched until the next call on d returns.
 * (New: the reference solves one model per newSimplex call, Source/simplex.c:27-73.)                */
int lpgpu_dense_upload_async(lpgpu_dense *d, const double *T, size_t ld, const int *basis);
/* synthetic feasible LP generated on the device; needs n = ns + m, R = 1 */
int lpgpu_dense_synth(lpgpu_dense *d, uint64_t seed, uint64_t lp_index, int ns);
/* resets iteration state (status RUNNING, iters 0, phase from opts->method) */
int lpgpu_dense_reset(lpgpu_dense *d, const lpgpu_opts *opts);
/* up to max_pivots more pivots; res->status == LPGPU_RUNNING when the budget ran out */
int lpgpu_dense_run(lpgpu_dense *d, int max_pivots, lpgpu_result *res);
/* one forced pivot on (row r, column k): pivot-row scaling + rank-1 elimination only */
int lpgpu_dense_pivot(lpgpu_dense *d, int r, int k);
int lpgpu_dense_download(lpgpu_dense *d, double *T, size_t ld, int *basis, int *pivot_log, double *x);
size_t lpgpu_dense_ld(const lpgpu_dense *d);           /* device leading dimension (doubles) */

/* One forced pivot on a CALLER-OWNED device tableau, allocating nothing: no cudaMalloc, cudaFree or allocating library
 * routine is called, now or at first use.  No counterpart in the reference; added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid it).  The
 * allocation-free variant of lpgpu_dense_pivot: the staged pivot column and scaled pivot row live in one fixed device
 * array inside the library (LPGPU_NOALLOC_MAX_ROWS + LPGPU_NOALLOC_MAX_COLS doubles = 1.25 MB per GPU that loads it).
 *   T_dev   device pointer, rows x ncols doubles with leading dimension ld >= ncols: the tableau with its objective
 *           row(s) (rows = m + R) and the b column (ncols = n + 1); updated in place, same bits as lpgpu_dense_pivot
 *   r, k    pivot row (0 <= r < rows) and tableau column (1 <= k < ncols)
 * Shapes beyond the maxima are rejected with LPGPU_EINVAL (lpgpu_last_error says so).  The two kernels are enqueued on the
 * context's stream; the call does not wait for them (lpgpu_sync).  Every call on a GPU shares the array: calls that
 * overlap on the GPU (two contexts / streams on one device) race -- keep them in one stream or order them.         */
#define LPGPU_NOALLOC_MAX_ROWS 32768
#define LPGPU_NOALLOC_MAX_COLS 131072
int lpgpu_pivot_noalloc(lpgpu_ctx *ctx, double *T_dev, size_t ld, int rows, int ncols, int r, int k);

/* ---- many small tableaux: one LP per CTA, resident in shared memory ---------------- */
/* T_all: n_lp tableaux back to back, each (m+R) x (n+1), ld = n+1.  Outputs may be NULL
 * except status.  pivot_log: n_lp * 2*log_cap ints.  T_all is not modified. */
int lpgpu_solve_batched(lpgpu_ctx *ctx, long n_lp, int m, int n, int R, const double *T_all,
                        const int *basis_all, const lpgpu_opts *opts, int *status, int *iters,
                        double *z, double *x, int *basis_out, int *pivot_log);

int lpgpu_batched_create(lpgpu_ctx *ctx, long n_lp, int m, int n, int R, lpgpu_batched **out);
void lpgpu_batched_destroy(lpgpu_batched *b);
int lpgpu_batched_upload(lpgpu_batched *b, const double *T_all, const int *basis_all);
int lpgpu_batched_synth(lpgpu_batched *b, uint64_t seed, uint64_t lp0, int ns);
/* one launch over all LPs; kernel_ms (may be NULL) = its device time from CUDA events */
int lpgpu_batched_run(lpgpu_batched *b, const lpgpu_opts *opts, float *kernel_ms);
/* any pointer may be NULL; T_final: n_lp tableaux as left by the kernel (parity tests) */
int lpgpu_batched_download(lpgpu_batched *b, int *status, int *iters, double *z, double *x,
                           int *basis_out, int *pivot_log, double *T_final);

/* ---- one large tableau, column-sharded over several GPUs ------------------------------ */
/* Each rank owns a contiguous block of the columns 1..n plus a replica of the b column, the
 * basis and the pivot state.  Per pivot the ranks exchange (a) one (reduced cost, global column)
 * candidate each -- the arg-max over those picks the entering column -- and (b) the entering
 * column itself, from the rank that owns it; the ratio test then runs redundantly on every
 * rank, so no (ratio, row) exchange is needed in this layout.
 * Methods: LPGPU_PLAIN and LPGPU_BIGM on every transport; LPGPU_TWOPHASE on the blocked path over the peer-memory
 * transport (phase switch inside the step kernel, drive-out pivots as forced steps).  LPGPU_VARIANT_BLAND is honoured.
 *   id128 != NULL : one rank per process, exchange = ncclAllGather over NVLink (libnccl is
 *                   dlopen'ed here).  Rank 0 makes the id with lpgpu_shard_unique_id and the
 *                   host ships its 128 bytes to the other ranks by any means it has.
 *   id128 == NULL : all `world` ranks live in this process on ctx's device and exchange by
 *                   device copies (single-GPU test vehicle for the N > 1 logic); `rank` ignored. */
typedef struct lpgpu_shard lpgpu_shard;
#define LPGPU_NCCL_ID_BYTES 128
int lpgpu_shard_unique_id(void *id128);
/* block of rank: global columns col0 .. col0+ncols-1 (1-based); blocks differ by at most one column */
int lpgpu_shard_columns(int n, int world, int rank, int *col0, int *ncols);
int lpgpu_shard_create(lpgpu_ctx *ctx, int rank, int world, const void *id128, int m, int n, int R,
                       lpgpu_shard **out);
/* Peer-memory transport: one rank per process like NCCL, but the two per-pivot exchanges are stores
 * into the other GPUs' memory over NVLink / NVSwitch from inside the select and elimination kernels
 * (sequence-numbered flags, no collective library on the path).  Protocol: create, export this rank's
 * 64-byte CUDA IPC handle, ship the handles of all ranks to all ranks (rank order), attach, then use
 * the shard like any other.  A peer that stops responding ends the run with LPGPU_ENCCL after 30 s.
 * Destroy only after EVERY rank has returned from its last lpgpu_shard_run (host barrier): until then
 * peers may still be storing flags into this rank's region. */
#define LPGPU_IPC_HANDLE_BYTES 64
int lpgpu_shard_create_p2p(lpgpu_ctx *ctx, int rank, int world, int m, int n, int R, lpgpu_shard **out);
int lpgpu_shard_p2p_export(lpgpu_shard *s, void *handle64);
int lpgpu_shard_p2p_attach(lpgpu_shard *s, const void *handles /* world x 64 bytes */);
void lpgpu_shard_destroy(lpgpu_shard *s);
/* T: the GLOBAL tableau (m+R) x (n+1) in host memory; each local rank copies b and its own block */
int lpgpu_shard_upload(lpgpu_shard *s, const double *T, size_t ld, const int *basis);
int lpgpu_shard_synth(lpgpu_shard *s, uint64_t seed, uint64_t lp_index, int ns);
int lpgpu_shard_reset(lpgpu_shard *s, const lpgpu_opts *opts);
/* collective: every rank must call it with the same max_pivots */
int lpgpu_shard_run(lpgpu_shard *s, int max_pivots, lpgpu_result *res);
/* writes b, the local ranks' column blocks and the objective cells into the GLOBAL layout of T
 * (other ranks' columns are left untouched); basis / pivot_log (global columns) / x (n) are replicas */
int lpgpu_shard_download(lpgpu_shard *s, double *T, size_t ld, int *basis, int *pivot_log, double *x);

/* ---- host-side helpers (no GPU work) --------------------------------------------- */
/* rel[i]: -2 "<=", 2 ">=", 3 "=" (Source/public.h:57).  Builds the canonical tableau of
 *   max c.x  s.t.  A x rel b, x >= 0
 * negating rows with b < 0 (Source/simplex.c:178-186); columns = structural, then one
 * slack/surplus per inequality row (Source/simplex.c:189-219), then one artificial per
 * ">=" / "=" row.  Returns n (total columns) or a negative code. */
int lpgpu_count_columns(int m, int ns, const double *b, const int *rel, int *n_art);
int lpgpu_build_tableau(int m, int ns, const double *A, size_t lda, const double *b, const double *c,
                        const int *rel, int R, double *T, size_t ld, int *basis, int *n_art);
/* the synthetic generator on the host (same bits as the device generator) */
void lpgpu_synth_host(uint64_t seed, uint64_t lp_index, int m, int ns, double *T, size_t ld, int *basis);

#ifdef __cplusplus
}
#endif
#endif /* LPGPU_H */
