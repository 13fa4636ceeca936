// dense.cu -- C-ABI entry points of the single-large-tableau path (include/lpgpu.h).
//
// Fills the gap the reference leaves between CreateSMatrix and RevokeSMatrix
// (Source/simplex.c:40-67): the pivot loop runs device-resident; the host only enqueues
// (select, eliminate) pairs and reads one small state record every `check_every` pivots.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <vector>

#include <cuda.h>

#include "dense_internal.h"
#include "dense_update.cuh"
#include "dense_block.cuh"
#include "dense_block_mma.cuh"
#include "dense_block_ring.cuh"
#include "dense_steps.cuh"

using namespace lpgpu;


namespace lpgpu {

// One thread-block cluster (8 CTAs x 1024 threads, distributed shared memory for the arg-min /
// arg-max exchange) decides a pivot and stages it for the elimination pass.
//   fast path  (st->k_next > 0): the entering column was priced one pivot ahead and its values
//              (and b) were extracted by the previous elimination pass into compact buffers, so
//              only the ratio test, the pivot-row scaling and the next pricing remain -- each a
//              single memory round trip spread over 8192 threads;
//   full path  : dense_select.cuh decide() over the tableau in HBM, run by CTA 0 and broadcast
//              (first pivot, phase changes, drive-out pivots, terminal statuses).
// Either way the cluster then prices the pivot AFTER this one (lookahead): the objective row(s)
// after elimination are sigma'_j = fma(-sigma_k, a_rj', sigma_j) -- the same operation on the
// same operands the elimination pass performs, so the choice is bit-identical to pricing the
// updated tableau.  forced_k > 0: skip the decision and pivot on (forced_r, forced_k).
__global__ void __cluster_dims__(SELECT_CLUSTER, 1, 1) __launch_bounds__(SELECT_THREADS)
dense_select_kernel(DenseParams P, int forced_r, int forced_k) {
    __shared__ Cand scratch[SELECT_THREADS / 32];
    __shared__ Cand slot_ratio[SELECT_CLUSTER], slot_look[SELECT_CLUSTER];
    __shared__ Decision dsh;
    cg::cluster_group cl = cg::this_cluster();
    const int rank = (int)cl.block_rank();
    const int gtid = rank * SELECT_THREADS + (int)threadIdx.x;
    constexpr int GSTRIDE = SELECT_CLUSTER * SELECT_THREADS;

    cluster_entry_arrive();     // DSMEM rule: no remote access before every CTA of the cluster has started (wait below)
    pdl_wait();                 // the elimination pass before this launch has completed
    pdl_launch_dependents();    // the next elimination pass may set itself up behind this kernel
    PivotState *st = P.st;
    int r = -1, k = 0, cur = st->cur, phase = st->phase;
    bool fast = false;
    cluster_entry_wait();

    if (forced_k > 0) {
        r = forced_r;
        k = forced_k;
    } else {
        // every thread of the cluster reads the same state; it is written only in the final section
        const int status0 = st->status, iters = st->iters, p10 = st->phase1_iters;
        const int budget = st->budget, knext = st->k_next;
        if (status0 != LPGPU_RUNNING) {
            cl.sync();                         // nobody still reads the state
            if (gtid == 0) st->do_pivot = 0;
            return;
        }
        Decision D;
        if (knext > 0 && phase != PH_DRIVEOUT) {
            // ratio test on the compact copies left by the previous elimination pass (coalesced)
            const double *col = P.pcol + (size_t)(cur ^ 1) * P.pcn;
            Cand c = cand_none();
            for (int i = gtid; i < P.m; i += GSTRIDE) {
                const double a = col[i];
                if (a > P.eps) {
                    Cand t;
                    t.p = 0.0;
                    t.q = -(P.bnext[i] / a);
                    t.idx = i;
                    if (cand_better(t, c)) c = t;
                }
            }
            const Cand l = cluster_best(cl, c, scratch, slot_ratio);
            D.status = status0; D.phase = phase; D.p1 = p10; D.k = knext; D.r = -1; D.have = false;
            if (l.idx < 0) D.status = LPGPU_UNBOUNDED;
            else { D.r = l.idx; D.have = true; }
            if (D.have && iters >= P.max_iter) { D.status = LPGPU_ITER_LIMIT; D.have = false; }
            fast = true;
        } else {
            if (rank == 0) {
                D = decide(P, scratch, status0, phase, iters, p10);
                if (threadIdx.x < SELECT_CLUSTER) *cl.map_shared_rank(&dsh, threadIdx.x) = D;
            }
            cl.sync();
            D = dsh;
        }
        // out of budget for this run call: the decision still settles terminal statuses, a
        // pending pivot is left for the next call (status stays RUNNING, lookahead kept)
        if (D.have && budget <= 0) D.have = false;
        r = D.r;
        k = D.k;
        phase = D.phase;
        if (D.have && fast) cur ^= 1;
        cl.sync();                             // all state reads are done before it is rewritten
        if (gtid == 0) {
            st->status = D.status;
            st->phase = D.phase;
            st->phase1_iters = (D.phase == PH_ONE) ? iters + (D.have ? 1 : 0) : D.p1;
            st->do_pivot = D.have ? 1 : 0;
            if (D.have) {
                st->r = r;
                st->k = k;
                st->cur = cur;
                st->iters = iters + 1;
                st->budget = budget - 1;
                if (iters < P.log_cap) { P.log[2 * iters] = k; P.log[2 * iters + 1] = r; }
            }
        }
        if (!D.have) return;
    }

    // ---- stage the pivot: pivot column (all rows incl. objective rows) and scaled pivot row
    double *pcol = P.pcol + (size_t)cur * P.pcn;
    const int rows = P.m + P.R;
    if (!fast)
        for (int i = gtid; i < rows; i += GSTRIDE) pcol[i] = P.T[(size_t)i * P.ld + (size_t)k];
    const double piv = P.T[(size_t)r * P.ld + (size_t)k];
    const double *xr = P.T + (size_t)r * P.ld;

    // lookahead pricing mode for the pivot after this one
    const bool look = (forced_k <= 0) && (phase != PH_DRIVEOUT) && (P.method != LPGPU_DUAL);   // dual: full decision every pivot
    const bool bigm = (P.method == LPGPU_BIGM);
    const int prow_obj = bigm ? P.m : ((phase == PH_ONE) ? P.m + 1 : P.m);
    const int jmax = (P.method == LPGPU_TWOPHASE && phase == PH_MAIN) ? P.n - P.n_art : P.n;
    const double *oc = P.T + (size_t)prow_obj * P.ld;
    const double *om = P.T + (size_t)(P.m + 1) * P.ld;            // Big-M row (R = 2 only)
    const double fc = -oc[k];
    const double fm = bigm ? -om[k] : 0.0;
    Cand c = cand_none();
    const double wk = P.wts ? P.wts[k] : 1.0;        // Devex: w_k before this pivot (rewritten only after the barrier below)
    for (size_t j = (size_t)gtid; j < P.ld; j += GSTRIDE) {
        const double pr = div_skip_zero(xr[j], piv);
        P.prow[j] = pr;
        double w = 1.0;
        if (P.wts && j >= 1 && j <= (size_t)P.n && j != (size_t)k) {       // every pivot updates the weights (forced and drive-out pivots too)
            w = devex_weight(P.wts[j], pr, wk, false);
            P.wts[j] = w;
        }
        if (look && j >= 1 && j <= (size_t)jmax) {
            const double q = __fma_rn(fc, pr, oc[j]);
            Cand t;
            t.idx = (int)j;
            if (bigm) {
                double p = __fma_rn(fm, pr, om[j]);
                if (fabs(p) <= P.eps) p = 0.0;
                t.p = P.bland ? 0.0 : p; t.q = P.bland ? 1.0 : q;
                if (P.wts) { t.q = (p == 0.0) ? (q * q) / w : q; t.p = (p * p) / w; }
                if ((p > 0.0 || (p == 0.0 && q > P.eps)) && cand_better(t, c)) c = t;
            } else {
                t.p = 0.0; t.q = P.bland ? 1.0 : P.wts ? (q * q) / w : q;
                if (q > P.eps && cand_better(t, c)) c = t;
            }
        }
    }
    int knew = 0;
    if (look) {
        c = cluster_best(cl, c, scratch, slot_look);
        knew = c.idx > 0 ? c.idx : 0;
    } else if (P.wts) cl.sync();                     // every thread has read w_k
    if (gtid == 0) {
        P.basis[r] = k;
        st->k_next = knew;
        if (P.wts) P.wts[k] = 1.0;                   // the entering column joins the basis
        if (forced_k > 0) { st->r = r; st->k = k; st->do_pivot = 1; }
    }
}

DenseParams make_params(const lpgpu_dense *d) {
    DenseParams P;
    P.T = d->T; P.ld = d->ld; P.m = d->m; P.n = d->n; P.R = d->R;
    P.basis = d->basis; P.prow = d->prow; P.pcol = d->pcol; P.pcn = d->pcn; P.bnext = d->bnext; P.st = d->st;
    P.log = d->log; P.log_cap = d->log_cap;
    P.eps = d->opts.eps; P.max_iter = d->opts.max_iter; P.method = d->opts.method; P.n_art = d->opts.n_art;
    P.bland = (d->opts.variant & LPGPU_VARIANT_BLAND) ? 1 : 0;
    P.wts = (d->opts.variant & LPGPU_VARIANT_DEVEX) ? d->wts : nullptr;
    return P;
}

// Launch with programmatic stream serialization (variant bit 0x400 turns it off): the kernel may become
// resident while its predecessor drains and blocks in pdl_wait() until that one has completed.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

// The same with the cluster shape as a launch attribute (the primal step kernel runs as 8 or 16 CTAs).
template <typename... KArgs, typename... Args>
static cudaError_t launch_cluster(bool pdl, int cluster, size_t smem, void (*kernel)(KArgs...), dim3 block, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster);
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

int launch_select(lpgpu_dense *d, int forced_r, int forced_k) {
    lpgpu_ctx *ctx = d->ctx;
    const bool pdl = !(d->opts.variant & LPGPU_VARIANT_NO_PDL) && !d->opts.time_kernels;
    LP_CUDA(ctx, launch_pdl(pdl, dense_select_kernel, dim3(SELECT_CLUSTER), dim3(SELECT_THREADS), ctx->stream,
                            make_params(d), forced_r, forced_k));
    ctx->launches++;
    return LPGPU_OK;
}

static int launch_update_impl(lpgpu_dense *d, double *cnext, void *const *peer_x, int n_peers, size_t peer_off);

int launch_update(lpgpu_dense *d, double *cnext) { return launch_update_impl(d, cnext, nullptr, 0, 0); }

int launch_update_peers(lpgpu_dense *d, void *const *peer_x, int n_peers, size_t peer_off) {
    return launch_update_impl(d, nullptr, peer_x, n_peers, peer_off);
}

static int launch_update_impl(lpgpu_dense *d, double *cnext, void *const *peer_x, int n_peers, size_t peer_off) {
    lpgpu_ctx *ctx = d->ctx;
    UpdateParams U;
    U.T = d->T; U.ld = d->ld; U.rows = d->m + d->R; U.prow = d->prow; U.colbuf = d->pcol; U.pcn = d->pcn; U.bnext = d->bnext; U.st = d->st; U.cnext = cnext;
    U.task_ctr = d->task_ctr; U.seq = d->upd_seq++;
    U.peer_x = peer_x; U.n_peers = n_peers; U.peer_off = peer_off;
    // variant: low nibble 0 = TMA-staged kernel with work stealing (default), 1 = TMA-staged with a
    //          static task stride, 2 = direct-load kernel;
    //          0x200 = do NOT alternate the sweep direction between pivots
    U.alternate = (d->opts.variant & LPGPU_VARIANT_NO_ALTERNATE) ? 0 : 1;
    const int kind = d->opts.variant & LPGPU_VARIANT_UPD_MASK;
    const bool pdl = !(d->opts.variant & LPGPU_VARIANT_NO_PDL) && !d->opts.time_kernels && cnext == nullptr && n_peers == 0;
    LP_CUDA(ctx, launch_pdl(pdl, kind == 2 ? dense_update_direct : kind == 1 ? dense_update_tma_staged : dense_update_dynamic, dim3(d->grid_update),
                            dim3(UPD_THREADS), ctx->stream, U));
    ctx->launches++;
    return LPGPU_OK;
}

int read_state(lpgpu_dense *d) {
    lpgpu_ctx *ctx = d->ctx;
    LP_CUDA(ctx, cudaMemcpyAsync(d->h_st, d->st, sizeof(PivotState), cudaMemcpyDeviceToHost, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

// ---- blocked path (dense_block.cuh): the default for the plain and Big-M methods; variant bit 0x800 selects the
// rank-1 path instead; pivots per block in variant bits 16..21 (0 = 32)
bool use_blocked(const lpgpu_dense *d) {
    return !(d->opts.variant & LPGPU_VARIANT_RANK1) &&
           (d->opts.method == LPGPU_PLAIN || d->opts.method == LPGPU_BIGM || d->opts.method == LPGPU_TWOPHASE ||
            d->opts.method == LPGPU_DUAL);
}

BlockBufs make_bufs(const lpgpu_dense *d) {
    BlockBufs B;
    B.Fs = d->Ft; B.cands = d->cands; B.Pm = d->Pm; B.objcur = d->objcur; B.bcur = d->bcur; B.rowstamp = d->rowstamp; B.steprow = d->steprow;
    B.Fall = d->Ft; B.Pall = d->Pm; B.stamp_prev = nullptr; B.info = nullptr; B.par = 0; B.pend = 0;
    B.l2keep = 0;
    if (d->pipe_on) {        // pipelined blocks: the set of this block's parity, the other set is the pending block's
        const bool odd = d->par != 0;
        B.Fs = odd ? d->Ft2 : d->Ft;
        B.Pm = odd ? d->Pm2 : d->Pm;
        B.rowstamp = odd ? d->rowstamp2 : d->rowstamp;   B.stamp_prev = odd ? d->rowstamp : d->rowstamp2;
        B.steprow = odd ? d->steprow2 : d->steprow;
        B.info = d->info; B.par = d->par; B.pend = d->pend ? 1 : 0;
        B.l2keep = (d->l2hint & 2) ? 1 : 0;
    }
    const int k = (d->opts.variant & LPGPU_VARIANT_BLOCK_K_MASK) >> 16;
    // pivots per block: 32; pipelined 28 (measured: the step kernels are the longer stage there, 56.5 k against 54.8 k pivots/s)
    B.K = (k >= 1 && k <= BLK_KMAX) ? k : (d->pipe_on ? 28 : BLK_KMAX);
    B.dbg = d->blk_dbg ? d->blk_dbg + 8 * (size_t)(d->blk_dbg_n % 4096) : nullptr;
    return B;
}

int launch_block_init(lpgpu_dense *d) {
    lpgpu_ctx *ctx = d->ctx;
    blk_init_kernel<<<ctx->sm_count, 1024, 0, ctx->stream>>>(make_params(d), make_bufs(d));
    LP_LAUNCH_CHECK(ctx);
    if (d->pipe_state == 1) {       // both sets of row stamps start clean
        LP_CUDA(ctx, cudaMemsetAsync(d->rowstamp, 0, d->pcn * sizeof(int), ctx->stream));
        LP_CUDA(ctx, cudaMemsetAsync(d->rowstamp2, 0, d->pcn * sizeof(int), ctx->stream));
    }
    return LPGPU_OK;
}

// ---- pipelined blocks: the pass of block b runs on a second stream, out of place (T -> T2), while the step kernels of
// block b+1 decide on T (the tableau BEFORE block b) through the slices of block b and their own.  All four methods.
// Default: by size -- the longer chains cost each step about 8 us, the hidden pass saves 2 x bytes / 4.3 TB/s per block,
// which pays from about 0.4 GB of tableau on (measured on 2 m = n tableaux: in place 1.27x / 1.12x faster at m = 1024 /
// 4096, pipelined 1.12x / 1.45x faster at m = 6144 / 8192).  variant bit 0x1000 forces pipelined blocks, 0x2000 (or no
// room for the second tableau buffer) the in-place order: steps, pass, steps, ...
static bool want_pipeline(const lpgpu_dense *d) {
    if (!use_blocked(d) || (d->opts.variant & LPGPU_VARIANT_INPLACE) || getenv("LPGPU_NO_PIPELINE")) return false;
    if (d->opts.variant & LPGPU_VARIANT_PIPELINED) return true;
    return ((size_t)d->m + d->R) * d->ld * sizeof(double) >= ((size_t)800 << 20);
}

typedef void (*steps_kernel_t)(DenseParams, BlockBufs, StepsArgs);
static steps_kernel_t steps_kernel(int R, int pr) {
    return pr == PR_DEVEX ? (R == 2 ? blk_steps_kernel<2, PR_DEVEX> : blk_steps_kernel<1, PR_DEVEX>)
         : pr == PR_BLAND ? (R == 2 ? blk_steps_kernel<2, PR_BLAND> : blk_steps_kernel<1, PR_BLAND>)
                          : (R == 2 ? blk_steps_kernel<2, PR_DANTZIG> : blk_steps_kernel<1, PR_DANTZIG>);
}

// the pricing rule of a tableau object as the kernels' template parameter
static int pricing_of(const lpgpu_dense *d) {
    return (d->opts.variant & LPGPU_VARIANT_DEVEX) ? PR_DEVEX : (d->opts.variant & LPGPU_VARIANT_BLAND) ? PR_BLAND : PR_DANTZIG;
}

// The persistent step kernel of pipelined blocks: grid, partition, shared memory, exchange slots, and the SMs the grid lands on
// while the device is idle (the pass leaves them free).  Unavailable (steps_state = -1) is not an error: the per-pivot launches
// of blk_step_kernel are used instead.
static int setup_steps(lpgpu_dense *d) {
    if (d->steps_state != 0) return LPGPU_OK;
    lpgpu_ctx *ctx = d->ctx;
    d->steps_state = -1;
    int G = d->green ? 16 : 32;       // under the green-context partition the pass leaves exactly 16 SMs
    { const char *e = getenv("LPGPU_STEP_CTAS"); if (e && atoi(e) > 0) G = atoi(e); }      // tuning aid
    G = std::max(1, std::min(G, std::min(STEPS_GMAX, ctx->sm_count / 2)));
    const int rpc = (int)((((size_t)d->m + G - 1) / G + 1) & ~(size_t)1);
    const int cpc = (int)(((d->ld + G - 1) / G + 1) & ~(size_t)1);
    const size_t budget = 200 * 1024, per_step = ((size_t)rpc + cpc) * sizeof(double);
    const int S = (int)std::min<size_t>(BLK_KMAX, budget / per_step);
    const size_t smem = (size_t)S * per_step;
    cudaError_t e = cudaSuccess;
    for (int R = 1; R <= 2 && e == cudaSuccess; ++R)
        for (int pr = 0; pr <= 2 && e == cudaSuccess; ++pr)
            e = cudaFuncSetAttribute(steps_kernel(R, pr), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(blk_steps_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int occ = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, steps_kernel(d->R, PR_DANTZIG), STEPS_TH, smem);
    unsigned *smids = nullptr;
    if (e == cudaSuccess && occ >= 1) e = cudaMalloc(&d->xch, 2 * STEPS_GMAX * sizeof(XSlot));
    if (e == cudaSuccess && occ >= 1) e = cudaMemsetAsync(d->xch, 0, 2 * STEPS_GMAX * sizeof(XSlot), ctx->stream);
    if (e == cudaSuccess && occ >= 1) e = cudaMalloc(&smids, STEPS_GMAX * sizeof(unsigned));
    if (e != cudaSuccess || occ < 1) { cudaFree(smids); cudaFree(d->xch); d->xch = nullptr; cudaGetLastError(); return LPGPU_OK; }
    LP_CUDA(ctx, cudaMemsetAsync(smids, 0xff, STEPS_GMAX * sizeof(unsigned), ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (d->stream2) LP_CUDA(ctx, cudaStreamSynchronize(d->stream2));
    blk_steps_probe_kernel<<<G, STEPS_TH, smem, ctx->stream>>>(smids);
    LP_LAUNCH_CHECK(ctx);
    unsigned h[STEPS_GMAX];
    LP_CUDA(ctx, cudaMemcpyAsync(h, smids, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(smids);
    for (unsigned &w : d->resv_steps) w = 0;
    int distinct = 0;
    if (!getenv("LPGPU_NO_RESERVE"))
        for (int q = 0; q < G; ++q) {
            const unsigned v = h[q];
            if (v < 256u && !((d->resv_steps[v >> 5] >> (v & 31u)) & 1u)) { d->resv_steps[v >> 5] |= 1u << (v & 31u); ++distinct; }
        }
    if (distinct != G)
        for (unsigned &w : d->resv_steps) w = 0;     // no reservation: still correct, the grid just starts when the pass has left room
    d->steps_G = G; d->steps_rpc = rpc; d->steps_cpc = cpc; d->steps_S = S; d->steps_smem = smem;
    d->xch_seq = 1;
    d->steps_state = 1;
    return LPGPU_OK;
}

static int launch_block_steps(lpgpu_dense *d, int nsteps, int first) {
    lpgpu_ctx *ctx = d->ctx;
    StepsArgs A;
    A.xch = static_cast<XSlot *>(d->xch);
    A.xbase = d->xch_seq;
    d->xch_seq += 2u * (unsigned)nsteps + 4u;        // at most one pricing, two exchanges per pivot and one closing decision
    A.nsteps = nsteps; A.first = first; A.smem_steps = d->steps_S; A.rpc = d->steps_rpc; A.cpc = d->steps_cpc;
    steps_kernel(d->R, pricing_of(d))<<<d->steps_G, STEPS_TH, d->steps_smem, ctx->stream>>>(make_params(d), make_bufs(d), A);
    LP_LAUNCH_CHECK(ctx);
    return LPGPU_OK;
}

static void free_pipeline(lpgpu_dense *d) {
    cudaFree(d->xch); d->xch = nullptr;
    cudaFree(d->T2); cudaFree(d->rowstamp2); cudaFree(d->steprow2); cudaFree(d->info); cudaFree(d->flush_flag);
    d->T2 = d->Ft2 = d->Pm2 = nullptr; d->rowstamp2 = d->steprow2 = nullptr; d->info = nullptr; d->flush_flag = nullptr;
    if (d->stream2) { cudaStreamDestroy(d->stream2); d->stream2 = nullptr; }
    if (d->ev_steps) { cudaEventDestroy(d->ev_steps); d->ev_steps = nullptr; }
    if (d->ev_flush) { cudaEventDestroy(d->ev_flush); d->ev_flush = nullptr; }
}

// SM partition through the driver's green contexts instead of the %smid probe: the SMs of the device are split into a group of
// 16 (one step cluster, 8-SM granularity, MAX_POTENTIAL_CLUSTER_SIZE so that a 16-CTA cluster fits it) and the rest; the
// pipelined pass runs on a stream of a green context over THE REST, so it can never occupy the step kernels' SMs and nothing
// has to be probed or vacated.  Entry points come from cudaGetDriverEntryPoint (the library does not link libcuda).
// Measured: 58.3 k against 56.6 k pivots/s with the probe + vacate reservation (tools/green_ab.sh).  LPGPU_GREEN=0, or the API
// being unavailable / failing, falls back to that reservation.
static cudaStream_t green_pass_stream(lpgpu_ctx *ctx) {
    const char *ge = getenv("LPGPU_GREEN");
    if ((ge && atoi(ge) == 0) || ctx->green_state < 0) return nullptr;
    typedef CUresult (*fn_devget)(CUdevice *, int);
    typedef CUresult (*fn_getres)(CUdevice, CUdevResource *, CUdevResourceType);
    typedef CUresult (*fn_split)(CUdevResource *, unsigned int *, const CUdevResource *, CUdevResource *, unsigned int, unsigned int);
    typedef CUresult (*fn_desc)(CUdevResourceDesc *, CUdevResource *, unsigned int);
    typedef CUresult (*fn_gcreate)(CUgreenCtx *, CUdevResourceDesc, CUdevice, unsigned int);
    typedef CUresult (*fn_gstream)(CUstream *, CUgreenCtx, unsigned int, int);
    auto entry = [](const char *name) -> void * {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint(name, &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { cudaGetLastError(); return nullptr; }
        return f;
    };
    fn_gstream gstream = (fn_gstream)entry("cuGreenCtxStreamCreate");
    if (ctx->green_state == 0) {
        ctx->green_state = -1;
        fn_devget devget = (fn_devget)entry("cuDeviceGet");
        fn_getres getres = (fn_getres)entry("cuDeviceGetDevResource");
        fn_split split = (fn_split)entry("cuDevSmResourceSplitByCount");
        fn_desc desc = (fn_desc)entry("cuDevResourceGenerateDesc");
        fn_gcreate gcreate = (fn_gcreate)entry("cuGreenCtxCreate");
        if (!devget || !getres || !split || !desc || !gcreate || !gstream) return nullptr;
        CUdevice dev;
        CUdevResource all, grp, rest;
        unsigned int ng = 1;
        CUdevResourceDesc rd;
        CUgreenCtx g;
        if (devget(&dev, ctx->device) != CUDA_SUCCESS || getres(dev, &all, CU_DEV_RESOURCE_TYPE_SM) != CUDA_SUCCESS) return nullptr;
        if (split(&grp, &ng, &all, &rest, CU_DEV_SM_RESOURCE_SPLIT_MAX_POTENTIAL_CLUSTER_SIZE, 16) != CUDA_SUCCESS || ng < 1) return nullptr;
        if (rest.sm.smCount < 64) return nullptr;
        if (desc(&rd, &rest, 1) != CUDA_SUCCESS || gcreate(&g, rd, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS) return nullptr;
        ctx->green_pass = g;
        ctx->green_sms = (int)rest.sm.smCount;
        ctx->green_state = 1;
        if (getenv("LPGPU_BLK_DEBUG")) fprintf(stderr, "[lpgpu] green context: step group %u SMs, pass %u SMs\n", grp.sm.smCount, rest.sm.smCount);
    }
    CUstream st = nullptr;
    if (!gstream || gstream(&st, (CUgreenCtx)ctx->green_pass, CU_STREAM_NON_BLOCKING, 0) != CUDA_SUCCESS) return nullptr;
    return (cudaStream_t)st;
}

// Buffers, stream and events of the pipelined path, once per tableau object.  Not having room for the second tableau
// buffer is not an error: the in-place order is used instead (pipe_state = -1).
static int setup_pipeline(lpgpu_dense *d) {
    if (d->pipe_state != 0) return LPGPU_OK;
    lpgpu_ctx *ctx = d->ctx;
    const size_t rows = (size_t)d->m + d->R;
    unsigned *smids = nullptr;
    cudaError_t e = cudaMalloc(&d->T2, rows * d->ld * sizeof(double));
    d->Ft2 = d->Ft + d->pcn * BLK_KMAX;             // second halves of the slice buffers (allocated with the object)
    d->Pm2 = d->Pm + (size_t)BLK_KMAX * d->ld;
    if (e == cudaSuccess) e = cudaMalloc(&d->rowstamp2, d->pcn * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&d->steprow2, BLK_KMAX * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&d->info, 2 * sizeof(BlkInfo));
    if (e == cudaSuccess) e = cudaMalloc(&d->flush_flag, sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&smids, STEP_CLUSTER_MAX * sizeof(unsigned));
    if (e == cudaSuccess) { d->stream2 = green_pass_stream(ctx); d->green = d->stream2 != nullptr; }
    if (e == cudaSuccess && !d->green) e = cudaStreamCreateWithFlags(&d->stream2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->ev_steps, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->ev_flush, cudaEventDisableTiming);
    if (e != cudaSuccess) {
        cudaFree(smids);
        free_pipeline(d);
        cudaGetLastError();
        d->pipe_state = -1;
        return LPGPU_OK;
    }
    LP_CUDA(ctx, cudaMemsetAsync(d->T2, 0, rows * d->ld * sizeof(double), ctx->stream));
    LP_CUDA(ctx, cudaMemsetAsync(d->rowstamp2, 0, d->pcn * sizeof(int), ctx->stream));
    LP_CUDA(ctx, cudaMemsetAsync(d->steprow2, 0, BLK_KMAX * sizeof(int), ctx->stream));
    LP_CUDA(ctx, cudaMemsetAsync(d->info, 0, 2 * sizeof(BlkInfo), ctx->stream));
    LP_CUDA(ctx, cudaMemsetAsync(d->flush_flag, 0, sizeof(int), ctx->stream));
    LP_CUDA(ctx, cudaMemsetAsync(smids, 0xff, STEP_CLUSTER_MAX * sizeof(unsigned), ctx->stream));
    // the SMs a step-kernel cluster is placed on while the device is idle: the pass leaves them free
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int cl = d->step_cl_max;
    LP_CUDA(ctx, launch_cluster(false, cl, 0, blk_probe_kernel, dim3(SELECT_THREADS), ctx->stream, smids));
    ctx->launches++;
    unsigned h[STEP_CLUSTER_MAX];
    LP_CUDA(ctx, cudaMemcpyAsync(h, smids, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(smids);
    for (unsigned &w : d->resv) w = 0;
    int distinct = 0;
    if (!getenv("LPGPU_NO_RESERVE"))
        for (int q = 0; q < cl; ++q) {
            const unsigned v = h[q];
            if (v < 256u && !((d->resv[v >> 5] >> (v & 31u)) & 1u)) { d->resv[v >> 5] |= 1u << (v & 31u); ++distinct; }
        }
    if (distinct != cl)
        for (unsigned &w : d->resv) w = 0;           // no reservation: still correct, the pass and the steps just take turns
    d->pipe_state = 1;
    return LPGPU_OK;
}

enum { PASS_RING = 0, PASS_DFMA = 1, PASS_DMMA = 2 };
static int pass_kind(const lpgpu_dense *d) {
    const int v = d->opts.variant;
    return (v & LPGPU_VARIANT_DMMA_PASS) ? PASS_DMMA : (v & LPGPU_VARIANT_DFMA_PASS) ? PASS_DFMA : PASS_RING;
}
static int launch_pass_kernel(lpgpu_dense *d, const FlushParams &U, cudaStream_t stream);

// Close the open block: its pass goes to stream2 (after this block's steps), the next block's steps wait for the
// pass before it (they read its output and reuse its slices).  ev0 / ev1: optional timing events around the pass.
static int close_block_pipe(lpgpu_dense *d, cudaEvent_t ev0, cudaEvent_t ev1) {
    lpgpu_ctx *ctx = d->ctx;
    const bool odd = d->par != 0;
    FlushParams U;
    U.Tsrc = d->pend ? d->T2 : d->T;
    U.T = d->pend ? d->T : d->T2;
    U.kc_ptr = &d->info[d->par].kc; U.seq_ptr = &d->info[d->par].seq;
    U.ld = d->ld; U.rows = d->m + d->R; U.Fs = odd ? d->Ft2 : d->Ft; U.pcn = d->pcn; U.Pm = odd ? d->Pm2 : d->Pm;
    U.rowstamp = odd ? d->rowstamp2 : d->rowstamp; U.steprow = odd ? d->steprow2 : d->steprow; U.st = d->st;
    U.task_ctr = d->task_ctr; U.seq = d->upd_seq++;
    U.rb = d->fl_rb; U.pf = d->fl_pf; U.ring_nd = d->ring_nd;
    const bool full = d->grid_flush >= ctx->sm_count * (pass_kind(d) == PASS_RING ? 1 : 2);   // a full persistent grid: every SM holds CTAs of the pass
    for (int w = 0; w < 8; ++w) U.resv[w] = (full && !d->green) ? (d->persist_on ? d->resv_steps[w] : d->resv[w]) : 0u;     // green context: nothing to vacate
    U.flag = d->flush_flag;
    U.l2first = (d->l2hint & 1) ? 1 : 0;
    LP_CUDA(ctx, cudaEventRecord(d->ev_steps, ctx->stream));
    if (d->pend) LP_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, d->ev_flush, 0));
    LP_CUDA(ctx, cudaStreamWaitEvent(d->stream2, d->ev_steps, 0));
    if (ev0) LP_CUDA(ctx, cudaEventRecord(ev0, d->stream2));
    {
        const int rcp = launch_pass_kernel(d, U, d->stream2);
        if (rcp != LPGPU_OK) return rcp;
    }
    const unsigned fx = (unsigned)std::max<size_t>(1, std::min<size_t>(64, (d->ld + 255) / 256));
    blk_fixrows_kernel<<<dim3(fx, BLK_KMAX), 256, 0, d->stream2>>>(U);
    LP_LAUNCH_CHECK(ctx);
    if (ev1) LP_CUDA(ctx, cudaEventRecord(ev1, d->stream2));
    LP_CUDA(ctx, cudaEventRecord(d->ev_flush, d->stream2));
    if (d->pend) std::swap(d->T, d->T2);         // T = the tableau before the block just closed, T2 = where its pass writes
    d->pend = true;
    d->par ^= 1;
    return LPGPU_OK;
}

// Bring d->T up to date: wait (on the stream, not on the host) for the pass in flight and make its output the tableau.
static int drain_pipe(lpgpu_dense *d) {
    lpgpu_ctx *ctx = d->ctx;
    if (!d->pend) return LPGPU_OK;
    LP_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, d->ev_flush, 0));
    std::swap(d->T, d->T2);
    d->pend = false;
    return LPGPU_OK;
}

typedef void (*step_kernel_t)(DenseParams, BlockBufs, int);
template <int BL>
static step_kernel_t step_kernel_bl(int R, bool pipe, int cl) {
    if (cl == 16)
        return pipe ? (R == 2 ? blk_step_kernel<2, true, 16, BL> : blk_step_kernel<1, true, 16, BL>)
                    : (R == 2 ? blk_step_kernel<2, false, 16, BL> : blk_step_kernel<1, false, 16, BL>);
    return pipe ? (R == 2 ? blk_step_kernel<2, true, 8, BL> : blk_step_kernel<1, true, 8, BL>)
                : (R == 2 ? blk_step_kernel<2, false, 8, BL> : blk_step_kernel<1, false, 8, BL>);
}
static step_kernel_t step_kernel(int R, bool pipe, int cl, int pr = PR_DANTZIG) {
    return pr == PR_DEVEX ? step_kernel_bl<PR_DEVEX>(R, pipe, cl) : pr == PR_BLAND ? step_kernel_bl<PR_BLAND>(R, pipe, cl) : step_kernel_bl<PR_DANTZIG>(R, pipe, cl);
}

// CTAs per cluster of the primal step kernel: 16 where the device can place such a cluster (a non-portable size,
// allowed per kernel), else 8; variant bit 0x8000 and LPGPU_STEP_CL8 keep 8.
static int probe_step_cluster(lpgpu_dense *d) {
    d->step_cl_max = SELECT_CLUSTER;
    if (getenv("LPGPU_STEP_CL8")) return LPGPU_OK;
    bool ok = true;
    for (int R = 1; R <= 2 && ok; ++R)
        for (int pipe = 0; pipe <= 1 && ok; ++pipe)
            for (int pr = 0; pr <= 2 && ok; ++pr)
                ok = cudaFuncSetAttribute(step_kernel(R, pipe != 0, 16, pr), cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
    if (ok) ok = cudaFuncSetAttribute(blk_probe_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
    if (ok) ok = cudaFuncSetAttribute(blk_dual_step_kernel<16, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
                 cudaFuncSetAttribute(blk_dual_step_kernel<16, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
    if (ok) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(16); cfg.blockDim = dim3(SELECT_THREADS / 2);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 16; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n = 0;
        ok = cudaOccupancyMaxActiveClusters(&n, step_kernel(d->R, true, 16), &cfg) == cudaSuccess && n >= 1;
    }
    cudaGetLastError();
    if (ok) d->step_cl_max = 16;
    cudaGetLastError();
    return LPGPU_OK;
}

static int step_cluster(const lpgpu_dense *d) { return (d->opts.variant & LPGPU_VARIANT_STEP_CL8) ? SELECT_CLUSTER : d->step_cl_max; }

static int launch_block_step(lpgpu_dense *d, int first) {
    lpgpu_ctx *ctx = d->ctx;
    const bool pdl = !(d->opts.variant & LPGPU_VARIANT_NO_PDL) && !d->opts.time_kernels;
    if (d->opts.method == LPGPU_DUAL) {
        const int cl = step_cluster(d);
        LP_CUDA(ctx, launch_cluster(pdl, cl, 0, cl == 16 ? (d->pipe_on ? blk_dual_step_kernel<16, true> : blk_dual_step_kernel<16, false>)
                                                         : (d->pipe_on ? blk_dual_step_kernel<8, true> : blk_dual_step_kernel<8, false>), dim3(cl == 16 ? SELECT_THREADS / 2 : SELECT_THREADS),
                                    ctx->stream, make_params(d), make_bufs(d), first));
    } else {
        const int cl = step_cluster(d);
        LP_CUDA(ctx, launch_cluster(pdl, cl, 0, step_kernel(d->R, d->pipe_on, cl, pricing_of(d)), dim3(cl == 16 ? SELECT_THREADS / 2 : SELECT_THREADS), ctx->stream,
                                    make_params(d), make_bufs(d), first));
    }
    if (d->blk_dbg) d->blk_dbg_n++;
    ctx->launches++;
    return LPGPU_OK;
}

// which pass kernel: default = the ring pass (dense_block_ring.cuh: TMA ring of slabs + DMMA chain, 0.36 ms per pass alone on the GPU at
// 8192 x 16384); LPGPU_VARIANT_DFMA_PASS = scalar DFMAs, software-pipelined in half tiles (round 1's default, 0.50 ms);
// LPGPU_VARIANT_DMMA_PASS = DMMA chain on tiles fetched with ordinary loads (0.66 ms).  All bit-identical.
static int launch_pass_kernel(lpgpu_dense *d, const FlushParams &U, cudaStream_t stream) {
    lpgpu_ctx *ctx = d->ctx;
    int grid = d->grid_flush;
    if (d->green && stream == d->stream2)        // the persistent grid fits the SMs of the green context it runs in
        grid = std::max(1, std::min(grid, grid / ctx->sm_count > 0 ? grid / ctx->sm_count * ctx->green_sms : ctx->green_sms));
    switch (pass_kind(d)) {
        case PASS_DMMA: blk_flush_mma_kernel<<<grid, FL_THREADS, MM_SMEM_BYTES, stream>>>(U); break;
        case PASS_DFMA: blk_flush_sp_kernel<<<grid, FL_THREADS, FL_SMEM_BYTES, stream>>>(U); break;
        default: blk_flush_ring_kernel<<<grid, RG_THREADS, RG_SMEM_BYTES, stream>>>(U); break;
    }
    LP_LAUNCH_CHECK(ctx);
    return LPGPU_OK;
}

int launch_block_flush(lpgpu_dense *d) {
    lpgpu_ctx *ctx = d->ctx;
    FlushParams U;
    U.T = d->T; U.ld = d->ld; U.rows = d->m + d->R; U.Fs = d->Ft; U.pcn = d->pcn; U.Pm = d->Pm; U.rowstamp = d->rowstamp; U.steprow = d->steprow; U.st = d->st;
    U.Tsrc = d->T; U.kc_ptr = &d->st->blk_t; U.seq_ptr = &d->st->blk_seq; U.flag = nullptr; U.l2first = 0;
    for (unsigned &w : U.resv) w = 0;
    U.task_ctr = d->task_ctr; U.seq = d->upd_seq++;
    U.rb = d->fl_rb; U.pf = d->fl_pf; U.ring_nd = d->ring_nd;
    {
        const int rcp = launch_pass_kernel(d, U, ctx->stream);
        if (rcp != LPGPU_OK) return rcp;
    }
    const unsigned fx = (unsigned)std::max<size_t>(1, std::min<size_t>(64, (d->ld + 255) / 256));
    blk_fixrows_kernel<<<dim3(fx, BLK_KMAX), 256, 0, ctx->stream>>>(U);
    LP_LAUNCH_CHECK(ctx);                   // (the macro counts the launch)
    return LPGPU_OK;
}

// Bring d->T up to date, in stream order: the open block is closed and applied, the pass in flight waited for.  A run call
// of the pipelined path leaves both pending (its results come from the eagerly kept objective rows / b column), so that the
// next run call continues the pipeline; everything else that touches the tableau calls this first.
int settle_dense(lpgpu_dense *d) {
    if (d->up_pending) {         // an asynchronous upload (lpgpu_dense_upload_async): everything on the context's stream comes after it
        const cudaError_t e = cudaStreamWaitEvent(d->ctx->stream, d->ev_up, 0);
        if (e != cudaSuccess) return fail(d->ctx, LPGPU_ECUDA, "settle: %s", cudaGetErrorString(e));
        d->up_pending = false;
    }
    if (d->blk_open) {
        const int rc = d->pipe_on ? close_block_pipe(d, nullptr, nullptr) : launch_block_flush(d);
        if (rc != LPGPU_OK) return rc;
        d->blk_open = false;
        d->blk_steps = 0;
    }
    return drain_pipe(d);
}

}  // namespace lpgpu

namespace {

__global__ void extract_x_kernel(const double *T, size_t ld, const int *basis, int m, int n, double *x) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const int j = basis[i];
        if (j >= 1 && j <= n) x[j - 1] = T[(size_t)i * ld];
    }
}

__global__ void fill_kernel(double *p, size_t n, double v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int size_grid(lpgpu_dense *d) {
    lpgpu_ctx *ctx = d->ctx;
    int occ0 = 0, occ1 = 0, occ3 = 0;
    LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ3, dense_update_dynamic, UPD_THREADS, 0));
    LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, dense_update_direct, UPD_THREADS, 0));
    LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, dense_update_tma_staged, UPD_THREADS, 0));
    const int kind = d->opts.variant & LPGPU_VARIANT_UPD_MASK;
    const int occ = std::max(1, kind == 2 ? occ0 : kind == 1 ? occ1 : occ3);
    const long nct = (long)((d->ld + UPD_TILE_COLS - 1) / UPD_TILE_COLS);
    const long nrb = (d->m + d->R + UPD_TILE_ROWS - 1) / UPD_TILE_ROWS;
    d->grid_update = (int)std::max(1L, std::min((long)ctx->sm_count * occ, nct * nrb));
    int occf = 0;
    const int pk = pass_kind(d);
    if (pk == PASS_RING) LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occf, blk_flush_ring_kernel, RG_THREADS, RG_SMEM_BYTES));
    else if (pk == PASS_DMMA) LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occf, blk_flush_mma_kernel, FL_THREADS, MM_SMEM_BYTES));
    else LP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occf, blk_flush_sp_kernel, FL_THREADS, FL_SMEM_BYTES));
    const long fct = (long)((d->ld + FL_COLS - 1) / FL_COLS);
    const char *erb = getenv("LPGPU_FL_RB"), *epf = getenv("LPGPU_FL_PF");      // tuning aids
    d->fl_rb = (erb && atoi(erb) > 0) ? atoi(erb) : pk == PASS_RING ? 2 * FL_RB : FL_RB;
    d->fl_pf = epf ? atoi(epf) : 1;
    { const char *end = getenv("LPGPU_RING_ND"); d->ring_nd = (end && atoi(end) > 0) ? atoi(end) : RG_ND; }     // tuning aid
    const long fsb = ((d->m + d->R + FL_ROWS - 1) / FL_ROWS + d->fl_rb - 1) / d->fl_rb;
    d->grid_flush = (int)std::max(1L, std::min((long)ctx->sm_count * std::max(1, occf), fct * fsb));
    d->flush_tasks = std::max(1L, fct * fsb);
    return LPGPU_OK;
}

// ---- one lpgpu_dense_run call: what its loops share
struct RunCall {
    lpgpu_dense *d = nullptr;
    int max_pivots = 0;
    bool timed = false, dirty = false;
    int launched = 0;           // pivots enqueued (and known to have been applied) so far
    int done_updates = 0;       // passes over the tableau
    int iters_before = 0;
    double upd_ms = 0.0, sel_ms = 0.0;
    std::vector<int> kinds;     // timed mode: 0 = step, 1 = pass, in launch order (events 2 i, 2 i + 1)
    bool pipe = false, persist = false;
    int K = BLK_KMAX;
};

// blocked path: enqueue one step launch (kind 0; the persistent kernel: nsteps pivots) or the pass that closes the block (kind 1)
int run_enqueue(RunCall &R, int kind, int first, int nsteps = 1) {
    lpgpu_dense *d = R.d;
    lpgpu_ctx *ctx = d->ctx;
    const size_t e = 2 * R.kinds.size();
    const bool pass_on_stream2 = kind && R.pipe;
    if (R.timed) {
        while (d->ev.size() < e + 2) {
            cudaEvent_t ev;
            LP_CUDA(ctx, cudaEventCreate(&ev));
            d->ev.push_back(ev);
        }
        if (!pass_on_stream2) LP_CUDA(ctx, cudaEventRecord(d->ev[e], ctx->stream));
    }
    int rc;
    if (pass_on_stream2) rc = close_block_pipe(d, R.timed ? d->ev[e] : nullptr, R.timed ? d->ev[e + 1] : nullptr);
    else rc = kind ? launch_block_flush(d) : (R.persist ? launch_block_steps(d, nsteps, first) : launch_block_step(d, first));
    if (rc != LPGPU_OK) return rc;
    if (R.timed) {
        if (!pass_on_stream2) LP_CUDA(ctx, cudaEventRecord(d->ev[e + 1], ctx->stream));
        R.kinds.push_back(kind);
    }
    return LPGPU_OK;
}

int run_flush(RunCall &R) {
    const int rc = run_enqueue(R, 1, 0);
    R.d->blk_open = false;
    R.d->blk_steps = 0;
    ++R.done_updates;
    return rc;
}

// the tableau in d->T is brought up to date (in stream order): open block closed, pass in flight waited for
int run_settle(RunCall &R) {
    if (R.d->blk_open) { const int rc = run_flush(R); if (rc != LPGPU_OK) return rc; }
    return R.pipe ? drain_pipe(R.d) : LPGPU_OK;
}

// timed mode: add up the event pairs recorded since the last call
int run_collect(RunCall &R) {
    lpgpu_dense *d = R.d;
    lpgpu_ctx *ctx = d->ctx;
    if (R.pipe && !R.kinds.empty()) LP_CUDA(ctx, cudaStreamSynchronize(d->stream2));
    for (size_t i = 0; i < R.kinds.size(); ++i) {
        float ms = 0.f;
        LP_CUDA(ctx, cudaEventElapsedTime(&ms, d->ev[2 * i], d->ev[2 * i + 1]));
        (R.kinds[i] ? R.upd_ms : R.sel_ms) += ms;
    }
    R.kinds.clear();
    return LPGPU_OK;
}

// two-phase: the drive-out pivots between the phases are decided by the rank-1 path's full decision (it scans tableau rows),
// a few pivots at a time, on a tableau brought up to date first; afterwards the blocked path re-reads its objective rows and
// b column and carries on in the main phase
int run_driveout(RunCall &R) {
    lpgpu_dense *d = R.d;
    while (d->h_st->status == LPGPU_RUNNING && d->h_st->phase == PH_DRIVEOUT && R.launched < R.max_pivots) {
        const int chunk = std::min(8, R.max_pivots - R.launched);
        for (int c = 0; c < chunk; ++c) {
            int rc = launch_select(d, -1, 0);
            if (rc == LPGPU_OK) rc = launch_update(d, nullptr);
            if (rc != LPGPU_OK) return rc;
        }
        R.launched += chunk;
        const int rc = read_state(d);
        if (rc != LPGPU_OK) return rc;
    }
    return launch_block_init(d);
}

// ---- blocked path: K step kernels (decide a pivot each on lazily updated slices), then one pass over the tableau that applies
// them all -- in place, or pipelined with the decisions of the next block
int run_blocked(RunCall &R) {
    lpgpu_dense *d = R.d;
    int rc = LPGPU_OK;
    // pipelined blocks (default by size; LPGPU_VARIANT_INPLACE = never): the pass of a block overlaps the decisions of the next one
    d->pipe_on = false;
    if (want_pipeline(d)) {
        rc = setup_pipeline(d);
        if (rc != LPGPU_OK) return rc;
        d->pipe_on = d->pipe_state == 1;
    }
    R.pipe = d->pipe_on;
    R.K = make_bufs(d).K;             // pivots per block (depends on the order chosen above)
    // pipelined blocks of the primal methods: one launch of the step cluster per pivot; LPGPU_VARIANT_STEP_PERSISTENT: one
    // persistent launch decides all the pivots of a block (dense_steps.cuh; same bits, measured slower: 42 k against 55 k pivots/s)
    d->persist_on = false;
    if (R.pipe && d->opts.method != LPGPU_DUAL && (d->opts.variant & LPGPU_VARIANT_STEP_PERSISTENT)) {
        rc = setup_steps(d);
        if (rc != LPGPU_OK) return rc;
        d->persist_on = d->steps_state == 1;
    }
    R.persist = d->persist_on;
    if (R.dirty) { rc = launch_block_init(d); if (rc != LPGPU_OK) return rc; }
    if (d->h_st->phase == PH_DRIVEOUT) { rc = run_settle(R); if (rc == LPGPU_OK) rc = run_driveout(R); if (rc != LPGPU_OK) return rc; }
    const int chunk_max = d->opts.check_every;
    while (R.launched < R.max_pivots && d->h_st->status == LPGPU_RUNNING) {
        const int chunk = std::min(chunk_max, R.max_pivots - R.launched);
        for (int c = 0; c < chunk;) {
            const int nst = R.persist ? std::min(chunk - c, R.K - d->blk_steps) : 1;     // persistent: all the pivots the block still has room for
            rc = run_enqueue(R, 0, d->blk_open ? 0 : 1, nst);
            if (rc != LPGPU_OK) return rc;
            d->blk_open = true;
            d->blk_steps += nst;
            c += nst;
            if (d->blk_steps >= R.K) { rc = run_flush(R); if (rc != LPGPU_OK) return rc; }
        }
        rc = read_state(d);
        if (rc == LPGPU_OK && R.timed) rc = run_collect(R);
        if (rc != LPGPU_OK) return rc;
        // launches after a phase switch (or a terminal status) were no-ops: only the applied pivots count
        R.launched += (d->h_st->status == LPGPU_RUNNING && d->h_st->phase == PH_DRIVEOUT) ? d->h_st->iters - R.iters_before : chunk;
        R.iters_before = d->h_st->iters;
        if (d->h_st->status != LPGPU_RUNNING) break;
        if (d->h_st->phase == PH_DRIVEOUT) {
            rc = run_settle(R);
            if (rc == LPGPU_OK) rc = run_driveout(R);
            if (rc != LPGPU_OK) return rc;
            R.iters_before = d->h_st->iters;
        }
    }
    // pipelined and still running: the open block and the pass in flight stay pending, so that the next run call continues the
    // pipeline (z comes from the eagerly kept objective rows; download / reset / upload / ... settle first)
    const bool keep_pending = R.pipe && !R.timed && d->h_st->status == LPGPU_RUNNING && d->h_st->phase != PH_DRIVEOUT;
    if (!keep_pending) { rc = run_settle(R); if (rc != LPGPU_OK) return rc; }
    if (d->h_st->status == LPGPU_RUNNING) {
        // decision only (budget is spent): settles a terminal status
        rc = (d->h_st->phase == PH_DRIVEOUT) ? launch_select(d, -1, 0)
             : R.persist ? launch_block_steps(d, 0, d->blk_open ? 0 : 1) : launch_block_step(d, d->blk_open ? 0 : 1);
        if (rc != LPGPU_OK) return rc;
    }
    rc = read_state(d);
    if (rc == LPGPU_OK && R.timed) rc = run_collect(R);
    if (rc != LPGPU_OK) return rc;
    // the decision-only step may have found a terminal status (or the end of phase one): nothing stays pending then
    if (d->h_st->status != LPGPU_RUNNING || d->h_st->phase == PH_DRIVEOUT) rc = run_settle(R);
    return rc;
}

// ---- rank-1 path: (select, eliminate) pairs, one pass over the tableau per pivot
int run_rank1(RunCall &R) {
    lpgpu_dense *d = R.d;
    lpgpu_ctx *ctx = d->ctx;
    const int chunk_max = d->opts.check_every;
    int rc = LPGPU_OK;
    while (d->ev.size() < (size_t)3 * chunk_max && R.timed) {
        cudaEvent_t e;
        LP_CUDA(ctx, cudaEventCreate(&e));
        d->ev.push_back(e);
    }
    while (R.launched < R.max_pivots) {
        const int chunk = std::min(chunk_max, R.max_pivots - R.launched);
        for (int c = 0; c < chunk; ++c) {
            if (R.timed) LP_CUDA(ctx, cudaEventRecord(d->ev[3 * c], ctx->stream));
            rc = launch_select(d, -1, 0);
            if (rc != LPGPU_OK) return rc;
            if (R.timed) LP_CUDA(ctx, cudaEventRecord(d->ev[3 * c + 1], ctx->stream));
            rc = launch_update(d, nullptr);
            if (rc != LPGPU_OK) return rc;
            if (R.timed) LP_CUDA(ctx, cudaEventRecord(d->ev[3 * c + 2], ctx->stream));
        }
        R.launched += chunk;
        rc = read_state(d);
        if (rc != LPGPU_OK) return rc;
        const int did = d->h_st->iters - R.iters_before;      // pivots applied in this chunk
        R.iters_before = d->h_st->iters;
        R.done_updates += did;
        if (R.timed) {
            for (int c = 0; c < did && c < chunk; ++c) {
                float a = 0.f, b = 0.f;
                LP_CUDA(ctx, cudaEventElapsedTime(&a, d->ev[3 * c], d->ev[3 * c + 1]));
                LP_CUDA(ctx, cudaEventElapsedTime(&b, d->ev[3 * c + 1], d->ev[3 * c + 2]));
                R.sel_ms += a;
                R.upd_ms += b;
            }
        }
        if (d->h_st->status != LPGPU_RUNNING) break;
    }
    if (d->h_st->status == LPGPU_RUNNING) {
        // budget spent: one decision-only select settles a terminal status (optimal, iteration limit, ...) without applying a
        // pivot, so a run of exactly the needed pivots reports it
        rc = launch_select(d, -1, 0);
        if (rc == LPGPU_OK) rc = read_state(d);
    }
    return rc;
}

bool opts_ok(const lpgpu_opts *o, int n, int R) {
    if (!o) return false;
    if (o->method != LPGPU_PLAIN && o->method != LPGPU_BIGM && o->method != LPGPU_TWOPHASE && o->method != LPGPU_DUAL) return false;
    if ((o->method == LPGPU_PLAIN || o->method == LPGPU_DUAL) != (R == 1)) return false;
    if (o->n_art < 0 || o->n_art > n || o->max_iter < 0 || o->log_cap < 0 || !(o->eps >= 0.0)) return false;
    if ((o->variant & (LPGPU_VARIANT_BLAND | LPGPU_VARIANT_DEVEX)) && o->method == LPGPU_DUAL) return false;      // pricing rules: primal methods
    if ((o->variant & LPGPU_VARIANT_BLAND) && (o->variant & LPGPU_VARIANT_DEVEX)) return false;
    return true;
}

}  // namespace

extern "C" {

int lpgpu_dense_create(lpgpu_ctx *ctx, int m, int n, int R, lpgpu_dense **out) {
    if (!ctx || !out) return LPGPU_EINVAL;
    if (m < 1 || n < 1 || (R != 1 && R != 2)) return fail(ctx, LPGPU_EINVAL, "dense_create: bad shape m=%d n=%d R=%d", m, n, R);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    lpgpu_dense *d = new (std::nothrow) lpgpu_dense();
    if (!d) return fail(ctx, LPGPU_ENOMEM, "dense_create: host allocation failed");
    d->ctx = ctx; d->m = m; d->n = n; d->R = R;
    d->ld = round_up((size_t)n + 1, 16);
    d->T = nullptr; d->basis = nullptr; d->prow = d->pcol = d->bnext = d->xbuf = d->wts = nullptr; d->la_dirty = true; d->st = nullptr; d->h_st = nullptr;
    d->log = nullptr; d->log_cap = 0; d->log_alloc = 0; d->task_ctr = nullptr; d->upd_seq = 0;
    d->Ft = d->Pm = d->objcur = d->bcur = nullptr; d->rowstamp = d->steprow = nullptr; d->cands = nullptr; d->blk_open = false; d->blk_steps = 0;
    d->pipe_state = 0; d->pipe_on = false; d->pend = false; d->par = 0;
    d->xch = nullptr; d->xch_seq = 1; d->steps_state = 0; d->steps_G = d->steps_rpc = d->steps_cpc = d->steps_S = 0; d->steps_smem = 0; d->persist_on = false;
    for (unsigned &w : d->resv_steps) w = 0;
    d->up_stream = nullptr; d->ev_up = d->ev_up_ready = nullptr; d->up_pending = false;
    { const char *eh = getenv("LPGPU_L2HINT"); d->l2hint = eh ? atoi(eh) : 3; }
    d->T2 = d->Ft2 = d->Pm2 = nullptr; d->rowstamp2 = d->steprow2 = nullptr; d->info = nullptr; d->flush_flag = nullptr;
    d->stream2 = nullptr; d->ev_steps = d->ev_flush = nullptr; d->green = false;
    for (unsigned &w : d->resv) w = 0;
    lpgpu_default_opts(&d->opts);
    d->opts.method = (R == 1) ? LPGPU_PLAIN : LPGPU_BIGM;
    const size_t rows = (size_t)m + R, pcn = round_up(rows, UPD_TILE_ROWS) + UPD_TILE_ROWS;
    cudaError_t e = cudaSuccess;
    if (e == cudaSuccess) e = cudaMalloc(&d->T, rows * d->ld * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->basis, (size_t)m * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&d->prow, d->ld * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->wts, d->ld * sizeof(double));
    d->pcn = pcn;
    if (e == cudaSuccess) e = cudaMalloc(&d->pcol, 2 * pcn * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->bnext, pcn * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->xbuf, (size_t)n * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->st, sizeof(PivotState));
    if (e == cudaSuccess) e = cudaMalloc(&d->task_ctr, 2 * sizeof(int));
    if (e == cudaSuccess) e = cudaMemsetAsync(d->task_ctr, 0, 2 * sizeof(int), ctx->stream);
    if (e == cudaSuccess) e = cudaMalloc(&d->Ft, 2 * pcn * BLK_KMAX * sizeof(double));              // two sets of slices (pipelined blocks)
    if (e == cudaSuccess) e = cudaMalloc(&d->Pm, 2 * (size_t)BLK_KMAX * d->ld * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->objcur, (size_t)R * d->ld * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->bcur, pcn * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d->rowstamp, pcn * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&d->steprow, BLK_KMAX * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&d->cands, STEP_CLUSTER_MAX * sizeof(KeyCand));
    if (e == cudaSuccess) e = cudaMemsetAsync(d->cands, 0xff, STEP_CLUSTER_MAX * sizeof(KeyCand), ctx->stream);
    d->blk_dbg = nullptr; d->blk_dbg_n = 0;
    if (e == cudaSuccess && getenv("LPGPU_BLK_DEBUG")) {
        e = cudaMalloc(&d->blk_dbg, 4096 * 8 * sizeof(unsigned long long));
        if (e == cudaSuccess) e = cudaMemsetAsync(d->blk_dbg, 0, 4096 * 8 * sizeof(unsigned long long), ctx->stream);
    }
    if (e == cudaSuccess) e = cudaMemsetAsync(d->steprow, 0, BLK_KMAX * sizeof(int), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->Ft, 0, 2 * pcn * BLK_KMAX * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->Pm, 0, 2 * (size_t)BLK_KMAX * d->ld * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->rowstamp, 0, pcn * sizeof(int), ctx->stream);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(blk_flush_sp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FL_SMEM_BYTES);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(blk_flush_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MM_SMEM_BYTES);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(blk_flush_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RG_SMEM_BYTES);
    if (e == cudaSuccess) e = cudaHostAlloc(&d->h_st, sizeof(PivotState), cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->T, 0, rows * d->ld * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->prow, 0, d->ld * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->pcol, 0, 2 * pcn * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->bnext, 0, pcn * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->basis, 0, (size_t)m * sizeof(int), ctx->stream);
    if (e != cudaSuccess) {
        lpgpu_dense_destroy(d);
        return fail(ctx, e == cudaErrorMemoryAllocation ? LPGPU_ENOMEM : LPGPU_ECUDA, "dense_create: %s", cudaGetErrorString(e));
    }
    int rc = size_grid(d);
    if (rc == LPGPU_OK) rc = probe_step_cluster(d);
    if (rc == LPGPU_OK) rc = lpgpu_dense_reset(d, &d->opts);
    if (rc != LPGPU_OK) { lpgpu_dense_destroy(d); return rc; }
    *out = d;
    return LPGPU_OK;
}

void lpgpu_dense_destroy(lpgpu_dense *d) {
    if (!d) return;
    cudaSetDevice(d->ctx->device);
    if (d->T && d->st) settle_dense(d);
    cudaStreamSynchronize(d->ctx->stream);
    if (d->stream2) cudaStreamSynchronize(d->stream2);
    free_pipeline(d);
    if (d->up_stream) { cudaStreamSynchronize(d->up_stream); cudaStreamDestroy(d->up_stream); }
    if (d->ev_up) cudaEventDestroy(d->ev_up);
    if (d->ev_up_ready) cudaEventDestroy(d->ev_up_ready);
    for (cudaEvent_t e : d->ev) cudaEventDestroy(e);
    if (d->blk_dbg) {       // development aid: mean time between the step kernel's phase stamps
        std::vector<unsigned long long> h(4096 * 8);
        cudaMemcpy(h.data(), d->blk_dbg, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        const long n = std::min<long>(d->blk_dbg_n, 4096);
        double acc[8] = {0}, gap = 0;
        long cnt = 0, gcnt = 0;
        for (long i = 0; i < n; ++i) {
            const unsigned long long *s = &h[8 * i];
            if (!s[6] || !s[0]) continue;
            for (int q = 1; q <= 6; ++q) acc[q] += (double)(s[q] - s[q - 1]);
            if (i + 1 < n && h[8 * (i + 1)]) { gap += (double)(h[8 * (i + 1)] - s[6]); ++gcnt; }
            ++cnt;
        }
        fprintf(stderr, "[lpgpu blk debug] step cluster max %d, pipeline state %d, reserved SM words %08x %08x %08x %08x %08x\n", d->step_cl_max, d->pipe_state,
                d->resv[0], d->resv[1], d->resv[2], d->resv[3], d->resv[4]);
        if (cnt) fprintf(stderr, "[lpgpu blk debug] %ld steps: pdl_wait %.2f | column+ratio loop %.2f | best %.2f | row+price loop %.2f | best %.2f | b+state %.2f | gap to next start %.2f us\n",
                         cnt, acc[1] / cnt / 1e3, acc[2] / cnt / 1e3, acc[3] / cnt / 1e3, acc[4] / cnt / 1e3, acc[5] / cnt / 1e3, acc[6] / cnt / 1e3, gcnt ? gap / gcnt / 1e3 : 0.0);
        cudaFree(d->blk_dbg);
    }
    cudaFree(d->T); cudaFree(d->basis); cudaFree(d->prow); cudaFree(d->wts); cudaFree(d->pcol); cudaFree(d->bnext); cudaFree(d->xbuf);
    cudaFree(d->st); cudaFree(d->log); cudaFree(d->task_ctr);
    cudaFree(d->Ft); cudaFree(d->Pm); cudaFree(d->objcur); cudaFree(d->bcur); cudaFree(d->rowstamp); cudaFree(d->steprow); cudaFree(d->cands);
    if (d->h_st) cudaFreeHost(d->h_st);
    delete d;
}

size_t lpgpu_dense_ld(const lpgpu_dense *d) { return d ? d->ld : 0; }

int lpgpu_dense_reset(lpgpu_dense *d, const lpgpu_opts *opts) {
    if (!d) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    if (!opts_ok(opts, d->n, d->R)) return fail(ctx, LPGPU_EINVAL, "dense_reset: bad options (method/R/n_art)");
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rcs = settle_dense(d); if (rcs != LPGPU_OK) return rcs; }
    d->opts = *opts;
    if (d->opts.check_every <= 0) d->opts.check_every = 64;
    // the buffer only ever grows (log_alloc); the ACTIVE limit is this call's opts->log_cap: the kernels log against it,
    // n_logged and the download length follow it (a caller's pivot_log holds 2 * opts->log_cap ints, no more)
    if (opts->log_cap > d->log_alloc) {
        LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(d->log);
        d->log = nullptr;
        d->log_alloc = 0;
        LP_CUDA(ctx, cudaMalloc(&d->log, (size_t)opts->log_cap * 2 * sizeof(int)));
        d->log_alloc = opts->log_cap;
    }
    d->log_cap = opts->log_cap;
    if (d->log_cap > 0) LP_CUDA(ctx, cudaMemsetAsync(d->log, 0xff, (size_t)d->log_cap * 2 * sizeof(int), ctx->stream));
    if (opts->variant & LPGPU_VARIANT_DEVEX) {            // reference framework: every weight starts at 1
        fill_kernel<<<64, 256, 0, ctx->stream>>>(d->wts, d->ld, 1.0);
        LP_LAUNCH_CHECK(ctx);
    }
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // h_st may still be in flight
    PivotState s;
    s.status = LPGPU_RUNNING; s.iters = 0; s.phase = (opts->method == LPGPU_TWOPHASE) ? PH_ONE : PH_MAIN;
    s.phase1_iters = 0; s.r = -1; s.k = 0; s.do_pivot = 0; s.budget = 0; s.k_next = 0; s.cur = 0;
    s.blk_t = 0; s.blk_seq = 0;
    d->blk_open = false; d->blk_steps = 0;
    d->la_dirty = true;      // the blocked path's objective rows / b column are re-read from the tableau
    *d->h_st = s;
    LP_CUDA(ctx, cudaMemcpyAsync(d->st, d->h_st, sizeof s, cudaMemcpyHostToDevice, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return size_grid(d);
}

int lpgpu_dense_upload(lpgpu_dense *d, const double *T, size_t ld, const int *basis) {
    if (!d || !T || !basis) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    if (ld < (size_t)d->n + 1) return fail(ctx, LPGPU_EINVAL, "dense_upload: ld %zu < n+1", ld);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rcs = settle_dense(d); if (rcs != LPGPU_OK) return rcs; }
    LP_CUDA(ctx, cudaMemcpy2DAsync(d->T, d->ld * sizeof(double), T, ld * sizeof(double), ((size_t)d->n + 1) * sizeof(double),
                                   (size_t)d->m + d->R, cudaMemcpyHostToDevice, ctx->stream));
    LP_CUDA(ctx, cudaMemcpyAsync(d->basis, basis, (size_t)d->m * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // the caller may reuse its buffers
    d->la_dirty = true;
    return LPGPU_OK;
}

// The same upload without blocking the host: the copy runs on a stream of its own, after everything enqueued for this
// tableau so far, and everything enqueued for it later (run, download, ...) comes after the copy.  With pinned host
// buffers the transfer of the NEXT problem into one tableau object overlaps lpgpu_dense_run on another one.  The host
// buffers must stay untouched until the next call on this object (run, reset, download, ...) returns.
int lpgpu_dense_upload_async(lpgpu_dense *d, const double *T, size_t ld, const int *basis) {
    if (!d || !T || !basis) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    if (ld < (size_t)d->n + 1) return fail(ctx, LPGPU_EINVAL, "dense_upload_async: ld %zu < n+1", ld);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rcs = settle_dense(d); if (rcs != LPGPU_OK) return rcs; }
    if (!d->up_stream) {
        LP_CUDA(ctx, cudaStreamCreateWithFlags(&d->up_stream, cudaStreamNonBlocking));
        LP_CUDA(ctx, cudaEventCreateWithFlags(&d->ev_up, cudaEventDisableTiming));
        LP_CUDA(ctx, cudaEventCreateWithFlags(&d->ev_up_ready, cudaEventDisableTiming));
    }
    LP_CUDA(ctx, cudaEventRecord(d->ev_up_ready, ctx->stream));              // the tableau's earlier work (passes, downloads) is done first
    LP_CUDA(ctx, cudaStreamWaitEvent(d->up_stream, d->ev_up_ready, 0));
    LP_CUDA(ctx, cudaMemcpy2DAsync(d->T, d->ld * sizeof(double), T, ld * sizeof(double), ((size_t)d->n + 1) * sizeof(double),
                                   (size_t)d->m + d->R, cudaMemcpyHostToDevice, d->up_stream));
    LP_CUDA(ctx, cudaMemcpyAsync(d->basis, basis, (size_t)d->m * sizeof(int), cudaMemcpyHostToDevice, d->up_stream));
    LP_CUDA(ctx, cudaEventRecord(d->ev_up, d->up_stream));
    d->up_pending = true;
    d->la_dirty = true;
    return LPGPU_OK;
}

int lpgpu_dense_synth(lpgpu_dense *d, uint64_t seed, uint64_t lp_index, int ns) {
    if (!d) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    if (d->R != 1 || ns < 1 || ns + d->m != d->n) return fail(ctx, LPGPU_EINVAL, "dense_synth: needs R=1 and n = ns + m");
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rcs = settle_dense(d); if (rcs != LPGPU_OK) return rcs; }
    dense_synth_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(d->T, d->ld, d->m, ns, d->basis, seed, lp_index);
    LP_LAUNCH_CHECK(ctx);
    d->la_dirty = true;
    return LPGPU_OK;
}


int lpgpu_dense_run(lpgpu_dense *d, int max_pivots, lpgpu_result *res) {
    if (!d || !res || max_pivots < 0) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    RunCall R;
    R.d = d; R.max_pivots = max_pivots; R.timed = d->opts.time_kernels != 0;
    // a pipelined run call leaves its last block(s) pending; anything but another pipelined run call needs the tableau itself
    int rc = LPGPU_OK;
    if (d->up_pending || !(use_blocked(d) && want_pipeline(d) && d->pipe_state == 1 && d->pipe_on && !R.timed)) rc = settle_dense(d);
    if (rc == LPGPU_OK) rc = read_state(d);
    if (rc != LPGPU_OK) return rc;
    d->h_st->budget = max_pivots;
    R.dirty = d->la_dirty;
    if (R.dirty) d->h_st->k_next = 0;             // never trust a lookahead across an upload / synth / forced pivot
    d->la_dirty = false;
    LP_CUDA(ctx, cudaMemcpyAsync(d->st, d->h_st, sizeof(PivotState), cudaMemcpyHostToDevice, ctx->stream));
    R.iters_before = d->h_st->iters;
    rc = use_blocked(d) ? run_blocked(R) : run_rank1(R);
    if (rc != LPGPU_OK) return rc;
    // objective cells: with blocks pending the tableau's own objective cells are behind, the eagerly kept objective rows hold the same bits
    double zc[2] = {0.0, 0.0};
    const double *zsrc = (d->pend || d->blk_open) ? d->objcur : d->T + (size_t)d->m * d->ld;
    LP_CUDA(ctx, cudaMemcpyAsync(&zc[0], zsrc, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (d->R == 2)
        LP_CUDA(ctx, cudaMemcpyAsync(&zc[1], zsrc + d->ld, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (d->h_st->status == LPGPU_STEP_TIMEOUT) return fail(ctx, LPGPU_ECUDA, "dense_run: the CTAs of the persistent step kernel lost each other (exchange timeout)");
    res->status = d->h_st->status;
    res->iters = d->h_st->iters;
    res->phase1_iters = (d->opts.method == LPGPU_TWOPHASE) ? d->h_st->phase1_iters : 0;
    res->phase = d->h_st->phase;
    res->z = -zc[0];
    res->zM = (d->R == 2) ? -zc[1] : 0.0;
    res->n_logged = std::min(res->iters, d->log_cap);
    res->update_launches = R.done_updates;
    res->update_ms = (float)R.upd_ms;
    res->select_ms = (float)R.sel_ms;
    return LPGPU_OK;
}

int lpgpu_dense_pivot(lpgpu_dense *d, int r, int k) {
    if (!d) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    if (r < 0 || r >= d->m || k < 1 || k > d->n) return fail(ctx, LPGPU_EINVAL, "dense_pivot: (r=%d,k=%d) out of range", r, k);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rcs = settle_dense(d); if (rcs != LPGPU_OK) return rcs; }
    d->la_dirty = true;
    int rc = launch_select(d, r, k);
    if (rc != LPGPU_OK) return rc;
    rc = launch_update(d, nullptr);
    if (rc != LPGPU_OK) return rc;
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

int lpgpu_dense_download(lpgpu_dense *d, double *T, size_t ld, int *basis, int *pivot_log, double *x) {
    if (!d) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = d->ctx;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rcs = settle_dense(d); if (rcs != LPGPU_OK) return rcs; }      // always: the basis / log copies must not race a pending asynchronous upload either
    if (T) {
        if (ld < (size_t)d->n + 1) return fail(ctx, LPGPU_EINVAL, "dense_download: ld %zu < n+1", ld);
        LP_CUDA(ctx, cudaMemcpy2DAsync(T, ld * sizeof(double), d->T, d->ld * sizeof(double), ((size_t)d->n + 1) * sizeof(double),
                                       (size_t)d->m + d->R, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (basis) LP_CUDA(ctx, cudaMemcpyAsync(basis, d->basis, (size_t)d->m * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (pivot_log && d->log_cap > 0) {
        int rc = read_state(d);
        if (rc != LPGPU_OK) return rc;
        const int nl = std::min(d->h_st->iters, d->log_cap);
        if (nl > 0) LP_CUDA(ctx, cudaMemcpyAsync(pivot_log, d->log, (size_t)nl * 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (x) {
        LP_CUDA(ctx, cudaMemsetAsync(d->xbuf, 0, (size_t)d->n * sizeof(double), ctx->stream));
        extract_x_kernel<<<std::max(1, std::min(1024, (d->m + 255) / 256)), 256, 0, ctx->stream>>>(d->T, d->ld, d->basis, d->m, d->n, d->xbuf);
        LP_LAUNCH_CHECK(ctx);
        LP_CUDA(ctx, cudaMemcpyAsync(x, d->xbuf, (size_t)d->n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    }
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

int lpgpu_solve_dense(lpgpu_ctx *ctx, int m, int n, int R, double *T, size_t ld, int *basis, const lpgpu_opts *opts,
                      int *pivot_log, double *x, lpgpu_result *res) {
    if (!ctx || !T || !basis || !opts || !res) return LPGPU_EINVAL;
    if (!opts_ok(opts, n, R)) return fail(ctx, LPGPU_EINVAL, "solve_dense: bad options (method/R/n_art)");
    // the tableau object (1 GB of device memory at the headline size) is kept in the context between calls of
    // the same shape: cudaMalloc / cudaFree of that size cost more than the upload itself
    lpgpu_dense *d = ctx->solve_cache;
    int rc = LPGPU_OK;
    if (!d || d->m != m || d->n != n || d->R != R) {
        if (d) lpgpu_dense_destroy(d);
        ctx->solve_cache = d = nullptr;
        rc = lpgpu_dense_create(ctx, m, n, R, &d);
        if (rc != LPGPU_OK) return rc;
        ctx->solve_cache = d;
    }
    rc = lpgpu_dense_reset(d, opts);
    if (rc == LPGPU_OK) rc = lpgpu_dense_upload(d, T, ld, basis);
    if (rc == LPGPU_OK) rc = lpgpu_dense_run(d, opts->max_iter, res);
    if (rc == LPGPU_OK) rc = lpgpu_dense_download(d, opts->results_only ? nullptr : T, ld, basis, pivot_log, x);
    if (rc != LPGPU_OK) {           // do not keep an object in an unknown state
        lpgpu_dense_destroy(d);
        ctx->solve_cache = nullptr;
    }
    return rc;
}

}  // extern "C"
