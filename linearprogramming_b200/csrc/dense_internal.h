// dense_internal.h -- what dense.cu shares with shard.cu (the column-sharded path reuses the
// single-GPU tableau object and its elimination launch for each rank's column block).
#pragma once

#include <vector>

#include "ctx.h"
#include "block_common.cuh"

struct lpgpu_dense {
    lpgpu_ctx *ctx;
    int m, n, R;
    size_t ld;
    double *T;
    int *basis;
    double *prow, *pcol, *bnext, *xbuf;   // pcol: 2 x pcn (colbuf)
    double *wts;            // Devex reference weights, ld doubles, all 1 after a reset (LPGPU_VARIANT_DEVEX)
    size_t pcn;
    bool la_dirty;          // tableau changed behind the lookahead's back (upload / synth / forced pivot)
    lpgpu::PivotState *st;
    lpgpu::PivotState *h_st;       // pinned mirror
    int *log;
    int log_cap;            // ACTIVE limit = opts.log_cap of the last reset (what the kernels log against, n_logged and the download length)
    int log_alloc;          // pivots the device buffer has room for (only ever grows)
    lpgpu_opts opts;
    int grid_update;
    int grid_flush;         // persistent grid of the blocked path's pass
    long flush_tasks;       // tasks of one pass
    int fl_rb, fl_pf;       // its row tiles per task and L2 prefetch switch
    int ring_nd;            // ring pass: data slots in use
    int *task_ctr;          // work-stealing counters of the elimination kernel (2 ints)
    unsigned upd_seq;       // elimination launches so far
    std::vector<cudaEvent_t> ev;
    // blocked path (dense_block.cuh): lazy slices of the pivots not yet applied to T
    double *Ft, *Pm, *objcur, *bcur;   // Ft: BLK_KMAX x pcn multipliers, Pm: BLK_KMAX x ld scaled pivot rows
    int *rowstamp, *steprow;
    lpgpu::KeyCand *cands;     // per-CTA pricing candidates handed from one step launch to the next
    unsigned long long *blk_dbg;   // LPGPU_BLK_DEBUG: phase stamps of the step kernel
    long blk_dbg_n;
    bool blk_open;          // step kernels were enqueued since the last flush
    int blk_steps;          // ... how many
    int l2hint;             // pipelined blocks: bit 0 = the pass streams the tableau with L2 evict-first, bit 1 = slices evict-last
    int step_cl_max;        // CTAs per cluster the primal step kernel may use: 16 where the device places such clusters, else 8
    // pipelined blocks (dense.cu): the pass of block b (stream2, T -> T2) runs while block b+1 is decided (ctx->stream)
    int pipe_state;         // 0 = not set up yet, 1 = ready, -1 = unavailable (second tableau buffer did not fit)
    bool pipe_on;           // this run call uses it
    bool pend;              // a block's pass is in flight: T = the tableau before that block, T2 = where the pass writes
    int par;                // parity of the open / next block: which set of slices, stamps and info it uses
    double *T2, *Ft2, *Pm2;
    int *rowstamp2, *steprow2;
    lpgpu::BlkInfo *info;   // [2]
    int *flush_flag;        // a pass's CTAs on reserved SMs leave once a worker CTA is known to exist
    bool green;             // stream2 belongs to the green context of the SMs the step kernels do not get (no reservation needed)
    cudaStream_t stream2;
    cudaEvent_t ev_steps, ev_flush;
    unsigned resv[8];       // SMs left to the step kernels while a pass runs (bit per %smid)
    // persistent step kernel (dense_steps.cuh): one launch of steps_G CTAs decides all the pivots of a block
    void *xch;              // XSlot[2][STEPS_GMAX]: the CTAs' all-to-all through L2
    unsigned xch_seq;       // next exchange sequence number (strictly increasing over the life of the object)
    int steps_state;        // 0 = not set up yet, 1 = ready, -1 = unavailable
    int steps_G, steps_rpc, steps_cpc, steps_S;
    size_t steps_smem;
    unsigned resv_steps[8]; // SMs that grid was placed on while the device was idle
    bool persist_on;        // this run call uses it
    // lpgpu_dense_upload_async: copy stream, "copy done" / "tableau free" events
    cudaStream_t up_stream;
    cudaEvent_t ev_up, ev_up_ready;
    bool up_pending;
};

namespace lpgpu {

DenseParams make_params(const lpgpu_dense *d);
// Enqueue the elimination pass.  cnext: where the pass leaves the next entering column
// (nullptr = the other half of the tableau's own ping-pong buffer).
int launch_update(lpgpu_dense *d, double *cnext);
// column-sharded, peer-memory transport: the next entering column goes to byte offset peer_off of every
// rank's exchange region (peer_x: device array of n_peers pointers)
int launch_update_peers(lpgpu_dense *d, void *const *peer_x, int n_peers, size_t peer_off);
int read_state(lpgpu_dense *d);
// blocked path (dense_block.cuh): buffers of the current block, objcur / bcur <- tableau, the pass that applies a block
BlockBufs make_bufs(const lpgpu_dense *d);
int launch_block_init(lpgpu_dense *d);
int launch_block_flush(lpgpu_dense *d);

}  // namespace lpgpu
