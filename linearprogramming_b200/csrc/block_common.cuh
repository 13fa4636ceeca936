// block_common.cuh -- device helpers shared by the blocked path of one tableau (dense_block.cuh) and its
// column-sharded form (shard.cu): buffers of a block, row stamps, arg-max on order-preserving keys.
#pragma once

#include "dense_select.cuh"

namespace lpgpu {

constexpr int BLK_KMAX = 32;           // pivots per block at most (shared-memory tile of the pass)
constexpr int STEP_CLUSTER_MAX = 16;   // the primal step kernel runs as a cluster of 8 (portable) or 16 CTAs

struct KeyCand;

// what the pass needs to know about a block once the step kernels have moved on to the next one
// (pipelined blocks: the pass of block b runs while block b+1 is being decided)
struct BlkInfo {
    int kc;              // pivots decided in the block
    int seq;             // its sequence number (the tag of its row stamps)
};

struct BlockBufs {
    double *Fs;          // BLK_KMAX x pcn: Fs[s*pcn + i] = a_{i,k_s} before step s (the multiplier of row i at step s)
    KeyCand *cands;      // SELECT_CLUSTER: per-CTA pricing candidates for the next pivot (st->k_next == -2)
    double *Pm;          // BLK_KMAX x ld: Pm[s*ld + j] = scaled pivot row of step s (col 0 = b_r')
    double *objcur;      // R x ld: objective rows as of the last decided pivot (row m first)
    double *bcur;        // pcn: b column as of the last decided pivot (rows < m)
    int *rowstamp;       // pcn: (block sequence << 8) | (s + 1) if the row was the pivot row of step s of that block
    int *steprow;        // BLK_KMAX: pivot row of step s of the current block
    int K;               // pivots per block, 1..BLK_KMAX
    // pipelined blocks (dense.cu, variant bit 0x1000 off): P.T is the tableau as it stood BEFORE the previous block,
    // whose pass is still running; the lazy chains run through that block's slices first, then this block's
    const double *Fall;  // both sets of slices, [2][BLK_KMAX] steps each: set `par` is this block's (= Fs / Pm), the other one
    const double *Pall;  // the previous block's
    const int *stamp_prev;   // the previous block's row stamps (pend != 0)
    BlkInfo *info;       // [2]: written for parity `par` by every step launch; [par ^ 1] = the previous block (nullptr: not pipelined)
    int par;             // parity of this block
    int pend;            // 1: the previous block has not been applied to P.T
    int l2keep;          // 1: the slices are read and written with the L2 evict-last priority (they are re-read every step
                         //    while the pass streams the whole tableau through the same L2)
    unsigned long long *dbg;   // LPGPU_BLK_DEBUG=1: 8 globaltimer stamps per launch slot (development aid), else nullptr
};

__device__ __forceinline__ int blk_row_start(int stamp, int seq) { return ((stamp >> 8) == seq) ? (stamp & 0xff) : 0; }

// ------------------------------------------------------------------ arg-max on order-preserving keys
// "maximise (p, q), ties to the lowest index" with redux.sync on 64-bit keys instead of shuffle trees
// of (double, double, int) triples: a warp level is 3 (one key) or 5 (two keys) reductions, a CTA level one
// barrier, the cluster level one DSMEM exchange.  Only the winning INDEX comes back (and the keys).
struct KeyCand {
    unsigned long long kp, kq;     // order-preserving images of p and q
    int idx;                       // -1 = none
};

__device__ __forceinline__ KeyCand keycand_none() {
    KeyCand c;
    c.kp = 0; c.kq = 0; c.idx = -1;
    return c;
}

// any double -> u64 with the same order; -0 and +0 map to the same key (they compare equal)
__device__ __forceinline__ unsigned long long ord_key(double x) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x + 0.0);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

__device__ __forceinline__ bool keycand_better(const KeyCand &a, const KeyCand &b) {
    if (a.idx < 0) return false;
    if (b.idx < 0) return true;
    if (a.kp != b.kp) return a.kp > b.kp;
    if (a.kq != b.kq) return a.kq > b.kq;
    return a.idx < b.idx;
}

template <bool TWO>
__device__ __forceinline__ KeyCand warp_key_best(KeyCand c) {
    constexpr unsigned FULL = 0xffffffffu;
    bool in = c.idx >= 0;
    if (__ballot_sync(FULL, in) == 0u) return keycand_none();
    KeyCand w;
    w.kp = 0;
    if (TWO) {
        const unsigned h = __reduce_max_sync(FULL, in ? (unsigned)(c.kp >> 32) : 0u);
        in = in && (unsigned)(c.kp >> 32) == h;
        const unsigned l = __reduce_max_sync(FULL, in ? (unsigned)c.kp : 0u);
        in = in && (unsigned)c.kp == l;
        w.kp = ((unsigned long long)h << 32) | l;
    }
    const unsigned h = __reduce_max_sync(FULL, in ? (unsigned)(c.kq >> 32) : 0u);
    in = in && (unsigned)(c.kq >> 32) == h;
    const unsigned l = __reduce_max_sync(FULL, in ? (unsigned)c.kq : 0u);
    in = in && (unsigned)c.kq == l;
    w.kq = ((unsigned long long)h << 32) | l;
    w.idx = (int)__reduce_min_sync(FULL, in ? (unsigned)c.idx : 0x7fffffffu);
    return w;
}

// CTA-wide: per-warp winners through shared memory, one barrier; every warp reduces the winners again, so
// the result is in all threads.  `scratch`: blockDim.x / 32 entries, used for ONE reduction per launch.
template <bool TWO>
__device__ __forceinline__ KeyCand block_key_best(KeyCand c, KeyCand *scratch) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
    c = warp_key_best<TWO>(c);
    if (lane == 0) scratch[warp] = c;
    __syncthreads();
    return warp_key_best<TWO>(lane < nwarp ? scratch[lane] : keycand_none());
}

// cluster-wide: every CTA pushes its winner into every CTA's `slot` (DSMEM), one cluster barrier
template <bool TWO, int CL = SELECT_CLUSTER>
__device__ __forceinline__ KeyCand cluster_key_best(cg::cluster_group &cl, KeyCand c, KeyCand *scratch, KeyCand *slot) {
    c = block_key_best<TWO>(c, scratch);
    if (threadIdx.x < CL) cl.map_shared_rank(slot, threadIdx.x)[cl.block_rank()] = c;
    cl.sync();
    const int lane = threadIdx.x & 31;
    return warp_key_best<TWO>(lane < CL ? slot[lane] : keycand_none());
}

// pricing candidate of column j from the (current) objective row values.  mode: 0 = cost row alone (plain,
// two-phase main phase), 1 = Big-M (M row first, cost row second), 2 = two-phase phase one (row m+1 alone)
enum { PRICE_COST = 0, PRICE_BIGM = 1, PRICE_PHASE1 = 2 };

// PR: pricing rule, a template parameter (the step kernels sit at their register limit and the default rule must not pay for
// the options).  PR_DANTZIG: largest reduced cost.  PR_BLAND: every eligible column gets the same key, the arg-max's tie-break
// (lowest column) decides.  PR_DEVEX: sigma_j^2 / w_j with the reference weight w_j >= 1 of the column (one rounded multiply,
// one IEEE division; under Big-M both levels are scaled, below a positive M part the cost breaks ties as it is) -- the
// oracle's price_single / price_bigm with pricing == 2.
enum { PR_DANTZIG = 0, PR_BLAND = 1, PR_DEVEX = 2 };

template <int PR = PR_DANTZIG>
__device__ __forceinline__ void blk_price(const DenseParams &P, int mode, int j, double oc, double om, double w, KeyCand &c) {
    KeyCand t;
    t.idx = j;
    if (mode == PRICE_BIGM) {
        double p = om;
        if (fabs(p) <= P.eps) p = 0.0;
        if (p > 0.0 || (p == 0.0 && oc > P.eps)) {
            if (PR == PR_DEVEX) {
                t.kq = ord_key((p == 0.0) ? (oc * oc) / w : oc);
                t.kp = ord_key((p * p) / w);
            } else {
                t.kp = (PR == PR_BLAND) ? 0ull : ord_key(p);
                t.kq = (PR == PR_BLAND) ? 1ull : ord_key(oc);
            }
            if (keycand_better(t, c)) c = t;
        }
    } else {
        const double q = (mode == PRICE_PHASE1) ? om : oc;
        if (q > P.eps) {
            t.kp = 0;
            t.kq = (PR == PR_BLAND) ? 1ull : (PR == PR_DEVEX) ? ord_key((q * q) / w) : ord_key(q);
            if (keycand_better(t, c)) c = t;
        }
    }
}

}  // namespace lpgpu
