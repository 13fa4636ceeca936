// shard.cu -- one large tableau column-sharded over several GPUs (include/lpgpu.h, lpgpu_shard_*).
//
// Rank g owns global columns col0_g .. col0_g + ncols_g - 1 of the SimplexMatrix-shaped tableau
// (Source/matrix.c:42-72) as a local tableau (m+R) x (1 + ncols_g) whose column 0 is a replica of
// the b column; basis and pivot state are replicated too.  Per pivot:
//
//   shard_select    ratio test on the exchanged entering column (redundant on every rank), pivot-row
//                   scaling of the local block, and LOOKAHEAD pricing of the local block for the pivot
//                   after this one -> one (p, q, global column) candidate per rank
//   exchange A      all-gather of the candidates (32 B per rank)
//   shard_resolve   arg-max over the gathered candidates (ties -> lowest global column)
//   dense_update    the single-GPU elimination kernel on the local block; the owner of the next
//                   entering column extracts its updated values during the pass
//   exchange B      all-gather of that column (8 * (m+R) bytes per rank; only the owner's is read)
//
// so a pivot costs one streaming pass over each rank's block plus two latency-bound exchanges.
// Exchanges go through ncclAllGather (one process per GPU, NVLink / NVSwitch) or, when every rank
// lives in one process on one device, through device copies (test vehicle for the N > 1 logic).
#include <algorithm>
#include <new>
#include <string.h>
#include <vector>

#include "dense_internal.h"
#include "shard_nccl.h"
#include "shard_p2p.cuh"

using namespace lpgpu;

namespace lpgpu {

constexpr int PAIR_DOUBLES = 4;

struct ShardState {
    int k_next;     // entering column of the pivot to decide next (global, 1-based; 0 = no candidate)
    int owner;      // rank that owns it
};

struct ShardParams {
    DenseParams P;           // the local tableau; P.n = local columns, P.basis holds GLOBAL columns
    int rank, world;
    int col0;                // global index of local column 1
    int n_global, n_art;     // artificial columns are the last n_art GLOBAL columns
    ShardState *sh;
    double *sendpair;        // 4 doubles: p, q, global column (-1 = none), 0
    const double *pairs;     // world x 4 doubles
    double *sendcol;         // pcn doubles: the next entering column if this rank owns it
    const double *gathered;  // world x pcn doubles
    // peer-memory transport (shard_p2p.cuh); peer_x == nullptr: stream-ordered exchange (NCCL / shared buffers)
    void *const *peer_x;     // world pointers: every rank's exchange region, own included
    unsigned iter;           // pipeline iteration, counted identically on every rank
    int post;                // 0: decide only, post nothing (the extra select that ends a run call)
    int ll;                  // 1: flag-in-payload exchange (blocked path over peer memory; shard_p2p.cuh), 0: data + release flag
    int forced_r;            // blocked path, two-phase drive-out: >= 0 = pivot in THIS row on its lowest eligible (non-artificial,
                             // |a| > eps) global column, found and agreed on inside the step kernel; -1 = the normal decision
};

constexpr unsigned long long XCHG_TIMEOUT_NS = 30ull * 1000 * 1000 * 1000;

// candidate of this rank -> every rank's pairs[parity][me], then the flags (release after the data)
__device__ __forceinline__ void post_pair(const ShardParams &S) {
    const int par = (int)(S.iter & 1u);
    for (int g = 0; g < S.world; ++g) {
        double *dst = xptr<double>(S.peer_x[g], XLayout::pairs(par)) + (size_t)S.rank * PAIR_DOUBLES;
#pragma unroll
        for (int q = 0; q < PAIR_DOUBLES; ++q) dst[q] = S.sendpair[q];
    }
    for (int g = 0; g < S.world; ++g)
        st_release_sys(xptr<unsigned>(S.peer_x[g], XLayout::pair_flag(par)) + S.rank, S.iter + 1u);
}

// The same, one lane per peer (call with the first warp of a CTA, all 32 lanes): the data stores, ONE system
// fence and the flag go out to all peers at once, instead of a release store per peer in sequence (each of
// which waits for an NVLink round trip).
__device__ __forceinline__ void post_pair_lanes(const ShardParams &S, const double (&v)[PAIR_DOUBLES]) {
    const int par = (int)(S.iter & 1u), g = (int)(threadIdx.x & 31);
    for (int gg = g; gg < S.world; gg += 32) {
        double *dst = xptr<double>(S.peer_x[gg], XLayout::pairs(par)) + (size_t)S.rank * PAIR_DOUBLES;
#pragma unroll
        for (int q = 0; q < PAIR_DOUBLES; ++q) dst[q] = v[q];
    }
    __threadfence_system();
    for (int gg = g; gg < S.world; gg += 32)
        st_release_sys(xptr<unsigned>(S.peer_x[gg], XLayout::pair_flag(par)) + S.rank, S.iter + 1u);
}

// block-wide wait for every rank's candidate of this iteration; false = a peer never posted
__device__ __forceinline__ bool wait_pairs(const ShardParams &S) {
    const int par = (int)(S.iter & 1u);
    int ok = 1;
    if (threadIdx.x < (unsigned)S.world)
        ok = spin_until(xptr<unsigned>(S.peer_x[S.rank], XLayout::pair_flag(par)) + threadIdx.x, S.iter + 1u, XCHG_TIMEOUT_NS);
    return __syncthreads_and(ok) != 0;
}

__device__ __forceinline__ void store_pair(double *dst, const Cand &c, int col0) {
    dst[0] = c.p;
    dst[1] = c.q;
    dst[2] = c.idx > 0 ? (double)(col0 + c.idx - 1) : -1.0;
    dst[3] = 0.0;
}

// flag-in-payload form of the same message: 5 words (p, q as two halves each, the global column) per sender
__device__ __forceinline__ void ll_post_pair_to(const ShardParams &S, int g, const double *v, unsigned iter) {
    unsigned long long *dst = xptr<unsigned long long>(S.peer_x[g], XLayout::ll_pairs((int)(iter & 1u), S.P.pcn)) + (size_t)S.rank * 8;
    ll_put_f64(dst, v[0], iter + 1u);
    ll_put_f64(dst + 2, v[1], iter + 1u);
    ll_put(dst + 4, (unsigned)(int)v[2], iter + 1u);
}
// this rank's candidate (4 doubles as store_pair leaves them) -> every rank; one thread
__device__ __forceinline__ void post_cand(const ShardParams &S, const double *v) {
    if (S.ll) { for (int g = 0; g < S.world; ++g) ll_post_pair_to(S, g, v, S.iter); }
    else post_pair(S);
}
// ... one lane per peer (first warp of a CTA, all 32 lanes)
__device__ __forceinline__ void post_cand_lanes(const ShardParams &S, const double (&v)[PAIR_DOUBLES]) {
    if (S.ll) { for (int g = (int)(threadIdx.x & 31); g < S.world; g += 32) ll_post_pair_to(S, g, v, S.iter); }
    else post_pair_lanes(S, v);
}

// arg-max over the gathered candidates; idx = GLOBAL column, owner through *owner
__device__ __forceinline__ Cand best_pair(const double *pairs, int world, int *owner) {
    Cand best = cand_none();
    int who = 0;
    for (int g = 0; g < world; ++g) {
        Cand t;
        t.p = __ldcg(pairs + g * PAIR_DOUBLES);          // written by other GPUs: L2 is the point of coherence
        t.q = __ldcg(pairs + g * PAIR_DOUBLES + 1);
        t.idx = (int)__ldcg(pairs + g * PAIR_DOUBLES + 2);
        if (t.idx > 0 && cand_better(t, best)) { best = t; who = g; }
    }
    *owner = who;
    return best;
}

// Bootstrap pricing straight from the tableau (first pivot after an upload / synth).
__global__ void __launch_bounds__(SELECT_THREADS) shard_price_kernel(ShardParams S) {
    __shared__ Cand scratch[SELECT_THREADS / 32];
    const DenseParams &P = S.P;
    Cand c;
    if (P.method == LPGPU_BIGM) c = price_bigm(P, scratch);
    else if (P.method == LPGPU_TWOPHASE && P.st->phase == PH_ONE) c = price_row(P, P.m + 1, P.n, scratch);
    else {
        // main phase of the two-phase method: the artificial columns (the last n_art GLOBAL columns) are not priced
        const int jmax = (P.method == LPGPU_TWOPHASE) ? max(0, min(P.n, S.n_global - S.n_art - S.col0 + 1)) : P.n;
        c = price_row(P, P.m, jmax, scratch);
    }
    if (threadIdx.x == 0) {
        store_pair(S.sendpair, c, S.col0);
        if (S.peer_x) post_cand(S, S.sendpair);
    }
}

// After exchange A.  gather = 1 (bootstrap): the owner also pulls the entering column out of its
// tableau (strided) and every rank refreshes its compact copy of b.
__global__ void __launch_bounds__(SELECT_THREADS) shard_resolve_kernel(ShardParams S, int gather) {
    const DenseParams &P = S.P;
    const int par = (int)(S.iter & 1u);
    const double *pairs = S.pairs;
    if (S.peer_x) {
        if (!wait_pairs(S)) {
            if (blockIdx.x == 0 && threadIdx.x == 0) { P.st->status = LPGPU_EXCHANGE_TIMEOUT; P.st->do_pivot = 0; }
            return;
        }
        pairs = xptr<double>(S.peer_x[S.rank], XLayout::pairs(par));
    }
    int owner = 0;
    const Cand best = best_pair(pairs, S.world, &owner);
    const int k = best.idx > 0 ? best.idx : 0;
    const bool mine = (k > 0) && (owner == S.rank);
    const int kl = mine ? k - S.col0 + 1 : 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        S.sh->k_next = k;
        S.sh->owner = owner;
        P.st->k_next = kl;
    }
    if (gather) {
        const int rows = P.m + P.R;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += gridDim.x * blockDim.x) {
            if (mine) {
                const double a = P.T[(size_t)i * P.ld + (size_t)kl];
                if (S.peer_x) {
                    for (int g = 0; g < S.world; ++g) xptr<double>(S.peer_x[g], XLayout::cols(par, P.pcn))[i] = a;
                } else {
                    S.sendcol[i] = a;
                }
            }
            P.bnext[i] = P.T[(size_t)i * P.ld];
        }
    }
}

// peer-memory transport: after the elimination pass (stream order) the OWNER of the entering column --
// the only rank that stored it -- tells every rank that the column of this iteration is complete in
// its region.  (Were every rank to post, a rank's own flag would overtake the owner's data.)
__global__ void shard_publish_kernel(ShardParams S) {
    if (S.sh->owner != S.rank) return;
    const int g = (int)threadIdx.x;
    if (g < S.world) {
        __threadfence_system();
        st_release_sys(xptr<unsigned>(S.peer_x[g], XLayout::col_flag((int)(S.iter & 1u))), S.iter + 1u);
    }
}

// any artificial (GLOBAL column > n_global - n_art) still basic at a positive level?  b from bnext.
__device__ __forceinline__ bool shard_art_positive(const ShardParams &S) {
    const int nreal = S.n_global - S.n_art;
    int f = 0;
    for (int i = (int)threadIdx.x; i < S.P.m; i += blockDim.x)
        if (S.P.basis[i] > nreal && S.P.bnext[i] > S.P.eps) f = 1;
    return __syncthreads_or(f) != 0;
}

// One cluster of 8 CTAs x 1024 threads per rank (same shape as dense_select_kernel's fast path).
__global__ void __cluster_dims__(SELECT_CLUSTER, 1, 1) __launch_bounds__(SELECT_THREADS)
shard_select_kernel(ShardParams S) {
    __shared__ Cand scratch[SELECT_THREADS / 32];
    __shared__ Cand slot_ratio[SELECT_CLUSTER], slot_look[SELECT_CLUSTER];
    cg::cluster_group cl = cg::this_cluster();
    const DenseParams &P = S.P;
    const int rank = (int)cl.block_rank();
    const int gtid = rank * SELECT_THREADS + (int)threadIdx.x;
    constexpr int GSTRIDE = SELECT_CLUSTER * SELECT_THREADS;

    cluster_entry_arrive();     // DSMEM rule: no remote access before every CTA of the cluster has started
    PivotState *st = P.st;
    const int status0 = st->status, iters = st->iters, budget = st->budget;
    const int k = S.sh->k_next, owner = S.sh->owner;
    cluster_entry_wait();
    if (status0 != LPGPU_RUNNING) {
        cl.sync();
        if (gtid == 0) {
            st->do_pivot = 0;
            if (S.peer_x && S.post) post_pair(S);      // peers still wait for a candidate every iteration
        }
        return;
    }
    int status = status0, r = -1;
    bool have = false;
    const double *col = S.gathered + (size_t)owner * P.pcn;
    if (S.peer_x) {
        // the entering column was stored into this rank's region by its owner during iteration iter-1
        const int cpar = (int)((S.iter - 1u) & 1u);
        col = xptr<double>(S.peer_x[S.rank], XLayout::cols(cpar, P.pcn));
        int ok = 1;
        if (threadIdx.x == 0 && k > 0)
            ok = spin_until(xptr<unsigned>(S.peer_x[S.rank], XLayout::col_flag(cpar)), S.iter, XCHG_TIMEOUT_NS);
        ok = __syncthreads_and(ok);
        if (!ok) {                                      // every CTA of the cluster times out alike
            cl.sync();
            if (gtid == 0) { st->status = LPGPU_EXCHANGE_TIMEOUT; st->do_pivot = 0; }
            return;
        }
    }
    if (k <= 0) {
        // nothing prices in: optimal, unless Big-M left an artificial basic at a positive level
        const bool bad = (P.method == LPGPU_BIGM) && shard_art_positive(S);
        status = bad ? LPGPU_INFEASIBLE : LPGPU_OPTIMAL;
    } else {
        Cand c = cand_none();
        for (int i = gtid; i < P.m; i += GSTRIDE) {
            const double a = __ldcg(col + i);
            if (a > P.eps) {
                Cand t;
                t.p = 0.0;
                t.q = -(P.bnext[i] / a);
                t.idx = i;
                if (cand_better(t, c)) c = t;
            }
        }
        const Cand l = cluster_best(cl, c, scratch, slot_ratio);
        if (l.idx < 0) status = LPGPU_UNBOUNDED;
        else { r = l.idx; have = true; }
        if (have && iters >= P.max_iter) { status = LPGPU_ITER_LIMIT; have = false; }
    }
    if (have && budget <= 0) have = false;      // pending pivot stays for the next run call
    cl.sync();                                  // all state reads are done before it is rewritten
    if (gtid == 0) {
        st->status = status;
        st->do_pivot = have ? 1 : 0;
        st->cur = 0;
        if (have) {
            st->r = r;
            st->k = k;
            st->iters = iters + 1;
            st->budget = budget - 1;
            if (iters < P.log_cap) { P.log[2 * iters] = k; P.log[2 * iters + 1] = r; }
            P.basis[r] = k;
        }
    }
    if (!have) {
        if (gtid == 0 && S.peer_x && S.post) post_pair(S);
        return;
    }

    // ---- stage the pivot: entering column (all rows) into the local pivot-column buffer, scaled
    // local slice of the pivot row, and the local lookahead candidate for the pivot after this one
    const int rows = P.m + P.R;
    for (int i = gtid; i < rows; i += GSTRIDE) P.pcol[i] = __ldcg(col + i);
    const double piv = __ldcg(col + r);
    const double *xr = P.T + (size_t)r * P.ld;
    const bool bigm = (P.method == LPGPU_BIGM);
    const double *oc = P.T + (size_t)P.m * P.ld;
    const double *om = P.T + (size_t)(P.m + 1) * P.ld;            // Big-M row (R = 2 only)
    const double fc = -__ldcg(col + P.m);
    const double fm = bigm ? -__ldcg(col + P.m + 1) : 0.0;
    Cand c = cand_none();
    for (size_t j = (size_t)gtid; j < P.ld; j += GSTRIDE) {
        const double pr = div_skip_zero(xr[j], piv);
        P.prow[j] = pr;
        if (j >= 1 && j <= (size_t)P.n) {
            const double q = __fma_rn(fc, pr, oc[j]);
            Cand t;
            t.idx = (int)j;
            if (bigm) {
                double p = __fma_rn(fm, pr, om[j]);
                if (fabs(p) <= P.eps) p = 0.0;
                t.p = P.bland ? 0.0 : p; t.q = P.bland ? 1.0 : q;
                if ((p > 0.0 || (p == 0.0 && q > P.eps)) && cand_better(t, c)) c = t;
            } else {
                t.p = 0.0; t.q = P.bland ? 1.0 : q;
                if (q > P.eps && cand_better(t, c)) c = t;
            }
        }
    }
    c = cluster_best(cl, c, scratch, slot_look);
    if (gtid < 32) {                                        // first warp of CTA 0: one lane per peer
        double v[PAIR_DOUBLES];
        store_pair(v, c, S.col0);
        if (gtid == 0) store_pair(S.sendpair, c, S.col0);
        if (S.peer_x) post_pair_lanes(S, v);
    }
}

// ------------------------------------------------------------------ blocked path, column-sharded (peer-memory transport)
// The blocked path of dense_block.cuh with the two decisions that need every rank's data exchanged over
// NVLink inside ONE cluster kernel per pivot and rank:
//   wait for every rank's pricing candidate of this pivot -> entering column k and its owner;
//   the OWNER runs its tableau column through the t pivots of the block (its own P, the replicated F) and
//   stores the result into every rank's exchange region, then the "column complete" flag;
//   every rank: column -> F[t], ratio test on its replica of b, its slice of row r lazily brought up to
//   date and scaled -> P[t], objective rows / b updated, local pricing -> candidate of the NEXT pivot posted.
// The pass that applies a block (blk_flush_kernel) is local: no exchange.  Every rank takes identical
// decisions from identical data, so no status has to be agreed on.
constexpr int SSTEP_PB = 8, SSTEP_CPT = 2;

// Two-phase (R = 2): phase one prices row m+1 over all columns, the main phase row m without the artificial columns (GLOBAL
// column > n_global - n_art); when phase one prices nothing in anywhere the kernel checks the artificials (infeasible) or
// sets the phase to drive-out and stops.  The drive-out pivots are forced steps (S.forced_r >= 0, one-step blocks on an
// up-to-date tableau, host loop of shard_run_blocked): every rank offers the lowest eligible column of its slice of that
// row, the lowest global column wins, its owner ships the column, and the row is pivoted without a ratio test.
// BL: Bland's rule (every eligible column gets the same key; the lowest global column wins the exchange).
template <int RR, bool BL>
__device__ __forceinline__ void shard_blk_step_body(const ShardParams &S, const BlockBufs &B, int first) {
    __shared__ KeyCand scr_b[SELECT_THREADS / 32], scr_c[SELECT_THREADS / 32];
    __shared__ KeyCand slot_b[SELECT_CLUSTER], slot_c[SELECT_CLUSTER];
    __shared__ double s_pk[BLK_KMAX], s_fr[BLK_KMAX];
    __shared__ int s_flag;
    cg::cluster_group cl = cg::this_cluster();
    const DenseParams &P = S.P;
    const int crank = (int)cl.block_rank();
    const int gtid = crank * SELECT_THREADS + (int)threadIdx.x;
    constexpr int GSTRIDE = SELECT_CLUSTER * SELECT_THREADS;

    auto stamp = [&](int slot) {        // LPGPU_BLK_DEBUG=1: phase times of a step (printed when the tableau object is destroyed)
        if (B.dbg && gtid == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            B.dbg[slot] = now;
        }
    };
    stamp(0);
    cluster_entry_arrive();     // DSMEM rule: no remote access before every CTA of the cluster has started
    PivotState *st = P.st;
    const int status0 = __ldcg(&st->status), iters = __ldcg(&st->iters), budget = __ldcg(&st->budget);
    const int t = first ? 0 : __ldcg(&st->blk_t), seq = __ldcg(&st->blk_seq) + (first ? 1 : 0);
    const bool bigm = (P.method == LPGPU_BIGM);
    const double *oc = B.objcur, *om = B.objcur + P.ld;
    const size_t pcn = P.pcn;
    const int rows = P.m + RR;
    const int phase0 = __ldcg(&st->phase);
    int phase = phase0, p1 = __ldcg(&st->phase1_iters);
    const bool forced = S.forced_r >= 0;
    const int pmode = bigm ? PRICE_BIGM : (phase0 == PH_ONE ? PRICE_PHASE1 : PRICE_COST);
    // local columns that may be priced: all of them, or (main phase of the two-phase method) the non-artificial ones
    const int jmax = (P.method == LPGPU_TWOPHASE && phase0 != PH_ONE) ? max(0, min(P.n, S.n_global - S.n_art - S.col0 + 1)) : P.n;
    cluster_entry_wait();

    // ---- every rank's candidate of this pivot (posted at the end of the previous step or by the bootstrap)
    const unsigned pit = S.iter - 1u;
    const int ppar = (int)(pit & 1u);
    if (forced && status0 == LPGPU_RUNNING) {
        // drive-out: this rank's offer = the lowest non-artificial column of its slice of row forced_r with |a| > eps
        // (the tableau is up to date: forced steps are one-step blocks).  Posted under the number the wait below expects.
        __shared__ KeyCand scr_f[SELECT_THREADS / 32];
        __shared__ KeyCand slot_f[SELECT_CLUSTER];
        const int jreal = max(0, min(P.n, S.n_global - S.n_art - S.col0 + 1));
        const double *xr = P.T + (size_t)S.forced_r * P.ld;
        KeyCand c = keycand_none();
        for (int j = 1 + gtid; j <= jreal; j += GSTRIDE)
            if (fabs(xr[j]) > P.eps) { c.kp = 0; c.kq = 1; c.idx = j; break; }
        c = cluster_key_best<false>(cl, c, scr_f, slot_f);          // equal keys: the lowest local column
        if (gtid < 32) {
            Cand out = cand_none();
            if (c.idx > 0) { out.idx = c.idx; out.p = 0.0; out.q = 1.0; }
            double v[PAIR_DOUBLES];
            store_pair(v, out, S.col0);
            ShardParams Sp = S;
            Sp.iter = pit;
            post_cand_lanes(Sp, v);
        }
    }
    __shared__ double s_cand[XMAXW * PAIR_DOUBLES];
    if (S.ll) {                 // flag-in-payload words: thread g collects rank g's candidate (a dead peer traps the kernel)
        if (threadIdx.x < (unsigned)S.world) {
            const unsigned long long *src = xptr<unsigned long long>(S.peer_x[S.rank], XLayout::ll_pairs(ppar, pcn)) + (size_t)threadIdx.x * 8;
            s_cand[threadIdx.x * PAIR_DOUBLES] = ll_wait_f64(src, pit + 1u);
            s_cand[threadIdx.x * PAIR_DOUBLES + 1] = ll_wait_f64(src + 2, pit + 1u);
            s_cand[threadIdx.x * PAIR_DOUBLES + 2] = (double)(int)ll_wait(src + 4, pit + 1u);
        }
        __syncthreads();
    } else {
        int ok = 1;
        if (threadIdx.x < (unsigned)S.world)
            ok = spin_until(xptr<unsigned>(S.peer_x[S.rank], XLayout::pair_flag(ppar)) + threadIdx.x, pit + 1u, XCHG_TIMEOUT_NS);
        if (!__syncthreads_and(ok)) {                       // every CTA of the cluster times out alike
            cl.sync();
            if (gtid == 0) { st->status = LPGPU_EXCHANGE_TIMEOUT; st->do_pivot = 0; }
            return;
        }
    }
    stamp(1);                   // every rank's candidate has arrived
    int owner = 0;
    Cand best = cand_none();
    if (S.ll) {
        for (int g = 0; g < S.world; ++g) {
            Cand c2;
            c2.p = s_cand[g * PAIR_DOUBLES]; c2.q = s_cand[g * PAIR_DOUBLES + 1]; c2.idx = (int)s_cand[g * PAIR_DOUBLES + 2];
            if (c2.idx > 0 && cand_better(c2, best)) { best = c2; owner = g; }
        }
    } else best = best_pair(xptr<double>(S.peer_x[S.rank], XLayout::pairs(ppar)), S.world, &owner);
    const int k = best.idx > 0 ? best.idx : 0;              // GLOBAL column
    const bool mine = (k > 0) && (owner == S.rank);
    const int kl = mine ? k - S.col0 + 1 : 0;
    const int par = (int)(S.iter & 1u);
    double *mycol = xptr<double>(S.peer_x[S.rank], XLayout::cols(par, pcn));

    int status = status0, r = -1;
    bool have = false;
    if (status0 == LPGPU_RUNNING && (forced || phase0 != PH_DRIVEOUT)) {
        if (k == 0 && forced) {
            // drive-out: no rank has an eligible entry in this row -- no pivot, the host tries the next artificial row
        } else if (k == 0) {
            // nothing prices in anywhere: optimal -- unless Big-M or phase one left an artificial basic at a positive level
            // (infeasible), or phase one is over (drive-out next: forced steps from the host loop)
            int f = 0;
            if (bigm || phase0 == PH_ONE) {
                const int nreal = S.n_global - S.n_art;
                for (int i = (int)threadIdx.x; i < P.m; i += blockDim.x)
                    if (__ldcg(P.basis + i) > nreal && __ldcg(B.bcur + i) > P.eps) f = 1;
            }
            f = __syncthreads_or(f);                                               // every CTA scans all rows: same answer
            if (f) status = LPGPU_INFEASIBLE;
            else if (phase0 == PH_ONE) { phase = PH_DRIVEOUT; p1 = iters; }
            else status = LPGPU_OPTIMAL;
        } else {
            // ---- the owner: column k as it stands after the t pivots of this block -> every rank's region
            if (mine) {
                if ((int)threadIdx.x < t) s_pk[threadIdx.x] = __ldcg(B.Pm + (size_t)threadIdx.x * P.ld + (size_t)kl);
                __syncthreads();
                for (int i = gtid; i < rows; i += GSTRIDE) {
                    double a;
                    if (i >= P.m) a = __ldcg(oc + (size_t)(i - P.m) * P.ld + kl);   // objective rows are kept eagerly: same bits
                    else {
                        const int s0 = blk_row_start(__ldcg(B.rowstamp + i), seq);
                        a = s0 ? s_pk[s0 - 1] : P.T[(size_t)i * P.ld + (size_t)kl];
                        for (int sb = s0; sb < t; sb += 8) {
                            double fv[8];
#pragma unroll
                            for (int u = 0; u < 8; ++u) fv[u] = (sb + u < t) ? B.Fs[(size_t)(sb + u) * pcn + i] : 0.0;
#pragma unroll
                            for (int u = 0; u < 8; ++u)
                                if (sb + u < t) a = __fma_rn(-fv[u], s_pk[sb + u], a);
                        }
                    }
                    if (S.ll) {                              // every element is its own message: no fence, no flag
                        for (int g = 0; g < S.world; ++g) ll_put_f64(xptr<unsigned long long>(S.peer_x[g], XLayout::ll_cols(par, pcn)) + 2 * (size_t)i, a, S.iter + 1u);
                    } else {
                        for (int g = 0; g < S.world; ++g) xptr<double>(S.peer_x[g], XLayout::cols(par, pcn))[i] = a;
                    }
                }
                if (!S.ll) {
                    __threadfence_system();                 // this thread's stores into the peers, before the flag below
                    cl.sync();
                    if (gtid < S.world) st_release_sys(xptr<unsigned>(S.peer_x[gtid], XLayout::col_flag(par)), S.iter + 1u);
                }
            }
            // ---- every rank: the column of this pivot has landed in its region
            stamp(2);           // (owner: column computed, stored to every rank and flagged)
            if (!S.ll) {
                if (threadIdx.x == 0) s_flag = spin_until(xptr<unsigned>(S.peer_x[S.rank], XLayout::col_flag(par)), S.iter + 1u, XCHG_TIMEOUT_NS) ? 1 : 0;
                __syncthreads();
                if (!s_flag) {
                    cl.sync();
                    if (gtid == 0) { st->status = LPGPU_EXCHANGE_TIMEOUT; st->do_pivot = 0; }
                    return;
                }
            }
            stamp(3);           // the column has arrived
            const unsigned long long *llcol = xptr<unsigned long long>(S.peer_x[S.rank], XLayout::ll_cols(par, pcn));
            KeyCand c = keycand_none();
            for (int i = gtid; i < rows; i += GSTRIDE) {
                const double a = S.ll ? ll_wait_f64(llcol + 2 * (size_t)i, S.iter + 1u) : __ldcg(mycol + i);
                B.Fs[(size_t)t * pcn + i] = a;
                if (i < P.m && a > P.eps) {
                    KeyCand x;
                    x.kp = 0;
                    x.kq = ord_key(-(__ldcg(B.bcur + i) / a));
                    x.idx = i;
                    if (keycand_better(x, c)) c = x;
                }
            }
            c = cluster_key_best<false>(cl, c, scr_b, slot_b);
            if (forced) {                                   // drive-out: the row is given, no ratio test
                if (iters >= P.max_iter) status = LPGPU_ITER_LIMIT;
                else if (budget > 0) { have = true; r = S.forced_r; }
            } else if (c.idx < 0) status = LPGPU_UNBOUNDED;
            else if (iters >= P.max_iter) status = LPGPU_ITER_LIMIT;
            else if (budget > 0) { have = true; r = c.idx; }
        }
    }
    if (!have) {
        cl.sync();                                          // all state reads are done
        if (gtid == 0) {
            st->status = status;
            st->do_pivot = 0;
            st->blk_t = t;
            st->blk_seq = seq;
            st->phase = phase;
            st->phase1_iters = p1;
            if (S.post && !forced) post_cand(S, S.sendpair);    // the same candidate again: peers expect one every iteration
        }
        return;
    }

    stamp(4);                   // ratio test done
    // ---- this rank's slice of row r, brought up to date and scaled; objective rows; local pricing
    // (the column of this pivot, as every thread of the cluster stored it into F[t] above, before the ratio test's cluster barrier)
    const double sig_c = __ldcg(B.Fs + (size_t)t * pcn + P.m);
    const double sig_m = (RR == 2) ? __ldcg(B.Fs + (size_t)t * pcn + P.m + 1) : 0.0;
    const double piv = __ldcg(B.Fs + (size_t)t * pcn + r);
    double av[SSTEP_CPT], ocv[SSTEP_CPT], omv[SSTEP_CPT];
    const size_t jb = 1 + (size_t)gtid;
#pragma unroll
    for (int cth = 0; cth < SSTEP_CPT; ++cth) {
        const size_t j = jb + (size_t)cth * GSTRIDE;
        av[cth] = ocv[cth] = omv[cth] = 0.0;
        if (j <= (size_t)P.n) {
            av[cth] = P.T[(size_t)r * P.ld + j];
            ocv[cth] = oc[j];
            if (RR == 2) omv[cth] = om[j];
        }
    }
    const int s0r = blk_row_start(__ldcg(B.rowstamp + r), seq);
    const double b_r = __ldcg(B.bcur + r);
    if ((int)threadIdx.x < t) s_fr[threadIdx.x] = __ldcg(B.Fs + (size_t)threadIdx.x * pcn + r);
    __syncthreads();
    double *pt = B.Pm + (size_t)t * P.ld;
    const double brp = div_skip_zero(b_r, piv);
    KeyCand c = keycand_none();
    if (gtid == 0) {                                        // column 0 (the replica of b and the z cells)
        pt[0] = brp;
        B.objcur[0] = __fma_rn(-sig_c, brp, oc[0]);
        if (RR == 2) B.objcur[P.ld] = __fma_rn(-sig_m, brp, om[0]);
    }
    for (size_t j0 = jb; j0 <= (size_t)P.n; j0 += (size_t)SSTEP_CPT * GSTRIDE) {
        if (j0 != jb) {
#pragma unroll
            for (int cth = 0; cth < SSTEP_CPT; ++cth) {
                const size_t j = j0 + (size_t)cth * GSTRIDE;
                if (j <= (size_t)P.n) { av[cth] = P.T[(size_t)r * P.ld + j]; ocv[cth] = oc[j]; if (RR == 2) omv[cth] = om[j]; }
            }
        }
        if (s0r > 0) {
#pragma unroll
            for (int cth = 0; cth < SSTEP_CPT; ++cth) {
                const size_t j = j0 + (size_t)cth * GSTRIDE;
                if (j <= (size_t)P.n) av[cth] = B.Pm[(size_t)(s0r - 1) * P.ld + j];
            }
        }
        for (int sb = s0r; sb < t; sb += SSTEP_PB) {
            double pv[SSTEP_CPT][SSTEP_PB];
#pragma unroll
            for (int cth = 0; cth < SSTEP_CPT; ++cth) {
                const size_t j = j0 + (size_t)cth * GSTRIDE;
#pragma unroll
                for (int u = 0; u < SSTEP_PB; ++u) pv[cth][u] = (sb + u < t && j <= (size_t)P.n) ? B.Pm[(size_t)(sb + u) * P.ld + j] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < SSTEP_PB; ++u)
                if (sb + u < t) {
#pragma unroll
                    for (int cth = 0; cth < SSTEP_CPT; ++cth) av[cth] = __fma_rn(-s_fr[sb + u], pv[cth][u], av[cth]);
                }
        }
#pragma unroll
        for (int cth = 0; cth < SSTEP_CPT; ++cth) {
            const size_t j = j0 + (size_t)cth * GSTRIDE;
            if (j <= (size_t)P.n) {
                const double pr = div_skip_zero(av[cth], piv);
                pt[j] = pr;
                const double qc = __fma_rn(-sig_c, pr, ocv[cth]);
                B.objcur[j] = qc;
                double qm = 0.0;
                if (RR == 2) {
                    qm = __fma_rn(-sig_m, pr, omv[cth]);
                    B.objcur[P.ld + j] = qm;
                }
                if (j <= (size_t)jmax) blk_price<BL ? PR_BLAND : PR_DANTZIG>(P, pmode, (int)j, qc, qm, 1.0, c);
            }
        }
    }
    for (int i = gtid; i < P.m; i += GSTRIDE)
        B.bcur[i] = (i == r) ? brp : __fma_rn(-B.Fs[(size_t)t * pcn + i], brp, B.bcur[i]);
    stamp(5);                   // row, objective rows, b column done
    // ---- this rank's candidate for the next pivot -> every rank
    c = bigm ? cluster_key_best<true>(cl, c, scr_c, slot_c) : cluster_key_best<false>(cl, c, scr_c, slot_c);
    if (gtid < 32 && !forced) {                             // first warp of CTA 0: one lane per peer (a forced step posts nothing:
        Cand out = cand_none();                             // the host re-prices once the drive-out is over)
        if (c.idx > 0) {
            out.idx = c.idx;
            out.p = 0.0;
            if (BL) out.q = 1.0;                            // Bland: equal keys, the lowest global column wins the exchange
            else if (pmode == PRICE_PHASE1) out.q = __ldcg(B.objcur + P.ld + c.idx);
            else {
                out.q = __ldcg(B.objcur + c.idx);           // written above, possibly by another CTA
                if (bigm) {
                    double p = __ldcg(B.objcur + P.ld + c.idx);
                    if (fabs(p) <= P.eps) p = 0.0;
                    out.p = p;
                }
            }
        }
        double v[PAIR_DOUBLES];
        store_pair(v, out, S.col0);
        if (gtid == 0) store_pair(S.sendpair, out, S.col0); // kept for the re-post of an iteration without a pivot
        post_cand_lanes(S, v);
    }
    if (gtid == 0) {
        P.basis[r] = k;
        B.rowstamp[r] = (seq << 8) | (t + 1);
        B.steprow[t] = r;
        st->status = status;
        st->do_pivot = 1;
        st->r = r;
        st->k = k;
        st->iters = iters + 1;
        st->budget = budget - 1;
        st->blk_t = t + 1;
        st->blk_seq = seq;
        st->phase = phase;
        st->phase1_iters = (phase0 == PH_ONE) ? iters + 1 : p1;
        if (iters < P.log_cap) { P.log[2 * iters] = k; P.log[2 * iters + 1] = r; }
    }
    stamp(6);
}

template <int RR, bool BL>
__global__ void __cluster_dims__(SELECT_CLUSTER, 1, 1) __launch_bounds__(SELECT_THREADS)
shard_blk_step_kernel(ShardParams S, BlockBufs B, int first) {
    shard_blk_step_body<RR, BL>(S, B, first);
}

__global__ void shard_synth_kernel(double *T, size_t ld, int m, int ns, int col0, int ncols, int *basis,
                                   uint64_t seed, uint64_t lp) {
    const size_t total = (size_t)(m + 1) * ld;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / ld), jl = (int)(e % ld);
        double v = 0.0;
        if (jl == 0) v = synth_elem(seed, lp, m, ns, i, 0);
        else if (jl <= ncols) v = synth_elem(seed, lp, m, ns, i, col0 + jl - 1);
        T[e] = v;
        if (jl == 0 && i < m) basis[i] = 1 + ns + i;
    }
}

__global__ void shard_extract_x_kernel(const double *T, size_t ld, const int *basis, int m, int n, double *x) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const int j = basis[i];
        if (j >= 1 && j <= n) x[j - 1] = T[(size_t)i * ld];
    }
}

}  // namespace lpgpu

// ------------------------------------------------------------------ host side
namespace {

struct ShardRank {
    lpgpu_dense *d;
    int rank, col0, ncols;
    ShardState *sh;
    double *sendpair, *pairs, *sendcol, *gathered;
};

}  // namespace

struct lpgpu_shard {
    lpgpu_ctx *ctx;
    int world, m, n, R;
    bool use_nccl;
    bool use_p2p;                     // peer-memory transport (shard_p2p.cuh)
    bool attached;                    // ... and the peers' regions are mapped
    ncclComm_t comm;
    void *xregion;                    // this rank's exchange region
    void **peer_host;                 // world mapped pointers (own = xregion), host copy
    void **peer_dev;                  // the same table in device memory
    unsigned iter;                    // pipeline iterations so far (identical on every rank)
    std::vector<ShardRank> rk;        // the ranks that live in this process (1 with NCCL, `world` without)
    double *pairs_buf, *gath_buf;     // exchange buffers (shared by all local ranks without NCCL)
    double *xbuf;
    size_t pcn;
    bool dirty;                       // tableau changed since the last lookahead (upload / synth / reset)
    bool blk_open;                    // blocked path: steps enqueued since the last pass over the tableau
    int blk_steps;
    bool drv_active;                  // two-phase: the drive-out's forced steps are under way (exchange numbering, see shard_run_blocked)
    lpgpu_opts opts;
};

namespace {

#define LP_NCCL(ctx, api, call)                                                                    \
    do {                                                                                           \
        ncclResult_t r__ = (call);                                                                 \
        if (r__ != ncclSuccess)                                                                    \
            return fail((ctx), LPGPU_ENCCL, "%s:%d: %s: %s", __FILE__, __LINE__, #call, (api)->GetErrorString(r__)); \
    } while (0)

void split_columns(int n, int world, int rank, int *col0, int *ncols) {
    const int base = n / world, extra = n % world;
    *ncols = base + (rank < extra ? 1 : 0);
    *col0 = 1 + rank * base + std::min(rank, extra);
}

ShardParams shard_params(const lpgpu_shard *s, const ShardRank &r) {
    ShardParams S;
    S.P = make_params(r.d);
    S.rank = r.rank; S.world = s->world; S.col0 = r.col0; S.n_global = s->n; S.n_art = s->opts.n_art;
    S.sh = r.sh; S.sendpair = r.sendpair; S.pairs = r.pairs; S.sendcol = r.sendcol; S.gathered = r.gathered;
    S.peer_x = s->use_p2p ? s->peer_dev : nullptr; S.iter = s->iter; S.post = 1; S.forced_r = -1;
    // blocked path over peer memory: flag-in-payload exchange (LPGPU_SHARD_FLAGS=1: data + release flag, as in round 1)
    S.ll = (s->use_p2p && !(s->opts.variant & LPGPU_VARIANT_RANK1) && !getenv("LPGPU_SHARD_FLAGS")) ? 1 : 0;
    return S;
}

// exchange A (which = 0): candidates; exchange B (which = 1): entering column.  Without NCCL the
// local ranks write straight into the shared gather buffers, so there is nothing to move.
int exchange(lpgpu_shard *s, int which) {
    if (s->use_p2p) {
        // candidates travel inside the select / price kernels; the column inside the elimination pass.
        // What is left is the "column complete" flag, after the pass in stream order.
        if (which == 1) {
            shard_publish_kernel<<<1, XMAXW, 0, s->ctx->stream>>>(shard_params(s, s->rk[0]));
            LP_LAUNCH_CHECK(s->ctx);
        }
        return LPGPU_OK;
    }
    if (!s->use_nccl) return LPGPU_OK;
    const NcclApi *api = nccl_api();
    const ShardRank &r = s->rk[0];
    if (which == 0)
        LP_NCCL(s->ctx, api, api->AllGather(r.sendpair, s->pairs_buf, PAIR_DOUBLES, ncclDouble, s->comm, s->ctx->stream));
    else
        LP_NCCL(s->ctx, api, api->AllGather(r.sendcol, s->gath_buf, s->pcn, ncclDouble, s->comm, s->ctx->stream));
    return LPGPU_OK;
}

int push_state(lpgpu_shard *s, const PivotState &st) {
    lpgpu_ctx *ctx = s->ctx;
    for (ShardRank &r : s->rk) {
        *r.d->h_st = st;
        LP_CUDA(ctx, cudaMemcpyAsync(r.d->st, r.d->h_st, sizeof(PivotState), cudaMemcpyHostToDevice, ctx->stream));
    }
    return LPGPU_OK;
}

int enqueue_select(lpgpu_shard *s, int post = 1) {
    for (ShardRank &r : s->rk) {
        ShardParams S = shard_params(s, r);
        S.post = post;
        shard_select_kernel<<<SELECT_CLUSTER, SELECT_THREADS, 0, s->ctx->stream>>>(S);
        LP_LAUNCH_CHECK(s->ctx);
    }
    return LPGPU_OK;
}

int enqueue_resolve(lpgpu_shard *s, int gather) {
    for (ShardRank &r : s->rk) {
        const int grid = gather ? std::max(1, std::min(64, (s->m + s->R + SELECT_THREADS - 1) / SELECT_THREADS)) : 1;
        shard_resolve_kernel<<<grid, gather ? SELECT_THREADS : XMAXW, 0, s->ctx->stream>>>(shard_params(s, r), gather);
        LP_LAUNCH_CHECK(s->ctx);
    }
    return LPGPU_OK;
}

int enqueue_update(lpgpu_shard *s) {
    int rc = LPGPU_OK;
    for (ShardRank &r : s->rk) {
        if (s->use_p2p) rc = launch_update_peers(r.d, s->peer_dev, s->world, XLayout::cols((int)(s->iter & 1u), s->pcn));
        else rc = launch_update(r.d, r.sendcol);
        if (rc != LPGPU_OK) return rc;
    }
    return rc;
}

int shard_create_common(lpgpu_ctx *ctx, int rank, int world, int transport, const void *id128, int m, int n, int R, lpgpu_shard **out);

}  // namespace

extern "C" {

int lpgpu_shard_unique_id(void *id128) {
    if (!id128) return LPGPU_EINVAL;
    const NcclApi *api = nccl_api();
    if (!api) return LPGPU_ENCCL;
    static_assert(sizeof(ncclUniqueId) == LPGPU_NCCL_ID_BYTES, "NCCL id size");
    ncclUniqueId id;
    if (api->GetUniqueId(&id) != ncclSuccess) return LPGPU_ENCCL;
    memcpy(id128, &id, sizeof id);
    return LPGPU_OK;
}

int lpgpu_shard_columns(int n, int world, int rank, int *col0, int *ncols) {
    if (n < 1 || world < 1 || world > n || rank < 0 || rank >= world || !col0 || !ncols) return LPGPU_EINVAL;
    split_columns(n, world, rank, col0, ncols);
    return LPGPU_OK;
}

int lpgpu_shard_create(lpgpu_ctx *ctx, int rank, int world, const void *id128, int m, int n, int R, lpgpu_shard **out) {
    return shard_create_common(ctx, rank, world, id128 ? 1 : 0, id128, m, n, R, out);
}

int lpgpu_shard_create_p2p(lpgpu_ctx *ctx, int rank, int world, int m, int n, int R, lpgpu_shard **out) {
    return shard_create_common(ctx, rank, world, 2, nullptr, m, n, R, out);
}

int lpgpu_shard_p2p_export(lpgpu_shard *s, void *handle64) {
    if (!s || !handle64) return LPGPU_EINVAL;
    if (!s->use_p2p) return fail(s->ctx, LPGPU_EINVAL, "shard_p2p_export: not a peer-memory shard");
    static_assert(sizeof(cudaIpcMemHandle_t) == LPGPU_IPC_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    LP_CUDA(s->ctx, cudaSetDevice(s->ctx->device));
    LP_CUDA(s->ctx, cudaIpcGetMemHandle(&h, s->xregion));
    memcpy(handle64, &h, sizeof h);
    return LPGPU_OK;
}

int lpgpu_shard_p2p_attach(lpgpu_shard *s, const void *handles) {
    if (!s || !handles) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = s->ctx;
    if (!s->use_p2p || s->attached) return fail(ctx, LPGPU_EINVAL, "shard_p2p_attach: not a peer-memory shard, or attached already");
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    const int me = s->rk[0].rank;
    for (int g = 0; g < s->world; ++g) {
        if (g == me) { s->peer_host[g] = s->xregion; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const unsigned char *>(handles) + (size_t)g * LPGPU_IPC_HANDLE_BYTES, sizeof h);
        LP_CUDA(ctx, cudaIpcOpenMemHandle(&s->peer_host[g], h, cudaIpcMemLazyEnablePeerAccess));
    }
    LP_CUDA(ctx, cudaMemcpyAsync(s->peer_dev, s->peer_host, (size_t)s->world * sizeof(void *), cudaMemcpyHostToDevice, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    s->attached = true;
    return LPGPU_OK;
}

}  // extern "C"

namespace {

int shard_create_common(lpgpu_ctx *ctx, int rank, int world, int transport, const void *id128, int m, int n, int R, lpgpu_shard **out) {
    if (!ctx || !out) return LPGPU_EINVAL;
    *out = nullptr;
    if (m < 1 || n < 1 || (R != 1 && R != 2) || world < 1 || world > n || world > 64)
        return fail(ctx, LPGPU_EINVAL, "shard_create: bad shape m=%d n=%d R=%d world=%d", m, n, R, world);
    if (transport != 0 && (rank < 0 || rank >= world)) return fail(ctx, LPGPU_EINVAL, "shard_create: rank %d of %d", rank, world);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    lpgpu_shard *s = new (std::nothrow) lpgpu_shard();
    if (!s) return fail(ctx, LPGPU_ENOMEM, "shard_create: host allocation failed");
    s->ctx = ctx; s->world = world; s->m = m; s->n = n; s->R = R;
    s->use_nccl = transport == 1; s->use_p2p = transport == 2; s->attached = false; s->comm = nullptr;
    s->xregion = nullptr; s->peer_host = nullptr; s->peer_dev = nullptr; s->iter = 0;
    s->pairs_buf = s->gath_buf = s->xbuf = nullptr; s->dirty = true; s->blk_open = false; s->blk_steps = 0; s->drv_active = false;
    lpgpu_default_opts(&s->opts);
    s->opts.method = (R == 1) ? LPGPU_PLAIN : LPGPU_BIGM;
    int rc = LPGPU_OK;
    const bool one_rank = transport != 0;
    const int first = one_rank ? rank : 0, last = one_rank ? rank + 1 : world;
    for (int g = first; g < last && rc == LPGPU_OK; ++g) {
        ShardRank r{};
        r.rank = g;
        split_columns(n, world, g, &r.col0, &r.ncols);
        rc = lpgpu_dense_create(ctx, m, r.ncols, R, &r.d);
        if (rc == LPGPU_OK) s->rk.push_back(r);
    }
    if (rc != LPGPU_OK) { lpgpu_shard_destroy(s); return rc; }
    s->pcn = s->rk[0].d->pcn;                   // depends on m + R only: the same on every rank
    cudaError_t e = cudaSuccess;
    // without NCCL every local rank sends straight into its slot of the shared gather buffers
    const size_t npairs = (size_t)world * PAIR_DOUBLES, ncol = (size_t)world * s->pcn;
    if (e == cudaSuccess) e = cudaMalloc(&s->pairs_buf, npairs * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&s->gath_buf, ncol * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&s->xbuf, (size_t)n * sizeof(double));
    if (e == cudaSuccess) e = cudaMemsetAsync(s->pairs_buf, 0, npairs * sizeof(double), ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(s->gath_buf, 0, ncol * sizeof(double), ctx->stream);
    for (ShardRank &r : s->rk) {
        if (e == cudaSuccess) e = cudaMalloc(&r.sh, sizeof(ShardState));
        if (e == cudaSuccess) e = cudaMemsetAsync(r.sh, 0, sizeof(ShardState), ctx->stream);
        r.pairs = s->pairs_buf;
        r.gathered = s->gath_buf;
        if (one_rank) {
            if (e == cudaSuccess) e = cudaMalloc(&r.sendpair, PAIR_DOUBLES * sizeof(double));
            if (e == cudaSuccess) e = cudaMalloc(&r.sendcol, s->pcn * sizeof(double));
            if (e == cudaSuccess) e = cudaMemsetAsync(r.sendcol, 0, s->pcn * sizeof(double), ctx->stream);
        } else {
            r.sendpair = s->pairs_buf + (size_t)r.rank * PAIR_DOUBLES;
            r.sendcol = s->gath_buf + (size_t)r.rank * s->pcn;
        }
    }
    if (e != cudaSuccess) {
        lpgpu_shard_destroy(s);
        return fail(ctx, e == cudaErrorMemoryAllocation ? LPGPU_ENOMEM : LPGPU_ECUDA, "shard_create: %s", cudaGetErrorString(e));
    }
    if (s->use_p2p) {
        const size_t xb = XLayout::bytes(s->pcn);
        s->peer_host = new (std::nothrow) void *[world]();
        if (e == cudaSuccess && !s->peer_host) e = cudaErrorMemoryAllocation;
        if (e == cudaSuccess) e = cudaMalloc(&s->xregion, xb);
        if (e == cudaSuccess) e = cudaMemsetAsync(s->xregion, 0, xb, ctx->stream);
        if (e == cudaSuccess) e = cudaMalloc(&s->peer_dev, (size_t)world * sizeof(void *));
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) {
            lpgpu_shard_destroy(s);
            return fail(ctx, e == cudaErrorMemoryAllocation ? LPGPU_ENOMEM : LPGPU_ECUDA, "shard_create_p2p: %s", cudaGetErrorString(e));
        }
        if (world == 1) {                 // nothing to map
            s->peer_host[0] = s->xregion;
            cudaMemcpy(s->peer_dev, s->peer_host, sizeof(void *), cudaMemcpyHostToDevice);
            s->attached = true;
        }
    }
    if (s->use_nccl) {
        const NcclApi *api = nccl_api();
        if (!api) { lpgpu_shard_destroy(s); return fail(ctx, LPGPU_ENCCL, "shard_create: libnccl.so.2 could not be loaded"); }
        ncclUniqueId id;
        memcpy(&id, id128, sizeof id);
        ncclResult_t nr = api->CommInitRank(&s->comm, world, id, rank);
        if (nr != ncclSuccess) {
            s->comm = nullptr;
            lpgpu_shard_destroy(s);
            return fail(ctx, LPGPU_ENCCL, "shard_create: ncclCommInitRank: %s", api->GetErrorString(nr));
        }
    }
    rc = lpgpu_shard_reset(s, &s->opts);
    if (rc != LPGPU_OK) { lpgpu_shard_destroy(s); return rc; }
    *out = s;
    return LPGPU_OK;
}

}  // namespace

extern "C" {

void lpgpu_shard_destroy(lpgpu_shard *s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    if (s->comm) { const NcclApi *api = nccl_api(); if (api) api->CommDestroy(s->comm); }
    for (ShardRank &r : s->rk) {
        lpgpu_dense_destroy(r.d);
        cudaFree(r.sh);
        if (s->use_nccl || s->use_p2p) { cudaFree(r.sendpair); cudaFree(r.sendcol); }
    }
    if (s->peer_host) {
        for (int g = 0; g < s->world; ++g)
            if (s->peer_host[g] && s->peer_host[g] != s->xregion) cudaIpcCloseMemHandle(s->peer_host[g]);
        delete[] s->peer_host;
    }
    cudaFree(s->xregion); cudaFree(s->peer_dev);
    cudaFree(s->pairs_buf); cudaFree(s->gath_buf); cudaFree(s->xbuf);
    delete s;
}

int lpgpu_shard_reset(lpgpu_shard *s, const lpgpu_opts *opts) {
    if (!s || !opts) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = s->ctx;
    if (opts->method == LPGPU_DUAL) return fail(ctx, LPGPU_EINVAL, "shard_reset: the column-sharded path runs the primal methods");
    if (opts->variant & LPGPU_VARIANT_DEVEX) return fail(ctx, LPGPU_EINVAL, "shard_reset: Devex pricing is not available on the column-sharded path (default rule or LPGPU_VARIANT_BLAND)");
    if (opts->method == LPGPU_TWOPHASE && !(s->use_p2p && !(opts->variant & LPGPU_VARIANT_RANK1)))
        return fail(ctx, LPGPU_EINVAL, "shard_reset: the two-phase method runs on the blocked path over the peer-memory transport (lpgpu_shard_create_p2p, no LPGPU_VARIANT_RANK1)");
    if ((opts->method == LPGPU_PLAIN) != (s->R == 1) || opts->n_art < 0 || opts->n_art > s->n)
        return fail(ctx, LPGPU_EINVAL, "shard_reset: bad options (method/R/n_art)");
    s->opts = *opts;
    if (s->opts.check_every <= 0) s->opts.check_every = 64;
    lpgpu_opts local = s->opts;
    local.n_art = 0;                     // artificial columns are tested against GLOBAL indices in shard kernels
    local.time_kernels = 0;
    for (ShardRank &r : s->rk) {
        int rc = lpgpu_dense_reset(r.d, &local);
        if (rc != LPGPU_OK) return rc;
    }
    s->dirty = true;
    s->blk_open = false;
    s->blk_steps = 0;
    s->drv_active = false;
    return LPGPU_OK;
}

int lpgpu_shard_upload(lpgpu_shard *s, const double *T, size_t ld, const int *basis) {
    if (!s || !T || !basis) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = s->ctx;
    if (ld < (size_t)s->n + 1) return fail(ctx, LPGPU_EINVAL, "shard_upload: ld %zu < n+1", ld);
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t rows = (size_t)s->m + s->R;
    for (ShardRank &r : s->rk) {
        lpgpu_dense *d = r.d;
        LP_CUDA(ctx, cudaMemcpy2DAsync(d->T, d->ld * sizeof(double), T, ld * sizeof(double), sizeof(double), rows,
                                       cudaMemcpyHostToDevice, ctx->stream));
        LP_CUDA(ctx, cudaMemcpy2DAsync(d->T + 1, d->ld * sizeof(double), T + r.col0, ld * sizeof(double),
                                       (size_t)r.ncols * sizeof(double), rows, cudaMemcpyHostToDevice, ctx->stream));
        LP_CUDA(ctx, cudaMemcpyAsync(d->basis, basis, (size_t)s->m * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    }
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    s->dirty = true;
    return LPGPU_OK;
}

int lpgpu_shard_synth(lpgpu_shard *s, uint64_t seed, uint64_t lp_index, int ns) {
    if (!s) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = s->ctx;
    if (s->R != 1 || ns < 1 || ns + s->m != s->n) return fail(ctx, LPGPU_EINVAL, "shard_synth: needs R=1 and n = ns + m");
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    for (ShardRank &r : s->rk) {
        shard_synth_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(r.d->T, r.d->ld, s->m, ns, r.col0, r.ncols, r.d->basis, seed, lp_index);
        LP_LAUNCH_CHECK(ctx);
    }
    s->dirty = true;
    return LPGPU_OK;
}

static int shard_finish(lpgpu_shard *s, lpgpu_result *res, int passes) {
    lpgpu_ctx *ctx = s->ctx;
    lpgpu_dense *d0 = s->rk[0].d;
    if (d0->h_st->status == LPGPU_EXCHANGE_TIMEOUT)
        return fail(ctx, LPGPU_ENCCL, "shard_run: a peer did not post its part of the exchange within %llu s", XCHG_TIMEOUT_NS / 1000000000ull);
    double zc[2] = {0.0, 0.0};
    LP_CUDA(ctx, cudaMemcpyAsync(&zc[0], d0->T + (size_t)s->m * d0->ld, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (s->R == 2)
        LP_CUDA(ctx, cudaMemcpyAsync(&zc[1], d0->T + (size_t)(s->m + 1) * d0->ld, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memset(res, 0, sizeof *res);
    res->status = d0->h_st->status;
    res->iters = d0->h_st->iters;
    res->z = -zc[0];
    res->zM = (s->R == 2) ? -zc[1] : 0.0;
    res->n_logged = std::min(res->iters, d0->log_cap);
    res->update_launches = passes;
    res->phase = d0->h_st->phase;
    res->phase1_iters = (s->opts.method == LPGPU_TWOPHASE) ? d0->h_st->phase1_iters : 0;
    return LPGPU_OK;
}

// Blocked path over the peer-memory transport: one cluster kernel per pivot (exchanges inside), one local
// pass over the rank's column block per K pivots.  The host loop is a replica on every rank.
static int shard_run_blocked(lpgpu_shard *s, int max_pivots, lpgpu_result *res) {
    lpgpu_ctx *ctx = s->ctx;
    ShardRank &r = s->rk[0];
    lpgpu_dense *d = r.d;
    int rc = read_state(d);
    if (rc != LPGPU_OK) return rc;
    PivotState st = *d->h_st;
    st.budget = max_pivots;
    st.cur = 0;
    if ((rc = push_state(s, st)) != LPGPU_OK) return rc;
    const int K = make_bufs(d).K;
    const bool bland = (s->opts.variant & LPGPU_VARIANT_BLAND) != 0;
    auto step = [&](int post, int forced_r = -1) -> int {
        ShardParams S = shard_params(s, r);
        S.post = post;
        S.forced_r = forced_r;
        const int first = s->blk_open ? 0 : 1;
        auto kern = s->R == 2 ? (bland ? shard_blk_step_kernel<2, true> : shard_blk_step_kernel<2, false>)
                              : (bland ? shard_blk_step_kernel<1, true> : shard_blk_step_kernel<1, false>);
        kern<<<SELECT_CLUSTER, SELECT_THREADS, 0, ctx->stream>>>(S, make_bufs(d), first);
        LP_LAUNCH_CHECK(ctx);
        if (d->blk_dbg) d->blk_dbg_n++;
        return LPGPU_OK;
    };
    int passes = 0;
    auto flush = [&]() -> int {
        const int rc2 = launch_block_flush(d);
        s->blk_open = false;
        s->blk_steps = 0;
        ++passes;
        return rc2;
    };
    auto bootstrap = [&]() -> int {
        // objective rows / b column out of the tableau, local pricing, candidate posted as iteration iter
        int rc2 = launch_block_init(d);
        if (rc2 != LPGPU_OK) return rc2;
        shard_price_kernel<<<1, SELECT_THREADS, 0, ctx->stream>>>(shard_params(s, r));
        LP_LAUNCH_CHECK(ctx);
        s->iter++;
        s->dirty = false;
        s->blk_open = false;
        s->blk_steps = 0;
        return LPGPU_OK;
    };
    int launched = 0;
    // Two-phase, between the phases: artificial variables still basic (at level zero) are driven out by FORCED steps -- the
    // lowest such row that has an eligible entry, on its lowest non-artificial column -- one-step blocks on an up-to-date
    // tableau, the offer / column exchange inside the step kernel.  The host loop is a replica on every rank (the basis is).
    // Exchange numbering: a forced step posts its offers under the number BEFORE its own and nothing under its own, so one
    // number is skipped on the way in and one on the way out; successive exchanges then keep alternating their slot parity.
    auto driveout = [&]() -> int {
        int rc2 = LPGPU_OK;
        if (s->blk_open && (rc2 = flush()) != LPGPU_OK) return rc2;
        if (!s->drv_active) { s->iter++; s->drv_active = true; }
        std::vector<int> hb((size_t)s->m);
        const int nreal = s->n - s->opts.n_art;
        bool pivoted = true;
        while (pivoted && launched < max_pivots && d->h_st->status == LPGPU_RUNNING) {
            pivoted = false;
            LP_CUDA(ctx, cudaMemcpyAsync(hb.data(), d->basis, hb.size() * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            for (int i = 0; i < s->m && !pivoted && d->h_st->status == LPGPU_RUNNING; ++i) {
                if (hb[(size_t)i] <= nreal) continue;
                const int before = d->h_st->iters;
                if ((rc2 = step(0, i)) != LPGPU_OK) return rc2;
                s->iter++;
                s->blk_open = true;
                if ((rc2 = flush()) != LPGPU_OK) return rc2;         // a one-step block (or nothing, if the row had no eligible entry)
                if ((rc2 = read_state(d)) != LPGPU_OK) return rc2;
                if (d->h_st->iters > before) { pivoted = true; ++launched; }
            }
        }
        if (!pivoted && d->h_st->status == LPGPU_RUNNING) {
            // no artificial row is left that can be pivoted on: the main phase starts (state is a replica: same write on every rank)
            PivotState ns = *d->h_st;
            ns.phase = PH_MAIN;
            ns.budget = std::max(0, max_pivots - launched);
            if ((rc2 = push_state(s, ns)) != LPGPU_OK) return rc2;
            s->iter++;
            s->drv_active = false;
            s->dirty = true;
        }
        return LPGPU_OK;
    };
    if (d->h_st->status == LPGPU_RUNNING && d->h_st->phase == PH_DRIVEOUT && (rc = driveout()) != LPGPU_OK) return rc;
    const int chunk_max = s->opts.check_every;
    while (launched < max_pivots && d->h_st->status == LPGPU_RUNNING && d->h_st->phase != PH_DRIVEOUT) {
        if (s->dirty && (rc = bootstrap()) != LPGPU_OK) return rc;
        const int chunk = std::min(chunk_max, max_pivots - launched);
        const int before = d->h_st->iters;
        for (int c = 0; c < chunk; ++c) {
            if ((rc = step(1)) != LPGPU_OK) return rc;
            s->iter++;
            s->blk_open = true;
            if (++s->blk_steps >= K && (rc = flush()) != LPGPU_OK) return rc;
        }
        if ((rc = read_state(d)) != LPGPU_OK) return rc;
        // steps after a phase switch (or a terminal status) were no-ops: only the applied pivots count
        launched += (d->h_st->status == LPGPU_RUNNING && d->h_st->phase == PH_DRIVEOUT) ? d->h_st->iters - before : chunk;
        if (d->h_st->status != LPGPU_RUNNING) break;
        if (d->h_st->phase == PH_DRIVEOUT && (rc = driveout()) != LPGPU_OK) return rc;
    }
    if (s->blk_open && (rc = flush()) != LPGPU_OK) return rc;
    if (d->h_st->status == LPGPU_RUNNING && d->h_st->phase != PH_DRIVEOUT) {
        // budget spent: one decision-only step (same iteration, nothing posted) settles a terminal status
        if (s->dirty && (rc = bootstrap()) != LPGPU_OK) return rc;
        if ((rc = step(0)) != LPGPU_OK) return rc;
        if ((rc = read_state(d)) != LPGPU_OK) return rc;
        // the decision may have ended phase one: nothing is pending then, the next run call starts with the drive-out
    }
    if ((rc = read_state(d)) != LPGPU_OK) return rc;
    return shard_finish(s, res, passes);
}

int lpgpu_shard_run(lpgpu_shard *s, int max_pivots, lpgpu_result *res) {
    if (!s || !res || max_pivots < 0) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = s->ctx;
    if (s->use_p2p && !s->attached) return fail(ctx, LPGPU_EINVAL, "shard_run: peer regions are not attached (lpgpu_shard_p2p_attach)");
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    lpgpu_dense *d0 = s->rk[0].d;
    int rc = read_state(d0);                    // the state is a replica on every rank
    if (rc != LPGPU_OK) return rc;
    PivotState st = *d0->h_st;
    st.budget = max_pivots;
    st.cur = 0;
    rc = push_state(s, st);
    if (rc != LPGPU_OK) return rc;
    // peer-memory transport, default: the blocked path (variant bit 0x800 selects the rank-1 pipeline below)
    if (s->use_p2p && !(s->opts.variant & LPGPU_VARIANT_RANK1)) return shard_run_blocked(s, max_pivots, res);
    if (s->dirty) {
        // bootstrap: price the tableau as it stands, fetch the entering column from its owner (strided)
        for (ShardRank &r : s->rk) {
            shard_price_kernel<<<1, SELECT_THREADS, 0, ctx->stream>>>(shard_params(s, r));
            LP_LAUNCH_CHECK(ctx);
        }
        if ((rc = exchange(s, 0)) != LPGPU_OK) return rc;
        if ((rc = enqueue_resolve(s, 1)) != LPGPU_OK) return rc;
        if ((rc = exchange(s, 1)) != LPGPU_OK) return rc;
        s->iter++;
        s->dirty = false;
    }
    const int chunk_max = s->opts.check_every;
    int launched = 0, done_updates = 0, iters_before = st.iters;
    while (launched < max_pivots) {
        const int chunk = std::min(chunk_max, max_pivots - launched);
        for (int c = 0; c < chunk; ++c) {
            if ((rc = enqueue_select(s)) != LPGPU_OK) return rc;
            if ((rc = exchange(s, 0)) != LPGPU_OK) return rc;
            if ((rc = enqueue_resolve(s, 0)) != LPGPU_OK) return rc;
            if ((rc = enqueue_update(s)) != LPGPU_OK) return rc;
            if ((rc = exchange(s, 1)) != LPGPU_OK) return rc;
            s->iter++;
        }
        launched += chunk;
        if ((rc = read_state(d0)) != LPGPU_OK) return rc;
        done_updates += d0->h_st->iters - iters_before;
        iters_before = d0->h_st->iters;
        if (d0->h_st->status != LPGPU_RUNNING) break;
    }
    if (d0->h_st->status == LPGPU_RUNNING) {
        // budget spent: one decision-only select settles a terminal status without applying a pivot
        if ((rc = enqueue_select(s, 0)) != LPGPU_OK) return rc;
        if ((rc = read_state(d0)) != LPGPU_OK) return rc;
    }
    if (d0->h_st->status == LPGPU_EXCHANGE_TIMEOUT)
        return fail(ctx, LPGPU_ENCCL, "shard_run: a peer did not post its part of the exchange within %llu s", XCHG_TIMEOUT_NS / 1000000000ull);
    double zc[2] = {0.0, 0.0};
    LP_CUDA(ctx, cudaMemcpyAsync(&zc[0], d0->T + (size_t)s->m * d0->ld, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (s->R == 2)
        LP_CUDA(ctx, cudaMemcpyAsync(&zc[1], d0->T + (size_t)(s->m + 1) * d0->ld, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memset(res, 0, sizeof *res);
    res->status = d0->h_st->status;
    res->iters = d0->h_st->iters;
    res->z = -zc[0];
    res->zM = (s->R == 2) ? -zc[1] : 0.0;
    res->n_logged = std::min(res->iters, d0->log_cap);
    res->update_launches = done_updates;
    return LPGPU_OK;
}

int lpgpu_shard_download(lpgpu_shard *s, double *T, size_t ld, int *basis, int *pivot_log, double *x) {
    if (!s) return LPGPU_EINVAL;
    lpgpu_ctx *ctx = s->ctx;
    LP_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t rows = (size_t)s->m + s->R;
    lpgpu_dense *d0 = s->rk[0].d;
    if (T) {
        if (ld < (size_t)s->n + 1) return fail(ctx, LPGPU_EINVAL, "shard_download: ld %zu < n+1", ld);
        LP_CUDA(ctx, cudaMemcpy2DAsync(T, ld * sizeof(double), d0->T, d0->ld * sizeof(double), sizeof(double), rows,
                                       cudaMemcpyDeviceToHost, ctx->stream));
        for (ShardRank &r : s->rk)
            LP_CUDA(ctx, cudaMemcpy2DAsync(T + r.col0, ld * sizeof(double), r.d->T + 1, r.d->ld * sizeof(double),
                                           (size_t)r.ncols * sizeof(double), rows, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (basis) LP_CUDA(ctx, cudaMemcpyAsync(basis, d0->basis, (size_t)s->m * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (pivot_log && d0->log_cap > 0) {
        int rc = read_state(d0);
        if (rc != LPGPU_OK) return rc;
        const int nl = std::min(d0->h_st->iters, d0->log_cap);
        if (nl > 0) LP_CUDA(ctx, cudaMemcpyAsync(pivot_log, d0->log, (size_t)nl * 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (x) {
        LP_CUDA(ctx, cudaMemsetAsync(s->xbuf, 0, (size_t)s->n * sizeof(double), ctx->stream));
        shard_extract_x_kernel<<<std::max(1, std::min(1024, (s->m + 255) / 256)), 256, 0, ctx->stream>>>(d0->T, d0->ld, d0->basis, s->m, s->n, s->xbuf);
        LP_LAUNCH_CHECK(ctx);
        LP_CUDA(ctx, cudaMemcpyAsync(x, s->xbuf, (size_t)s->n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    }
    LP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LPGPU_OK;
}

}  // extern "C"
