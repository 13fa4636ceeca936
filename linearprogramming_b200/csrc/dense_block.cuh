// dense_block.cuh -- delayed ("blocked") application of the rank-1 eliminations of ONE large tableau.
//
// The rank-1 path (dense_update.cuh) reads and writes the whole tableau once per pivot: 16 B per
// element per pivot, which is what bounds it (it runs at the measured HBM copy peak).  Here up to
// BLK_KMAX pivots are DECIDED on lazily updated slices and APPLIED to the tableau in one pass:
//
//   per pivot t of a block (blk_step_kernel, one 8-CTA cluster):
//     entering column k_t   from the eagerly maintained objective row(s)  objcur
//     column k_t, current   = T[:,k_t] run through the pivots s < t of this block      -> F[:,t]
//     ratio test            on the eagerly maintained b column  bcur                   -> r_t
//     row r_t, current      = T[r_t,:] run through the pivots s < t, divided by a_rk   -> P[t,:]
//     objcur, bcur          updated with (F[:,t], P[t,:]);  next entering column priced
//   per block (the pass: blk_flush_ring_kernel, persistent grid):
//     a_ij <- fma(-F[i,K-1], P[K-1,j], ... fma(-F[i,0], P[0,j], a_ij))   every element, once
//     (a row that was the pivot row of step s restarts from P[s,j] and runs the steps after s)
//
// Every element goes through exactly the operations of the rank-1 path, in the same order, on the
// same operands (F[i,s] IS the element (i,k_s) at step s, produced by the same chain), so tableau,
// basis and pivot log are bit-identical to it -- only the HBM traffic per pivot drops from
// 16 B/element to 16/K B/element.  At K = 32 the pass is about as long on the FP64 pipe as on HBM.
// Reference: the loop missing from newSimplex (Source/simplex.c:40-67); layout Source/matrix.c:42-72.
#pragma once

#include "block_common.cuh"

namespace lpgpu {

// objcur / bcur <- tableau (after an upload, a synth, a forced pivot or a run of the rank-1 path)
__global__ void blk_init_kernel(DenseParams P, BlockBufs B) {
    const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gs = (size_t)gridDim.x * blockDim.x;
    for (size_t e = gt; e < (size_t)P.R * P.ld; e += gs) B.objcur[e] = P.T[(size_t)P.m * P.ld + e];
    for (size_t i = gt; i < (size_t)P.m; i += gs) B.bcur[i] = P.T[i * P.ld];
    for (size_t i = gt; i < (size_t)(P.m + P.R); i += gs) B.rowstamp[i] = 0;
}

// Which SMs a step-kernel cluster lands on (same cluster shape and CTA size, so one CTA per SM): run once on an idle
// device; the pass of the pipelined path leaves exactly these SMs free.
__global__ void __launch_bounds__(SELECT_THREADS) blk_probe_kernel(unsigned *smids) {      // cluster shape and CTA size: launch attributes
    cg::cluster_group cl = cg::this_cluster();
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (threadIdx.x == 0) smids[cl.block_rank()] = smid;
    cl.sync();
}

// One pivot of a block: decision + lazy slices + eager objective rows / b column (PLAIN and BIGM).
// first != 0: the tableau was flushed (or initialised) since the last step: a new block starts.
// The kernel is a chain of dependent memory round trips (~1 us each), so every phase first issues ALL
// the loads it can (tableau column / row, stamps, b, objective row) and only then waits; the second
// cluster-wide reduction (next pivot's pricing) is left to the next launch: each CTA leaves its best
// candidate in `cands` (st->k_next = -2) and the next step reduces the 8 of them on entry.
constexpr int STEP_PB = 8;             // earlier pivot rows per batch of loads in the row phase (64 registers per thread)
constexpr int STEP_CPT = 2;            // columns per thread in flight in the row phase (covers n <= 2 * 8192 in one go)

// CL = CTAs in the cluster (launch attribute): 8 x 1024 threads, or 16 x 512 threads (a non-portable cluster size).  The
// lazy chains are bounded by LATENCY: the multipliers / pivot-row entries of the earlier steps come from L2 in batches,
// one dependent round trip per batch (0.85 us; twice that while a pass is streaming the tableau).  Half the threads per
// CTA means twice the registers per thread, i.e. batches of 32 instead of 8: a whole block's chain in one round trip.
template <int CL> struct StepShape {
    static constexpr int TH = (CL == 16) ? SELECT_THREADS / 2 : SELECT_THREADS;   // threads per CTA
    static constexpr int FB = (CL == 16) ? 32 : 8;              // earlier steps per batch of loads, column phase
    static constexpr int PB = (CL == 16) ? 16 : STEP_PB;        // ... row phase (x STEP_CPT columns per thread)
};

template <int RR, bool PIPE, int CL, int BL>       // BL: pricing rule (PR_DANTZIG / PR_BLAND / PR_DEVEX, block_common.cuh)
__global__ void __launch_bounds__(StepShape<CL>::TH)
blk_step_kernel(DenseParams P, BlockBufs B, int first) {
    constexpr int TH = StepShape<CL>::TH, FB = StepShape<CL>::FB, PB = StepShape<CL>::PB;
    // one scratch / slot array per reduction of a launch: no reuse, so no extra barriers
    __shared__ KeyCand scr_a[TH / 32], scr_b[TH / 32], scr_c[TH / 32], scr_d[TH / 32];
    __shared__ KeyCand slot_a[CL], slot_b[CL], slot_d[CL];
    __shared__ double s_pk[2 * BLK_KMAX], s_fr[2 * BLK_KMAX];      // the previous (pending) block's steps first, then this block's
    cg::cluster_group cl = cg::this_cluster();
    const int rank = (int)cl.block_rank();
    const int gtid = rank * TH + (int)threadIdx.x;
    constexpr int GSTRIDE = CL * TH;

    auto stamp = [&](int slot) {
        if (B.dbg && gtid == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            B.dbg[slot] = now;
        }
    };
    stamp(0);
    cluster_entry_arrive();     // DSMEM rule: no remote access before every CTA of the cluster has started (wait below)
    pdl_wait();
    pdl_launch_dependents();
    stamp(1);
    PivotState *st = P.st;
    // every thread of the cluster reads the same state; it is rewritten only after a cluster barrier
    const int status0 = st->status, iters = st->iters, budget = st->budget;
    const int t = first ? 0 : st->blk_t, seq = st->blk_seq + (first ? 1 : 0);
    const int kn = st->k_next;
    cluster_entry_wait();
    // pipelined blocks: P.T is the tableau before the previous block (its pass is still running); the lazy chains
    // run through that block's Kp steps (combined step index c < Kp) and then this block's t steps (c - Kp)
    // (PIPE = false: the in-place order, Kp = 0 at compile time and the code below is the plain one-block chain)
    const int Kp = (PIPE && B.pend) ? B.info[B.par ^ 1].kc : 0;
    const int seqp = (PIPE && B.pend) ? B.info[B.par ^ 1].seq : -1;      // the pending block's own sequence number (not seq - 1: a
                                                                         // decision-only launch between two run calls also counts one up)
    const int tc = Kp + t;
    const uint64_t pol = l2_policy(B.l2keep ? 2 : 0);
    const int offP = PIPE ? (B.par ^ 1) * BLK_KMAX : 0, offC = PIPE ? B.par * BLK_KMAX - Kp : 0;     // combined step c -> row of Fall / Pall
    auto srow = [&](int cc) -> size_t { return (size_t)((cc < Kp ? offP : offC) + cc); };
    KeyCand cin = keycand_none();                  // the previous step's per-CTA candidates, requested together with the state
    if ((threadIdx.x & 31) < CL) {
        const int ln = threadIdx.x & 31;
        cin.kp = __ldcg(&B.cands[ln].kp); cin.kq = __ldcg(&B.cands[ln].kq); cin.idx = __ldcg(&B.cands[ln].idx);
    }
    const bool bigm = (P.method == LPGPU_BIGM);
    const double *oc = B.objcur, *om = B.objcur + P.ld;          // om valid for R == 2 only
    const size_t pcn = P.pcn;
    // two-phase: phase one prices row m+1 over all columns, the main phase row m without the artificial columns;
    // the drive-out pivots between them are decided by the rank-1 path on an up-to-date tableau (host loop)
    const int phase0 = st->phase;
    const int pmode = bigm ? PRICE_BIGM : (phase0 == PH_ONE ? PRICE_PHASE1 : PRICE_COST);
    const int jmax = (P.method == LPGPU_TWOPHASE && phase0 == PH_MAIN) ? P.n - P.n_art : P.n;
    const bool two_keys = bigm;

    int status = status0, k = 0, r = -1, phase = phase0, p1 = st->phase1_iters;
    bool have = false;
    double sig_c = 0.0, sig_m = 0.0, wk = 1.0;
    if (status0 == LPGPU_RUNNING && phase0 != PH_DRIVEOUT) {
        // ---- entering column: priced by the previous step (one candidate per CTA), or from scratch
        if (kn > 0) k = kn;
        else if (kn == -2) {
            const KeyCand c = two_keys ? warp_key_best<true>(cin) : warp_key_best<false>(cin);
            k = c.idx > 0 ? c.idx : 0;
        } else {
            KeyCand c = keycand_none();
            for (int j = 1 + gtid; j <= jmax; j += GSTRIDE) blk_price<BL>(P, pmode, j, oc[j], (RR == 2) ? om[j] : 0.0, (BL == PR_DEVEX) ? P.wts[j] : 1.0, c);
            c = cluster_key_best<true, CL>(cl, c, scr_a, slot_a);
            k = c.idx > 0 ? c.idx : 0;
        }
        if (k == 0) {
            // nothing prices in: optimal -- unless Big-M or phase one left an artificial basic at a positive level
            // (infeasible), or phase one is over (drive-out next: the host loop hands over to the rank-1 path)
            KeyCand a = keycand_none();
            if (bigm || phase0 == PH_ONE) {
                const int nreal = P.n - P.n_art;
                for (int i = gtid; i < P.m; i += GSTRIDE)
                    if (P.basis[i] > nreal && B.bcur[i] > P.eps) { a.idx = i; break; }
            }
            a = cluster_key_best<false, CL>(cl, a, scr_d, slot_d);
            if (a.idx >= 0) status = LPGPU_INFEASIBLE;
            else if (phase0 == PH_ONE) { phase = PH_DRIVEOUT; p1 = iters; }
            else status = LPGPU_OPTIMAL;
        } else {
            // ---- column k as it stands after the t pivots of this block, and the ratio test on it
            sig_c = oc[k];                                        // read before the row phase rewrites objcur
            sig_m = (RR == 2) ? om[k] : 0.0;
            if (BL == PR_DEVEX) wk = P.wts[k];                    // Devex: w_k before this pivot (rewritten in the row phase, after a cluster barrier)
            if (gtid < RR) B.Fs[(size_t)t * pcn + P.m + gtid] = oc[(size_t)gtid * P.ld + k];   // objective rows: kept eagerly, same bits
            // first row of this thread: everything it needs is requested before anyone waits
            int i = gtid;
            double tv = 0.0, bv = 0.0;
            int rs = 0, rsp = 0;
            if (i < P.m) { tv = P.T[(size_t)i * P.ld + (size_t)k]; rs = B.rowstamp[i]; bv = B.bcur[i]; if (Kp) rsp = B.stamp_prev[i]; }
            if ((int)threadIdx.x < tc) {
                const int cc = (int)threadIdx.x;
                s_pk[cc] = ld_hint(B.Pall + srow(cc) * P.ld + (size_t)k, pol);
            }
            __syncthreads();
            KeyCand c = keycand_none();
            for (; i < P.m; i += GSTRIDE) {
                if (i != gtid) { tv = P.T[(size_t)i * P.ld + (size_t)k]; rs = B.rowstamp[i]; bv = B.bcur[i]; if (Kp) rsp = B.stamp_prev[i]; }
                // where this row's chain starts: after its last turn as a pivot row (this block, else the pending one), else at the tableau
                const int s0 = blk_row_start(rs, seq);
                const int c0 = s0 ? Kp + s0 : (Kp ? blk_row_start(rsp, seqp) : 0);
                double a = c0 ? s_pk[c0 - 1] : tv;
                for (int sb = c0; sb < tc; sb += FB) {            // multipliers of FB earlier steps in flight (coalesced over rows)
                    double fv[FB];
#pragma unroll
                    for (int u = 0; u < FB; ++u) {
                        const int cc = sb + u;
                        fv[u] = (cc < tc) ? ld_hint(B.Fall + srow(cc) * pcn + i, pol) : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < FB; ++u)
                        if (sb + u < tc) a = __fma_rn(-fv[u], s_pk[sb + u], a);
                }
                st_hint(B.Fs + (size_t)t * pcn + i, a, pol);
                if (a > P.eps) {
                    KeyCand x;
                    x.kp = 0;
                    x.kq = ord_key(-(bv / a));                  // arg max -theta == arg min theta
                    x.idx = i;
                    if (keycand_better(x, c)) c = x;
                }
            }
            stamp(2);
            c = cluster_key_best<false, CL>(cl, c, scr_b, slot_b);
            stamp(3);
            if (c.idx < 0) status = LPGPU_UNBOUNDED;
            else if (iters >= P.max_iter) status = LPGPU_ITER_LIMIT;
            else if (budget > 0) { have = true; r = c.idx; }
        }
    }
    if (!have) {
        cl.sync();                               // all state reads are done
        if (gtid == 0) {
            st->status = status;
            st->do_pivot = 0;
            st->blk_t = t;
            st->blk_seq = seq;
            st->phase = phase;
            st->phase1_iters = p1;
            if (status == LPGPU_RUNNING) st->k_next = (phase == PH_DRIVEOUT) ? 0 : k;   // budget ran out: the priced column stays valid
            if (B.info) { B.info[B.par].kc = t; B.info[B.par].seq = seq; }
        }
        return;
    }

    // ---- row r as it stands after the t pivots of this block, scaled; objective rows and next pricing
    double av[STEP_CPT], ocv[STEP_CPT], omv[STEP_CPT], wv[STEP_CPT];
    const size_t jb = 1 + (size_t)gtid;           // columns 1..n; column 0 (b_r') is brp, padding columns stay zero
#pragma unroll
    for (int cth = 0; cth < STEP_CPT; ++cth) {
        const size_t j = jb + (size_t)cth * GSTRIDE;
        av[cth] = ocv[cth] = omv[cth] = 0.0;
        wv[cth] = 1.0;
        if (j <= (size_t)P.n) {
            av[cth] = P.T[(size_t)r * P.ld + j];
            ocv[cth] = oc[j];
            if (RR == 2) omv[cth] = om[j];
            if (BL == PR_DEVEX) wv[cth] = P.wts[j];
        }
    }
    const int s0r = blk_row_start(B.rowstamp[r], seq);
    const int c0r = s0r ? Kp + s0r : (Kp ? blk_row_start(B.stamp_prev[r], seqp) : 0);
    const double b_r = B.bcur[r];
    const double piv = __ldcg(B.Fs + (size_t)t * pcn + r);                   // written above, possibly by another CTA
    if ((int)threadIdx.x < tc) {
        const int cc = (int)threadIdx.x;
        s_fr[cc] = ld_hint(B.Fall + srow(cc) * pcn + r, pol);
    }
    __syncthreads();
    // the scaled pivot row the chain restarts from, if this row was a pivot row in this block or the pending one
    const double *rrow = c0r > 0 ? B.Pall + srow(c0r - 1) * P.ld : nullptr;
    double *pt = B.Pm + (size_t)t * P.ld;
    const double brp = div_skip_zero(b_r, piv);       // = P[t][0]: column 0 of the row is the b column
    KeyCand c = keycand_none();
    if (gtid == 0) {                              // column 0
        pt[0] = brp;
        B.objcur[0] = __fma_rn(-sig_c, brp, oc[0]);
        if (RR == 2) B.objcur[P.ld] = __fma_rn(-sig_m, brp, om[0]);
    }
    for (size_t j0 = jb; j0 <= (size_t)P.n; j0 += (size_t)STEP_CPT * GSTRIDE) {
        if (j0 != jb) {
#pragma unroll
            for (int cth = 0; cth < STEP_CPT; ++cth) {
                const size_t j = j0 + (size_t)cth * GSTRIDE;
                if (j <= (size_t)P.n) { av[cth] = P.T[(size_t)r * P.ld + j]; ocv[cth] = oc[j]; if (RR == 2) omv[cth] = om[j]; if (BL == PR_DEVEX) wv[cth] = P.wts[j]; }
            }
        }
        if (rrow) {                               // the row was a pivot row earlier: restart from its scaled row
#pragma unroll
            for (int cth = 0; cth < STEP_CPT; ++cth) {
                const size_t j = j0 + (size_t)cth * GSTRIDE;
                if (j <= (size_t)P.n) av[cth] = ld_hint(rrow + j, pol);
            }
        }
        for (int sb = c0r; sb < tc; sb += PB) {         // PB earlier pivot rows x STEP_CPT columns in flight
            double pv[STEP_CPT][PB];
#pragma unroll
            for (int cth = 0; cth < STEP_CPT; ++cth) {
                const size_t j = j0 + (size_t)cth * GSTRIDE;
#pragma unroll
                for (int u = 0; u < PB; ++u) {
                    const int cc = sb + u;
                    pv[cth][u] = (cc < tc && j <= (size_t)P.n) ? ld_hint(B.Pall + srow(cc) * P.ld + j, pol) : 0.0;
                }
            }
#pragma unroll
            for (int u = 0; u < PB; ++u)
                if (sb + u < tc) {
#pragma unroll
                    for (int cth = 0; cth < STEP_CPT; ++cth) av[cth] = __fma_rn(-s_fr[sb + u], pv[cth][u], av[cth]);
                }
        }
#pragma unroll
        for (int cth = 0; cth < STEP_CPT; ++cth) {
            const size_t j = j0 + (size_t)cth * GSTRIDE;
            if (j <= (size_t)P.n) {
                const double pr = div_skip_zero(av[cth], piv);
                st_hint(pt + j, pr, pol);
                const double qc = __fma_rn(-sig_c, pr, ocv[cth]);
                B.objcur[j] = qc;
                double qm = 0.0;
                if (RR == 2) {
                    qm = __fma_rn(-sig_m, pr, omv[cth]);
                    B.objcur[P.ld + j] = qm;
                }
                double w = 1.0;
                if (BL == PR_DEVEX) {
                    w = devex_weight(wv[cth], pr, wk, j == (size_t)k);
                    P.wts[j] = w;
                }
                if (j <= (size_t)jmax) blk_price<BL>(P, pmode, (int)j, qc, qm, w, c);
            }
        }
    }
    // ---- this CTA's best candidate for the next pivot; the next launch reduces the SELECT_CLUSTER of them
    c = two_keys ? block_key_best<true>(c, scr_c) : block_key_best<false>(c, scr_c);
    if (threadIdx.x == 0) B.cands[rank] = c;
    if (PIPE) {      // ... and every CTA of the cluster learns them too (split barrier: nobody waits here), see the prefetch below
        if (threadIdx.x < CL) cl.map_shared_rank(slot_a, threadIdx.x)[rank] = c;
        cluster_entry_arrive();
    }
    stamp(4);
    // ---- b column
    for (int i = gtid; i < P.m; i += GSTRIDE)
        B.bcur[i] = (i == r) ? brp : __fma_rn(-B.Fs[(size_t)t * pcn + i], brp, B.bcur[i]);
    stamp(5);

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
        st->k_next = -2;
        st->phase1_iters = (phase0 == PH_ONE) ? iters + 1 : p1;
        st->blk_t = t + 1;
        st->blk_seq = seq;
        if (B.info) { B.info[B.par].kc = t + 1; B.info[B.par].seq = seq; }
        if (iters < P.log_cap) { P.log[2 * iters] = k; P.log[2 * iters + 1] = r; }
    }
    if (PIPE) {
        // Pipelined blocks: the pass streams the tableau at the HBM peak beside this kernel, and the next launch's first act is
        // a strided read of the entering column from HBM -- several microseconds in that queue.  The column is known now
        // (the best of the CL candidates): one 32-byte sector per row starts its way into L2 before this launch retires.
        cluster_entry_wait();
        const KeyCand w = two_keys ? warp_key_best<true>((threadIdx.x & 31) < CL ? slot_a[threadIdx.x & 31] : keycand_none())
                                   : warp_key_best<false>((threadIdx.x & 31) < CL ? slot_a[threadIdx.x & 31] : keycand_none());
        if (w.idx > 0)
            for (int i = gtid; i < P.m; i += GSTRIDE) asm volatile("prefetch.global.L2 [%0];" ::"l"(P.T + (size_t)i * P.ld + (size_t)w.idx));
    }
    stamp(6);
}

// ------------------------------------------------------------------ dual simplex step (LPGPU_DUAL, R = 1)
// Same block machinery, decisions in the dual order: leaving row = most negative b (from bcur; the candidates
// of the NEXT pivot are left per CTA at the end of a step, as in the primal kernel), row r brought up to
// date lazily, entering column = min sigma_j / a_rj over a_rj < -eps on it, column k brought up to date
// lazily -> F[t], row scaled by a_rk = F[t][r] -> P[t], objective row and b updated.
template <int CL, bool PIPE>        // PIPE: pipelined blocks -- the chains run through the pending block's steps first (see blk_step_kernel)
__global__ void __launch_bounds__(StepShape<CL>::TH)
blk_dual_step_kernel(DenseParams P, BlockBufs B, int first) {
    constexpr int TH = StepShape<CL>::TH, FB = StepShape<CL>::FB, PB = StepShape<CL>::PB;      // cluster shape: see blk_step_kernel
    __shared__ KeyCand scr_a[TH / 32], scr_b[TH / 32], scr_c[TH / 32];
    __shared__ KeyCand slot_a[CL], slot_b[CL];
    __shared__ double s_pk[2 * BLK_KMAX], s_fr[2 * BLK_KMAX];      // the pending block's steps first, then this block's
    cg::cluster_group cl = cg::this_cluster();
    const int rank = (int)cl.block_rank();
    const int gtid = rank * TH + (int)threadIdx.x;
    constexpr int GSTRIDE = CL * TH;

    cluster_entry_arrive();
    pdl_wait();
    pdl_launch_dependents();
    PivotState *st = P.st;
    const int status0 = st->status, iters = st->iters, budget = st->budget;
    const int t = first ? 0 : st->blk_t, seq = st->blk_seq + (first ? 1 : 0);
    const int kn = st->k_next;                       // -2: per-CTA leaving-row candidates are in B.cands
    cluster_entry_wait();
    KeyCand cin = keycand_none();
    if ((threadIdx.x & 31) < CL) {
        const int ln = threadIdx.x & 31;
        cin.kp = __ldcg(&B.cands[ln].kp); cin.kq = __ldcg(&B.cands[ln].kq); cin.idx = __ldcg(&B.cands[ln].idx);
    }
    const double *oc = B.objcur;
    const size_t pcn = P.pcn;
    // pipelined blocks: P.T is the tableau before the pending block; combined step index c < Kp = the pending block's step c
    const int Kp = (PIPE && B.pend) ? B.info[B.par ^ 1].kc : 0;
    const int seqp = (PIPE && B.pend) ? B.info[B.par ^ 1].seq : -1;
    const int tc = Kp + t;
    const int offP = PIPE ? (B.par ^ 1) * BLK_KMAX : 0, offC = PIPE ? B.par * BLK_KMAX - Kp : 0;
    auto srow = [&](int cc) -> size_t { return (size_t)((cc < Kp ? offP : offC) + cc); };
    auto row_start = [&](int i) -> int {       // where row i's chain starts: after its last turn as a pivot row (this block, else the pending one)
        const int s0 = blk_row_start(B.rowstamp[i], seq);
        return s0 ? Kp + s0 : (Kp ? blk_row_start(B.stamp_prev[i], seqp) : 0);
    };

    int status = status0, k = 0, r = -1;
    bool have = false;
    double av[STEP_CPT], ocv[STEP_CPT];
    const size_t jb = 1 + (size_t)gtid;
    if (status0 == LPGPU_RUNNING) {
        // ---- leaving row: most negative b_i < -eps, ties to the lowest row
        if (kn == -2) r = warp_key_best<false>(cin).idx;
        else {
            KeyCand c = keycand_none();
            for (int i = gtid; i < P.m; i += GSTRIDE) {
                const double b = B.bcur[i];
                if (b < -P.eps) {
                    KeyCand x;
                    x.kp = 0; x.kq = ord_key(-b); x.idx = i;
                    if (keycand_better(x, c)) c = x;
                }
            }
            r = cluster_key_best<false, CL>(cl, c, scr_a, slot_a).idx;
        }
        if (r < 0) status = LPGPU_OPTIMAL;
        else {
            // ---- row r as it stands after the t pivots of this block; entering column on it
            const int s0r = row_start(r);
#pragma unroll
            for (int cth = 0; cth < STEP_CPT; ++cth) {
                const size_t j = jb + (size_t)cth * GSTRIDE;
                av[cth] = ocv[cth] = 0.0;
                if (j <= (size_t)P.n) { av[cth] = P.T[(size_t)r * P.ld + j]; ocv[cth] = oc[j]; }
            }
            if ((int)threadIdx.x < tc) s_fr[threadIdx.x] = B.Fall[srow((int)threadIdx.x) * pcn + r];
            __syncthreads();
            KeyCand c = keycand_none();
            // n <= STEP_CPT * GSTRIDE columns are kept in registers across the column phase; wider tableaux recompute
            for (size_t j0 = jb; j0 <= (size_t)P.n; j0 += (size_t)STEP_CPT * GSTRIDE) {
                double aw[STEP_CPT], ow[STEP_CPT];
#pragma unroll
                for (int cth = 0; cth < STEP_CPT; ++cth) {
                    const size_t j = j0 + (size_t)cth * GSTRIDE;
                    aw[cth] = ow[cth] = 0.0;
                    if (j0 == jb) { aw[cth] = av[cth]; ow[cth] = ocv[cth]; }
                    else if (j <= (size_t)P.n) { aw[cth] = P.T[(size_t)r * P.ld + j]; ow[cth] = oc[j]; }
                    if (s0r > 0 && j <= (size_t)P.n) aw[cth] = B.Pall[srow(s0r - 1) * P.ld + j];
                }
                for (int sb = s0r; sb < tc; sb += PB) {
                    double pv[STEP_CPT][PB];
#pragma unroll
                    for (int cth = 0; cth < STEP_CPT; ++cth) {
                        const size_t j = j0 + (size_t)cth * GSTRIDE;
#pragma unroll
                        for (int u = 0; u < PB; ++u) pv[cth][u] = (sb + u < tc && j <= (size_t)P.n) ? B.Pall[srow(sb + u) * P.ld + j] : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < PB; ++u)
                        if (sb + u < tc) {
#pragma unroll
                            for (int cth = 0; cth < STEP_CPT; ++cth) aw[cth] = __fma_rn(-s_fr[sb + u], pv[cth][u], aw[cth]);
                        }
                }
#pragma unroll
                for (int cth = 0; cth < STEP_CPT; ++cth) {
                    const size_t j = j0 + (size_t)cth * GSTRIDE;
                    if (j0 == jb) av[cth] = aw[cth];
                    if (j <= (size_t)P.n && aw[cth] < -P.eps) {
                        KeyCand x;
                        x.kp = 0; x.kq = ord_key(-(ow[cth] / aw[cth])); x.idx = (int)j;
                        if (keycand_better(x, c)) c = x;
                    }
                }
            }
            c = cluster_key_best<false, CL>(cl, c, scr_b, slot_b);
            if (c.idx < 0) status = LPGPU_INFEASIBLE;
            else if (iters >= P.max_iter) status = LPGPU_ITER_LIMIT;
            else if (budget > 0) { have = true; k = c.idx; }
        }
    }
    if (!have) {
        cl.sync();
        if (gtid == 0) {
            st->status = status;
            st->do_pivot = 0;
            st->blk_t = t;
            st->blk_seq = seq;
            st->k_next = 0;                       // the leaving row is looked for again from bcur
            if (B.info) { B.info[B.par].kc = t; B.info[B.par].seq = seq; }
        }
        return;
    }

    // ---- column k as it stands after the t pivots of this block -> F[t]
    const double sig = oc[k];
    if (gtid == 0) B.Fs[(size_t)t * pcn + P.m] = sig;
    if ((int)threadIdx.x < tc) s_pk[threadIdx.x] = B.Pall[srow((int)threadIdx.x) * P.ld + (size_t)k];
    __syncthreads();
    for (int i = gtid; i < P.m; i += GSTRIDE) {
        const int s0 = row_start(i);
        double a = s0 ? s_pk[s0 - 1] : P.T[(size_t)i * P.ld + (size_t)k];
        for (int sb = s0; sb < tc; sb += FB) {
            double fv[FB];
#pragma unroll
            for (int u = 0; u < FB; ++u) fv[u] = (sb + u < tc) ? B.Fall[srow(sb + u) * pcn + i] : 0.0;
#pragma unroll
            for (int u = 0; u < FB; ++u)
                if (sb + u < tc) a = __fma_rn(-fv[u], s_pk[sb + u], a);
        }
        B.Fs[(size_t)t * pcn + i] = a;
    }
    const double b_r = B.bcur[r];
    const int s0r = row_start(r);
    cl.sync();                                    // F[t] is complete (the pivot element among it)
    const double piv = __ldcg(B.Fs + (size_t)t * pcn + r);

    // ---- scaled row -> P[t], objective row, b column, next leaving-row candidate
    double *pt = B.Pm + (size_t)t * P.ld;
    const double brp = div_skip_zero(b_r, piv);
    if (gtid == 0) {
        pt[0] = brp;
        B.objcur[0] = __fma_rn(-sig, brp, oc[0]);
    }
    for (size_t j0 = jb; j0 <= (size_t)P.n; j0 += (size_t)STEP_CPT * GSTRIDE) {
#pragma unroll
        for (int cth = 0; cth < STEP_CPT; ++cth) {
            const size_t j = j0 + (size_t)cth * GSTRIDE;
            if (j > (size_t)P.n) continue;
            double a, o;
            if (j0 == jb) { a = av[cth]; o = ocv[cth]; }
            else {                                // wide tableau: the row slice again (rare; same chain, same bits)
                a = s0r ? B.Pall[srow(s0r - 1) * P.ld + j] : P.T[(size_t)r * P.ld + j];
                for (int s = s0r; s < tc; ++s) a = __fma_rn(-s_fr[s], B.Pall[srow(s) * P.ld + j], a);
                o = oc[j];
            }
            const double pr = div_skip_zero(a, piv);
            pt[j] = pr;
            B.objcur[j] = __fma_rn(-sig, pr, o);
        }
    }
    KeyCand c = keycand_none();
    for (int i = gtid; i < P.m; i += GSTRIDE) {
        const double nb = (i == r) ? brp : __fma_rn(-B.Fs[(size_t)t * pcn + i], brp, B.bcur[i]);
        B.bcur[i] = nb;
        if (nb < -P.eps) {
            KeyCand x;
            x.kp = 0; x.kq = ord_key(-nb); x.idx = i;
            if (keycand_better(x, c)) c = x;
        }
    }
    c = block_key_best<false>(c, scr_c);
    if (threadIdx.x == 0) B.cands[rank] = c;
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
        st->k_next = -2;
        st->blk_t = t + 1;
        st->blk_seq = seq;
        if (B.info) { B.info[B.par].kc = t + 1; B.info[B.par].seq = seq; }
        if (iters < P.log_cap) { P.log[2 * iters] = k; P.log[2 * iters + 1] = r; }
    }
}

// ------------------------------------------------------------------ the pass that applies a block
// CTA tile = FL_ROWS x FL_COLS, 4 warps stacked over the rows; a thread owns 16 rows x 4 columns
// (two 128-bit accesses per row, 64 accumulator registers).  Per step of the chain a thread needs two
// pivot-row pairs (per-lane LDS.128: 4 shared-memory wavefronts per warp each) and 16 multipliers that
// are warp-uniform (a broadcast LDS.128 = 1 wavefront serves two steps of a row): 16 wavefronts per 64
// DFMAs, so the FP64 pipe, not shared memory, is the busier unit (a 2 x 8 thread tile measured 4x
// slower: 18 wavefronts per 16 DFMAs).  Steps are taken in pairs.
// A task = one column tile x FL_RB consecutive row tiles: the K staged pivot-row segments (K x 1 KB,
// TMA bulk copies) are reused FL_RB times, the multipliers (64 rows x K, one 16 KB bulk copy) are
// double buffered.  Tasks are claimed from a global counter (protocol of dense_update_dynamic).
// Rows that were pivot rows inside the block are NOT special-cased here: blk_fixrows_kernel rewrites
// them afterwards.
constexpr int FL_THREADS = 128;
constexpr int FL_COLS = 128;
constexpr int FL_ROWS = 64;
constexpr int FL_TR = 16;              // rows per thread (x 4 columns: two 128-bit accesses 64 columns apart)
constexpr int FL_RG = 4;               // rows whose multipliers are in flight together
constexpr int FL_RB = 4;

struct FlushParams {
    double *T;                // where the pass writes
    const double *Tsrc;       // where it reads: == T in place, the other tableau buffer with pipelined blocks
    const int *kc_ptr;        // pivots in the block (st->blk_t in place, info[par].kc pipelined)
    const int *seq_ptr;       // the block's sequence number
    unsigned resv[8];         // bit smid set: that SM is left to the step kernels of the next block (CTAs landing there exit)
    int *flag;                // >= seq + 1 once a CTA of this launch runs on an SM that is not reserved
    int l2first;              // 1: the tableau goes through L2 with the evict-first priority (blk_flush_sp_kernel)
    size_t ld;
    int rows;                 // m + R
    const double *Fs;         // BLK_KMAX x pcn
    size_t pcn;
    const double *Pm;         // BLK_KMAX x ld
    const int *rowstamp;
    const int *steprow;       // pivot row of step s of the current block
    const PivotState *st;
    int *task_ctr;
    unsigned seq;
    int rb;                   // row tiles per task (FL_RB; LPGPU_FL_RB overrides it for tuning)
    int pf;                   // 1: HBM -> L2 prefetch of the next tile
    int ring_nd;              // ring pass: data slots in use, 2 .. RG_ND (fewer = fewer bytes queued in front of the step kernels' requests)
};

constexpr size_t FL_SMEM_BYTES = (size_t)BLK_KMAX * FL_COLS * 8 + 2 * (size_t)FL_ROWS * BLK_KMAX * 8;

// where a tile's operands live
struct FlTile {
    size_t j0;          // first column of the tile
    int ncols;          // columns in it (a multiple of 16)
    int rt;             // row tile index
};

// ------------------------------------------------------------------ the scalar pass, software-pipelined in half tiles (round 1's default; LPGPU_VARIANT_DFMA_PASS)
// A CTA that loads its whole 64 x 128 tile (32 LDG.128 per thread), waits, runs the chain on it and stores it: memory
// idles while the CTA computes and the FP64 pipe idles while it loads, and three CTAs per SM in arbitrary phase only
// partly cover for one another (ncu: DRAM 50 %, FP64 pipe 53 % busy).  Here a thread's 16 rows are two halves of 8:
// while the chain runs on one half (32 accumulator registers) the loads of the other half -- of this tile, or of the
// first half of the CTA's next tile -- are in flight, so every CTA keeps both the memory system and the FP64 pipe busy
// all the time.  Registers: the same 64 accumulators.
constexpr int FL_HR = FL_TR / 2;       // rows per half

template <int NR>
__device__ __forceinline__ void fl_load_rows(const FlushParams &U, const FlTile &t, int row0, int jc, double2 (&acc)[NR][2], uint64_t pol) {
    const double *tp = U.Tsrc + (size_t)row0 * U.ld + t.j0 + jc;
    const bool ok0 = jc < t.ncols, ok1 = jc + 64 < t.ncols;
#pragma unroll
    for (int rr = 0; rr < NR; ++rr) {
        const bool rowok = row0 + rr < U.rows;
        acc[rr][0] = (ok0 && rowok) ? ld_stream2_hint(tp + (size_t)rr * U.ld, pol) : make_double2(0.0, 0.0);
        acc[rr][1] = (ok1 && rowok) ? ld_stream2_hint(tp + (size_t)rr * U.ld + 64, pol) : make_double2(0.0, 0.0);
    }
}

template <int NR>
__device__ __forceinline__ void fl_store_rows(const FlushParams &U, const FlTile &t, int row0, int jc, double2 (&acc)[NR][2], uint64_t pol) {
    double *tp = U.T + (size_t)row0 * U.ld + t.j0 + jc;
    const bool ok0 = jc < t.ncols, ok1 = jc + 64 < t.ncols;
#pragma unroll
    for (int rr = 0; rr < NR; ++rr) {
        const bool rowok = row0 + rr < U.rows;
        if (ok0 && rowok) st_stream2_hint(tp + (size_t)rr * U.ld, acc[rr][0], pol);
        if (ok1 && rowok) st_stream2_hint(tp + (size_t)rr * U.ld + 64, acc[rr][1], pol);
    }
}

// the chain of kc steps on NR rows x 4 columns of one thread (fw: the multipliers of its first row, pw: its column pair)
template <int NR>
__device__ __forceinline__ void fl_chain_rows(int kc, const double *fw, const double *pw, double2 (&acc)[NR][2]) {
    int s = 0;
    for (; s + 1 < kc; s += 2) {
        double2 p[2][2];                                  // [step][column pair]
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int q = 0; q < 2; ++q) p[e][q] = *reinterpret_cast<const double2 *>(pw + (size_t)(s + e) * FL_COLS + 64 * q);
#pragma unroll
        for (int rg = 0; rg < NR; rg += FL_RG) {
            double2 f[2][FL_RG / 2];
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int u = 0; u < FL_RG / 2; ++u) f[e][u] = *reinterpret_cast<const double2 *>(fw + (size_t)(s + e) * FL_ROWS + rg + 2 * u);
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int u = 0; u < FL_RG / 2; ++u)
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        acc[rg + 2 * u][q].x = __fma_rn(-f[e][u].x, p[e][q].x, acc[rg + 2 * u][q].x);
                        acc[rg + 2 * u][q].y = __fma_rn(-f[e][u].x, p[e][q].y, acc[rg + 2 * u][q].y);
                        acc[rg + 2 * u + 1][q].x = __fma_rn(-f[e][u].y, p[e][q].x, acc[rg + 2 * u + 1][q].x);
                        acc[rg + 2 * u + 1][q].y = __fma_rn(-f[e][u].y, p[e][q].y, acc[rg + 2 * u + 1][q].y);
                    }
            asm volatile("" ::: "memory");
        }
    }
    if (s < kc) {
        double2 p[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) p[q] = *reinterpret_cast<const double2 *>(pw + (size_t)s * FL_COLS + 64 * q);
#pragma unroll
        for (int rr = 0; rr < NR; ++rr) {
            const double f = fw[(size_t)s * FL_ROWS + rr];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                acc[rr][q].x = __fma_rn(-f, p[q].x, acc[rr][q].x);
                acc[rr][q].y = __fma_rn(-f, p[q].y, acc[rr][q].y);
            }
        }
    }
}

__global__ void __launch_bounds__(FL_THREADS, 3) blk_flush_sp_kernel(FlushParams U) {
    extern __shared__ __align__(128) unsigned char fl_smem[];
    double *sP = reinterpret_cast<double *>(fl_smem);                                 // [BLK_KMAX][FL_COLS]
    double *sF = sP + (size_t)BLK_KMAX * FL_COLS;                                     // [2][BLK_KMAX][FL_ROWS]
    __shared__ __align__(8) uint64_t bar_p, bar_f[2];
    __shared__ long s_task[2];

    if (threadIdx.x == 0) {
        mbar_init(&bar_p, 1);
        mbar_init(&bar_f[0], 1);
        mbar_init(&bar_f[1], 1);
        fence_mbar_init();
    }
    pdl_wait();
    int *ctr = U.task_ctr + (U.seq & 1u);
    if (blockIdx.x == 0 && threadIdx.x == 0) U.task_ctr[(U.seq & 1u) ^ 1u] = 0;
    if (U.flag) {       // pipelined blocks: see blk_flush_ring_kernel
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        const int want = (int)(U.seq + 1u);
        if ((U.resv[(smid >> 5) & 7u] >> (smid & 31u)) & 1u) {
            __shared__ int s_leave;
            if (threadIdx.x == 0) {
                unsigned long long t0, now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
                bool leave = false;
                do {
                    leave = (*(volatile int *)U.flag - want) >= 0;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                } while (!leave && now - t0 < 20000ull);
                s_leave = leave ? 1 : 0;
            }
            __syncthreads();
            if (s_leave) return;
        } else if (threadIdx.x == 0) atomicMax(U.flag, want);
    }
    const int kc = *U.kc_ptr;
    if (kc <= 0 && U.Tsrc == U.T) return;

    const int nct = (int)((U.ld + FL_COLS - 1) / FL_COLS);
    const int nrt = (U.rows + FL_ROWS - 1) / FL_ROWS;
    const int nsb = (nrt + U.rb - 1) / U.rb;
    const long ntasks = (long)nct * nsb;
    const int lane = threadIdx.x & 31, wr = threadIdx.x >> 5;
    const int jc = 2 * lane;
    const uint32_t FBYTES = (uint32_t)(kc * FL_ROWS * 8);
    auto issue_f = [&](int rt, int fb) {
        mbar_arrive_expect_tx(&bar_f[fb], FBYTES);
        for (int s = 0; s < kc; ++s)
            tma_load_1d(sF + ((size_t)fb * BLK_KMAX + s) * FL_ROWS, U.Fs + (size_t)s * U.pcn + (size_t)rt * FL_ROWS, (uint32_t)(FL_ROWS * 8), &bar_f[fb]);
    };
    auto tile_of = [&](long tt, int k) {             // k-th tile of task tt
        FlTile t;
        t.j0 = (size_t)(tt % nct) * FL_COLS;
        t.ncols = (int)min((size_t)FL_COLS, U.ld - t.j0);
        t.rt = (int)(tt / nct) * U.rb + k;
        return t;
    };

    if (threadIdx.x == 0) s_task[0] = (long)atomicAdd(ctr, 1);
    __syncthreads();
    uint32_t par_p = 0, par_f = 0;
    const uint64_t pol = l2_policy(U.l2first ? 1 : 0);
    double2 acc0[FL_HR][2], acc1[FL_HR][2];            // the two halves of a thread's 16 rows
    bool pre = false;                                  // acc0 already holds (is being loaded with) the first half of the coming tile
    for (int it = 0;; ++it) {
        const long tt = s_task[it & 1];
        if (tt >= ntasks) break;
        FlTile t = tile_of(tt, 0);
        const int rt0 = t.rt, rt1 = min(rt0 + U.rb, nrt);
        if (threadIdx.x == 0) {
            s_task[(it & 1) ^ 1] = (long)atomicAdd(ctr, 1);            // next task; visible after the barriers below
            mbar_arrive_expect_tx(&bar_p, (uint32_t)(kc * t.ncols * 8));
            for (int s = 0; s < kc; ++s) tma_load_1d(sP + (size_t)s * FL_COLS, U.Pm + (size_t)s * U.ld + t.j0, (uint32_t)(t.ncols * 8), &bar_p);
            issue_f(rt0, 0);
        }
        for (int rt = rt0; rt < rt1; ++rt) {
            const int fb = (rt - rt0) & 1;
            t.rt = rt;
            const int row0 = rt * FL_ROWS + FL_TR * wr;
            if (!pre) fl_load_rows<FL_HR>(U, t, row0, jc, acc0, pol);
            fl_load_rows<FL_HR>(U, t, row0 + FL_HR, jc, acc1, pol);      // in flight while the chain runs on the first half
            if (threadIdx.x == 0 && rt + 1 < rt1) issue_f(rt + 1, fb ^ 1);
            if (rt == rt0) { mbar_wait(&bar_p, par_p); par_p ^= 1; }
            mbar_wait(&bar_f[fb], (par_f >> fb) & 1u);
            par_f ^= 1u << fb;
            const double *fw = sF + (size_t)fb * BLK_KMAX * FL_ROWS + FL_TR * wr;
            fl_chain_rows<FL_HR>(kc, fw, sP + jc, acc0);
            fl_store_rows<FL_HR>(U, t, row0, jc, acc0, pol);
            // the tile after this one (same task, or the first tile of the next task once its claim is visible):
            // its first half is requested now and arrives while the chain runs on the second half
            FlTile nx = t;
            pre = true;
            if (rt + 1 < rt1) nx.rt = rt + 1;
            else {
                const long nt = s_task[(it & 1) ^ 1];
                pre = (rt > rt0) && nt < ntasks;       // the claim was published before tile rt0's closing barrier
                if (pre) nx = tile_of(nt, 0);
            }
            if (pre) fl_load_rows<FL_HR>(U, nx, nx.rt * FL_ROWS + FL_TR * wr, jc, acc0, pol);
            fl_chain_rows<FL_HR>(kc, fw + FL_HR, sP + jc, acc1);
            fl_store_rows<FL_HR>(U, t, row0 + FL_HR, jc, acc1, pol);
            __syncthreads();       // sF[fb] may be refilled (two tiles ahead), and at the last tile sP and s_task
        }
    }
}

// Rows that were a pivot row inside the block: restart from the scaled pivot row of their LAST such step
// and run the steps after it.  grid = (column chunks, BLK_KMAX); block y = step s.
__global__ void __launch_bounds__(256) blk_fixrows_kernel(FlushParams U) {
    const int kc = *U.kc_ptr, bseq = *U.seq_ptr, s0 = (int)blockIdx.y;
    if (s0 >= kc) return;
    const int r = U.steprow[s0];
    if (blk_row_start(U.rowstamp[r], bseq) != s0 + 1) return;      // the row was a pivot row again later in the block
    __shared__ double s_f[BLK_KMAX];
    if ((int)threadIdx.x < kc) s_f[threadIdx.x] = U.Fs[(size_t)threadIdx.x * U.pcn + r];
    __syncthreads();
    for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < U.ld; j += (size_t)gridDim.x * blockDim.x) {
        double a = U.Pm[(size_t)s0 * U.ld + j];
        for (int sb = s0 + 1; sb < kc; sb += 16) {            // 16 later pivot rows in flight: one round trip per batch, not per step
            double pv[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) pv[u] = (sb + u < kc) ? U.Pm[(size_t)(sb + u) * U.ld + j] : 0.0;
#pragma unroll
            for (int u = 0; u < 16; ++u)
                if (sb + u < kc) a = __fma_rn(-s_f[sb + u], pv[u], a);
        }
        U.T[(size_t)r * U.ld + j] = a;
    }
}

}  // namespace lpgpu
