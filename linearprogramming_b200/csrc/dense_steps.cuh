// dense_steps.cuh -- the decisions of a whole block of pivots in ONE persistent launch (pipelined blocks, primal methods).
//
// blk_step_kernel (dense_block.cuh) is one launch of one 16-CTA cluster per pivot.  With pipelined blocks every launch
// re-reads, from L2, the lazily applied slices of the pending block AND of its own block for its rows and columns:
// (Kp + t) * 8 * (m + n) bytes per pivot -- 9 MB on average at 8192 x 16384, 590 KB per SM, which is what bounds it
// (about 64 B / clock / SM out of L2; measured 17.9 us per pivot beside the pass, profiles/r2_step_phases_start.txt).
//
// Here G CTAs (32 by default, one per SM, NO cluster: they exchange through L2) stay resident for all the pivots of a
// block.  CTA c owns a contiguous range of rows (column phase, ratio test, b column) and of columns (row phase,
// objective rows, pricing):
//   * the PENDING block's slices of its rows / columns (F: Kp x rows, P: Kp x columns; constant for the whole block)
//     are brought into shared memory once, by TMA bulk copies, at the start of the launch;
//   * the current block's slices (t <= 31 steps) come from L2 in ONE batch per phase, requested together with the
//     tableau column / row, the stamps, b and the objective row -- one memory round trip per phase;
//   * the two arg-max reductions of a pivot (ratio test; pricing of the next pivot) are a CTA reduction on
//     order-preserving keys followed by an all-to-all through sequence-numbered 64-byte slots in global memory
//     (st.release.gpu / ld.acquire.gpu): one lane per peer CTA polls, one warp reduces.  The ratio winner carries
//     its pivot element and b_r along, so nobody goes back to memory for them.
// Per pivot: 2 memory round trips + 2 exchanges + 2 short chains out of shared memory / registers, whatever the
// length of the chains.  Same operations on the same operands in the same order as blk_step_kernel (and so as the
// rank-1 path and the oracle): every blocked parity test runs through this kernel.
//
// All G CTAs must be resident together (they wait for each other).  They are: the pass of the previous block leaves
// the SMs this grid was placed on while the device was idle (the reservation of dense.cu), and where that fails a CTA
// of this launch simply starts once the pass has finished -- waiting costs time, never the result.  Polls are bounded:
// a CTA that never hears from a peer reports LPGPU_STEP_TIMEOUT through the pivot state instead of hanging the GPU.
// Reference: the loop missing from newSimplex (Source/simplex.c:40-67).
#pragma once

#include "dense_block.cuh"

namespace lpgpu {

constexpr int STEPS_TH = 512;
constexpr int STEPS_GMAX = 64;                  // CTAs of the persistent grid at most (two polls per lane)
constexpr int LPGPU_STEP_TIMEOUT = 6;           // internal status: a peer CTA never posted
constexpr unsigned STEPS_SPIN_MAX = 1u << 23;   // polls before giving up (a few seconds)

struct __align__(64) XSlot {
    unsigned long long kp, kq;
    int idx, aux;
    double v0, v1;
    unsigned flag;
    unsigned pad[5];
};

struct StepsArgs {
    XSlot *xch;          // [2][STEPS_GMAX]
    unsigned xbase;      // sequence number of this launch's first exchange (host: strictly increasing, never reused)
    int nsteps;          // pivots this launch may apply (room in the block and in the run budget); 0 = decision only
    int first;           // 1: a new block starts with this launch
    int smem_steps;      // pending steps whose slices fit the CTA's dynamic shared memory
    int rpc, cpc;        // rows / tableau columns per CTA (both even)
};

__device__ __forceinline__ double ldcg_f64(const double *p) { return __ldcg(p); }

__device__ __forceinline__ void xslot_post(XSlot *s, const KeyCand &c, double v0, double v1, unsigned seq) {
    __stcg(&s->kp, c.kp); __stcg(&s->kq, c.kq); __stcg(&s->idx, c.idx); __stcg(&s->v0, v0); __stcg(&s->v1, v1);
    __threadfence();       // this CTA's earlier stores (slices, objective row, b, stamps) and the payload, before the flag
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&s->flag), "r"(seq) : "memory");
}

__device__ __forceinline__ bool xslot_wait(const XSlot *s, unsigned seq) {
    for (unsigned spin = 0; spin < STEPS_SPIN_MAX; ++spin) {
        unsigned f;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(f) : "l"(&s->flag) : "memory");
        if (f == seq) return true;
    }
    return false;
}

struct StepsShared {
    KeyCand scr[2][STEPS_TH / 32];
    double scrv[2][2 * (STEPS_TH / 32)];
    KeyCand res;
    double res_v0, res_v1;
    int res_ok;
};

// CTA-wide best candidate WITH the two payload doubles of the winner; result in every thread.  `set`: which scratch set
// (successive reductions alternate, a set is reused only after several barriers).
template <bool TWO>
__device__ __forceinline__ KeyCand block_best_payload(KeyCand c, double &v0, double &v1, StepsShared &sh, int set) {
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    KeyCand w = warp_key_best<TWO>(c);
    unsigned who = __ballot_sync(FULL, c.idx >= 0 && c.idx == w.idx);
    int sl = who ? __ffs(who) - 1 : 0;
    const double a0 = __shfl_sync(FULL, v0, sl), a1 = __shfl_sync(FULL, v1, sl);
    if (lane == 0) { sh.scr[set][warp] = w; sh.scrv[set][2 * warp] = a0; sh.scrv[set][2 * warp + 1] = a1; }
    __syncthreads();
    constexpr int NW = STEPS_TH / 32;
    const KeyCand x = lane < NW ? sh.scr[set][lane] : keycand_none();
    const double b0 = lane < NW ? sh.scrv[set][2 * lane] : 0.0, b1 = lane < NW ? sh.scrv[set][2 * lane + 1] : 0.0;
    w = warp_key_best<TWO>(x);
    who = __ballot_sync(FULL, x.idx >= 0 && x.idx == w.idx);
    sl = who ? __ffs(who) - 1 : 0;
    v0 = __shfl_sync(FULL, b0, sl);
    v1 = __shfl_sync(FULL, b1, sl);
    return w;
}

// All-to-all of the CTAs' winners: thread 0 posts this CTA's (valid in every thread), warp 0 polls every CTA's slot of
// this sequence number and reduces; the grid-wide winner and its payload come back in every thread.  ok = false: timeout.
template <bool TWO>
__device__ __forceinline__ KeyCand grid_best(const StepsArgs &A, unsigned seq, KeyCand mine, double &v0, double &v1, StepsShared &sh, bool &ok) {
    constexpr unsigned FULL = 0xffffffffu;
    const int G = (int)gridDim.x, lane = threadIdx.x & 31;
    XSlot *slots = A.xch + (size_t)(seq & 1u) * STEPS_GMAX;
    if (threadIdx.x == 0) xslot_post(slots + blockIdx.x, mine, v0, v1, seq);
    if (threadIdx.x < 32) {
        KeyCand best = keycand_none();
        double w0 = 0.0, w1 = 0.0;
        int good = 1;
        for (int base = 0; base < G; base += 32) {
            KeyCand x = keycand_none();
            double b0 = 0.0, b1 = 0.0;
            const int s = base + lane;
            if (s < G) {
                if (xslot_wait(slots + s, seq)) {
                    x.kp = __ldcg(&slots[s].kp); x.kq = __ldcg(&slots[s].kq); x.idx = __ldcg(&slots[s].idx);
                    b0 = __ldcg(&slots[s].v0); b1 = __ldcg(&slots[s].v1);
                } else good = 0;
            }
            const KeyCand w = warp_key_best<TWO>(x);
            const unsigned who = __ballot_sync(FULL, x.idx >= 0 && x.idx == w.idx);
            const int sl = who ? __ffs(who) - 1 : 0;
            const double c0 = __shfl_sync(FULL, b0, sl), c1 = __shfl_sync(FULL, b1, sl);
            if (keycand_better(w, best)) { best = w; w0 = c0; w1 = c1; }
        }
        good = __all_sync(FULL, good);
        if (lane == 0) { sh.res = best; sh.res_v0 = w0; sh.res_v1 = w1; sh.res_ok = good; }
    }
    __syncthreads();
    const KeyCand r = sh.res;
    v0 = sh.res_v0; v1 = sh.res_v1;
    ok = sh.res_ok != 0;
    __syncthreads();       // sh.res may be rewritten by the next exchange
    return r;
}

template <int RR, int BL>          // BL: pricing rule (PR_DANTZIG / PR_BLAND / PR_DEVEX, block_common.cuh)
__global__ void __launch_bounds__(STEPS_TH, 1) blk_steps_kernel(DenseParams P, BlockBufs B, StepsArgs A) {
    extern __shared__ __align__(128) unsigned char steps_smem[];
    __shared__ StepsShared sh;
    __shared__ double s_pk[2 * BLK_KMAX], s_fr[2 * BLK_KMAX];
    __shared__ __align__(8) uint64_t bar;
    constexpr int TH = STEPS_TH;
    const int tid = (int)threadIdx.x, cta = (int)blockIdx.x;

    PivotState *st = P.st;
    // every CTA reads the same state; it is rewritten (by CTA 0) only after the last exchange of the launch
    int status = st->status, iters = st->iters, budget = st->budget, phase = st->phase, p1 = st->phase1_iters;
    int t = A.first ? 0 : st->blk_t;
    const int seq = st->blk_seq + (A.first ? 1 : 0);
    int k = st->k_next;
    const bool pend = B.pend != 0;
    const int Kp = pend ? B.info[B.par ^ 1].kc : 0;
    const int seqp = pend ? B.info[B.par ^ 1].seq : -1;
    const int offP = (B.par ^ 1) * BLK_KMAX, offC = B.par * BLK_KMAX - Kp;       // combined step c -> row of Fall / Pall
    auto srow = [&](int cc) -> size_t { return (size_t)((cc < Kp ? offP : offC) + cc); };
    const uint64_t pol = l2_policy(B.l2keep ? 2 : 0);
    const size_t pcn = P.pcn;

    // ---- what this CTA owns
    const int row0 = cta * A.rpc;
    const int nr = max(0, min(A.rpc, P.m - row0));                     // rows row0 .. row0 + nr - 1
    const long col0 = (long)cta * A.cpc;
    const int nc = (int)max(0L, min((long)A.cpc, (long)P.ld - col0));  // tableau columns col0 .. col0 + nc - 1 (padding columns included)
    const int S = min(Kp, A.smem_steps);                               // pending steps held in shared memory
    double *sFp = reinterpret_cast<double *>(steps_smem);              // [S][rpc]
    double *sPp = sFp + (size_t)A.smem_steps * A.rpc;                  // [S][cpc]

    if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    __syncthreads();
    const bool active0 = (status == LPGPU_RUNNING && phase != PH_DRIVEOUT);
    if (active0 && S > 0 && tid == 0) {
        const uint32_t nrb = (uint32_t)(((nr + 1) & ~1) * 8), ncb = (uint32_t)(nc * 8);
        mbar_arrive_expect_tx(&bar, (uint32_t)S * (nrb + ncb));
        for (int c = 0; c < S; ++c) {
            if (nrb) tma_load_1d(sFp + (size_t)c * A.rpc, B.Fall + srow(c) * pcn + row0, nrb, &bar);
            if (ncb) tma_load_1d(sPp + (size_t)c * A.cpc, B.Pall + srow(c) * P.ld + col0, ncb, &bar);
        }
    }
    bool staged = !(active0 && S > 0);          // the pending slices have arrived (waited for once, before their first use)

    const bool bigm = (P.method == LPGPU_BIGM);
    double *oc = B.objcur, *om = B.objcur + P.ld;
    unsigned xs = A.xbase;
    int applied = 0, last_r = -1, last_k = 0;
    bool ok = true;

    if (active0) {
        if (k <= 0) {
            // ---- no priced column handed over (first launch after an upload / a drive-out / a per-step kernel): price the objective row(s)
            const int pmode = bigm ? PRICE_BIGM : (phase == PH_ONE ? PRICE_PHASE1 : PRICE_COST);
            const int jmax = (P.method == LPGPU_TWOPHASE && phase == PH_MAIN) ? P.n - P.n_art : P.n;
            KeyCand c = keycand_none();
            for (int lj = tid; lj < nc; lj += TH) {
                const long j = col0 + lj;
                if (j >= 1 && j <= jmax) blk_price<BL>(P, pmode, (int)j, __ldcg(oc + j), (RR == 2) ? __ldcg(om + j) : 0.0, (BL == PR_DEVEX) ? __ldcg(P.wts + j) : 1.0, c);
            }
            double d0 = 0.0, d1 = 0.0;
            c = bigm ? block_best_payload<true>(c, d0, d1, sh, 1) : block_best_payload<false>(c, d0, d1, sh, 1);
            c = bigm ? grid_best<true>(A, xs++, c, d0, d1, sh, ok) : grid_best<false>(A, xs++, c, d0, d1, sh, ok);
            k = c.idx > 0 ? c.idx : 0;
        }
        while (ok) {
            const int pmode = bigm ? PRICE_BIGM : (phase == PH_ONE ? PRICE_PHASE1 : PRICE_COST);
            const int jmax = (P.method == LPGPU_TWOPHASE && phase == PH_MAIN) ? P.n - P.n_art : P.n;
            const int tc = Kp + t;
            if (k == 0) {
                // nothing prices in: optimal -- unless Big-M or phase one left an artificial basic at a positive level
                // (infeasible), or phase one is over (drive-out next: the host loop hands over to the rank-1 path)
                KeyCand a = keycand_none();
                if (bigm || phase == PH_ONE) {
                    const int nreal = P.n - P.n_art;
                    for (int li = tid; li < nr; li += TH) {
                        const int i = row0 + li;
                        if (__ldcg(P.basis + i) > nreal && __ldcg(B.bcur + i) > P.eps) { a.idx = i; break; }
                    }
                }
                double d0 = 0.0, d1 = 0.0;
                a = block_best_payload<false>(a, d0, d1, sh, 0);
                a = grid_best<false>(A, xs++, a, d0, d1, sh, ok);
                if (!ok) break;
                if (a.idx >= 0) status = LPGPU_INFEASIBLE;
                else if (phase == PH_ONE) { phase = PH_DRIVEOUT; p1 = iters; }
                else status = LPGPU_OPTIMAL;
                break;
            }
            // ================= column k as it stands after the tc pending steps, and the ratio test on it
            const double sig_c = __ldcg(oc + k);                       // read before the row phase rewrites the objective rows
            const double sig_m = (RR == 2) ? __ldcg(om + k) : 0.0;
            const double wk = (BL == PR_DEVEX) ? __ldcg(P.wts + k) : 1.0;   // Devex: w_k before this pivot (rewritten in the row phase, after the exchange)
            if (tid < tc) s_pk[tid] = ld_hint(B.Pall + srow(tid) * P.ld + (size_t)k, pol);
            if (cta == 0 && tid < RR) B.Fs[(size_t)t * pcn + P.m + tid] = __ldcg(oc + (size_t)tid * P.ld + k);   // objective rows: kept eagerly, same bits
            KeyCand cr = keycand_none();
            double cr_a = 0.0, cr_b = 0.0;
            for (int li0 = 0; li0 < nr; li0 += TH) {                   // one trip unless the CTA owns more than TH rows
                const int li = li0 + tid, i = row0 + li;
                const bool mine = li < nr;
                double tv = 0.0, bv = 0.0, fcur[BLK_KMAX];
                int rs = 0, rsp = 0;
                if (mine) {
                    tv = P.T[(size_t)i * P.ld + (size_t)k];
                    rs = __ldcg(B.rowstamp + i);
                    bv = __ldcg(B.bcur + i);
                    if (Kp) rsp = B.stamp_prev[i];
#pragma unroll
                    for (int u = 0; u < BLK_KMAX; ++u) fcur[u] = (u < t) ? ld_hint(B.Fs + (size_t)u * pcn + i, pol) : 0.0;
                }
                if (li0 == 0) {
                    __syncthreads();                                    // s_pk is complete
                    if (!staged) { mbar_wait(&bar, 0); staged = true; }
                }
                if (mine) {
                    // where this row's chain starts: after its last turn as a pivot row (this block, else the pending one), else at the tableau
                    const int s0 = blk_row_start(rs, seq);
                    const int c0 = s0 ? Kp + s0 : (Kp ? blk_row_start(rsp, seqp) : 0);
                    double a = c0 ? s_pk[c0 - 1] : tv;
                    for (int cc = c0; cc < Kp; ++cc) {
                        const double f = (cc < S) ? sFp[(size_t)cc * A.rpc + li] : ld_hint(B.Fall + srow(cc) * pcn + i, pol);
                        a = __fma_rn(-f, s_pk[cc], a);
                    }
                    const int u0 = max(c0 - Kp, 0);
#pragma unroll
                    for (int u = 0; u < BLK_KMAX; ++u)
                        if (u >= u0 && u < t) a = __fma_rn(-fcur[u], s_pk[Kp + u], a);
                    st_hint(B.Fs + (size_t)t * pcn + i, a, pol);
                    if (a > P.eps) {
                        KeyCand x;
                        x.kp = 0;
                        x.kq = ord_key(-(bv / a));                     // arg max -theta == arg min theta
                        x.idx = i;
                        if (keycand_better(x, cr)) { cr = x; cr_a = a; cr_b = bv; }
                    }
                }
            }
            if (nr <= 0) {                                             // a CTA without rows still takes part in the barriers above
                __syncthreads();
                if (!staged) { mbar_wait(&bar, 0); staged = true; }
            }
            cr = block_best_payload<false>(cr, cr_a, cr_b, sh, 0);
            cr = grid_best<false>(A, xs++, cr, cr_a, cr_b, sh, ok);
            if (!ok) break;
            if (cr.idx < 0) { status = LPGPU_UNBOUNDED; break; }
            if (iters >= P.max_iter) { status = LPGPU_ITER_LIMIT; break; }
            if (applied >= A.nsteps || budget <= 0) break;             // decision only: the priced column stays valid
            const int r = cr.idx;
            const double piv = cr_a, b_r = cr_b;

            // ================= row r as it stands after the tc pending steps, scaled; objective rows; next pricing; b column
            if (tid < tc) s_fr[tid] = ld_hint(B.Fall + srow(tid) * pcn + r, pol);
            const int s0r = blk_row_start(__ldcg(B.rowstamp + r), seq);
            const int c0r = s0r ? Kp + s0r : (Kp ? blk_row_start(B.stamp_prev[r], seqp) : 0);
            const double brp = div_skip_zero(b_r, piv);                // = P[t][0]: column 0 of the row is the b column
            double *pt = B.Pm + (size_t)t * P.ld;
            KeyCand cp = keycand_none();
            for (int lj0 = 0; lj0 < nc; lj0 += TH) {
                const int lj = lj0 + tid;
                const long j = col0 + lj;
                const bool mine = lj < nc && j >= 1 && j <= (long)P.n;
                double av = 0.0, ocv = 0.0, omv = 0.0, wv = 1.0, pcur[BLK_KMAX];
                if (mine) {
                    av = P.T[(size_t)r * P.ld + (size_t)j];
                    ocv = __ldcg(oc + j);
                    if (RR == 2) omv = __ldcg(om + j);
                    if (BL == PR_DEVEX) wv = __ldcg(P.wts + j);
#pragma unroll
                    for (int u = 0; u < BLK_KMAX; ++u) pcur[u] = (u < t) ? ld_hint(B.Pm + (size_t)u * P.ld + (size_t)j, pol) : 0.0;
                }
                if (lj0 == 0) __syncthreads();                          // s_fr is complete
                if (mine) {
                    double a = av;
                    // a row that was a pivot row (in this block or the pending one) restarts from its scaled row P[c0r - 1]
                    for (int cc = max(c0r - 1, 0); cc < Kp; ++cc) {
                        const double pv = (cc < S) ? sPp[(size_t)cc * A.cpc + lj] : ld_hint(B.Pall + srow(cc) * P.ld + (size_t)j, pol);
                        a = (cc == c0r - 1) ? pv : __fma_rn(-s_fr[cc], pv, a);
                    }
#pragma unroll
                    for (int u = 0; u < BLK_KMAX; ++u)
                        if (u < t) {
                            const int cc = Kp + u;
                            if (cc == c0r - 1) a = pcur[u];
                            else if (cc >= c0r) a = __fma_rn(-s_fr[cc], pcur[u], a);
                        }
                    const double pr = div_skip_zero(a, piv);
                    st_hint(pt + j, pr, pol);
                    const double qc = __fma_rn(-sig_c, pr, ocv);
                    __stcg(B.objcur + j, qc);
                    double qm = 0.0;
                    if (RR == 2) {
                        qm = __fma_rn(-sig_m, pr, omv);
                        __stcg(B.objcur + P.ld + j, qm);
                    }
                    double w = 1.0;
                    if (BL == PR_DEVEX) {
                        w = devex_weight(wv, pr, wk, j == (long)k);
                        __stcg(P.wts + j, w);
                    }
                    if (j <= (long)jmax) blk_price<BL>(P, pmode, (int)j, qc, qm, w, cp);
                }
            }
            if (nc <= 0) __syncthreads();
            if (cta == 0 && tid == 0) {                                // column 0 and the bookkeeping of this pivot
                pt[0] = brp;
                __stcg(B.objcur, __fma_rn(-sig_c, brp, __ldcg(oc)));
                if (RR == 2) __stcg(B.objcur + P.ld, __fma_rn(-sig_m, brp, __ldcg(om)));
                P.basis[r] = k;
                __stcg(B.rowstamp + r, (seq << 8) | (t + 1));
                B.steprow[t] = r;
                if (iters < P.log_cap) { P.log[2 * iters] = k; P.log[2 * iters + 1] = r; }
            }
            // ---- b column of this CTA's rows (the multipliers were stored above by the same threads)
            for (int li = tid; li < nr; li += TH) {
                const int i = row0 + li;
                const double f = __ldcg(B.Fs + (size_t)t * pcn + i), bv = __ldcg(B.bcur + i);
                __stcg(B.bcur + i, (i == r) ? brp : __fma_rn(-f, brp, bv));
            }
            double d0 = 0.0, d1 = 0.0;
            cp = bigm ? block_best_payload<true>(cp, d0, d1, sh, 1) : block_best_payload<false>(cp, d0, d1, sh, 1);
            cp = bigm ? grid_best<true>(A, xs++, cp, d0, d1, sh, ok) : grid_best<false>(A, xs++, cp, d0, d1, sh, ok);
            if (!ok) break;
            last_r = r; last_k = k;
            iters += 1; budget -= 1; applied += 1; t += 1;
            if (phase == PH_ONE) p1 = iters;
            k = cp.idx > 0 ? cp.idx : 0;
            if (applied >= A.nsteps) break;                            // the next launch (or the closing decision) goes on from k
        }
    }
    if (!ok) status = LPGPU_STEP_TIMEOUT;
    if ((cta == 0 || !ok) && tid == 0) {
        st->status = status;
        st->phase = phase;
        st->phase1_iters = p1;
        st->iters = iters;
        st->budget = budget;
        st->blk_t = t;
        st->blk_seq = seq;
        st->do_pivot = applied > 0 ? 1 : 0;
        if (applied > 0) { st->r = last_r; st->k = last_k; }
        st->k_next = (status == LPGPU_RUNNING && phase != PH_DRIVEOUT && k > 0) ? k : 0;
        if (B.info) { B.info[B.par].kc = t; B.info[B.par].seq = seq; }
    }
}

// which SMs the persistent grid lands on while the device is idle (same CTA size and shared memory): the pass leaves them free
__global__ void __launch_bounds__(STEPS_TH, 1) blk_steps_probe_kernel(unsigned *smids) {
    extern __shared__ __align__(128) unsigned char steps_smem[];
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (threadIdx.x == 0) { steps_smem[0] = 0; smids[blockIdx.x] = smid; }
}

}  // namespace lpgpu
