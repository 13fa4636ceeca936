/*
 * lp_oracle.h -- CPU ORACLE for the dense-tableau simplex pivot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, the smoke check in
 * __graft_entry__.py and the cpu_baseline / --impl reference legs of bench.py may
 * link, load or call anything in oracle/.  The product (liblpgpu.so and
 * linearprogramming_b200/) never routes through it and has no CPU fallback.
 *
 * PARITY STATUS: "parity unpinned" for the pivot loop itself.  The reference
 * (SomeBottle/LinearProgramming) builds the tableau (Source/matrix.c:19-91) and frees
 * it again (Source/simplex.c:40,67); the Big-M / two-phase switch is two empty cases
 * (Source/simplex.c:57-63) and AddBigM is an empty body (Source/simplex.c:79-82).  There
 * is no pricing, ratio test, elimination, status or iteration count to compare against.
 * What IS pinned by running the reference's own code (oracle/_ref, tests/golden):
 *   - the exact-rational arithmetic restated in lpo_rat_* (Source/numOprts.c:37-129,
 *     137-294 for the non-M case) against the known-answer vectors of SURVEY.md App. A;
 *   - the initial tableau (layout, column order, row negation, slack naming, basis) that
 *     Source/simplex.c:91-260 + Source/matrix.c:19-91 produce, against App. B dumps.
 * The pivot rule below ("Appendix D rule") is defined by THIS repository and stated in
 * DESIGN.md; the GPU path must reproduce it bit for bit.
 *
 * Tableau layout (follows SimplexMatrix, Source/matrix.h:13-21: cMatrix[i][0] = b_i,
 * cMatrix[i][j+1] = a_ij), row-major with leading dimension ld >= n+1:
 *   rows 0..m-1     constraint rows   [ b_i | a_i1 .. a_in ]
 *   row  m          objective row     [ -z  | sigma_1 .. sigma_n ]   sigma_j = c_j - z_j
 *   row  m+1 (R=2)  M-row / phase-1   [ -zM | sigmaM_1 .. sigmaM_n ]
 * The problem is always "max" (Source/simplex.c:99-106), every b_i >= 0
 * (Source/simplex.c:178-186) and the tableau is canonical: basis[i] (1-based column) is a
 * unit column e_i and has zero entries in the objective rows.
 */
#ifndef LP_ORACLE_H
#define LP_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { LPO_OPTIMAL = 0, LPO_UNBOUNDED = 1, LPO_INFEASIBLE = 2, LPO_ITER_LIMIT = 3,
       LPO_RUNNING = 4, LPO_OVERFLOW = 5 };
enum { LPO_PLAIN = 0, LPO_BIGM = 1, LPO_TWOPHASE = 2, LPO_DUAL = 3 };
/* phases of the state machine */
enum { LPO_PH_MAIN = 0, LPO_PH_ONE = 1, LPO_PH_DRIVEOUT = 2 };

typedef struct {
    double eps;      /* 1e-9: pricing / ratio / snap tolerance                      */
    int max_iter;    /* iteration cap (-> LPO_ITER_LIMIT)                            */
    int method;      /* LPO_PLAIN (R=1) | LPO_BIGM (R=2) | LPO_TWOPHASE (R=2) | LPO_DUAL (R=1) */
    int n_art;       /* artificial columns = the LAST n_art columns of the tableau   */
    int threads;     /* OpenMP threads for the elimination (results do not depend)   */
    int pricing;     /* 0 = largest reduced cost, ties -> lowest column (Dantzig);
                      * 1 = Bland: the LOWEST eligible column (with the lowest-row ratio tie-break: no cycling).
                      * 2 = Devex: largest sigma_j^2 / w_j over the eligible columns, ties -> lowest column; reference
                      *     weights w_j as devex_update() in lp_oracle.c maintains them.
                      * fp64 primal methods only (the reference prices nothing)  */
} lpo_opts;

typedef struct {
    int status;
    int iters;         /* pivots performed (incl. phase 1 and drive-out pivots)      */
    int phase1_iters;  /* pivots performed before the phase switch (two-phase)       */
    int phase;         /* state-machine phase on return                              */
    double z;          /* objective value  = -T[m][0]                                 */
    double zM;         /* M-part / phase-1 = -T[m+1][0] (0 when R = 1)                */
} lpo_result;

/* ---- fp64 path ---------------------------------------------------------------- */

/* Run the pivot loop on a canonical tableau in place.  pivot_log receives
 * (entering column j [1-based], leaving row r [0-based]) pairs, at most log_cap of them.
 * Returns 0, or -1 on bad arguments. */
int lpo_solve_f64(int m, int n, int R, double *T, size_t ld, int *basis,
                  const lpo_opts *opts, int *pivot_log, int log_cap, lpo_result *res);

/* One rank-1 pivot on (r, k): pivot-row scaling + elimination of every other row
 * (incl. objective rows).  Exposed so tests can compare a single update pass. */
void lpo_pivot_f64(int rows, int n, double *T, size_t ld, int r, int k, int threads);

/* x[0..n-1] from basis and b column (x[j-1] = b_i for basis[i] = j, else 0). */
void lpo_extract_x(int m, int n, const double *T, size_t ld, const int *basis, double *x);

/* Batched: n_lp tableaux packed back to back, each (m+R) x (n+1) with ld = n+1. */
int lpo_solve_batched_f64(long n_lp, int m, int n, int R, double *T_all, int *basis_all,
                          const lpo_opts *opts, int *status, int *iters, double *z,
                          double *x /* n_lp*n or NULL */, int *pivot_log, int log_cap);

/* ---- tableau construction (host logic restated; compared with the product's) ---- */

/* rel[i]: -2 "<=", 2 ">=", 3 "=" (codes of Source/public.h:57).  Builds a canonical
 * tableau for  max c.x  s.t.  A x rel b,  x >= 0,  with b >= 0 enforced by negating rows
 * (Source/simplex.c:178-186).  Column order: structural 1..ns, then one slack/surplus
 * column per inequality row in row order (Source/simplex.c:189-219), then one artificial
 * column per row that lacks a unit column (">=" and "=" rows), in row order.
 * Returns n (total columns) and fills *n_art; T must hold (m+R) * ld doubles with
 * ld >= n+1 where n <= ns + 2m.  R = 1 requires n_art == 0 on return (else -1). */
int lpo_build_tableau(int m, int ns, const double *A, size_t lda, const double *b,
                      const double *c, const int *rel, int R, double *T, size_t ld,
                      int *basis, int *n_art);
int lpo_count_columns(int m, int ns, const double *b, const int *rel, int *n_art);

/* ---- synthetic workloads ------------------------ */

uint64_t lpo_mix64(uint64_t z);
/* uniform in [lo, hi): fma(u, hi-lo, lo), u = (hash >> 11) * 2^-53 */
double lpo_uniform(uint64_t seed, uint64_t lp, uint64_t elem, double lo, double hi);
/* max c.x, A x <= b, A ~ U(0.1,1.1), b ~ U(1,2)*ns, c ~ U(0.1,1.1); n = ns + m columns,
 * slack basis.  Writes an (m+1) x (n+1) canonical tableau (R = 1). */
void lpo_synth_feasible(uint64_t seed, uint64_t lp, int m, int ns, double *T, size_t ld,
                        int *basis);

/* ---- exact rational path (numOprts.c semantics, no symbolic M) -------------------- */

typedef struct { long num; long den; int valid; } lpo_rat;

lpo_rat lpo_rat_make(long num, long den);
lpo_rat lpo_rat_add(lpo_rat a, lpo_rat b);  /* Source/numOprts.c:78-129  FractionAdd */
lpo_rat lpo_rat_sub(lpo_rat a, lpo_rat b);  /* Source/numOprts.c:204-208 NSub        */
lpo_rat lpo_rat_mul(lpo_rat a, lpo_rat b);  /* Source/numOprts.c:37-67   FractionMul */
lpo_rat lpo_rat_div(lpo_rat a, lpo_rat b);  /* Source/numOprts.c:267-283 NDiv        */
lpo_rat lpo_rat_neg(lpo_rat a);             /* Source/numOprts.c:290-294 NInv        */
double  lpo_rat_to_double(lpo_rat a);       /* Source/basicFuncs.c:298-313 Decimalize */
long    lpo_gcd(long a, long b);            /* Source/basicFuncs.c:123-138 */
long    lpo_lcm(long a, long b);            /* Source/basicFuncs.c:146-158 (-1 = overflow) */
int     lpo_rat_cmp(lpo_rat a, lpo_rat b);  /* exact, via 128-bit cross products */
/* string -> rational, the non-M subset of Source/basicFuncs.c:165-291 Fractionize
 * ("3/4", "-2.45", "5"), including its decimal truncation quirk. */
lpo_rat lpo_rat_parse(const char *s);

/* Same rule as lpo_solve_f64 with eps = 0, on rationals.  T is (m+R) x ld lpo_rat. */
int lpo_solve_rat(int m, int n, int R, lpo_rat *T, size_t ld, int *basis,
                  const lpo_opts *opts, int *pivot_log, int log_cap, lpo_result *res);

#ifdef __cplusplus
}
#endif
#endif
