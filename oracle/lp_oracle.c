/*
 * lp_oracle.c -- CPU ORACLE (test infrastructure; see lp_oracle.h for the rules on who
 * may call it and for the "parity unpinned" statement).
 *
 * Build: gcc -O2 -std=c11 -fopenmp -ffp-contract=off -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off matters: every multiply-add below that must round once is written as
 * an explicit fma(), everything else must round per operation, exactly as the CUDA path
 * (__fma_rn / __ddiv_rn / plain compares) does.
 */
#include "lp_oracle.h"

#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================================= fp64 path */

typedef struct {
    int m, n, R;
    double *T;
    size_t ld;
    int *basis;
    double eps;
    int method, n_art, threads, pricing;
    double *w;       /* pricing == 2: Devex reference weights, w[1..n] (>= 1 always) */
} lpo_ctx;

#define AT(c, i, j) ((c)->T[(size_t)(i) * (c)->ld + (size_t)(j)])

/* Entering column (Appendix D): largest reduced cost > eps, ties -> lowest column.
 * The reference has no pricing code (Source/simplex.c:57-63 are empty cases); the
 * candidate range follows its column order (Source/matrix.c:50-58). Returns 0 if none. */
static int price_single(const lpo_ctx *c, int row, int jmax) {
    int best = 0;
    double bv = 0.0;
    for (int j = 1; j <= jmax; ++j) {
        double v = AT(c, row, j);
        if (c->pricing == 2) {                   /* Devex: largest sigma_j^2 / w_j (one multiply, one division) */
            if (v > c->eps) {
                const double sc = (v * v) / c->w[j];
                if (best == 0 || sc > bv) { best = j; bv = sc; }
            }
            continue;
        }
        if (v > c->eps && (best == 0 || v > bv)) {
            best = j; bv = v;
            if (c->pricing == 1) break;          /* Bland: the lowest eligible column */
        }
    }
    return best;
}

/* Big-M pricing: (sigmaM, sigma) compared lexicographically, |sigmaM| <= eps snapped to 0.
 * Stands in for the symbolic aM+b comparison the reference never defined
 * (Source/numOprts.h:15-37 has the representation, no ordering). */
static int price_bigm(const lpo_ctx *c) {
    int best = 0;
    double bp = 0.0, bq = 0.0;
    for (int j = 1; j <= c->n; ++j) {
        double p = AT(c, c->m + 1, j), q = AT(c, c->m, j);
        if (fabs(p) <= c->eps) p = 0.0;
        int eligible = (p > 0.0) || (p == 0.0 && q > c->eps);
        if (!eligible) continue;
        if (c->pricing == 2) {                   /* Devex: both levels scaled by the weight; below an M part the cost breaks ties as it is */
            q = (p == 0.0) ? (q * q) / c->w[j] : q;
            p = (p * p) / c->w[j];
        }
        if (best == 0 || p > bp || (p == bp && q > bq)) { best = j; bp = p; bq = q; }
        if (c->pricing == 1) break;              /* Bland: the lowest eligible column */
    }
    return best;
}

/* Leaving row: min b_i / a_ik over a_ik > eps, ties -> lowest row; -1 = unbounded. */
static int ratio_test(const lpo_ctx *c, int k) {
    int best = -1;
    double bt = 0.0;
    for (int i = 0; i < c->m; ++i) {
        double a = AT(c, i, k);
        if (a > c->eps) {
            double t = AT(c, i, 0) / a;
            if (best < 0 || t < bt) { best = i; bt = t; }
        }
    }
    return best;
}

/* Dual simplex (the reference only has the menu stub, Source/router.c:32-34, and the dual = 1
 * branch of LPStandardize, Source/simplex.c:177-178).  Rule defined here: leaving row = most negative
 * b_i < -eps, ties -> lowest row; entering column = min sigma_j / a_rj over a_rj < -eps, ties ->
 * lowest column (one IEEE division each); no such column -> the row proves infeasibility. */
static int dual_leave(const lpo_ctx *c) {
    int best = -1;
    double bv = 0.0;
    for (int i = 0; i < c->m; ++i) {
        double b = AT(c, i, 0);
        if (b < -c->eps && (best < 0 || b < bv)) { best = i; bv = b; }
    }
    return best;
}

static int dual_enter(const lpo_ctx *c, int r) {
    int best = 0;
    double bt = 0.0;
    for (int j = 1; j <= c->n; ++j) {
        double a = AT(c, r, j);
        if (a < -c->eps) {
            double t = AT(c, c->m, j) / a;
            if (best == 0 || t < bt) { best = j; bt = t; }
        }
    }
    return best;
}

void lpo_pivot_f64(int rows, int n, double *T, size_t ld, int r, int k, int threads) {
    double *prow = (double *)malloc(sizeof(double) * (size_t)(n + 1));
    double *pcol = (double *)malloc(sizeof(double) * (size_t)rows);
    const double piv = T[(size_t)r * ld + (size_t)k];
    for (int j = 0; j <= n; ++j) prow[j] = T[(size_t)r * ld + (size_t)j] / piv;
    for (int i = 0; i < rows; ++i) pcol[i] = T[(size_t)i * ld + (size_t)k];
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static) if (threads > 1)
    for (int i = 0; i < rows; ++i) {
        double *row = T + (size_t)i * ld;
        if (i == r) {
            for (int j = 0; j <= n; ++j) row[j] = prow[j];
        } else {
            const double f = -pcol[i];
            for (int j = 0; j <= n; ++j) row[j] = fma(f, prow[j], row[j]);
        }
    }
    free(prow);
    free(pcol);
}

/* Devex reference-framework weights (Forrest & Goldfarb; pricing == 2).  All weights start at 1.  After the pivot on
 * (r, k), with a_rj' the scaled pivot row:  w_j <- max(w_j, (a_rj' * a_rj') * w_k)  for every column j != k (two rounded
 * multiplies, in this order), then  w_k <- 1  (the entering column joins the basis; when it leaves again the rule above gives
 * it max(1, w_q / a_rq^2), Devex's weight of the leaving variable).  Every pivot of a primal method updates them, the
 * drive-out pivots between the phases too; the weights are NOT reset at the phase switch. */
static void devex_update(lpo_ctx *c, int r, int k) {
    const double wk = c->w[k];
    for (int j = 1; j <= c->n; ++j) {
        if (j == k) continue;
        const double pr = AT(c, r, j);
        const double t = pr * pr;
        const double u = t * wk;
        if (u > c->w[j]) c->w[j] = u;
    }
    c->w[k] = 1.0;
}

static int art_basic_positive(const lpo_ctx *c) {
    for (int i = 0; i < c->m; ++i)
        if (c->basis[i] > c->n - c->n_art && AT(c, i, 0) > c->eps) return 1;
    return 0;
}

/* Drive-out step at the phase switch: lowest row whose basic variable is artificial and
 * that has a non-artificial entry with |a_ij| > eps; lowest such column. */
static int driveout_find(const lpo_ctx *c, int *r, int *k) {
    const int nreal = c->n - c->n_art;
    for (int i = 0; i < c->m; ++i) {
        if (c->basis[i] <= nreal) continue;
        for (int j = 1; j <= nreal; ++j)
            if (fabs(AT(c, i, j)) > c->eps) { *r = i; *k = j; return 1; }
    }
    return 0;
}

int lpo_solve_f64(int m, int n, int R, double *T, size_t ld, int *basis,
                  const lpo_opts *opts, int *pivot_log, int log_cap, lpo_result *res) {
    if (m < 1 || n < 1 || (R != 1 && R != 2) || !T || !basis || !opts || !res) return -1;
    if (ld < (size_t)n + 1) return -1;
    if ((opts->method == LPO_PLAIN || opts->method == LPO_DUAL) != (R == 1)) return -1;
    if (opts->n_art < 0 || opts->n_art > n) return -1;
    if (opts->pricing != 0 && ((opts->pricing != 1 && opts->pricing != 2) || opts->method == LPO_DUAL)) return -1;
    lpo_ctx c = { m, n, R, T, ld, basis, opts->eps, opts->method, opts->n_art,
                  opts->threads > 0 ? opts->threads : 1, opts->pricing, NULL };
    if (c.pricing == 2) {
        c.w = (double *)malloc(sizeof(double) * (size_t)(n + 1));
        if (!c.w) return -1;
        for (int j = 0; j <= n; ++j) c.w[j] = 1.0;
    }
    int phase = (opts->method == LPO_TWOPHASE) ? LPO_PH_ONE : LPO_PH_MAIN;
    int status = LPO_RUNNING, iters = 0, p1 = 0;
    while (status == LPO_RUNNING) {
        int r = -1, k = 0;
        if (c.method == LPO_DUAL) {
            r = dual_leave(&c);
            if (r < 0) { status = LPO_OPTIMAL; break; }
            k = dual_enter(&c, r);
            if (k == 0) { status = LPO_INFEASIBLE; break; }
        } else if (phase == LPO_PH_DRIVEOUT) {
            if (!driveout_find(&c, &r, &k)) { phase = LPO_PH_MAIN; continue; }
        } else {
            if (c.method == LPO_BIGM) k = price_bigm(&c);
            else if (phase == LPO_PH_ONE) k = price_single(&c, m + 1, n);
            else k = price_single(&c, m, n - (c.method == LPO_TWOPHASE ? c.n_art : 0));
            if (k == 0) {
                if (phase == LPO_PH_ONE) {
                    if (art_basic_positive(&c)) { status = LPO_INFEASIBLE; break; }
                    phase = LPO_PH_DRIVEOUT;
                    p1 = iters;
                    continue;
                }
                if (c.method == LPO_BIGM && art_basic_positive(&c)) status = LPO_INFEASIBLE;
                else status = LPO_OPTIMAL;
                break;
            }
            r = ratio_test(&c, k);
            if (r < 0) { status = LPO_UNBOUNDED; break; }
        }
        if (iters >= opts->max_iter) { status = LPO_ITER_LIMIT; break; }
        lpo_pivot_f64(m + R, n, T, ld, r, k, c.threads);
        if (c.pricing == 2) devex_update(&c, r, k);
        basis[r] = k;
        if (pivot_log && iters < log_cap) { pivot_log[2 * iters] = k; pivot_log[2 * iters + 1] = r; }
        ++iters;
    }
    if (phase == LPO_PH_ONE) p1 = iters;
    free(c.w);
    res->status = status;
    res->iters = iters;
    res->phase1_iters = (opts->method == LPO_TWOPHASE) ? p1 : 0;
    res->phase = phase;
    res->z = -T[(size_t)m * ld];
    res->zM = (R == 2) ? -T[(size_t)(m + 1) * ld] : 0.0;
    return 0;
}

void lpo_extract_x(int m, int n, const double *T, size_t ld, const int *basis, double *x) {
    for (int j = 0; j < n; ++j) x[j] = 0.0;
    for (int i = 0; i < m; ++i)
        if (basis[i] >= 1 && basis[i] <= n) x[basis[i] - 1] = T[(size_t)i * ld];
}

int lpo_solve_batched_f64(long n_lp, int m, int n, int R, double *T_all, int *basis_all,
                          const lpo_opts *opts, int *status, int *iters, double *z,
                          double *x, int *pivot_log, int log_cap) {
    if (n_lp < 0 || !T_all || !basis_all || !opts) return -1;
    const size_t ld = (size_t)n + 1, tsz = (size_t)(m + R) * ld;
    int threads = opts->threads > 0 ? opts->threads : 1;
    lpo_opts inner = *opts;
    inner.threads = 1;
    int bad = 0;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16) if (threads > 1)
    for (long p = 0; p < n_lp; ++p) {
        lpo_result r;
        int rc = lpo_solve_f64(m, n, R, T_all + (size_t)p * tsz, ld, basis_all + (size_t)p * m,
                               &inner, pivot_log ? pivot_log + (size_t)p * 2 * log_cap : NULL,
                               log_cap, &r);
        if (rc != 0) {
#pragma omp atomic write
            bad = 1;
            continue;
        }
        if (status) status[p] = r.status;
        if (iters) iters[p] = r.iters;
        if (z) z[p] = r.z;
        if (x) lpo_extract_x(m, n, T_all + (size_t)p * tsz, ld, basis_all + (size_t)p * m,
                             x + (size_t)p * n);
    }
    return bad ? -1 : 0;
}

/* ============================================================ tableau construction */

int lpo_count_columns(int m, int ns, const double *b, const int *rel, int *n_art) {
    int n = ns, na = 0;
    for (int i = 0; i < m; ++i) {
        int rl = rel[i];
        if (b[i] < 0 && rl != 3) rl = -rl;
        if (rl == -2) n += 1;
        else if (rl == 2) { n += 1; na += 1; }
        else if (rl == 3) na += 1;
        else return -1;
    }
    if (n_art) *n_art = na;
    return n + na;
}

int lpo_build_tableau(int m, int ns, const double *A, size_t lda, const double *b,
                      const double *c, const int *rel, int R, double *T, size_t ld,
                      int *basis, int *n_art) {
    int na = 0;
    const int n = lpo_count_columns(m, ns, b, rel, &na);
    if (n < 0 || ld < (size_t)n + 1) return -1;
    if (R == 1 && na != 0) return -1;
    for (size_t e = 0; e < (size_t)(m + R) * ld; ++e) T[e] = 0.0;
    int next_slack = ns + 1;          /* slack / surplus columns in row order           */
    int next_art = n - na + 1;        /* artificial columns trail, in row order         */
    for (int i = 0; i < m; ++i) {
        double *row = T + (size_t)i * ld;
        int rl = rel[i];
        const int neg = b[i] < 0;     /* Source/simplex.c:178-186: make b_i >= 0        */
        row[0] = neg ? -b[i] : b[i];
        for (int j = 0; j < ns; ++j) {
            double a = A[(size_t)i * lda + (size_t)j];
            row[1 + j] = neg ? -a : a;
        }
        if (neg && rl != 3) rl = -rl;
        if (rl == -2) { row[next_slack] = 1.0; basis[i] = next_slack++; }
        else if (rl == 2) { row[next_slack++] = -1.0; row[next_art] = 1.0; basis[i] = next_art++; }
        else { row[next_art] = 1.0; basis[i] = next_art++; }
    }
    double *obj = T + (size_t)m * ld;
    for (int j = 0; j < ns; ++j) obj[1 + j] = c[j];
    if (R == 2) {
        /* cost -M on artificials: M-row starts at -1 on artificial columns and is priced
         * out by adding every artificial-basic row, rows ascending, one add per element. */
        double *om = T + (size_t)(m + 1) * ld;
        for (int j = n - na + 1; j <= n; ++j) om[j] = -1.0;
        for (int i = 0; i < m; ++i) {
            if (basis[i] <= n - na) continue;
            const double *row = T + (size_t)i * ld;
            for (int j = 0; j <= n; ++j) om[j] = om[j] + row[j];
        }
    }
    if (n_art) *n_art = na;
    return n;
}

/* ============================================================== synthetic workloads */

uint64_t lpo_mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

double lpo_uniform(uint64_t seed, uint64_t lp, uint64_t elem, double lo, double hi) {
    uint64_t h = lpo_mix64(lpo_mix64(lpo_mix64(seed) ^ lp) ^ elem);
    double u = (double)(h >> 11) * 0x1.0p-53;
    return fma(u, hi - lo, lo);
}

void lpo_synth_feasible(uint64_t seed, uint64_t lp, int m, int ns, double *T, size_t ld,
                        int *basis) {
    const int n = ns + m;
    const uint64_t nA = (uint64_t)m * (uint64_t)ns;
    for (int i = 0; i <= m; ++i) {
        double *row = T + (size_t)i * ld;
        for (int j = 0; j <= n; ++j) row[j] = 0.0;
    }
    for (int i = 0; i < m; ++i) {
        double *row = T + (size_t)i * ld;
        row[0] = lpo_uniform(seed, lp, nA + (uint64_t)i, 1.0, 2.0) * (double)ns;
        for (int j = 0; j < ns; ++j)
            row[1 + j] = lpo_uniform(seed, lp, (uint64_t)i * (uint64_t)ns + (uint64_t)j, 0.1, 1.1);
        row[1 + ns + i] = 1.0;
        basis[i] = 1 + ns + i;
    }
    double *obj = T + (size_t)m * ld;
    for (int j = 0; j < ns; ++j)
        obj[1 + j] = lpo_uniform(seed, lp, nA + (uint64_t)m + (uint64_t)j, 0.1, 1.1);
}

/* ============================================================== exact rational path */

static lpo_rat rat_invalid(void) { lpo_rat r = { 0, 0, 0 }; return r; }

lpo_rat lpo_rat_make(long num, long den) { lpo_rat r = { num, den, den != 0 }; return r; }

/* Euclid on absolute values, result >= 0 (Source/basicFuncs.c:123-138). */
long lpo_gcd(long a, long b) {
    a = labs(a);
    b = labs(b);
    while (b) { long t = a % b; a = b; b = t; }
    return a;
}

/* |a| / gcd * |b|, -1 when the product overflows (Source/basicFuncs.c:146-158). */
long lpo_lcm(long a, long b) {
    long g = lpo_gcd(a, b), out;
    a = labs(a);
    b = labs(b);
    if (g == 0) return 0;
    if (__builtin_mul_overflow(a / g, b, &out)) return -1;
    return out;
}

/* Fraction product with cross-reduction BEFORE multiplying and no sign normalisation of
 * the denominator (Source/numOprts.c:37-67); overflow or a zero denominator -> invalid. */
lpo_rat lpo_rat_mul(lpo_rat a, lpo_rat b) {
    if (!a.valid || !b.valid || a.den == 0 || b.den == 0) return rat_invalid();
    long g1 = lpo_gcd(a.num, b.den), g2 = lpo_gcd(a.den, b.num);
    long an = a.num / g1, bd = b.den / g1, ad = a.den / g2, bn = b.num / g2;
    lpo_rat r = { 0, 0, 1 };
    if (__builtin_mul_overflow(an, bn, &r.num) || __builtin_mul_overflow(ad, bd, &r.den))
        return rat_invalid();
    return r;
}

/* Fraction sum over the LCM of the denominators, reduced AFTER adding
 * (Source/numOprts.c:78-129); the result's denominator is positive. */
lpo_rat lpo_rat_add(lpo_rat a, lpo_rat b) {
    if (!a.valid || !b.valid || a.den == 0 || b.den == 0) return rat_invalid();
    long l = lpo_lcm(a.den, b.den), pa, pb, s;
    if (l == -1) return rat_invalid();
    if (__builtin_mul_overflow(a.num, l / a.den, &pa)) return rat_invalid();
    if (__builtin_mul_overflow(b.num, l / b.den, &pb)) return rat_invalid();
    if (__builtin_add_overflow(pa, pb, &s)) return rat_invalid();
    long g = lpo_gcd(s, l);
    lpo_rat r = { s / g, l / g, 1 };
    return r;
}

lpo_rat lpo_rat_neg(lpo_rat a) { a.num = -a.num; return a; }

lpo_rat lpo_rat_sub(lpo_rat a, lpo_rat b) { return lpo_rat_add(a, lpo_rat_neg(b)); }

/* Multiply by the reciprocal (Source/numOprts.c:267-283): a zero divisor puts 0 in the
 * denominator, which the product rejects. */
lpo_rat lpo_rat_div(lpo_rat a, lpo_rat b) {
    long t = b.num;
    b.num = b.den;
    b.den = t;
    return lpo_rat_mul(a, b);
}

double lpo_rat_to_double(lpo_rat a) {
    if (!a.valid || a.den == 0) return 0.0;
    return (double)a.num / (double)a.den;
}

int lpo_rat_cmp(lpo_rat a, lpo_rat b) {
    __int128 l = (__int128)a.num * (__int128)b.den, r = (__int128)b.num * (__int128)a.den;
    int flip = ((a.den < 0) != (b.den < 0));
    int c = (l > r) - (l < r);
    return flip ? -c : c;
}

static int rat_sign(lpo_rat a) {
    int s = (a.num > 0) - (a.num < 0);
    return a.den < 0 ? -s : s;
}

lpo_rat lpo_rat_parse(const char *s) {
    char buf[128];
    size_t len = strlen(s);
    if (len + 2 >= sizeof buf) return rat_invalid();
    memcpy(buf, s, len + 1);
    char *end;
    char *slash = strchr(buf, '/');
    if (slash) {                       /* "p/q": both integers, p != 0, q > 0 after reduce */
        *slash = '\0';
        if (buf[0] == '\0' || slash[1] == '\0') return rat_invalid();
        long p = strtol(buf, &end, 10);
        if (*end) return rat_invalid();
        long q = strtol(slash + 1, &end, 10);
        if (*end || p == 0) return rat_invalid();
        long g = lpo_gcd(p, q);
        if (g == 0) return rat_invalid();
        p /= g;
        q /= g;
        if (q <= 0) return rat_invalid();
        return lpo_rat_make(p, q);
    }
    if (len == 0 || (len == 1 && (buf[0] == '+' || buf[0] == '-'))) {   /* bare sign = 1 */
        buf[len] = '1';
        buf[len + 1] = '\0';
    }
    char *dot = strchr(buf, '.');
    if (dot) {                         /* decimal: digits after '.' fix the denominator    */
        double v = strtod(buf, &end);
        if (*end || dot[1] == '\0') return rat_invalid();
        size_t digits = strlen(dot + 1);
        long q = (long)pow(10.0, (double)digits);
        long p = (long)(v * (double)q);    /* truncates: "0.29" -> 28/100, as the reference */
        long g = lpo_gcd(p, q);
        if (g == 0) return rat_invalid();
        return lpo_rat_make(p / g, q / g);
    }
    long p = strtol(buf, &end, 10);
    if (*end) return rat_invalid();
    return lpo_rat_make(p, 1);
}

#define RAT(T, ld, i, j) ((T)[(size_t)(i) * (ld) + (size_t)(j)])

static int rat_price_single(const lpo_rat *T, size_t ld, int row, int jmax) {
    int best = 0;
    for (int j = 1; j <= jmax; ++j) {
        lpo_rat v = RAT(T, ld, row, j);
        if (rat_sign(v) > 0 && (best == 0 || lpo_rat_cmp(v, RAT(T, ld, row, best)) > 0)) best = j;
    }
    return best;
}

static int rat_price_bigm(const lpo_rat *T, size_t ld, int m, int n) {
    int best = 0;
    for (int j = 1; j <= n; ++j) {
        lpo_rat p = RAT(T, ld, m + 1, j), q = RAT(T, ld, m, j);
        int sp = rat_sign(p);
        if (!(sp > 0 || (sp == 0 && rat_sign(q) > 0))) continue;
        if (best == 0) { best = j; continue; }
        int cp = lpo_rat_cmp(p, RAT(T, ld, m + 1, best));
        if (cp > 0 || (cp == 0 && lpo_rat_cmp(q, RAT(T, ld, m, best)) > 0)) best = j;
    }
    return best;
}

int lpo_solve_rat(int m, int n, int R, lpo_rat *T, size_t ld, int *basis,
                  const lpo_opts *opts, int *pivot_log, int log_cap, lpo_result *res) {
    if (m < 1 || n < 1 || (R != 1 && R != 2) || !T || !basis || !opts || !res) return -1;
    if ((opts->method == LPO_PLAIN) != (R == 1)) return -1;
    const int nreal = n - opts->n_art;
    int phase = (opts->method == LPO_TWOPHASE) ? LPO_PH_ONE : LPO_PH_MAIN;
    int status = LPO_RUNNING, iters = 0, p1 = 0;
    lpo_rat *prow = (lpo_rat *)malloc(sizeof(lpo_rat) * (size_t)(n + 1));
    while (status == LPO_RUNNING) {
        int r = -1, k = 0;
        if (phase == LPO_PH_DRIVEOUT) {
            for (int i = 0; i < m && r < 0; ++i) {
                if (basis[i] <= nreal) continue;
                for (int j = 1; j <= nreal; ++j)
                    if (rat_sign(RAT(T, ld, i, j)) != 0) { r = i; k = j; break; }
            }
            if (r < 0) { phase = LPO_PH_MAIN; continue; }
        } else {
            if (opts->method == LPO_BIGM) k = rat_price_bigm(T, ld, m, n);
            else if (phase == LPO_PH_ONE) k = rat_price_single(T, ld, m + 1, n);
            else k = rat_price_single(T, ld, m, opts->method == LPO_TWOPHASE ? nreal : n);
            if (k == 0) {
                int artpos = 0;
                for (int i = 0; i < m; ++i)
                    if (basis[i] > nreal && rat_sign(RAT(T, ld, i, 0)) > 0) artpos = 1;
                if (phase == LPO_PH_ONE) {
                    if (artpos) { status = LPO_INFEASIBLE; break; }
                    phase = LPO_PH_DRIVEOUT;
                    p1 = iters;
                    continue;
                }
                status = (opts->method == LPO_BIGM && artpos) ? LPO_INFEASIBLE : LPO_OPTIMAL;
                break;
            }
            lpo_rat bt = rat_invalid();
            for (int i = 0; i < m; ++i) {
                lpo_rat a = RAT(T, ld, i, k);
                if (rat_sign(a) <= 0) continue;
                lpo_rat t = lpo_rat_div(RAT(T, ld, i, 0), a);
                if (!t.valid) { status = LPO_OVERFLOW; break; }
                if (r < 0 || lpo_rat_cmp(t, bt) < 0) { r = i; bt = t; }
            }
            if (status != LPO_RUNNING) break;
            if (r < 0) { status = LPO_UNBOUNDED; break; }
        }
        if (iters >= opts->max_iter) { status = LPO_ITER_LIMIT; break; }
        /* one NDiv per pivot-row element, then one NMul + one NSub per other element
         *  */
        lpo_rat piv = RAT(T, ld, r, k);
        int ok = 1;
        for (int j = 0; j <= n; ++j) { prow[j] = lpo_rat_div(RAT(T, ld, r, j), piv); ok &= prow[j].valid; }
        for (int i = 0; i < m + R && ok; ++i) {
            if (i == r) continue;
            lpo_rat f = RAT(T, ld, i, k);
            for (int j = 0; j <= n; ++j) {
                lpo_rat v = lpo_rat_sub(RAT(T, ld, i, j), lpo_rat_mul(f, prow[j]));
                ok &= v.valid;
                RAT(T, ld, i, j) = v;
            }
        }
        for (int j = 0; j <= n; ++j) RAT(T, ld, r, j) = prow[j];
        if (!ok) { status = LPO_OVERFLOW; break; }
        basis[r] = k;
        if (pivot_log && iters < log_cap) { pivot_log[2 * iters] = k; pivot_log[2 * iters + 1] = r; }
        ++iters;
    }
    free(prow);
    if (phase == LPO_PH_ONE) p1 = iters;
    res->status = status;
    res->iters = iters;
    res->phase1_iters = (opts->method == LPO_TWOPHASE) ? p1 : 0;
    res->phase = phase;
    res->z = -lpo_rat_to_double(RAT(T, ld, m, 0));
    res->zM = (R == 2) ? -lpo_rat_to_double(RAT(T, ld, m + 1, 0)) : 0.0;
    return 0;
}
