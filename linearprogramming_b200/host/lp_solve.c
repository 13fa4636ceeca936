/*
 * lp_solve.c -- what newSimplex (Source/simplex.c:27-73) would do after CreateSMatrix if the reference
 * had a pivot loop: add artificial columns where a row lacks a unit column (the empty AddBigM,
 * Source/simplex.c:79-82), solve on the GPU through the C ABI (include/lpgpu.h) and map the solution
 * back to the model's own variables: x(unr) = x'' - x' (Source/simplex.c:110-169), x(<=0) = -x'
 * (Source/simplex.c:343-354), objective constant re-added (Source/matrix.c:23-28), sign restored for
 * min (Source/simplex.c:99-106).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/lpgpu.h"
#include "lp_frontend.h"

static int seterr(char *err, size_t n, const char *msg) {
    if (err && n) { strncpy(err, msg, n - 1); err[n - 1] = '\0'; }
    return -1;
}

static double column_value(const lpf_model *mo, const double *x, const char *name) {
    for (int j = 0; j < mo->n; ++j)
        if (strcmp(mo->names[j], name) == 0) return x[j];
    return 0.0;
}

int lpf_solve_model(void *ctx, const lpf_model *mo, int method_if_needed, lpf_solution *sol, char *err, size_t errlen) {
    if (!ctx || !mo || !mo->valid || !sol) return seterr(err, errlen, "bad arguments");
    memset(sol, 0, sizeof *sol);
    int lacking = 0;
    for (int i = 0; i < mo->m; ++i) lacking += (mo->basis[i] == 0);
    int method = LPGPU_PLAIN;
    if (method_if_needed == LPGPU_DUAL) {
        if (lacking) return seterr(err, errlen, "Dual simplex needs a slack in every row (no equality rows)");
        for (int j = 0; j < mo->n; ++j)
            if (mo->costs[j].num > 0) return seterr(err, errlen, "Dual simplex needs a dual feasible start (every cost <= 0 in the max form)");
        method = LPGPU_DUAL;
    } else if (lacking) {
        if (method_if_needed != LPGPU_BIGM && method_if_needed != LPGPU_TWOPHASE)
            return seterr(err, errlen, "Artificial variables are needed: choose Big M (1) or Two-phase (2)");
        method = method_if_needed;
    }
    double *T = NULL;
    int *basis = NULL, n = 0, n_art = 0;
    if (lpf_build_tableau(mo, method == LPGPU_DUAL ? LPGPU_PLAIN : method, &T, &basis, &n, &n_art) != 0) return seterr(err, errlen, "cannot build the tableau");
    const int R = (method == LPGPU_PLAIN || method == LPGPU_DUAL) ? 1 : 2;
    lpgpu_opts o;
    lpgpu_default_opts(&o);
    o.method = method;
    o.n_art = n_art;
    {   /* LPGPU_PRICING=bland: the lowest eligible column enters (anti-cycling); =devex: reference-weight pricing; primal methods */
        const char *pe = getenv("LPGPU_PRICING");
        if (pe && !strcmp(pe, "bland") && method != LPGPU_DUAL) o.variant |= LPGPU_VARIANT_BLAND;
        if (pe && !strcmp(pe, "devex") && method != LPGPU_DUAL) o.variant |= LPGPU_VARIANT_DEVEX;
    }
    o.log_cap = 16 * (mo->m + n) + 64;
    {   /* "message, never hang": a cycling model ends with ITERATION LIMIT (LPGPU_MAX_ITER overrides the cap) */
        const char *me = getenv("LPGPU_MAX_ITER");
        o.max_iter = (me && atol(me) > 0) ? (int)(atol(me) > 0x7fffffffL ? 0x7fffffffL : atol(me)) : 50 * (mo->m + n) + 1000;
    }
    int *log = (int *)malloc(sizeof(int) * 2 * (size_t)o.log_cap);
    double *x = (double *)calloc((size_t)n, sizeof(double));
    lpgpu_result res;
    int rc = lpgpu_solve_dense((lpgpu_ctx *)ctx, mo->m, n, R, T, (size_t)n + 1, basis, &o, log, x, &res);
    if (rc != LPGPU_OK) {
        seterr(err, errlen, lpgpu_last_error((lpgpu_ctx *)ctx));
        free(T); free(basis); free(log); free(x);
        return -1;
    }
    sol->status = res.status;
    sol->iters = res.iters;
    sol->method = method;
    const double zstd = res.z + (double)mo->constant.num / (double)mo->constant.den;
    sol->z = mo->sense == 1 ? zstd : -zstd;
    sol->n_named = mo->nuser;
    sol->names = (char (*)[8])calloc((size_t)(mo->nuser ? mo->nuser : 1), 8);
    sol->values = (double *)calloc((size_t)(mo->nuser ? mo->nuser : 1), sizeof(double));
    for (int v = 0; v < mo->nuser; ++v) {
        const lpf_var *e = &mo->vars[v];
        strcpy(sol->names[v], e->name);
        if (e->relation == 0 && e->former[0]) sol->values[v] = column_value(mo, x, e->former) - column_value(mo, x, e->latter);
        else if (e->relation < 0) sol->values[v] = -column_value(mo, x, e->name);
        else sol->values[v] = column_value(mo, x, e->name);
    }
    sol->n_log = res.n_logged;
    sol->pivot_log = log;
    free(T); free(basis); free(x);
    return 0;
}

int lpf_solve_file(void *ctx, const char *path, int method_if_needed, lpf_solution *sol, char *err, size_t errlen) {
    lpf_model mo;
    if (lpf_parse_file_mode(path, method_if_needed == LPGPU_DUAL, &mo) != 0) return seterr(err, errlen, mo.error);
    int rc = lpf_solve_model(ctx, &mo, method_if_needed, sol, err, errlen);
    lpf_free(&mo);
    return rc;
}

void lpf_free_solution(lpf_solution *sol) {
    if (!sol) return;
    free(sol->names); free(sol->values); free(sol->pivot_log);
    memset(sol, 0, sizeof *sol);
}
