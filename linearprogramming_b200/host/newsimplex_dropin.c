/*
 * newsimplex_dropin.c -- drop-in replacement for the reference's newSimplex (Source/simplex.c:27-73).
 *
 * Compiled against the reference's own headers (public.h, simplex.h, matrix.h; -I/root/reference/Source
 * at build time, nothing is copied) and linked with the reference's unmodified main.c, router.c,
 * dataReader.c, basicFuncs.c, hashTable.c, numOprts.c, matrix.c and with simplex.c built with
 * -DnewSimplex=ref_newSimplex_unused, so LPStandardize / LPAlign stay the reference's
 * (Source/simplex.h:9-11) and only newSimplex (Source/simplex.h:13) is ours.  Same signature, same
 * caller (Entry, Source/router.c:28-31), same ownership (works on a CopyModel, frees it), same error
 * convention (messages, never aborts).
 *
 * Where the reference frees the matrix without solving (Source/simplex.c:40,67; empty cases :57-63;
 * empty AddBigM :79-82) this version converts the SimplexMatrix to the dense fp64 layout with
 * Decimalize (Source/basicFuncs.c:298-313), adds one artificial column per row that lacks a unit
 * column, calls lpgpu_solve_dense (include/lpgpu.h) and prints the pivots and the solution in the
 * model's own variables.
 *
 * Environment switches for non-interactive runs (the reference is menu driven):
 *   LPGPU_BATCH=1          skip the "Press Enter" pauses of this function
 *   LPGPU_METHOD=bigm|twophase   answer the "1. Big M / 2. Two-phase" prompt
 *   LPGPU_METHOD=dual            dual simplex: standardise with the reference's LPStandardize(model, 1)
 *                                (Source/simplex.c:177-178; the menu entry "2. Dual Simplex Method" of
 *                                Source/router.c:32-34 calls nothing) and pivot with LPGPU_DUAL
 *   LPGPU_PRICING=bland          Bland's rule (anti-cycling), primal methods
 *   LPGPU_MAX_ITER=<n>           pivot cap (default 50 * (m + n) + 1000: "message, never hang")
 *   LPGPU_PRINT_TABLEAU=1        print the textbook tableau  C_B | X_B | b | a_ij  with the sigma_j row after
 *                                every iteration (the layout SimplexMatrix's fields imply: basicCosts = C_B,
 *                                basicVars = X_B, b in column 0, Source/matrix.h:13-21); models up to 64 columns
 */
#include "public.h"
#include "simplex.h"
#include "matrix.h"

#include "lpgpu.h"

static void pause_if_interactive(void) {
    if (getenv("LPGPU_BATCH")) return;
    PAUSE;
}

static int is_one(Number v) { return v.valid && v.constant == NULL && v.numerator == v.denominator && v.denominator != 0; }
static int is_zero(Number v) { return v.valid && v.constant == NULL && v.numerator == 0; }

static const char *status_text(int s) {
    switch (s) {
        case LPGPU_OPTIMAL: return "OPTIMAL";
        case LPGPU_UNBOUNDED: return "UNBOUNDED";
        case LPGPU_INFEASIBLE: return "INFEASIBLE";
        case LPGPU_ITER_LIMIT: return "ITERATION LIMIT";
        default: return "?";
    }
}

static int choose_method(void) {
    const char *e = getenv("LPGPU_METHOD");
    if (e && (!strcmp(e, "bigm") || !strcmp(e, "1"))) return LPGPU_BIGM;
    if (e && (!strcmp(e, "twophase") || !strcmp(e, "2"))) return LPGPU_TWOPHASE;
    short int chosen = 0;
    while (1) {     /* the reference's prompt, Source/simplex.c:43-55 */
        printf("Artificial variables are needed due to the lack of Identity Matrix.\n");
        printf("Choose one method to continue:\n");
        printf("\t1. Big M Method.\n");
        printf("\t2. Two-phase Method.\n");
        printf("Type in the correspond number:");
        fflush(stdout);
        chosen = (short int)(ReadChar() - '0');
        if (chosen >= 1 && chosen <= 2) break;
    }
    return chosen == 1 ? LPGPU_BIGM : LPGPU_TWOPHASE;
}

/* pivot cap of the host front ends: a cycling model ends with ITERATION LIMIT and a hint, never with a hang */
static int pivot_cap(int m, int n) {
    const char *e = getenv("LPGPU_MAX_ITER");
    if (e && atol(e) > 0) return (int)(atol(e) > 0x7fffffffL ? 0x7fffffffL : atol(e));
    return 50 * (m + n) + 1000;
}

#define PRINT_TABLEAU_MAX_COLS 64

static void column_name(char *out, size_t cap, const SimplexMatrix *sm, int n0, int col) {
    if (col >= 1 && col <= n0) snprintf(out, cap, "%s", sm->varNames[col - 1]);
    else snprintf(out, cap, "a%d", col - n0);
}

/* The textbook tableau after `iter` pivots:  C_B | X_B | b | a_i1 .. a_in, then the sigma_j row(s) with -z in the
 * b column.  cost[j-1] = c_j of the standard form; artificial columns cost -M (Big M) or 0 in the cost row. */
static void print_tableau(const SimplexMatrix *sm, int m, int n, int n0, int R, int method, const double *T, size_t ld,
                          const int *basis, const double *cost, int iter, int enter, int leave) {
    char nm[16];
    int i, j;
    if (iter == 0) printf("\n> Initial tableau\n\n");
    else {
        column_name(nm, sizeof nm, sm, n0, enter);
        printf("\n> Tableau after iteration %d (%s enters, row %d leaves)\n\n", iter, nm, leave + 1);
    }
    printf("\t%8s %6s %12s |", "C_B", "X_B", "b");
    for (j = 1; j <= n; j++) { column_name(nm, sizeof nm, sm, n0, j); printf(" %10s", nm); }
    printf("\n\t%8s %6s %12s |", "", "c_j", "");
    for (j = 1; j <= n; j++) {
        if (j > n0 && method == LPGPU_BIGM) printf(" %10s", "-M");
        else printf(" %10.6g", j <= n0 ? cost[j - 1] : 0.0);
    }
    printf("\n");
    for (i = 0; i < m; i++) {
        const int bj = basis[i];
        column_name(nm, sizeof nm, sm, n0, bj);
        if (bj > n0 && method == LPGPU_BIGM) printf("\t%8s", "-M");
        else printf("\t%8.6g", (bj >= 1 && bj <= n0) ? cost[bj - 1] : 0.0);
        printf(" %6s %12.6g |", nm, T[(size_t)i * ld]);
        for (j = 1; j <= n; j++) printf(" %10.6g", T[(size_t)i * ld + (size_t)j]);
        printf("\n");
    }
    printf("\t%8s %6s %12.6g |", "sigma_j", "-z", T[(size_t)m * ld]);
    for (j = 1; j <= n; j++) printf(" %10.6g", T[(size_t)m * ld + (size_t)j]);
    printf("\n");
    if (R == 2) {
        printf("\t%8s %6s %12.6g |", method == LPGPU_BIGM ? "sigma^M" : "sigma^I", method == LPGPU_BIGM ? "-z_M" : "-w", T[(size_t)(m + 1) * ld]);
        for (j = 1; j <= n; j++) printf(" %10.6g", T[(size_t)(m + 1) * ld + (size_t)j]);
        printf("\n");
    }
}

/* LPGPU_PRINT_TABLEAU=1: the same solve, one pivot per lpgpu_dense_run call on a device-resident tableau
 * (include/lpgpu.h), the tableau read back and printed after every iteration.  Same pivots, same bits as the
 * one-shot call; meant for small models. */
static int solve_with_tableau_print(lpgpu_ctx *ctx, const SimplexMatrix *sm, int m, int n, int n0, int R, double *T, size_t ld,
                                    int *basis, const lpgpu_opts *o, int *plog, double *x, lpgpu_result *res) {
    lpgpu_dense *d = NULL;
    double *cost = (double *)calloc((size_t)n0 + 1, sizeof(double));
    int j, rc = lpgpu_dense_create(ctx, m, n, R, &d);
    for (j = 0; j < n0; j++) cost[j] = T[(size_t)m * ld + 1 + (size_t)j];
    if (rc == LPGPU_OK) rc = lpgpu_dense_reset(d, o);
    if (rc == LPGPU_OK) rc = lpgpu_dense_upload(d, T, ld, basis);
    if (rc == LPGPU_OK) print_tableau(sm, m, n, n0, R, o->method, T, ld, basis, cost, 0, 0, 0);
    memset(res, 0, sizeof *res);
    res->status = LPGPU_RUNNING;
    int done = 0;
    while (rc == LPGPU_OK && res->status == LPGPU_RUNNING && done < o->max_iter) {
        rc = lpgpu_dense_run(d, 1, res);
        if (rc != LPGPU_OK || res->iters == done) break;           /* terminal status without a pivot */
        done = res->iters;
        rc = lpgpu_dense_download(d, T, ld, basis, plog, NULL);
        if (rc == LPGPU_OK && done <= o->log_cap) print_tableau(sm, m, n, n0, R, o->method, T, ld, basis, cost, done, plog[2 * (done - 1)], plog[2 * (done - 1) + 1]);
    }
    if (rc == LPGPU_OK && res->status == LPGPU_RUNNING) rc = lpgpu_dense_run(d, 0, res);   /* settles OPTIMAL / ITER_LIMIT */
    if (rc == LPGPU_OK) rc = lpgpu_dense_download(d, T, ld, basis, plog, x);
    if (d) lpgpu_dense_destroy(d);
    free(cost);
    return rc;
}

void newSimplex(LPModel *model) {
    LPModel modelCopy = CopyModel(model);
    const short int sense = model->objective.type;          /* 1 max, -1 min, before standardising */
    const char *menv = getenv("LPGPU_METHOD");
    const short int dual = (menv && !strcmp(menv, "dual")) ? 1 : 0;
    LPStandardize(&modelCopy, dual);
    printf("\n---------------\n> Standard Form\n\n");
    PrintModel(&modelCopy);
    pause_if_interactive();
    LPAlign(&modelCopy);
    PrintModel(&modelCopy);
    pause_if_interactive();
    if (!modelCopy.valid) {
        printf("ERROR occurred during the Standardization and the Alignment :( \n");
        FreeModel(&modelCopy);
        return;
    }
    /* objective constant, before CreateSMatrix strips it (Source/matrix.c:23-28) */
    double constant = 0.0;
    size_t i, j;
    for (i = 0; i < modelCopy.objective.rightLen; i++)
        if (strlen(modelCopy.objective.right[i]->variable) == 0) constant = Decimalize(modelCopy.objective.right[i]->coefficient);

    /* Source/matrix.c:86 writes through lack[lackPtr] instead of (*lack)[lackPtr]: hand it one slot per
     * row so that the stray writes stay inside this frame; the list itself is rebuilt exactly below. */
    const size_t rowsMax = modelCopy.stLen + 1;
    size_t *scratch = (size_t *)calloc(rowsMax, sizeof(size_t));
    size_t **slots = (size_t **)calloc(rowsMax, sizeof(size_t *));
    for (i = 0; i < rowsMax; i++) slots[i] = &scratch[i];
    short int valid = 1;
    SimplexMatrix sm = CreateSMatrix(&modelCopy, slots, &valid);
    const int m = (int)sm.basicLen, n0 = (int)sm.ofLen;
    size_t *lackFromRef = slots[0] != &scratch[0] ? slots[0] : NULL;       /* calloc'ed by CreateSMatrix */

    /* exact unit-column test on zero-cost columns (Source/matrix.c:59-78 accepts (1, 1/2)) */
    int *basis = (int *)calloc((size_t)m, sizeof(int));
    for (j = 0; j < (size_t)n0; j++) {
        if (!is_zero(*sm.ofCosts[j])) continue;
        int one = -1, ok = 1;
        for (i = 0; i < (size_t)m && ok; i++) {
            Number a = *sm.cMatrix[i][j + 1];
            if (is_zero(a)) continue;
            if (is_one(a) && one < 0) one = (int)i;
            else ok = 0;
        }
        if (ok && one >= 0) basis[one] = (int)j + 1;
    }
    int n_art = 0;
    for (i = 0; i < (size_t)m; i++) n_art += (basis[i] == 0);
    int method = LPGPU_PLAIN;
    if (dual) {
        int ok = (n_art == 0);
        for (j = 0; j < (size_t)n0 && ok; j++)
            if (Decimalize(*sm.ofCosts[j]) > 0) ok = 0;
        if (!ok) {
            printf("The Dual Simplex Method needs a slack in every row and no positive cost in the max form.\n");
            pause_if_interactive();
            free(basis); free(scratch); free(slots);
            if (lackFromRef != NULL) free(lackFromRef);
            RevokeSMatrix(&sm);
            FreeModel(&modelCopy);
            return;
        }
        method = LPGPU_DUAL;
    } else if (n_art) method = choose_method();
    const int R = (method == LPGPU_PLAIN || method == LPGPU_DUAL) ? 1 : 2, n = n0 + n_art;
    const size_t ld = (size_t)n + 1;

    /* SimplexMatrix -> dense fp64 tableau (include/lpgpu.h layout) */
    double *T = (double *)calloc((size_t)(m + R) * ld, sizeof(double));
    int next_art = n0 + 1;
    for (i = 0; i < (size_t)m; i++) {
        for (j = 0; j <= (size_t)n0; j++) T[i * ld + j] = Decimalize(*sm.cMatrix[i][j]);
        if (!basis[i]) { T[i * ld + (size_t)next_art] = 1.0; basis[i] = next_art++; }
    }
    for (j = 0; j < (size_t)n0; j++) T[(size_t)m * ld + 1 + j] = Decimalize(*sm.ofCosts[j]);
    if (R == 2) {
        double *om = T + (size_t)(m + 1) * ld;
        for (j = (size_t)n0 + 1; j <= (size_t)n; j++) om[j] = -1.0;
        for (i = 0; i < (size_t)m; i++) {
            if (basis[i] <= n0) continue;
            for (j = 0; j <= (size_t)n; j++) om[j] += T[i * ld + j];
        }
    }

    lpgpu_ctx *ctx = NULL;
    int rc = lpgpu_create(&ctx, 0, NULL);
    if (rc != LPGPU_OK) {
        printf("ERROR: no sm_100 CUDA device available (lpgpu_create returned %d); there is no CPU fallback.\n", rc);
    } else {
        lpgpu_opts o;
        lpgpu_default_opts(&o);
        o.method = method;
        o.n_art = n_art;
        {   /* LPGPU_PRICING=bland: the lowest eligible column enters (anti-cycling); =devex: reference-weight pricing; primal methods */
            const char *pe = getenv("LPGPU_PRICING");
            if (pe && !strcmp(pe, "bland") && method != LPGPU_DUAL) o.variant |= LPGPU_VARIANT_BLAND;
            if (pe && !strcmp(pe, "devex") && method != LPGPU_DUAL) o.variant |= LPGPU_VARIANT_DEVEX;
        }
        o.log_cap = 16 * (m + n) + 64;
        o.max_iter = pivot_cap(m, n);
        int *plog = (int *)calloc((size_t)o.log_cap * 2, sizeof(int));
        double *x = (double *)calloc((size_t)n, sizeof(double));
        lpgpu_result res;
        const char *pt = getenv("LPGPU_PRINT_TABLEAU");
        if (pt && atoi(pt) > 0 && n <= PRINT_TABLEAU_MAX_COLS) rc = solve_with_tableau_print(ctx, &sm, m, n, n0, R, T, ld, basis, &o, plog, x, &res);
        else rc = lpgpu_solve_dense(ctx, m, n, R, T, ld, basis, &o, plog, x, &res);
        if (rc != LPGPU_OK) {
            printf("ERROR: GPU solve failed (%d): %s\n", rc, lpgpu_last_error(ctx));
        } else {
            if (method == LPGPU_DUAL) printf("\n---------------\n> Dual Simplex Method\n\n");
            else printf("\n---------------\n> Primal Simplex Method (%s)\n\n",
                        method == LPGPU_PLAIN ? "identity basis" : method == LPGPU_BIGM ? "Big M" : "Two-phase");
            for (int p = 0; p < res.n_logged; p++) {
                const int col = plog[2 * p], row = plog[2 * p + 1];
                printf("\tIteration %d: ", p + 1);
                if (col <= n0) printf("%s enters", sm.varNames[col - 1]);
                else printf("artificial a%d enters", col - n0);
                printf(", row %d leaves\n", row + 1);
            }
            printf("\n> %s after %d iteration(s)\n", status_text(res.status), res.iters);
            if (res.status == LPGPU_ITER_LIMIT)
                printf("\tThe pivot cap of %d was reached (a degenerate model may cycle under largest-coefficient pricing):\n"
                       "\ttry LPGPU_PRICING=bland, or raise the cap with LPGPU_MAX_ITER.\n", o.max_iter);
            if (res.status == LPGPU_OPTIMAL) {
                const double zstd = res.z + constant;
                printf("\t%s z = %.10g\n", sense == 1 ? "max" : "min", sense == 1 ? zstd : -zstd);
                /* back to the model's own variables: x(unr) = x'' - x' (Source/simplex.c:110-169),
                 * x(<=0) = -x' (Source/simplex.c:343-354) */
                size_t nv = 0;
                short int vv = 1;
                VarItem **items = GetVarItems(&nv, &vv);
                for (size_t v = 0; v < nv; v++) {
                    VarItem *it = items[v];
                    int generated = 0;
                    for (size_t w = 0; w < nv; w++)
                        if (!strcmp(items[w]->formerX, it->keyName) || !strcmp(items[w]->latterX, it->keyName)) generated = 1;
                    double val = 0.0;
                    int found = 0;
                    if (it->relation == 0 && strlen(it->formerX) > 0) {
                        for (j = 0; j < (size_t)n0; j++) {
                            if (!strcmp(sm.varNames[j], it->formerX)) val += x[j];
                            if (!strcmp(sm.varNames[j], it->latterX)) val -= x[j];
                        }
                        found = 1;
                    } else {
                        for (j = 0; j < (size_t)n0; j++)
                            if (!strcmp(sm.varNames[j], it->keyName)) { val = it->relation < 0 ? -x[j] : x[j]; found = 1; }
                    }
                    if (found && !generated) printf("\t%s = %.10g\n", it->keyName, val);
                }
                free(items);
            }
        }
        free(plog);
        free(x);
        lpgpu_destroy(ctx);
    }
    pause_if_interactive();
    free(T);
    free(basis);
    free(scratch);
    free(slots);
    if (lackFromRef != NULL) free(lackFromRef);
    RevokeSMatrix(&sm);
    FreeModel(&modelCopy);
}
