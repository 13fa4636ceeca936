/*
 * lp_cli.c -- non-interactive mirror of the reference's main.c + router.c flow (Source/main.c:4-45,
 * Source/router.c:13-42): read a model file, print the standard form, solve with the primal simplex
 * on the GPU, print the solution.  The interactive prompts (PAUSE, menu digits, "1. Big M / 2.
 * Two-phase", Source/simplex.c:43-55) become flags:
 *     lp_cli model.txt [--method bigm|twophase|dual] [--json] [--parse-only] [--device N]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/lpgpu.h"
#include "lp_frontend.h"

static const char *status_name(int s) {
    static const char *n[] = { "OPTIMAL", "UNBOUNDED", "INFEASIBLE", "ITER_LIMIT", "RUNNING" };
    return (s >= 0 && s <= 4) ? n[s] : "?";
}

static void print_rat(FILE *o, lpf_rat r) {
    if (r.den == 1) fprintf(o, "%ld", r.num);
    else fprintf(o, "%ld/%ld", r.num, r.den);
}

static void print_model_json(const lpf_model *m) {
    printf("{\"sense\":%d,\"m\":%d,\"n\":%d,\"constant\":[%ld,%ld],\"names\":[", m->sense, m->m, m->n, m->constant.num, m->constant.den);
    for (int j = 0; j < m->n; ++j) printf("%s\"%s\"", j ? "," : "", m->names[j]);
    printf("],\"inverted\":[");
    for (int j = 0; j < m->n; ++j) printf("%s%d", j ? "," : "", m->inverted[j]);
    printf("],\"costs\":[");
    for (int j = 0; j < m->n; ++j) printf("%s[%ld,%ld]", j ? "," : "", m->costs[j].num, m->costs[j].den);
    printf("],\"rows\":[");
    for (int i = 0; i < m->m; ++i) {
        printf("%s[", i ? "," : "");
        for (int j = 0; j <= m->n; ++j) {
            lpf_rat a = m->rows[(size_t)i * (size_t)(m->n + 1) + j];
            printf("%s[%ld,%ld]", j ? "," : "", a.num, a.den);
        }
        printf("]");
    }
    printf("],\"basis\":[");
    for (int i = 0; i < m->m; ++i) printf("%s%d", i ? "," : "", m->basis[i]);
    printf("]}\n");
}

static void print_model_text(const lpf_model *m) {
    printf("> Standard Form (max)\n\tz' =");
    for (int j = 0; j < m->n; ++j) { printf(" + "); print_rat(stdout, m->costs[j]); printf("[%s%s]", m->names[j], m->inverted[j] ? "'" : ""); }
    printf(" + "); print_rat(stdout, m->constant); printf("\nSubject to:\n");
    for (int i = 0; i < m->m; ++i) {
        printf("\t");
        for (int j = 0; j < m->n; ++j) {
            lpf_rat a = m->rows[(size_t)i * (size_t)(m->n + 1) + 1 + j];
            if (a.num == 0) continue;
            printf(" + "); print_rat(stdout, a); printf("[%s%s]", m->names[j], m->inverted[j] ? "'" : "");
        }
        printf(" = "); print_rat(stdout, m->rows[(size_t)i * (size_t)(m->n + 1)]);
        if (m->basis[i]) printf("\t(basic: %s)\n", m->names[m->basis[i] - 1]);
        else printf("\t(needs an artificial variable)\n");
    }
}

int main(int argc, char **argv) {
    const char *path = NULL;
    int method = LPGPU_BIGM, json = 0, parse_only = 0, device = 0;
    for (int a = 1; a < argc; ++a) {
        if (!strcmp(argv[a], "--method") && a + 1 < argc) {
            ++a;
            if (!strcmp(argv[a], "bigm") || !strcmp(argv[a], "1")) method = LPGPU_BIGM;
            else if (!strcmp(argv[a], "twophase") || !strcmp(argv[a], "2")) method = LPGPU_TWOPHASE;
            else if (!strcmp(argv[a], "dual")) method = LPGPU_DUAL;      /* the reference's menu item 2, Source/router.c:32-34 */
            else { fprintf(stderr, "unknown method %s\n", argv[a]); return 2; }
        } else if (!strcmp(argv[a], "--json")) json = 1;
        else if (!strcmp(argv[a], "--parse-only")) parse_only = 1;
        else if (!strcmp(argv[a], "--device") && a + 1 < argc) device = atoi(argv[++a]);
        else path = argv[a];
    }
    if (!path) { printf("Missing argument. \nUsage: \n\t%s <filename> [--method bigm|twophase|dual] [--json] [--parse-only]\n\n", argv[0]); return 2; }
    lpf_model mo;
    if (lpf_parse_file_mode(path, method == LPGPU_DUAL, &mo) != 0) {
        if (json) printf("{\"parsed\":0,\"error\":\"%s\"}\n", mo.error);
        else printf("%s\nError occurred when parsing the model.\n", mo.error);
        return 1;
    }
    if (parse_only) {
        if (json) print_model_json(&mo); else print_model_text(&mo);
        lpf_free(&mo);
        return 0;
    }
    lpgpu_ctx *ctx = NULL;
    int rc = lpgpu_create(&ctx, device, NULL);
    if (rc != LPGPU_OK) { fprintf(stderr, "lpgpu_create failed (%d): no sm_100 CUDA device; there is no CPU fallback\n", rc); lpf_free(&mo); return 3; }
    lpf_solution sol;
    char err[512] = "";
    if (lpf_solve_model(ctx, &mo, method, &sol, err, sizeof err) != 0) { fprintf(stderr, "solve failed: %s\n", err); lpgpu_destroy(ctx); lpf_free(&mo); return 4; }
    if (json) {
        printf("{\"status\":\"%s\",\"iters\":%d,\"method\":%d,\"z\":%.17g,\"x\":{", status_name(sol.status), sol.iters, sol.method, sol.z);
        for (int v = 0; v < sol.n_named; ++v) printf("%s\"%s\":%.17g", v ? "," : "", sol.names[v], sol.values[v]);
        printf("},\"log\":[");
        for (int p = 0; p < sol.n_log; ++p) printf("%s[%d,%d]", p ? "," : "", sol.pivot_log[2 * p], sol.pivot_log[2 * p + 1]);
        printf("]}\n");
    } else {
        print_model_text(&mo);
        if (sol.method == LPGPU_DUAL) printf("\n> Dual Simplex Method\n");
        else printf("\n> Primal Simplex Method (%s)\n", sol.method == LPGPU_PLAIN ? "identity basis" : sol.method == LPGPU_BIGM ? "Big M" : "Two-phase");
        for (int p = 0; p < sol.n_log; ++p) {
            const int col = sol.pivot_log[2 * p], row = sol.pivot_log[2 * p + 1];
            printf("\titeration %d: column %d enters, row %d leaves\n", p + 1, col, row);
        }
        printf("> %s after %d iteration(s)\n", status_name(sol.status), sol.iters);
        if (sol.status == LPGPU_ITER_LIMIT) printf("\tpivot cap reached (a degenerate model may cycle): try LPGPU_PRICING=bland or raise LPGPU_MAX_ITER\n");
        if (sol.status == LPGPU_OPTIMAL) {
            printf("\t%s z = %.10g\n", mo.sense == 1 ? "max" : "min", sol.z);
            for (int v = 0; v < sol.n_named; ++v) printf("\t%s = %.10g\n", sol.names[v], sol.values[v]);
        }
    }
    lpf_free_solution(&sol);
    lpgpu_destroy(ctx);
    lpf_free(&mo);
    return 0;
}
