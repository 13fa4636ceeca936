/*
 * lp_frontend.h -- host-side C front end: the reference's model text format to a dense tableau.
 *
 * Keeps the input format of Source/testdata.txt / Source/dataReader.c reachable without the
 * reference's sources:  OF { max|min : z = ... }  ST { row ; row ; x1>=0 ; x2<=0 }  with integer,
 * fraction (3/4) or decimal (2.45) coefficients and variable names "letter + up to two digits"
 * (Source/testdata.txt:13-25).  The pipeline restates, in new code, what the reference does between
 * reading the file and returning from CreateSMatrix:
 *   parse + per-row GCD simplification   (Source/dataReader.c:148-235, 301-432)
 *   move terms / merge like terms / extract sign rows      (Source/dataReader.c:45-140)
 *   standard form: min -> max(-z), x(unr) -> x'' - x', b < 0 -> negate row, slack / surplus,
 *   sort columns, x <= 0 -> x' = -x                          (Source/simplex.c:91-230)
 *   align rows to the objective's column order              (Source/simplex.c:238-260)
 *   tableau  b | A, costs, exact unit-column basis           (Source/matrix.c:19-91)
 * and then adds what the reference stops short of: artificial columns (Source/simplex.c:79-82 is an
 * empty AddBigM), the objective rows and the solve through the C ABI of include/lpgpu.h, plus the
 * read-back (undo x'' - x', undo inversion, add the stripped constant, flip sign for min).
 * tests/test_frontend.py pins the tableau against dumps produced by the reference's own code.
 *
 * Deliberate divergences from reference defects: decimals are converted
 * exactly ("0.29" = 29/100; Source/basicFuncs.c:264 truncates to 28/100); "0/5" is accepted as 0
 * (Source/basicFuncs.c:210 rejects it); the initial basis is detected with an exact unit-column
 * test on zero-cost columns (Source/matrix.c:59-78 sums truncated integers and accepts (1, 1/2)).
 */
#ifndef LP_FRONTEND_H
#define LP_FRONTEND_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { long num, den; } lpf_rat;          /* den > 0 */

typedef struct {
    char name[8];
    int relation;        /* 2: x >= 0, -2: x <= 0, 0: unrestricted (Source/hashTable.h:13-23)      */
    char former[8];      /* unrestricted x = former - latter after standardisation                  */
    char latter[8];
} lpf_var;

typedef struct {
    int valid;
    char error[256];
    int sense;           /* 1 = max, -1 = min, as written in the file                               */
    lpf_rat constant;    /* objective constant of the STANDARD form (negated for min), stripped     */
    int m, n;            /* rows, columns of the standard form (structural + split + slack/surplus) */
    char (*names)[8];    /* n column names in tableau order                                         */
    int *inverted;       /* n flags: column stands for x' = -x                                      */
    lpf_rat *costs;      /* n                                                                       */
    lpf_rat *rows;       /* m x (n+1): b | a                                                        */
    int *basis;          /* m: 1-based zero-cost unit column of the row, 0 if the row lacks one     */
    int nvars;
    int nuser;           /* vars[0..nuser) are the model's own variables, the rest are generated    */
    lpf_var *vars;       /* dictionary of every variable incl. generated ones                       */
} lpf_model;

/* parse + standardise + align.  Returns 0 and model->valid = 1, or -1 with model->error set. */
int lpf_parse_string(const char *text, lpf_model *model);
int lpf_parse_file(const char *path, lpf_model *model);
/* dual = 1: the standard form the dual simplex starts from (Source/simplex.c:177-178, LPStandardize
 * with dual = 1): ">=" rows are negated into "<=" instead of making b >= 0. */
int lpf_parse_string_mode(const char *text, int dual, lpf_model *model);
int lpf_parse_file_mode(const char *path, int dual, lpf_model *model);
void lpf_free(lpf_model *model);

/* Dense fp64 canonical tableau (layout of include/lpgpu.h).  method: 0 plain (fails if a row lacks a
 * unit column), 1 Big-M, 2 two-phase; rows lacking a unit column get one trailing artificial column
 * each, in row order.  *T is malloc'ed: (m+R) x (*n_total+1), ld = *n_total+1.  Returns 0 / -1. */
int lpf_build_tableau(const lpf_model *model, int method, double **T, int **basis, int *n_total, int *n_art);

typedef struct {
    int status;          /* LPGPU_* status                                                          */
    int iters;
    int method;          /* method actually used (plain when an identity basis exists)             */
    double z;            /* objective value in the ORIGINAL sense, constant included                 */
    int n_named;         /* user-visible variables                                                  */
    char (*names)[8];
    double *values;      /* their values in the original variables                                  */
    int n_log;
    int *pivot_log;      /* (entering column, leaving row) pairs                                    */
} lpf_solution;

/* Parse `path`, build the tableau, solve it on the GPU through liblpgpu.so and map the solution back.
 * method_if_needed: 1 Big-M or 2 two-phase, used only when artificials are required; 3 = dual simplex
 * (what the reference's menu item 2 announces, Source/router.c:32-34): needs every row to own a slack
 * and every cost <= 0 in the max form, else -1 with a message.
 * ctx: an lpgpu_ctx* (void* here to keep this header free of lpgpu.h).  Returns 0 / -1. */
int lpf_solve_file(void *ctx, const char *path, int method_if_needed, lpf_solution *sol, char *err, size_t errlen);
int lpf_solve_model(void *ctx, const lpf_model *model, int method_if_needed, lpf_solution *sol, char *err, size_t errlen);
void lpf_free_solution(lpf_solution *sol);

#ifdef __cplusplus
}
#endif
#endif
