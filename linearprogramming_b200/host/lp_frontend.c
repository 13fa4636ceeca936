/*
 * lp_frontend.c -- model text -> standard form -> dense tableau (see lp_frontend.h for the map to
 * the reference's files).  Plain C, no GPU code; the solve itself lives in lp_solve.c.
 */
#include "lp_frontend.h"

#include <ctype.h>
#include <errno.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ small rationals */

static long gcd_l(long a, long b) {
    if (a < 0) a = -a;
    if (b < 0) b = -b;
    while (b) { long t = a % b; a = b; b = t; }
    return a;
}

static int rat_norm(__int128 num, __int128 den, lpf_rat *out) {
    if (den == 0) return -1;
    if (den < 0) { num = -num; den = -den; }
    __int128 a = num < 0 ? -num : num, b = den;
    while (b) { __int128 t = a % b; a = b; b = t; }
    if (a > 1) { num /= a; den /= a; }
    if (num > (__int128)0x7fffffffffffffffLL || num < -(__int128)0x7fffffffffffffffLL ||
        den > (__int128)0x7fffffffffffffffLL) return -1;
    out->num = (long)num;
    out->den = (long)den;
    return 0;
}

static int rat_add(lpf_rat a, lpf_rat b, lpf_rat *out) {
    return rat_norm((__int128)a.num * b.den + (__int128)b.num * a.den, (__int128)a.den * b.den, out);
}

static lpf_rat rat_neg(lpf_rat a) { a.num = -a.num; return a; }
static lpf_rat rat_int(long v) { lpf_rat r = { v, 1 }; return r; }
static double rat_val(lpf_rat a) { return (double)a.num / (double)a.den; }   /* Decimalize, basicFuncs.c:298-313 */
static int rat_sign(lpf_rat a) { return (a.num > 0) - (a.num < 0); }

/* "5", "-7", "3/4", "2.45", "+", "-" (bare sign = 1)  ->  rational.  Returns 0 / -1. */
static int rat_parse(const char *s, lpf_rat *out) {
    char buf[96];
    size_t len = strlen(s);
    if (len + 2 >= sizeof buf) return -1;
    memcpy(buf, s, len + 1);
    if (len == 0 || (len == 1 && (buf[0] == '+' || buf[0] == '-'))) { buf[len] = '1'; buf[len + 1] = '\0'; }
    char *end, *slash = strchr(buf, '/'), *dot = strchr(buf, '.');
    if (slash) {
        *slash = '\0';
        if (!buf[0] || !slash[1] || dot) return -1;
        errno = 0;
        long p = strtol(buf, &end, 10);
        if (*end || errno == ERANGE) return -1;
        long q = strtol(slash + 1, &end, 10);
        if (*end || q <= 0 || errno == ERANGE) return -1;
        return rat_norm(p, q, out);
    }
    if (dot) {
        /* exact decimal: digits after the point fix the power of ten */
        int neg = 0;
        const char *p = buf;
        if (*p == '+' || *p == '-') { neg = (*p == '-'); ++p; }
        __int128 num = 0, den = 1;
        int seen_digit = 0, after = 0;
        for (; *p; ++p) {
            if (*p == '.') { if (after) return -1; after = 1; continue; }
            if (!isdigit((unsigned char)*p)) return -1;
            seen_digit = 1;
            num = num * 10 + (*p - '0');
            if (after) den *= 10;
            if (num > (__int128)1 << 100 || den > (__int128)1 << 100) return -1;
        }
        if (!seen_digit) return -1;
        return rat_norm(neg ? -num : num, den, out);
    }
    errno = 0;
    long v = strtol(buf, &end, 10);
    if (*end || errno == ERANGE) return -1;     /* a coefficient that does not fit a long is an error, not LONG_MAX */
    *out = rat_int(v);
    return 0;
}

/* ------------------------------------------------------------------ terms and formulas */

typedef struct {
    lpf_rat c;
    char var[8];          /* "" = constant term */
    int inverted;
} term_t;

typedef struct {
    term_t *t;
    int n, cap;
} terms_t;

typedef struct {
    terms_t left, right;
    int rel;              /* -2 <=, -1 <, 1 >, 2 >=, 3 =   (codes of Source/public.h:57) */
} formula_t;

static void terms_push(terms_t *v, term_t x, int pos) {
    if (v->n == v->cap) {
        v->cap = v->cap ? v->cap * 2 : 8;
        v->t = (term_t *)realloc(v->t, sizeof(term_t) * (size_t)v->cap);
    }
    if (pos < 0 || pos >= v->n) v->t[v->n++] = x;
    else {
        memmove(v->t + pos + 1, v->t + pos, sizeof(term_t) * (size_t)(v->n - pos));
        v->t[pos] = x;
        v->n++;
    }
}

static void terms_remove(terms_t *v, int pos) {
    memmove(v->t + pos, v->t + pos + 1, sizeof(term_t) * (size_t)(v->n - pos - 1));
    v->n--;
}

static void formula_free(formula_t *f) { free(f->left.t); free(f->right.t); memset(f, 0, sizeof *f); }

static int fail(lpf_model *m, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(m->error, sizeof m->error, fmt, ap);
    va_end(ap);
    m->valid = 0;
    return -1;
}

/* variable names: a letter followed by digits only; only the first three characters count
 * (Source/dataReader.c:332-333, Source/basicFuncs.c:374-377) */
static int valid_var(const char *s) {
    if (!isalpha((unsigned char)s[0])) return 0;
    for (const char *p = s + 1; *p; ++p)
        if (!isdigit((unsigned char)*p)) return 0;
    return 1;
}

/* One statement without blanks, e.g. "-2x1+x2>=-1/3" or "z=x1+x2+6+1/3", into terms on both sides.
 * A term is [sign][number][variable]; a missing number means 1. */
static int parse_formula(const char *s, formula_t *f, lpf_model *m) {
    memset(f, 0, sizeof *f);
    int side = 0;
    size_t i = 0, len = strlen(s);
    while (i < len) {
        char ch = s[i];
        if (ch == '<' || ch == '>' || ch == '=') {
            if (side) return fail(m, "more than one relational operator in \"%s\"", s);
            int code = (ch == '=') ? 3 : (ch == '<' ? -1 : 1);
            ++i;
            if (ch != '=' && i < len && s[i] == '=') { code *= 2; ++i; }
            f->rel = code;
            side = 1;
            continue;
        }
        /* a term: sign, number, variable */
        char num[96], var[32];
        size_t nn = 0, nv = 0;
        if (ch == '+' || ch == '-') { num[nn++] = ch; ++i; }
        while (i < len && (isdigit((unsigned char)s[i]) || s[i] == '.' || s[i] == '/') && nn + 1 < sizeof num) num[nn++] = s[i++];
        num[nn] = '\0';
        while (i < len && !strchr("+-<>=", s[i]) && nv + 1 < sizeof var) var[nv++] = s[i++];
        var[nv] = '\0';
        if (nn == 0 && nv == 0) return fail(m, "empty term in \"%s\"", s);
        term_t t;
        memset(&t, 0, sizeof t);
        if (rat_parse(num, &t.c) != 0) return fail(m, "bad coefficient \"%s\" in \"%s\"", num, s);
        if (nv) {
            var[3] = '\0';                       /* at most three characters are kept */
            if (!valid_var(var)) return fail(m, "Invalid variable name: %s", var);
            strcpy(t.var, var);
        } else if (nn == 1 && (num[0] == '+' || num[0] == '-')) {
            return fail(m, "dangling sign in \"%s\"", s);
        }
        terms_push(side ? &f->right : &f->left, t, -1);
    }
    if (!f->rel || f->left.n < 1 || f->right.n < 1) return fail(m, "Formula invalid: \"%s\"", s);
    /* divide every numerator by their common GCD and every denominator by theirs
     * (Source/dataReader.c:395-432) */
    long gn = 0, gd = 0;
    for (int sd = 0; sd < 2; ++sd) {
        terms_t *v = sd ? &f->right : &f->left;
        for (int k = 0; k < v->n; ++k) { gn = gcd_l(gn, v->t[k].c.num); gd = gcd_l(gd, v->t[k].c.den); }
    }
    if (gn > 1 || gd > 1)
        for (int sd = 0; sd < 2; ++sd) {
            terms_t *v = sd ? &f->right : &f->left;
            for (int k = 0; k < v->n; ++k) {
                if (gn > 1) v->t[k].c.num /= gn;
                if (gd > 1) v->t[k].c.den /= gd;
            }
        }
    return 0;
}

/* merge terms with the same variable name, drop the ones that cancel (Source/basicFuncs.c:338-366) */
static int combine_like(terms_t *v, lpf_model *m) {
    for (int j = 0; j < v->n; ++j) {
        for (int k = j + 1; k < v->n; ++k)
            if (strcmp(v->t[j].var, v->t[k].var) == 0) {
                if (rat_add(v->t[j].c, v->t[k].c, &v->t[j].c) != 0) return fail(m, "coefficient overflow while combining terms");
                terms_remove(v, k);
                --k;
            }
        if (v->t[j].c.num == 0) { terms_remove(v, j); --j; }
    }
    return 0;
}

/* ------------------------------------------------------------------ variable dictionary */

typedef struct {
    lpf_var *v;
    int n, cap;
    long max_x_sub;       /* largest subscript among names starting with 'x' (Source/hashTable.c:229-233) */
} dict_t;

static lpf_var *dict_get(dict_t *d, const char *name) {
    for (int i = 0; i < d->n; ++i)
        if (strcmp(d->v[i].name, name) == 0) return &d->v[i];
    return NULL;
}

static lpf_var *dict_put(dict_t *d, const char *name, int relation) {
    lpf_var *e = dict_get(d, name);
    if (!e) {
        if (d->n == d->cap) {
            d->cap = d->cap ? d->cap * 2 : 16;
            d->v = (lpf_var *)realloc(d->v, sizeof(lpf_var) * (size_t)d->cap);
        }
        e = &d->v[d->n++];
        memset(e, 0, sizeof *e);
        strncpy(e->name, name, sizeof e->name - 1);
        if (name[0] == 'x') {
            long sub = strtol(name + 1, NULL, 10);
            if (sub > d->max_x_sub) d->max_x_sub = sub;
        }
    }
    e->relation = relation;
    return e;
}

/* column order of the standard form: first letter, then numeric subscript
 * (Source/simplex.c:316-325); the constant term (empty name) sorts first */
static int var_cmp(const char *a, const char *b) {
    if (a[0] != b[0]) return (unsigned char)a[0] > (unsigned char)b[0] ? 1 : -1;
    long sa = strlen(a) > 1 ? strtol(a + 1, NULL, 10) : 0, sb = strlen(b) > 1 ? strtol(b + 1, NULL, 10) : 0;
    return (sa > sb) - (sa < sb);
}

static void terms_sort(terms_t *v) {
    for (int i = 1; i < v->n; ++i) {
        term_t x = v->t[i];
        int j = i - 1;
        while (j >= 0 && var_cmp(v->t[j].var, x.var) > 0) { v->t[j + 1] = v->t[j]; --j; }
        v->t[j + 1] = x;
    }
}

static void new_x_name(dict_t *d, char *out) {
    snprintf(out, 8, "x%d", (int)(++d->max_x_sub % 100000));
    dict_put(d, out, 2);
}

/* ------------------------------------------------------------------ text -> statements */

typedef struct {
    formula_t objective;
    int have_objective;
    int sense;
    formula_t *st;
    int nst, cap;
    int nuser;
} raw_model;

static void raw_free(raw_model *r) {
    if (r->have_objective) formula_free(&r->objective);
    for (int i = 0; i < r->nst; ++i) formula_free(&r->st[i]);
    free(r->st);
    memset(r, 0, sizeof *r);
}

/* The section name is the blank-delimited token in front of '{'; inside the braces blanks are
 * dropped and ';' separates statements; text outside the two sections is ignored
 * (Source/dataReader.c:26-37,174-213). */
static int split_sections(const char *text, raw_model *r, lpf_model *m) {
    const char *p = text;
    char token[64], last[64] = "";
    size_t nt = 0;
    while (*p) {
        if (*p == '{') {
            token[nt] = '\0';
            if (nt == 0) strcpy(token, last);        /* "OF {" : the name ended at the blank */
            last[0] = '\0';
            int section = strcmp(token, "OF") == 0 ? 1 : strcmp(token, "ST") == 0 ? 2 : 0;
            nt = 0;
            const char *q = strchr(p, '}');
            if (!q) q = p + strlen(p);
            char *stmt = (char *)malloc((size_t)(q - p) + 1);
            size_t ns = 0;
            for (const char *c = p + 1; c <= q; ++c) {
                if (c < q && *c != ';') {
                    if (!isspace((unsigned char)*c)) stmt[ns++] = *c;
                    continue;
                }
                stmt[ns] = '\0';
                if (ns > 0 && section == 1) {
                    char *colon = strchr(stmt, ':');
                    if (!colon) { free(stmt); return fail(m, "Objective function invalid."); }
                    *colon = '\0';
                    int sense = strcmp(stmt, "max") == 0 ? 1 : strcmp(stmt, "min") == 0 ? -1 : 0;
                    if (!sense) { free(stmt); return fail(m, "Objective function invalid."); }   /* e.g. "ma:" (testdata.txt:2) */
                    if (r->have_objective) { free(stmt); return fail(m, "There can be only ONE Objective function!"); }
                    if (parse_formula(colon + 1, &r->objective, m) != 0) { formula_free(&r->objective); free(stmt); return -1; }
                    r->have_objective = 1;
                    r->sense = sense;
                    const formula_t *f = &r->objective;
                    if (f->rel != 3) { free(stmt); return fail(m, "Wrong relational operator in Objective function!"); }
                    if (f->left.n != 1 || f->left.t[0].c.num != f->left.t[0].c.den) { free(stmt); return fail(m, "Non-standard Objective function!"); }
                } else if (ns > 0 && section == 2) {
                    if (r->nst == r->cap) {
                        r->cap = r->cap ? r->cap * 2 : 8;
                        r->st = (formula_t *)realloc(r->st, sizeof(formula_t) * (size_t)r->cap);
                    }
                    if (parse_formula(stmt, &r->st[r->nst], m) != 0) { formula_free(&r->st[r->nst]); free(stmt); return -1; }
                    r->nst++;
                }
                ns = 0;
            }
            free(stmt);
            p = *q ? q + 1 : q;
            continue;
        }
        if (isspace((unsigned char)*p)) {
            if (nt) { token[nt] = '\0'; strcpy(last, token); }
            nt = 0;
        } else if (nt + 1 < sizeof token) token[nt++] = *p;
        ++p;
    }
    if (!r->have_objective) return fail(m, "MISSING DATA: Objective Function not found.");
    if (r->nst == 0) return fail(m, "MISSING DATA: Constraints not found.");
    return 0;
}

/* ------------------------------------------------------------------ normalise, standardise, align */

/* constants to the right, variables to the left, like terms merged, sign rows "x >= 0" / "x <= 0"
 * pulled out into the dictionary (Source/dataReader.c:45-140) */
static int normalise(raw_model *r, dict_t *d, lpf_model *m) {
    if (combine_like(&r->objective.right, m) != 0) return -1;
    for (int i = 0; i < r->nst; ++i) {
        formula_t *f = &r->st[i];
        for (int j = 0; j < f->left.n; ++j)
            if (f->left.t[j].var[0] == '\0') {
                term_t t = f->left.t[j];
                t.c = rat_neg(t.c);
                terms_remove(&f->left, j--);
                terms_push(&f->right, t, -1);
            }
        for (int j = 0; j < f->right.n; ++j)
            if (f->right.t[j].var[0] != '\0') {
                term_t t = f->right.t[j];
                t.c = rat_neg(t.c);
                terms_remove(&f->right, j--);
                terms_push(&f->left, t, -1);
            }
        if (f->left.n < 1 || f->right.n < 1) return fail(m, "LPModel invalid due to the incomplete CONSTRAINT (ST Line: %d).", i + 1);
        if (combine_like(&f->left, m) != 0) return -1;
        for (int j = f->right.n - 1; j > 0; --j) {
            if (rat_add(f->right.t[0].c, f->right.t[j].c, &f->right.t[0].c) != 0) return fail(m, "right-hand side overflow (ST Line: %d)", i + 1);
            f->right.n--;
        }
        if (f->left.n < 1) return fail(m, "No term left in the left hand side of the CONSTRAINT (ST Line: %d) after combining similar terms.", i + 1);
        for (int j = 0; j < f->left.n; ++j)
            if (!dict_get(d, f->left.t[j].var)) dict_put(d, f->left.t[j].var, 0);      /* unrestricted until told otherwise */
        if (f->left.n == 1 && f->left.t[0].c.num == f->left.t[0].c.den && f->right.t[0].c.num == 0 && abs(f->rel) == 2) {
            dict_put(d, f->left.t[0].var, f->rel);
            formula_free(f);
            memmove(&r->st[i], &r->st[i + 1], sizeof(formula_t) * (size_t)(r->nst - i - 1));
            r->nst--;
            --i;
        }
    }
    int ofvars = 0;
    for (int j = 0; j < r->objective.right.n; ++j) {
        const char *v = r->objective.right.t[j].var;
        if (!v[0]) continue;
        ++ofvars;
        if (!dict_get(d, v)) return fail(m, "Mismatch in the number of variables in the Objective Function and Constraints.");
    }
    if (ofvars != d->n) return fail(m, "Mismatch in the number of variables in the Objective Function and Constraints.");
    if (r->nst < 1) return fail(m, "MISSING DATA: Constraints not found.");
    return 0;
}

static void negate_terms(terms_t *v) {
    for (int j = 0; j < v->n; ++j) v->t[j].c = rat_neg(v->t[j].c);
}

/* min -> max(-z); x(unr) -> x'' - x'; b < 0 -> negate the row; slack / surplus; sort; x <= 0 -> -x'
 * (Source/simplex.c:91-230).  dual = 1 is the reference's variant for the dual simplex
 * (Source/simplex.c:177-178): rows are NOT made b >= 0, instead every ">=" row is negated into "<=" so
 * that each inequality row gets a +1 slack. */
static void standardise(raw_model *r, dict_t *d, int dual) {
    terms_t *of = &r->objective.right;
    if (r->sense != 1) negate_terms(of);
    r->nuser = d->n;                                /* everything added below is generated */
    for (int i = 0; i < of->n; ++i) {
        lpf_var *e = of->t[i].var[0] ? dict_get(d, of->t[i].var) : NULL;
        if (!e || e->relation != 0) continue;
        char target[8], former[8], latter[8];
        strcpy(target, of->t[i].var);
        new_x_name(d, former);
        new_x_name(d, latter);
        e = dict_get(d, target);                    /* the dictionary may have moved */
        strcpy(e->former, former);
        strcpy(e->latter, latter);
        term_t neg = of->t[i];
        strcpy(of->t[i].var, former);
        strcpy(neg.var, latter);
        neg.c = rat_neg(neg.c);
        terms_push(of, neg, i + 1);
        ++i;
        for (int s = 0; s < r->nst; ++s) {
            terms_t *l = &r->st[s].left;
            for (int k = 0; k < l->n; ++k)
                if (strcmp(l->t[k].var, target) == 0) {
                    term_t n2 = l->t[k];
                    strcpy(l->t[k].var, former);
                    strcpy(n2.var, latter);
                    n2.c = rat_neg(n2.c);
                    terms_push(l, n2, k + 1);
                    ++k;
                }
        }
    }
    for (int s = 0; s < r->nst; ++s) {
        formula_t *f = &r->st[s];
        if ((!dual && rat_sign(f->right.t[0].c) < 0) || (dual && f->rel > 0 && f->rel != 3)) {
            f->right.t[0].c = rat_neg(f->right.t[0].c);
            negate_terms(&f->left);
            if (f->rel != 3) f->rel = -f->rel;
        }
        if (f->rel != 3) {
            term_t sl;
            memset(&sl, 0, sizeof sl);
            new_x_name(d, sl.var);
            sl.c = rat_int(0);
            terms_push(of, sl, -1);
            sl.c = rat_int(f->rel > 0 ? -1 : 1);          /* surplus for > / >=, slack for < / <= */
            terms_push(&f->left, sl, -1);
            f->rel = 3;
        }
        terms_sort(&f->left);
    }
    terms_sort(of);
    for (int s = 0; s <= r->nst; ++s) {
        terms_t *v = s < r->nst ? &r->st[s].left : of;
        for (int k = 0; k < v->n; ++k) {
            lpf_var *e = v->t[k].var[0] ? dict_get(d, v->t[k].var) : NULL;
            if (e && e->relation < 0) { v->t[k].c = rat_neg(v->t[k].c); v->t[k].inverted = 1; }
        }
    }
}

static int finish_model(raw_model *r, dict_t *d, lpf_model *m) {
    terms_t *of = &r->objective.right;
    m->constant = rat_int(0);
    for (int j = 0; j < of->n; ++j)
        if (!of->t[j].var[0]) { m->constant = of->t[j].c; terms_remove(of, j--); }      /* Source/matrix.c:23-28 */
    m->sense = r->sense;
    m->m = r->nst;
    m->n = of->n;
    m->names = (char (*)[8])calloc((size_t)m->n, 8);
    m->inverted = (int *)calloc((size_t)m->n, sizeof(int));
    m->costs = (lpf_rat *)calloc((size_t)m->n, sizeof(lpf_rat));
    m->rows = (lpf_rat *)calloc((size_t)m->m * (size_t)(m->n + 1), sizeof(lpf_rat));
    m->basis = (int *)calloc((size_t)m->m, sizeof(int));
    for (int j = 0; j < m->n; ++j) {
        strcpy(m->names[j], of->t[j].var);
        m->inverted[j] = of->t[j].inverted;
        m->costs[j] = of->t[j].c;
    }
    /* align: every row gets every column in objective order, zeros where absent (simplex.c:238-260) */
    for (int i = 0; i < m->m; ++i) {
        lpf_rat *row = m->rows + (size_t)i * (size_t)(m->n + 1);
        row[0] = r->st[i].right.t[0].c;
        for (int j = 0; j < m->n; ++j) row[1 + j] = rat_int(0);
        terms_t *l = &r->st[i].left;
        for (int k = 0; k < l->n; ++k) {
            int j = 0;
            while (j < m->n && strcmp(m->names[j], l->t[k].var) != 0) ++j;
            if (j == m->n) return fail(m, "internal: constraint variable %s not in the objective", l->t[k].var);
            row[1 + j] = l->t[k].c;
        }
    }
    /* initial basis: exact unit columns with zero cost; the last such column of a row wins, as the
     * column scan of Source/matrix.c:50-78 would leave it */
    for (int j = 0; j < m->n; ++j) {
        if (m->costs[j].num != 0) continue;
        int one = -1, ok = 1;
        for (int i = 0; i < m->m && ok; ++i) {
            lpf_rat a = m->rows[(size_t)i * (size_t)(m->n + 1) + 1 + j];
            if (a.num == 0) continue;
            if (a.num == a.den && one < 0) one = i;
            else ok = 0;
        }
        if (ok && one >= 0) m->basis[one] = j + 1;
    }
    m->nvars = d->n;
    m->nuser = r->nuser;
    m->vars = (lpf_var *)malloc(sizeof(lpf_var) * (size_t)(d->n ? d->n : 1));
    memcpy(m->vars, d->v, sizeof(lpf_var) * (size_t)d->n);
    m->valid = 1;
    return 0;
}

int lpf_parse_string(const char *text, lpf_model *m) { return lpf_parse_string_mode(text, 0, m); }

int lpf_parse_string_mode(const char *text, int dual, lpf_model *m) {
    memset(m, 0, sizeof *m);
    if (!text) return fail(m, "no input");
    raw_model r;
    dict_t d;
    memset(&r, 0, sizeof r);
    memset(&d, 0, sizeof d);
    int rc = split_sections(text, &r, m);
    if (rc == 0) rc = normalise(&r, &d, m);
    if (rc == 0) { standardise(&r, &d, dual); rc = finish_model(&r, &d, m); }
    raw_free(&r);
    free(d.v);
    if (rc != 0) { char e[256]; memcpy(e, m->error, sizeof e); lpf_free(m); memcpy(m->error, e, sizeof e); }
    return rc;
}

int lpf_parse_file(const char *path, lpf_model *m) { return lpf_parse_file_mode(path, 0, m); }

int lpf_parse_file_mode(const char *path, int dual, lpf_model *m) {
    memset(m, 0, sizeof *m);
    FILE *f = fopen(path, "rb");
    if (!f) return fail(m, "File not readable or does not exist.");
    long sz = -1;
    if (fseek(f, 0, SEEK_END) == 0) sz = ftell(f);
    if (sz < 0 || fseek(f, 0, SEEK_SET) != 0) { fclose(f); return fail(m, "File not readable or does not exist."); }   /* unseekable path */
    char *buf = (char *)malloc((size_t)sz + 1);
    if (!buf) { fclose(f); return fail(m, "Out of memory while reading the model file."); }
    size_t got = fread(buf, 1, (size_t)sz, f);
    buf[got] = '\0';
    fclose(f);
    int rc = lpf_parse_string_mode(buf, dual, m);
    free(buf);
    return rc;
}

void lpf_free(lpf_model *m) {
    if (!m) return;
    free(m->names); free(m->inverted); free(m->costs); free(m->rows); free(m->basis); free(m->vars);
    memset(m, 0, sizeof *m);
}

/* ------------------------------------------------------------------ dense tableau */

int lpf_build_tableau(const lpf_model *mo, int method, double **Tout, int **bout, int *n_total, int *n_art) {
    if (!mo || !mo->valid || !Tout || !bout || method < 0 || method > 2) return -1;
    const int m = mo->m, n0 = mo->n, R = method == 0 ? 1 : 2;
    if (m < 1 || n0 < 1) return -1;
    int na = 0;
    for (int i = 0; i < m; ++i) na += (mo->basis[i] == 0);
    if (method == 0 && na) return -1;
    const int n = n0 + na;
    const size_t ld = (size_t)n + 1;
    double *T = (double *)calloc((size_t)(m + R) * ld, sizeof(double));
    int *basis = (int *)calloc((size_t)m, sizeof(int));
    int next_art = n0 + 1;
    for (int i = 0; i < m; ++i) {
        const lpf_rat *row = mo->rows + (size_t)i * (size_t)(n0 + 1);
        for (int j = 0; j <= n0; ++j) T[(size_t)i * ld + j] = rat_val(row[j]);
        if (mo->basis[i]) basis[i] = mo->basis[i];
        else { T[(size_t)i * ld + next_art] = 1.0; basis[i] = next_art++; }
    }
    for (int j = 0; j < n0; ++j) T[(size_t)m * ld + 1 + j] = rat_val(mo->costs[j]);
    if (R == 2) {
        /* artificial cost -M: M-row = -1 on artificial columns, priced out by adding every
         * artificial-basic row, rows ascending (same order as lpgpu_build_tableau) */
        double *om = T + (size_t)(m + 1) * ld;
        for (int j = n0 + 1; j <= n; ++j) om[j] = -1.0;
        for (int i = 0; i < m; ++i) {
            if (basis[i] <= n0) continue;
            for (int j = 0; j <= n; ++j) om[j] += T[(size_t)i * ld + j];
        }
    }
    *Tout = T;
    *bout = basis;
    if (n_total) *n_total = n;
    if (n_art) *n_art = na;
    return 0;
}
