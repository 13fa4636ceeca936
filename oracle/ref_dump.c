/*
 * ref_dump.c -- harness around the UNMODIFIED reference front end (test infrastructure).
 *
 * Compiled by oracle/Makefile against the reference sources where they lie
 * (/root/reference/Source, never copied) into oracle/_ref/ref_dump.  It runs
 *   Parser -> LPTrans -> CopyModel -> LPStandardize(.,dual) -> LPAlign -> CreateSMatrix
 * (Source/main.c:16-19, Source/simplex.c:27-40) on one model file and writes what the
 * reference built as JSON, so that tests/golden/ can pin this repository's own front end
 * and tableau layout against real reference behaviour.
 *
 * usage: ref_dump <model.txt> <out.json>
 */
#include "public.h"
#include "simplex.h"
#include "matrix.h"

static void put_num(FILE *o, Number v) {
    fprintf(o, "[%ld,%ld,%d]", v.numerator, v.denominator, (int)(v.valid && v.constant == NULL));
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s model.txt out.json [dual]\n", argv[0]); return 2; }
    const short dual = (argc > 3 && argv[3][0] == '1') ? 1 : 0;   /* LPStandardize's second argument */
    FILE *in = fopen(argv[1], "r");
    FILE *o = fopen(argv[2], "w");
    if (!in || !o) { fprintf(stderr, "cannot open files\n"); return 2; }
    InitVarDict();
    LPModel parsed = Parser(in);
    if (parsed.valid) LPTrans(&parsed);
    fclose(in);
    if (!parsed.valid) { fprintf(o, "{\"parsed\":0}\n"); fclose(o); return 0; }

    LPModel mc = CopyModel(&parsed);
    const int sense = parsed.objective.type;        /* 1 max, -1 min, before standardising */
    LPStandardize(&mc, dual);
    LPAlign(&mc);
    if (!mc.valid) { fprintf(o, "{\"parsed\":1,\"standard\":0}\n"); fclose(o); return 0; }

    /* objective constant that CreateSMatrix strips (Source/matrix.c:23-28) */
    Number constant = Fractionize("0");
    for (size_t i = 0; i < mc.objective.rightLen; i++)
        if (strlen(mc.objective.right[i]->variable) == 0) constant = mc.objective.right[i]->coefficient;

    /* Source/matrix.c:86 writes through lack[lackPtr] instead of (*lack)[lackPtr]; give it
     * a slot per row so that the write stays inside this harness' memory. */
    size_t scratch[256];
    size_t *slots[256];
    for (int i = 0; i < 256; i++) { scratch[i] = (size_t)-1; slots[i] = &scratch[i]; }
    short int valid = 1;
    SimplexMatrix sm = CreateSMatrix(&mc, slots, &valid);

    fprintf(o, "{\"parsed\":1,\"standard\":1,\"sense\":%d,\"m\":%zu,\"n\":%zu,\"valid\":%d,\n",
            sense, sm.basicLen, sm.ofLen, (int)valid);
    fprintf(o, " \"constant\":");
    put_num(o, constant);
    fprintf(o, ",\n \"names\":[");
    for (size_t j = 0; j < sm.ofLen; j++) fprintf(o, "%s\"%s\"", j ? "," : "", sm.varNames[j]);
    fprintf(o, "],\n \"inverted\":[");
    for (size_t j = 0; j < sm.ofLen; j++) fprintf(o, "%s%d", j ? "," : "", (int)mc.objective.right[j]->inverted);
    fprintf(o, "],\n \"costs\":[");
    for (size_t j = 0; j < sm.ofLen; j++) { if (j) fprintf(o, ","); put_num(o, *sm.ofCosts[j]); }
    fprintf(o, "],\n \"rows\":[");
    for (size_t i = 0; i < sm.basicLen; i++) {
        fprintf(o, "%s\n  [", i ? "," : "");
        for (size_t j = 0; j <= sm.ofLen; j++) { if (j) fprintf(o, ","); put_num(o, *sm.cMatrix[i][j]); }
        fprintf(o, "]");
    }
    fprintf(o, "],\n \"basis\":[");
    for (size_t i = 0; i < sm.basicLen; i++) {
        if (sm.basicVars[i]) fprintf(o, "%s\"%s\"", i ? "," : "", sm.basicVars[i]);
        else fprintf(o, "%snull", i ? "," : "");
    }
    fprintf(o, "],\n \"vars\":[");
    size_t nv = 0;
    short int vv = 1;
    VarItem **items = GetVarItems(&nv, &vv);
    for (size_t i = 0; i < nv; i++)
        fprintf(o, "%s{\"name\":\"%s\",\"relation\":%d,\"former\":\"%s\",\"latter\":\"%s\"}", i ? "," : "",
                items[i]->keyName, (int)items[i]->relation, items[i]->formerX, items[i]->latterX);
    free(items);
    fprintf(o, "]}\n");
    fclose(o);
    return 0;
}
