/*
 * ref_kat.c -- known-answer harness around the UNMODIFIED reference arithmetic (test
 * infrastructure).  Compiled by oracle/Makefile against /root/reference/Source
 * (numOprts.c, basicFuncs.c, hashTable.c, dataReader.c for the `constants` globals) into
 * oracle/_ref/ref_kat.  Reads lines "op a b" on stdin where op is one of
 * add sub mul div parse dec and a, b are strings accepted by Fractionize
 * (Source/basicFuncs.c:165-291); writes "num den valid hasM [double]" per line.
 * Used by oracle/gen_golden.py to produce tests/golden/rational_kat.json
 * (SURVEY.md Appendix A).
 */
#include "public.h"

int main(void) {
    /* register the constant M the way Parser does (Source/dataReader.c:165-173) */
    constants = (Constant *)calloc(CONSTANTS_LEN_PER_ALLOC, sizeof(Constant));
    Constant big = { .name = 'M', .relation = 1, .val = Fractionize("0") };
    constants[constantsNum++] = big;

    char op[16], a[64], b[64];
    while (scanf("%15s %63s %63s", op, a, b) == 3) {
        Number x = Fractionize(a), y = Fractionize(b), r = x;
        if (!strcmp(op, "add")) r = NAdd(x, y);
        else if (!strcmp(op, "sub")) r = NSub(x, y);
        else if (!strcmp(op, "mul")) r = NMul(x, y);
        else if (!strcmp(op, "div")) r = NDiv(x, y);
        else if (!strcmp(op, "parse") || !strcmp(op, "dec")) r = x;
        fprintf(stderr, "%ld %ld %d %d %.17g\n", r.numerator, r.denominator, (int)r.valid,
                (int)(r.constant != NULL),
                (r.valid && r.constant == NULL && r.denominator) ? Decimalize(r) : 0.0);
    }
    return 0;
}
