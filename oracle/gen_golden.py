#!/usr/bin/env python3
"""Generate tests/golden/*.json from the UNMODIFIED reference (test infrastructure).

Runs only in the build container, where /root/reference exists: `make -C oracle ref`
compiles the reference's own sources in place into oracle/_ref/ref_dump and
oracle/_ref/ref_kat; this script drives them and commits what they print, so that the
golden vectors travel to the GPU box (where /root/reference does not exist).

  frontend_dumps.json  Parser -> LPTrans -> LPStandardize -> LPAlign -> CreateSMatrix
                       (Source/main.c:16-19, Source/simplex.c:27-40) per model file in
                       tests/golden/models/  -- SURVEY.md Appendix B.
  rational_kat.json    NAdd/NSub/NMul/NDiv/Fractionize/Decimalize known answers
                       (Source/numOprts.c:137-294, Source/basicFuncs.c:165-313)
                       -- SURVEY.md Appendix A.
"""
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, "tests", "golden")

KAT = [
    # (op, a, b)
    ("parse", "1/3", "0"), ("parse", "6/4", "0"), ("parse", "-2.45", "0"), ("parse", "0.1", "0"),
    ("parse", "0.29", "0"), ("parse", "0", "0"), ("parse", "0/5", "0"), ("parse", "5", "0"),
    ("parse", "-7", "0"), ("parse", "2.5", "0"), ("parse", "-", "0"), ("parse", "+", "0"),
    ("parse", "10/4", "0"), ("parse", "-3/9", "0"), ("parse", "3/-4", "0"), ("parse", "1.250", "0"),
    ("parse", "12x", "0"), ("parse", "3M", "0"), ("parse", "-M", "0"), ("parse", "3M/4", "0"),
    ("add", "1/3", "1/2"), ("mul", "2/3", "5/4"), ("sub", "51/14", "1/6"), ("div", "1", "-2"),
    ("div", "5", "0"), ("div", "0", "7"), ("mul", "4611686018427387904", "4"),
    ("add", "1/4294967311", "1/4294967357"), ("add", "-1/2", "1/3"), ("sub", "1/3", "1/3"),
    ("mul", "0", "5/7"), ("mul", "-3/7", "14/9"), ("div", "51/14", "9/14"), ("div", "-1/3", "2"),
    ("add", "9223372036854775807", "1"), ("sub", "-9223372036854775807", "2"),
    ("mul", "3037000500", "3037000500"), ("mul", "-3037000500", "3037000500"),
    ("add", "17/3", "19/3"), ("sub", "0", "5/8"), ("div", "7/3", "7/3"), ("add", "1/6", "1/10"),
    ("mul", "6/35", "35/6"), ("div", "2/4", "4/8"), ("sub", "1/1000000007", "1/998244353"),
    ("add", "3M", "1"), ("mul", "3M", "2"), ("mul", "0", "M"),
]


def run_kat():
    exe = os.path.join(HERE, "_ref", "ref_kat")
    inp = "".join(f"{op} {a} {b}\n" for op, a, b in KAT)
    p = subprocess.run([exe], input=inp, capture_output=True, text=True, check=True)
    lines = [l for l in p.stderr.splitlines() if l.strip()]
    assert len(lines) == len(KAT), (len(lines), len(KAT))
    out = []
    for (op, a, b), l in zip(KAT, lines):
        num, den, valid, has_m, dec = l.split()
        out.append({"op": op, "a": a, "b": b, "num": int(num), "den": int(den),
                    "valid": int(valid), "has_m": int(has_m), "dec": float(dec)})
    return out


def run_dumps(dual=0):
    """dual = 1: the reference's LPStandardize(model, dual = 1) branch (Source/simplex.c:177-178)."""
    exe = os.path.join(HERE, "_ref", "ref_dump")
    mdir = os.path.join(GOLD, "models")
    out = {}
    for name in sorted(os.listdir(mdir)):
        if not name.endswith(".txt"):
            continue
        with tempfile.NamedTemporaryFile(suffix=".json") as tf:
            subprocess.run([exe, os.path.join(mdir, name), tf.name, str(dual)], capture_output=True, check=True)
            out[name] = json.load(open(tf.name))
    return out


def main():
    if not os.path.isdir("/root/reference/Source"):
        sys.exit("reference checkout absent: golden vectors can only be regenerated in the build container")
    subprocess.run(["make", "-C", HERE, "ref"], check=True, capture_output=True)
    os.makedirs(GOLD, exist_ok=True)
    with open(os.path.join(GOLD, "rational_kat.json"), "w") as f:
        json.dump(run_kat(), f, indent=1)
    with open(os.path.join(GOLD, "frontend_dumps.json"), "w") as f:
        json.dump(run_dumps(), f, indent=1)
    with open(os.path.join(GOLD, "frontend_dumps_dual.json"), "w") as f:
        json.dump(run_dumps(dual=1), f, indent=1)
    print("wrote", GOLD)


if __name__ == "__main__":
    main()
