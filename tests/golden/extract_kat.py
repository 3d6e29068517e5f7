#!/usr/bin/env python3
"""Extract the reference's known-answer distance vectors into tests/golden/kat.json.

Source: the *_General tests of /root/reference/tests/ailego/math/
{euclidean_distance_matrix,inner_product_matrix}_{int8,fp32,fp16}_test.cc
(hard-coded operands + EXPECT_FLOAT_EQ expected values).  Only numbers are taken.
Run in the build container (needs /root/reference); the JSON is committed.
"""
import json
import os
import re
import sys

REF = os.environ.get("ZVEC_REFERENCE", "/root/reference")
FILES = {
    "int8": ["euclidean_distance_matrix_int8_test.cc", "inner_product_matrix_int8_test.cc"],
    "fp32": ["euclidean_distance_matrix_fp32_test.cc", "inner_product_matrix_fp32_test.cc"],
    "fp16": ["euclidean_distance_matrix_fp16_test.cc", "inner_product_matrix_fp16_test.cc"],
}
FUNCS = {"EuclideanDistance": "L2", "SquaredEuclideanDistance": "L2SQ",
         "InnerProductDistance": "IP_POS", "MinusInnerProductDistance": "IP"}


def nums(s):
    s = re.sub(r"\(\s*int8_t\s*\)", "", s)
    out = []
    for tok in s.split(","):
        tok = tok.strip().rstrip("fF")
        if tok:
            out.append(float(tok))
    return out


def test_bodies(text):
    for m in re.finditer(r"\bTEST\s*\(\s*(\w+)\s*,\s*(\w+)\s*\)\s*\{", text):
        i = m.end()
        depth = 1
        while depth and i < len(text):
            depth += {"{": 1, "}": -1}.get(text[i], 0)
            i += 1
        yield m.group(2), text[m.end():i - 1], text[:m.start()].count("\n") + 1


def main():
    cases = []
    for dtype, files in FILES.items():
        for fn in files:
            path = os.path.join(REF, "tests/ailego/math", fn)
            if not os.path.exists(path):
                continue
            text = open(path).read()
            for tname, body, line in test_bodies(text):
                if not tname.endswith("_General"):
                    continue
                vecs = {}
                for m in re.finditer(r"FixedVector<\s*\w+\s*,\s*(\d+)\s*>\s*([^;]*);", body):
                    n = int(m.group(1))
                    for d in re.finditer(r"(\w+)\s*\{([^}]*)\}", m.group(2)):
                        v = nums(d.group(2))
                        vecs[d.group(1)] = (v + [0.0] * n)[:n]
                for m in re.finditer(r"\b(?:int8_t|float)\s+(\w+)\s*\[\s*\]\s*=\s*\{([^}]*)\}", body):
                    vecs[m.group(1)] = nums(m.group(2))
                for m in re.finditer(r"EXPECT_FLOAT_EQ\(\s*(-?[\d.]+(?:e[-+]?\d+)?)f?\s*,\s*(\w+)\s*\((.*?)\)\s*\)\s*;",
                                     body, re.S):
                    func = m.group(2)
                    if func not in FUNCS:
                        continue
                    args = m.group(3)
                    names = re.findall(r"Cast\(\s*(\w+)\s*\)", args) or [a.strip() for a in args.split(",")]
                    sizes = [int(x) for x in re.findall(r"FixedVector<\s*\w+\s*,\s*(\d+)\s*>::Cast", args)]
                    if len(names) != 2 or names[0] not in vecs or names[1] not in vecs:
                        continue
                    a, b = vecs[names[0]], vecs[names[1]]
                    n = sizes[0] if sizes else len(a)
                    if len(a) < n or len(b) < n:
                        continue  # e.g. the 2048-d arrays whose tails are elided
                    cases.append({"dtype": dtype, "metric": FUNCS[func], "a": a[:n], "b": b[:n],
                                  "expected": float(m.group(1)), "source": "tests/ailego/math/%s:%d (%s)" % (fn, line, tname)})
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kat.json")
    json.dump(cases, open(out, "w"), indent=0)
    by = {}
    for c in cases:
        by[(c["dtype"], c["metric"])] = by.get((c["dtype"], c["metric"]), 0) + 1
    print(len(cases), "cases", by, file=sys.stderr)


if __name__ == "__main__":
    main()
