"""Regenerates tests/golden/*.json.  Run from the repo root:  python tests/golden/make_golden.py

kat_v1.json  -- known answers derived BY HAND from the reference source
                (Helpers/RagIndex.cs:59-67,124-135): expected values below are written as
                closed-form Python float expressions, not produced by either oracle.
synth_v1.json -- regression vectors: the synthetic generator's first values and the numpy
                oracle's answers on a small seeded corpus (pins C oracle == numpy oracle ==
                CUDA path on identical inputs).  The reference itself cannot run here
                (no dotnet), so these are NOT reference outputs: parity stays "unpinned".
"""
import json
import math
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import oracle_np as onp  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
EPS = 1e-9


def f32(x):
    return float(np.float32(x))


def kat():
    cases = []

    def add(name, rows, q, k, idx, scores, why, pins=None):
        cases.append(dict(name=name, rows=rows, query=q, top_k=k, expect_idx=idx,
                          expect_score=scores, why=why))
        if pins:   # which of the C#-specific facts the equivalence argument rests on this case exercises
            cases[-1]["pins_csharp_fact"] = pins

    one = 1.0 / (1.0 * 1.0 + EPS)
    add("onehot_orthogonal", [[1, 0, 0], [0, 1, 0], [0, 0, 1]], [0, 1, 0], 3,
        [1, 0, 2], [one, 0.0, 0.0],
        "dot=1,na=nb=1 -> 1/(1+1e-9); orthogonal rows score +0.0 and keep insertion order (:64 stable)")
    add("ties_lower_index_first", [[1, 1], [2, 2], [1, 1], [3, 3]], [1, 1], 4,
        None, None, "parallel rows: scores computed below; equal scores keep lower index first")
    # rows 0 and 2 are bit-identical -> exactly equal scores; fill expectations by formula
    def cos_formula(a, b):
        dot = na = nb = 0.0
        for x, y in zip(a, b):
            dot += f32(np.float32(x) * np.float32(y))
            na += f32(np.float32(x) * np.float32(x))
            nb += f32(np.float32(y) * np.float32(y))
        return dot / (math.sqrt(na) * math.sqrt(nb) + EPS)
    c = cases[-1]
    sc = [cos_formula(c["query"], r) for r in c["rows"]]
    order = sorted(range(4), key=lambda i: (-sc[i], i))
    c["expect_idx"], c["expect_score"] = order, [sc[i] for i in order]
    assert order.index(0) < order.index(2)
    c["pins_csharp_fact"] = "System.Linq ordering is stable; Take(k) after it keeps the first k in source order among equal keys"

    add("topk_zero_gives_one", [[1, 0], [0, 1]], [0, 1], 0, [1], [one], "Math.Max(1, topK) (:65)")
    add("topk_negative_gives_one", [[1, 0], [0, 1]], [1, 0], -7, [0], [one], "Math.Max(1, topK) (:65)")
    add("topk_larger_than_n", [[1, 0], [0, 1]], [1, 0], 10, [0, 1], [one, 0.0], "Take(k) over N<k rows returns N")
    add("empty_index", [], [1, 0], 5, [], [], "no chunks -> empty list (Program.cs:74,117)")
    add("negative_zero_products_sum_to_plus_zero", [[-1.0, -2.0], [3.0, 4.0]], [0.0, 0.0], 2, [0, 1], [0.0, 0.0],
        "every product of row 0 is -0.0f, but the accumulator starts at +0.0 (`double dot = 0`, :127) and +0.0 + -0.0 = +0.0; "
        "na = 0 -> +0.0 / 1e-9 = +0.0 with a POSITIVE sign for both rows (found by the fuzz test in tests/test_oracle.py)")
    add("zero_row_scores_plus_zero", [[0, 0], [1, 0], [0, 0]], [1, 0], 3, [1, 0, 2], [one, 0.0, 0.0],
        "0/(0+1e-9) = +0.0 exactly (:134)")
    add("zero_query", [[1, 0], [0, 1], [1, 1]], [0, 0], 2, [0, 1], [0.0, 0.0], "all scores +0.0 -> insertion order")
    add("dim_mismatch_all_minus_one", [[1, 0, 0], [0, 1, 0], [0, 0, 1]], [1, 0], 2, [0, 1], [-1.0, -1.0],
        "a.Length != b.Length -> -1 (:126); stable order keeps rows 0,1")
    add("opposite_row_last", [[-2, 0], [0, 1], [1, 0]], [1, 0], 3, [2, 1, 0],
        [one, 0.0, -2.0 / (1.0 * 2.0 + EPS)], "cos=-1 ranks last")
    x = 14.0
    y = 77.0
    add("small_integers", [[4, 5, 6]], [1, 2, 3], 1, [0], [32.0 / (math.sqrt(x) * math.sqrt(y) + EPS)],
        "dot=32, na=14, nb=77, all exact")
    t = 1.0 + 2.0 ** -12
    p = 1.0 + 2.0 ** -11          # fp32(t*t): exact 1+2^-11+2^-24 ties-to-even -> 1+2^-11
    add("fp32_product_rounding", [[t]], [t], 1, [0], [p / (math.sqrt(p) * math.sqrt(p) + EPS)],
        "a[i]*b[i] is rounded to fp32 BEFORE widening (:130); an fp64 product would give 1+2^-11+2^-24",
        pins="float * float is evaluated in fp32 on .NET (x64 RyuJIT), then widened by `+=` into the double accumulator")
    big = 2.0 ** 30
    na_ = (2.0 ** 60 + 1.0) + 2.0 ** 60     # sequential fp64: 2^60+1 -> 2^60, then +2^60 -> 2^61
    add("sequential_fp64_order", [[big, 1.0, big]], [big, 1.0, -big], 1, [0],
        [0.0 / (math.sqrt(na_) * math.sqrt(na_) + EPS)],
        "dot: (2^60 + 1) rounds to 2^60 in fp64, then -2^60 -> 0 exactly; any other order gives 1",
        pins="the loop adds in index order into ONE fp64 accumulator (no vectorisation, no reassociation)")
    # four terms, so that a pairwise tree ((t0+t1)+(t2+t3)) and the sequential loop (((t0+t1)+t2)+t3) disagree:
    # products 1, 2^60, -2^60, 1.  sequential: 1+2^60 -> 2^60; -2^60 -> 0; +1 -> 1.   pairwise: 2^60 + (-2^60 + 1 -> -2^60) = 0
    n4 = ((1.0 + 2.0 ** 60) + 2.0 ** 60) + 1.0          # na = nb, sequential: 2^60, 2^61, 2^61
    add("sequential_not_pairwise_sum", [[1.0, big, -big, 1.0]], [1.0, big, big, 1.0], 1, [0],
        [1.0 / (math.sqrt(n4) * math.sqrt(n4) + EPS)],
        "dot = ((1 + 2^60) - 2^60) + 1 = 1 when summed in index order; a pairwise (SIMD-style) tree gives 0",
        pins="the loop adds in index order into ONE fp64 accumulator (no vectorisation, no reassociation)")
    # stability: eight bit-identical rows among others, fewer requested than there are duplicates.  An unstable
    # selection (heap, quickselect, parallel partial sort) returns SOME five of the eight; LINQ returns the first five.
    dup = [0.6, 0.8]
    rows_st = [[1, 0], dup, [0, 1], dup, dup, [-1, 0], dup, dup, [0.8, 0.6], dup, dup, dup]
    sc_dup = cos_formula(dup, dup)
    add("stable_sort_first_of_many_equal_keys", rows_st, dup, 5, [1, 3, 4, 6, 7], [sc_dup] * 5,
        "OrderByDescending is a STABLE sort (:64): equal keys keep source order, so Take(5) is the five lowest-numbered duplicates",
        pins="System.Linq ordering is stable; Take(k) after it keeps the first k in source order among equal keys")
    add("scale_invariance", [[3, 4], [0.75, 1.0], [300, 400]], [3, 4], 3, None, None,
        "positive scaling leaves cosine unchanged up to rounding; equal-score rows keep index order")
    c = cases[-1]
    sc = [cos_formula(c["query"], r) for r in c["rows"]]
    order = sorted(range(3), key=lambda i: (-sc[i], i))
    c["expect_idx"], c["expect_score"] = order, [sc[i] for i in order]
    add("nan_row_sorts_last", [[float("nan"), 1], [0, 1], [1, 0]], [0, 1], 3, [1, 2, 0], [one, 0.0, "nan"],
        "double.CompareTo: NaN is below every number -> last under OrderByDescending")
    return cases


def synth():
    out = {}
    rows = onp.synth_rows(0, 4, 8, 1234)
    out["generator"] = dict(seed=1234, row0=0, n_rows=4, dim=8,
                            f32_hex=[[np.float32(v).view(np.uint32).item() for v in r] for r in rows])
    rb = onp.synth_rows(5, 2, 8, 99, round_bf16=True)
    out["generator_bf16"] = dict(seed=99, row0=5, n_rows=2, dim=8,
                                 f32_hex=[[np.float32(v).view(np.uint32).item() for v in r] for r in rb])
    cases = []
    for (n, d, k, cs, qs) in [(257, 64, 5, 1234, 5678), (1000, 768, 10, 1234, 5678), (333, 100, 6, 7, 8), (4096, 96, 100, 11, 12)]:
        corpus = onp.synth_rows(0, n, d, cs)
        # adversarial inserts: duplicates, zero row, +q and -q rows
        q = onp.synth_rows(0, 1, d, qs)[0]
        corpus[7] = corpus[42]
        corpus[n - 3] = corpus[42]
        corpus[13] = 0.0
        corpus[21] = q
        corpus[22] = -q
        idx, sc = onp.query(corpus, q, k)
        cases.append(dict(n=n, dim=d, top_k=k, corpus_seed=cs, query_seed=qs,
                          inserts="rows 7,n-3 := row 42; row 13 := 0; row 21 := q; row 22 := -q",
                          expect_idx=[int(i) for i in idx], expect_score_hex=[float(s).hex() for s in sc]))
    out["queries"] = cases
    return out


if __name__ == "__main__":
    with open(os.path.join(HERE, "kat_v1.json"), "w") as f:
        json.dump(kat(), f, indent=1)
    with open(os.path.join(HERE, "synth_v1.json"), "w") as f:
        json.dump(synth(), f, indent=1)
    print("wrote kat_v1.json, synth_v1.json")
