"""Generates tests/golden/kat.json: hand-checkable known-answer vectors for the hot path.

The reference ships no tests or golden vectors, so these pins are authored here
from the formulas in SURVEY.md 3 with an INDEPENDENT pure-Python implementation (textbook
sum p*ln(p/(pa*pb)) over the observed-state sub-table with +pc per observed cell), not with
the oracle. Run: python tests/golden/make_golden.py
"""
import json
import math
import os

GAP = 4


def textbook_mi(col_i, col_j, w, pc=0.5):
    Pi = sorted(set(col_i)); Pj = sorted(set(col_j))
    C = {(a, b): pc for a in Pi for b in Pj}
    for a, b, ws in zip(col_i, col_j, w):
        C[(a, b)] += ws
    Z = sum(C.values())
    pa = {a: sum(C[(a, b)] for b in Pj) / Z for a in Pi}
    pb = {b: sum(C[(a, b)] for a in Pi) / Z for b in Pj}
    return sum((v / Z) * math.log((v / Z) / (pa[a] * pb[b])) for (a, b), v in C.items())


def literal_wo_gaps(col_i, col_j, w, pc=0.5):
    """The reference's literal 'MI without gaps' (SURVEY 3.3 quirk): second normalisation pass
    on the pass-1 normalised table with the gap bit removed from both masks."""
    Pi = sorted(set(col_i)); Pj = sorted(set(col_j))
    T = [[0.0] * 5 for _ in range(5)]          # T[b][a]
    for a, b, ws in zip(col_i, col_j, w):
        T[b][a] += ws
    def one_pass(T, Pi, Pj):
        for b in Pj:
            for a in Pi:
                T[b][a] += pc
        Z = sum(T[b][a] for b in Pj for a in Pi)
        joint = icond = 0.0
        v = [0.0] * 5
        for b in Pj:
            for a in range(5):
                T[b][a] /= Z
            joint += sum(T[b][a] * math.log(T[b][a]) for a in Pi)
            rs = sum(T[b])
            icond += rs * math.log(rs)
            for a in range(5):
                v[a] += T[b][a]
        jcond = sum(v[a] * math.log(v[a]) for a in Pi)
        return joint - icond - jcond
    mi1 = one_pass(T, Pi, Pj)
    mi2 = one_pass(T, [a for a in Pi if a != GAP], [b for b in Pj if b != GAP])
    return mi1, mi2


def main():
    kat = {"mi": [], "wo_gaps": []}
    cases = [
        ("toy_identical_2x2", [0, 2], [0, 2], [1.0, 1.0]),
        ("independent_4", [0, 0, 2, 2], [1, 3, 1, 3], [1.0, 1.0, 1.0, 1.0]),
        ("mask_A_gap_vs_CGT", [0, 0, 4, 4, 0, 4], [2, 3, 1, 2, 3, 1], [1.0] * 6),
        ("weighted", [0, 1, 1, 0, 2], [3, 3, 2, 2, 2], [0.5, 0.25, 1.0, 0.125, 2.0]),
        ("with_gaps", [0, 1, 4, 0, 1, 4, 0], [2, 2, 2, 4, 3, 4, 3], [1.0, 0.5, 0.5, 1.0, 0.25, 0.25, 1.0]),
    ]
    for name, ci, cj, w in cases:
        kat["mi"].append({"name": name, "col_i": ci, "col_j": cj, "w": w, "pc": 0.5, "mi": textbook_mi(ci, cj, w)})
    # closed form for the 2x2 toy (SURVEY 8c pin 1)
    toy = 2 * (3 / 8) * math.log(3 / 8) + 2 * (1 / 8) * math.log(1 / 8) - 2 * (2 * 0.5 * math.log(0.5))
    kat["toy_closed_form"] = toy
    # duplicate-sample == weight-2 equivalence
    kat["dup_equiv"] = {"a": {"col_i": [0, 1, 1, 2], "col_j": [3, 3, 3, 0], "w": [1.0, 1.0, 1.0, 1.0]},
                        "b": {"col_i": [0, 1, 2], "col_j": [3, 3, 0], "w": [1.0, 2.0, 1.0]}}
    for name, ci, cj, w in cases[2:]:
        m1, m2 = literal_wo_gaps(ci, cj, w)
        kat["wo_gaps"].append({"name": name, "col_i": ci, "col_j": cj, "w": w, "pc": 0.5, "mi": m1, "mi_wo_gaps": m2})
    # ARACNE pins (SURVEY 8c pin 6): edges (i,j,w) sorted desc; expected direct flags
    kat["aracne"] = [
        {"name": "triangle", "edges": [[0, 1, 0.3], [1, 2, 0.2], [0, 2, 0.1]], "direct": [1, 1, 0]},
        {"name": "tie_top", "edges": [[0, 1, 0.2], [1, 2, 0.2], [0, 2, 0.1]], "direct": [1, 1, 0]},
        {"name": "tie_min", "edges": [[0, 1, 0.3], [0, 2, 0.1], [1, 2, 0.1]], "direct": [1, 1, 1]},
        {"name": "path_no_triangle", "edges": [[0, 1, 0.3], [1, 2, 0.2], [2, 3, 0.1]], "direct": [1, 1, 1]},
        # 4-clique with weights 0.6..0.1: (0,1).6 (2,3).5 (0,2).4 (1,3).3 (0,3).2 (1,2).1
        # triangles: 012: .6,.4,.1 -> (1,2) ; 013: .6,.3,.2 -> (0,3) ; 023: .4,.5,.2 -> (0,3) ; 123: .1,.3,.5 -> (1,2)
        {"name": "clique4", "edges": [[0, 1, 0.6], [2, 3, 0.5], [0, 2, 0.4], [1, 3, 0.3], [0, 3, 0.2], [1, 2, 0.1]],
         "direct": [1, 1, 1, 1, 0, 0]},
    ]
    # sort tie-break pin (SURVEY 8c pin 7): equal weights ordered by (node2<<32)|node1
    kat["sort"] = {"edges": [[5, 9, 0.5], [1, 9, 0.5], [7, 8, 0.5], [0, 1, 0.7], [2, 3, 0.1]],
                   "sorted": [[0, 1, 0.7], [7, 8, 0.5], [1, 9, 0.5], [5, 9, 0.5], [2, 3, 0.1]]}
    # quartile index pin (SURVEY 8c pin 8): L=10 -> sorted[2], sorted[6]
    colmax = [0.9, 0.1, 0.5, 0.3, 0.7, 0.2, 0.8, 0.4, 0.6, 1.0]
    s = sorted(colmax); q1, q3 = s[2], s[6]
    kat["quartile"] = {"colmax": colmax, "outlier": q3 + 1.5 * (q3 - q1), "extreme": q3 + 3.0 * (q3 - q1)}
    # writer pins (SURVEY 8c pin 9): pos = original positions (0-based), base 1, circular genome of 100
    kat["writer"] = {
        "pos": [3, 10, 50, 98], "n_original_positions": 100,
        "edges": [[0, 3, 0.75], [1, 2, 0.5000004], [0, 1, 0.25]], "direct": [1, 0, 1],
        "edges_text": "4 99 5 1 0.750000\n11 51 40 0 0.500000\n4 11 7 1 0.250000\n",
        "wo_gaps": [[0, 3, 0.6]], "outlier": 0.4, "extreme": 0.7,
        "outliers_text": "4 99 5 1 0.750000 0.600000 20.0 1\n11 51 40 0 0.500000 0.500000 0.0 0\n",
    }
    # neighbour rule pin (SURVEY 8c pin 5) is constructed in the test itself (needs L=64/65 arrays)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kat.json")
    with open(out, "w") as f:
        json.dump(kat, f, indent=1)
    print("wrote", out)


if __name__ == "__main__":
    main()
