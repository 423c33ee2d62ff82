"""Oracle vs the hand-authored known-answer vectors (tests/golden/kat.json) and properties.

The reference has no tests / golden vectors: these pins are this repo's own.
"""
import json
import math
import os

import numpy as np
import pytest

import oracle.oracle as orc
from spydrpick_b200.synth import generate_alignment

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "kat.json")))
EPS = 2.220446049250313e-16


def _cols(case):
    s = np.stack([np.array(case["col_i"], dtype=np.uint8), np.array(case["col_j"], dtype=np.uint8)], axis=1)
    return s, np.array(case["w"], dtype=np.float64)


def _edges(lst):
    e = np.zeros(len(lst), dtype=orc.EDGE_DTYPE)
    for k, (i, j, w) in enumerate(lst):
        e[k] = (i, j, w)
    return e


@pytest.mark.parametrize("case", KAT["mi"], ids=lambda c: c["name"])
def test_mi_known_answers(case):
    s, w = _cols(case)
    assert orc.mi_pair(s, w, 0, 1, case["pc"]) == pytest.approx(case["mi"], rel=1e-12, abs=1e-14)


def test_toy_closed_form():
    assert KAT["mi"][0]["mi"] == pytest.approx(KAT["toy_closed_form"], rel=1e-13)
    assert KAT["toy_closed_form"] == pytest.approx(0.130812, abs=1e-6)


def test_mi_symmetry_and_nonneg():
    s, _, _ = generate_alignment(60, 40, seed=11, n_planted=0)
    w = np.random.default_rng(0).uniform(0.1, 1.0, 60)
    for i, j in [(0, 1), (3, 17), (5, 39)]:
        a = orc.mi_pair(s, w, i, j); b = orc.mi_pair(s, w, j, i)
        assert a == pytest.approx(b, rel=1e-12, abs=1e-15)
        assert a > -1e-12


def test_duplicate_sample_equals_weight_two():
    a, b = KAT["dup_equiv"]["a"], KAT["dup_equiv"]["b"]
    sa, wa = _cols(a); sb, wb = _cols(b)
    assert orc.mi_pair(sa, wa, 0, 1) == pytest.approx(orc.mi_pair(sb, wb, 0, 1), rel=1e-13)


def test_sample_permutation_invariance():
    s, _, _ = generate_alignment(80, 30, seed=5, n_planted=0)
    w = np.random.default_rng(1).uniform(0.1, 1.0, 80)
    p = np.random.default_rng(2).permutation(80)
    assert orc.mi_pair(s, w, 2, 9) == pytest.approx(orc.mi_pair(s[p], w[p], 2, 9), rel=1e-11)


@pytest.mark.parametrize("case", KAT["wo_gaps"], ids=lambda c: c["name"])
def test_mi_wo_gaps_literal_quirk(case):
    s, w = _cols(case)
    mi, wo = orc.mi_pair(s, w, 0, 1, case["pc"], wo_gaps=True)
    assert mi == pytest.approx(case["mi"], rel=1e-12)
    assert wo == pytest.approx(case["mi_wo_gaps"], rel=1e-12)


def test_neighbour_rule_lastblock_quirk():
    # SURVEY 8c pin 5: L=64, ident = min_identical-1, last block fully identical
    L = 64
    a = np.zeros(L, dtype=np.uint8); b = np.zeros(L, dtype=np.uint8)
    b[:10] = 1                                   # 10 mismatches in the FIRST block -> ident = 54
    s = np.stack([a, b])
    assert list(orc.neighbour_counts(s, 55, quirks=orc.QUIRK_LASTBLOCK)) == [2, 2]    # 54 + 32 >= 55
    assert list(orc.neighbour_counts(s, 55, quirks=False)) == [1, 1]
    assert list(orc.neighbour_counts(s, 87, quirks=orc.QUIRK_LASTBLOCK)) == [1, 1]    # 54 + 32 < 87
    # L=65: no quirk regardless of the flag
    s65 = np.concatenate([s, np.zeros((2, 1), dtype=np.uint8)], axis=1)          # ident = 55
    assert list(orc.neighbour_counts(s65, 56, quirks=orc.QUIRK_LASTBLOCK)) == [1, 1]
    assert list(orc.neighbour_counts(s65, 55, quirks=orc.QUIRK_LASTBLOCK)) == [2, 2]
    # gap == gap counts as identical
    g = np.full((2, 40), 4, dtype=np.uint8)
    assert list(orc.neighbour_counts(g, 40, quirks=False)) == [2, 2]
    # the reference's sample-0 double count (pinned against the compiled reference in tests/test_ref_pins.py):
    # counts[0] = 1 + 2 (self pair) + 2 * #neighbours of 0, every neighbour of sample 0 carries one extra count
    assert list(orc.neighbour_counts(g, 40, quirks=True)) == [5, 3]
    three = np.zeros((3, 40), dtype=np.uint8); three[2] = 1                      # sample 2 is nobody's neighbour
    assert list(orc.neighbour_counts(three, 36, quirks=True)) == [5, 3, 1]


def test_neighbour_counts_bruteforce():
    s, _, _ = generate_alignment(50, 96, seed=9, n_planted=0)
    thr = int(0.8 * 96)
    ident = (s[:, None, :] == s[None, :, :]).sum(-1)
    last = (s[:, None, 64:] == s[None, :, 64:]).sum(-1)
    exp_q = ((ident + last) >= thr).sum(1)       # includes self
    exp = (ident >= thr).sum(1)
    assert np.array_equal(orc.neighbour_counts(s, thr, orc.QUIRK_LASTBLOCK), exp_q.astype(np.uint64))
    sim0 = (ident + last)[0] >= thr                                        # sample 0 vs everyone, itself included, once more
    exp_ref = exp_q.copy(); exp_ref += sim0; exp_ref[0] += sim0.sum()
    assert np.array_equal(orc.neighbour_counts(s, thr, True), exp_ref.astype(np.uint64))
    assert np.array_equal(orc.neighbour_counts(s, thr, False), exp.astype(np.uint64))
    w = orc.weights_from_counts(exp)
    assert np.allclose(w, 1.0 / exp)


@pytest.mark.parametrize("case", KAT["aracne"], ids=lambda c: c["name"])
def test_aracne_known_answers(case):
    e = _edges(case["edges"])
    assert np.array_equal(orc.sort_edges(e), e)            # fixtures are pre-sorted
    for bs in (1, 2, 16384):
        assert list(orc.aracne_blocked(e, EPS, bs)) == case["direct"], bs
    assert list(orc.aracne_closed(e, EPS)) == case["direct"]


def test_aracne_blocked_equals_closed_random():
    rng = np.random.default_rng(3)
    for trial in range(5):
        nn = 40
        iu = np.triu_indices(nn, 1)
        pick = rng.random(len(iu[0])) < 0.3
        w = np.round(rng.random(pick.sum()), 2 if trial % 2 else 6)     # rounded -> plenty of exact ties
        e = np.zeros(pick.sum(), dtype=orc.EDGE_DTYPE)
        e["i"] = iu[0][pick]; e["j"] = iu[1][pick]; e["mi"] = w
        e = orc.sort_edges(e)
        ref = orc.aracne_closed(e, EPS)
        for bs in (1, 7, 64, 16384):
            assert np.array_equal(orc.aracne_blocked(e, EPS, bs), ref)
        assert np.array_equal(orc.aracne_blocked(e, 0.0, 16384), orc.aracne_closed(e, 0.0))
        assert ref.sum() < len(ref)


def test_sort_tiebreak():
    e = orc.sort_edges(_edges(KAT["sort"]["edges"]))
    assert [[int(a), int(b), float(c)] for a, b, c in e] == KAT["sort"]["sorted"]


def test_quartile_index():
    o, x = orc.outlier_thresholds(KAT["quartile"]["colmax"])
    assert o == pytest.approx(KAT["quartile"]["outlier"], rel=1e-14)
    assert x == pytest.approx(KAT["quartile"]["extreme"], rel=1e-14)


def test_writers_exact_bytes():
    W = KAT["writer"]
    e = _edges(W["edges"])
    txt = orc.format_edges(e, W["direct"], W["pos"], 1, False, W["n_original_positions"])
    assert txt == W["edges_text"]
    out = orc.format_outliers(e, W["direct"], _edges(W["wo_gaps"]), W["pos"], W["outlier"], W["extreme"], 1, False,
                              W["n_original_positions"], 0)
    assert out == W["outliers_text"]


def test_scan_matches_pairwise_and_colmax():
    s, pos, _ = generate_alignment(40, 50, seed=21, n_planted=0)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 50)))
    thr = 0.05
    r = orc.mi_scan(s, w, pos, threshold=thr, ld_threshold=0, circle_size=50)
    masks = orc.state_masks(s)
    full = np.zeros((50, 50))
    for i in range(50):
        for j in range(i + 1, 50):
            full[i, j] = full[j, i] = orc.mi_pair(s, w, i, j, masks=masks)
    exp = {(i, j) for i in range(50) for j in range(i + 1, 50) if thr < full[i, j]}
    got = {(int(e["i"]), int(e["j"])) for e in r["edges"]}
    assert got == exp
    np.fill_diagonal(full, -np.inf)
    assert np.allclose(r["colmax"], full.max(1), rtol=1e-12)
    gapcols = (masks & 16) != 0
    expwo = {(i, j) for (i, j) in exp if gapcols[i] or gapcols[j]}
    assert {(int(e["i"]), int(e["j"])) for e in r["edges_wo_gaps"]} == expwo


def test_threshold_pairs_rule():
    # SpydrPick.hpp:258-282
    assert orc.determine_threshold_pairs(0, 10**9, 0.999) == 100000          # 100000*0.001 = 100 -> not < 100
    assert orc.determine_threshold_pairs(0, 10**9, 0.9995) == 200000
    assert orc.determine_threshold_pairs(0, 1000, 0.5) == 100                # small-alignment safeguard
    assert orc.determine_threshold_pairs(777, 10**9, 0.5) == 777
    pi, pj = orc.sample_pairs(500, 99, seed=4)
    assert (pi < pj).all() and pj.max() <= 99
    assert len({(a, b) for a, b in zip(pi, pj)}) == 500


def test_filter_rule():
    n = 100
    s = np.zeros((n, 4), dtype=np.uint8)
    s[:1, 0] = 1            # minor allele 1% -> kept (>= 0.01)
    s[:50, 2] = 1; s[50:66, 2] = 4   # gap 16% -> dropped
    s[:50, 3] = 2; s[50:65, 3] = 4   # gap 15% -> kept
    assert list(orc.filter_columns(s, 0.01, 0.15)) == [True, False, False, True]
