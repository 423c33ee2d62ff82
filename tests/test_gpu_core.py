"""GPU parity tests (through the C-ABI) for K1 (Hamming), the fp64 MI kernels (K2b/K2r/K2c and the
exhaustive engine), edge sort and K3 (ARACNE), each against the CPU oracle on the same inputs.
Integer / index / flag results must be bit-exact; fp64 MI within 1e-12 relative (re-association only)."""
import json
import os

import numpy as np
import pytest

import oracle.oracle as orc
from spydrpick_b200 import capi
from spydrpick_b200 import Context, EDGE_DTYPE
from spydrpick_b200.synth import generate_alignment

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "kat.json")))
EPS = 2.220446049250313e-16



def assert_edges_match(e_got, e_exp, rtol=1e-10, atol=1e-13):
    """Same edge set, MI equal within tolerance (joined by edge id); e_got validly sorted by
    (weight desc, id asc). Near-exact MI ties (duplicated column patterns) may be ordered differently
    by two fp64 implementations that associate the weight sums differently, so order is checked on
    each side's own values, not across sides."""
    assert len(e_got) == len(e_exp)
    idg = (e_got["j"].astype(np.uint64) << np.uint64(32)) | e_got["i"].astype(np.uint64)
    ide = (e_exp["j"].astype(np.uint64) << np.uint64(32)) | e_exp["i"].astype(np.uint64)
    og, oe = np.argsort(idg), np.argsort(ide)
    assert np.array_equal(idg[og], ide[oe]), "edge sets differ"
    assert np.allclose(e_got["mi"][og], e_exp["mi"][oe], rtol=rtol, atol=atol)
    d = np.diff(e_got["mi"])
    assert (d <= 0).all(), "not sorted by weight desc"
    ties = d == 0
    assert (idg[1:][ties] > idg[:-1][ties]).all(), "ties not ordered by id asc"


@pytest.fixture(scope="module")
def ctx():
    c = Context()
    yield c
    c.close()


def _edge_set(e):
    return {(int(a), int(b)) for a, b in zip(e["i"], e["j"])}


def _edges(lst):
    e = np.zeros(len(lst), dtype=EDGE_DTYPE)
    for k, (i, j, w) in enumerate(lst):
        e[k] = (i, j, w)
    return e


# ---------------------------------------------------------------- K1
@pytest.mark.parametrize("n,L", [(50, 96), (130, 257), (64, 64), (200, 1000), (333, 1024), (7, 33)])
@pytest.mark.parametrize("quirk", [True, False, orc.QUIRK_LASTBLOCK, orc.QUIRK_SAMPLE0])
def test_neighbour_counts_bit_exact(ctx, n, L, quirk):
    s, _, _ = generate_alignment(n, L, seed=100 + n + L, n_planted=0)
    s[n // 3] = s[0]                                           # a copy of sample 0: the sample-0 double count has neighbours to work on
    ctx.set_alignment(s)
    for frac in (0.9, 0.7, 0.99):
        thr = int(frac * float(L))
        got = ctx.neighbour_counts(thr, quirk)
        exp = orc.neighbour_counts(s, thr, quirk)
        assert np.array_equal(got, exp), (n, L, frac, quirk)


def test_neighbour_rule_pins(ctx):
    L = 64
    a = np.zeros(L, dtype=np.uint8); b = np.zeros(L, dtype=np.uint8); b[:10] = 1
    s = np.stack([a, b])
    ctx.set_alignment(s)
    assert list(ctx.neighbour_counts(55, orc.QUIRK_LASTBLOCK)) == [2, 2]
    assert list(ctx.neighbour_counts(55, False)) == [1, 1]
    assert list(ctx.neighbour_counts(87, orc.QUIRK_LASTBLOCK)) == [1, 1]
    g = np.full((2, 40), 4, dtype=np.uint8)
    ctx.set_alignment(g)
    assert list(ctx.neighbour_counts(40, False)) == [2, 2]
    # the reference's sample-0 double count (pinned against the compiled reference, tests/test_ref_pins.py)
    assert list(ctx.neighbour_counts(40, True)) == [5, 3]
    three = np.zeros((3, 40), dtype=np.uint8); three[2] = 1
    ctx.set_alignment(three)
    assert list(ctx.neighbour_counts(36, True)) == [5, 3, 1]


def test_reweight_matches_counts_and_oracle(ctx):
    s, _, _ = generate_alignment(150, 640, seed=31)
    s[5] = s[0]
    ctx.set_alignment(s)
    w, cnt = ctx.reweight(int(0.9 * 640), True)
    exp = orc.neighbour_counts(s, int(0.9 * 640), True)
    assert np.array_equal(cnt, exp) and np.array_equal(w, orc.weights_from_counts(exp))
    w2, _ = ctx.reweight(int(0.9 * 640), True, rescale=True)
    assert np.array_equal(w2, orc.weights_from_counts(exp, rescale=True))     # (w * n) / n_eff, bit for bit


# ---------------------------------------------------------------- fp64 MI kernels
@pytest.mark.parametrize("case", KAT["mi"], ids=lambda c: c["name"])
def test_mi_pairs_known_answers(ctx, case):
    s = np.stack([np.array(case["col_i"], dtype=np.uint8), np.array(case["col_j"], dtype=np.uint8)], axis=1)
    ctx.set_alignment(s)
    ctx.set_weights(np.array(case["w"]))
    got = ctx.mi_pairs([0], [1], case["pc"])[0]
    assert got == pytest.approx(case["mi"], rel=1e-12, abs=1e-14)


def test_mi_pairs_vs_oracle(ctx):
    s, _, _ = generate_alignment(300, 500, seed=7)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 500)))
    ctx.set_alignment(s); ctx.set_weights(w)
    pi, pj = orc.sample_pairs(4000, 499, seed=3)
    got = ctx.mi_pairs(pi, pj)
    exp = orc.mi_pairs(s, w, pi, pj)
    assert np.allclose(got, exp, rtol=1e-11, atol=1e-14)
    # symmetry
    got2 = ctx.mi_pairs(pj, pi)
    assert np.allclose(got, got2, rtol=1e-11, atol=1e-14)


@pytest.mark.parametrize("n,L,linear,ld", [(120, 300, False, 0), (77, 130, True, 40), (64, 96, False, 25)])
def test_exhaustive_scan_vs_oracle(ctx, n, L, linear, ld):
    s, pos, _ = generate_alignment(n, L, seed=n * L)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * L)))
    ctx.set_alignment(s, pos); ctx.set_weights(w)
    thr = 0.04
    exp = orc.mi_scan(s, w, pos, threshold=thr, ld_threshold=ld, linear_genome=linear, circle_size=L)
    got = ctx.mi_scan(thr, ld_threshold=ld, linear_genome=linear, engine=1, sort_desc=True)
    e_exp = orc.sort_edges(exp["edges"]); e_got = got["edges"]
    assert len(e_got) == len(e_exp) > 0
    assert_edges_match(e_got, e_exp, rtol=1e-11, atol=1e-14)
    assert np.allclose(got["colmax"], exp["colmax"], rtol=1e-11, atol=1e-14)
    wo_exp = {(int(a), int(b)): m for a, b, m in exp["edges_wo_gaps"]}
    wo_got = {(int(a), int(b)): m for a, b, m in got["edges_wo_gaps"]}
    assert wo_exp.keys() == wo_got.keys()
    for k in wo_exp:
        assert wo_got[k] == pytest.approx(wo_exp[k], rel=1e-10, abs=1e-13)
    assert ctx.outlier_thresholds(got["colmax"]) == pytest.approx(orc.outlier_thresholds(exp["colmax"]), rel=1e-10)


def test_exhaustive_scan_partition_union(ctx):
    s, pos, _ = generate_alignment(60, 200, seed=5)
    ctx.set_alignment(s, pos)
    whole = ctx.mi_scan(0.05, engine=1)
    parts = [ctx.mi_scan(0.05, engine=1, part_index=k, n_parts=3) for k in range(3)]
    assert sum(p["n_edges"] for p in parts) == whole["n_edges"]
    u = set().union(*[_edge_set(p["edges"]) for p in parts])
    assert u == _edge_set(whole["edges"])
    cm = np.max(np.stack([p["colmax"] for p in parts]), axis=0)
    assert np.array_equal(cm, whole["colmax"])


def test_threshold_estimator_equals_oracle(ctx):
    """spk_determine_mi_threshold uses the reference's sampler (mt19937(seed + iteration), duplicates rejected) exactly as the
    oracle restates it, so the two must agree to the last bits of the fp64 MI values."""
    s, _, _ = generate_alignment(150, 400, seed=8)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 400)))
    ctx.set_alignment(s); ctx.set_weights(w)
    for n_values, pairs, its, seed in ((4000, 0, 3, 5), (800, 5000, 10, 77), (20000, 0, 4, 1)):
        thr = ctx.determine_mi_threshold(n_values, threshold_pairs=pairs, iterations=its, seed=seed)
        exp = orc.determine_mi_threshold(s, w, n_values, threshold_pairs=pairs, iterations=its, seed=seed)
        assert thr == pytest.approx(exp, rel=1e-12, abs=1e-15), (n_values, pairs, its, seed)
    with pytest.raises(capi.SpkError):
        ctx.determine_mi_threshold(400 * 399 // 2)               # every pair wanted: the reference indexes out of range here


def test_threshold_estimator_device_sampler(ctx):
    """the device sampler draws DISTINCT pairs from a keyed permutation of the pair index space: same quantile statistics"""
    s, _, _ = generate_alignment(150, 400, seed=8)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 400)))
    ctx.set_alignment(s); ctx.set_weights(w)
    full = np.sort(orc.mi_scan(s, w, threshold=-1.0)["edges"]["mi"])
    exact = full[len(full) - 4000]                                # the exact threshold that saves 4000 pairs
    thr = np.array([ctx.determine_mi_threshold(4000, iterations=10, seed=sd, sampler="device") for sd in range(6)])
    ref = np.array([ctx.determine_mi_threshold(4000, iterations=10, seed=sd) for sd in range(6)])
    assert abs(thr.mean() - exact) <= 3 * max(ref.std(), 1e-4) + abs(ref.mean() - exact)
    assert len(set(thr.tolist())) > 1                             # the seed matters
    # small alignment: pairs / 10 samples, all distinct by construction even when that is a large share of all pairs
    s2, _, _ = generate_alignment(60, 40, seed=3)
    ctx.set_alignment(s2)
    t2 = ctx.determine_mi_threshold(100, iterations=5, seed=9, sampler="device")
    f2 = np.sort(orc.mi_scan(s2, np.ones(60), threshold=-1.0)["edges"]["mi"])
    assert f2[0] <= t2 <= f2[-1]


# ---------------------------------------------------------------- sort + K3
def test_sort_edges_bit_exact(ctx):
    rng = np.random.default_rng(1)
    e = np.zeros(5000, dtype=EDGE_DTYPE)
    e["i"] = rng.integers(0, 300, 5000); e["j"] = e["i"] + 1 + rng.integers(0, 300, 5000)
    e["mi"] = np.round(rng.random(5000), 2)            # many ties -> id tie-break exercised
    got = ctx.sort_edges(e); exp = orc.sort_edges(e)
    assert got.tobytes() == exp.tobytes()
    k = _edges(KAT["sort"]["edges"])
    assert [[int(a), int(b), float(c)] for a, b, c in ctx.sort_edges(k)] == KAT["sort"]["sorted"]


@pytest.mark.parametrize("runs", [[5000], [1, 4999], [1700, 0, 1650, 1650], [700, 600, 900, 800, 500, 400, 600, 500], [1000, 1000, 3000]])
def test_merge_sorted_runs_equals_full_sort(ctx, runs):
    """spk_merge_sorted_runs_device (multi-GPU: every rank sorts its own edges, the gathered list is merged) gives
    the list Graph::sort gives, bit for bit, for empty, odd and unequal runs."""
    import torch
    rng = np.random.default_rng(len(runs))
    n = sum(runs)
    e = np.zeros(n, dtype=EDGE_DTYPE)
    e["i"] = rng.integers(0, 300, n); e["j"] = e["i"] + 1 + rng.integers(0, 300, n)
    e["mi"] = np.round(rng.random(n), 2)               # many ties -> id tie-break exercised
    _, first = np.unique((e["j"].astype(np.uint64) << np.uint64(32)) | e["i"], return_index=True)
    e = e[np.sort(first)]                              # distinct edges (a graph has no duplicate ids)
    runs = list(runs); runs[-1] -= n - len(e); n = len(e)
    parts, o = [], 0
    for r in runs:
        parts.append(orc.sort_edges(e[o:o + r])); o += r
    cat = np.concatenate(parts)
    d = torch.from_numpy(cat.view(np.uint8).copy()).cuda()
    ctx.merge_sorted_runs_device(d.data_ptr(), runs)
    got = d.cpu().numpy().view(EDGE_DTYPE)
    assert got.tobytes() == orc.sort_edges(e).tobytes()


@pytest.mark.parametrize("case", KAT["aracne"], ids=lambda c: c["name"])
def test_aracne_known_answers(ctx, case):
    assert list(ctx.aracne(_edges(case["edges"]), EPS)) == case["direct"]


@pytest.mark.parametrize("trial", range(4))
def test_aracne_vs_oracle_random(ctx, trial):
    rng = np.random.default_rng(30 + trial)
    nn = [40, 200, 600, 1500][trial]
    m = [300, 6000, 30000, 100000][trial]
    a = rng.integers(0, nn, m); b = rng.integers(0, nn, m)
    keep = a != b
    lo = np.minimum(a, b)[keep]; hi = np.maximum(a, b)[keep]
    uniq = np.unique(lo.astype(np.uint64) << 32 | hi.astype(np.uint64))
    e = np.zeros(len(uniq), dtype=EDGE_DTYPE)
    e["i"] = (uniq >> 32).astype(np.uint32); e["j"] = (uniq & 0xffffffff).astype(np.uint32)
    e["mi"] = np.round(rng.random(len(uniq)), 3 if trial % 2 else 12)
    e = orc.sort_edges(e)
    for thr in (EPS, 0.0, 0.05):
        exp = orc.aracne_closed(e, thr)
        assert np.array_equal(ctx.aracne(e, thr), exp)
    if trial < 3:
        assert np.array_equal(orc.aracne_blocked(e, EPS, 16384), orc.aracne_closed(e, EPS))
    # edge-range partition (multi-GPU sharding of K3)
    f = np.zeros(len(e), dtype=np.uint8)
    for k in range(3):
        fk = ctx.aracne(e, EPS, part_index=k, n_parts=3)
        lo_, hi_ = len(e) * k // 3, len(e) * (k + 1) // 3
        f[lo_:hi_] = fk[lo_:hi_]
    assert np.array_equal(f, orc.aracne_closed(e, EPS))
    # a caller-supplied list that is NOT sorted by weight: the predicate is order independent (the device path
    # detects it and orders the adjacency lists itself)
    perm = rng.permutation(len(e))
    es = e[perm]
    for thr in (EPS, 0.0):
        assert np.array_equal(ctx.aracne(es, thr), orc.aracne_closed(e, thr)[perm])


def test_neighbour_counts_partitions_sum_to_full(ctx):
    """K1 sharding (spk_neighbour_counts_part): partial counts over 1, 2, 3 and 5 partitions add up to the full result."""
    s, pos, _ = generate_alignment(300, 1000, seed=21)
    ctx.set_alignment(s, pos)
    full = ctx.neighbour_counts(int(0.9 * 1000), True)
    assert np.array_equal(full, orc.neighbour_counts(s, int(0.9 * 1000), True))
    for parts in (2, 3, 5):
        tot = np.zeros_like(full)
        for k in range(parts):
            tot += ctx.neighbour_counts(int(0.9 * 1000), True, part_index=k, n_parts=parts)
        assert np.array_equal(tot, full)
    from spydrpick_b200 import SpkError
    with pytest.raises(SpkError):
        ctx.neighbour_counts(900, True, part_index=3, n_parts=3)


def test_aracne_many_nodes_fallback(ctx):
    """node ids beyond what the shared-memory neighbourhood bitmap holds: generic search kernel + separate offset table"""
    rng = np.random.default_rng(77)
    hubs = rng.integers(0, 2_000_000, 300)
    a = rng.choice(hubs, 40000); b = rng.choice(hubs, 40000)
    keep = a != b
    lo = np.minimum(a, b)[keep]; hi = np.maximum(a, b)[keep]
    uniq = np.unique(lo.astype(np.uint64) << 32 | hi.astype(np.uint64))
    e = np.zeros(len(uniq), dtype=EDGE_DTYPE)
    e["i"] = (uniq >> 32).astype(np.uint32); e["j"] = (uniq & 0xffffffff).astype(np.uint32)
    e["mi"] = np.round(rng.random(len(uniq)), 4)
    e = orc.sort_edges(e)
    for thr in (EPS, 0.0):
        assert np.array_equal(ctx.aracne(e, thr), orc.aracne_closed(e, thr))


def test_error_behaviour(ctx):
    from spydrpick_b200 import SpkError
    c2 = Context()
    with pytest.raises(SpkError):
        c2.neighbour_counts(10)                      # no alignment
    s, _, _ = generate_alignment(20, 40, seed=2, n_planted=0)
    c2.set_alignment(s)
    with pytest.raises(SpkError):
        c2.mi_pairs([0], [40])                       # column out of range
    with pytest.raises(SpkError):
        c2.mi_scan(0.1, pseudocount=0.0)
    c2.close()


def test_set_alignment_device_equals_host_upload(ctx):
    """spk_set_alignment_device (states already in HBM, e.g. all-gathered row shards) = spk_set_alignment."""
    import torch
    from spydrpick_b200 import SpkError
    s, pos, _ = generate_alignment(90, 333, seed=5)
    ctx.set_alignment(s, pos)
    c_host = ctx.neighbour_counts(int(0.9 * 333), True)
    ctx.set_weights(1.0 / c_host.astype(np.float64))
    r_host = ctx.mi_scan(0.05)
    d = torch.from_numpy(s).cuda()
    ctx.set_alignment_device(d.data_ptr(), 90, 333, pos)
    del d
    c_dev = ctx.neighbour_counts(int(0.9 * 333), True)
    assert np.array_equal(c_host, c_dev)
    ctx.set_weights(1.0 / c_dev.astype(np.float64))
    r_dev = ctx.mi_scan(0.05)
    assert r_dev["edges"].tobytes() == r_host["edges"].tobytes() and len(r_host["edges"]) > 0
    assert np.array_equal(r_dev["colmax"], r_host["colmax"])
    with pytest.raises(SpkError):
        ctx.set_alignment_device(s.ctypes.data, 90, 333, pos)       # pageable host pointer is rejected, not copied


# ---------------------------------------------------------------- column filter pre-pass
@pytest.mark.parametrize("n,L", [(37, 5), (150, 1003), (300, 4096)])
def test_column_filter_vs_oracle(ctx, n, L):
    """spk_column_state_counts / spk_set_alignment_filtered (Alignment_filter.hpp:250-339): counts bit-exact, keep mask =
    the oracle's, and the context afterwards equals one given the host-filtered matrix."""
    rng = np.random.default_rng(L)
    raw = rng.integers(0, 4, size=(n, L)).astype(np.uint8)
    inv = rng.random(L) < 0.3
    raw[:, inv] = rng.integers(0, 4, size=int(inv.sum()))[None, :]                      # invariant columns
    rare = rng.random(L) < 0.2
    raw[:, rare] = np.where(rng.random((n, int(rare.sum()))) < 0.004, 1, 0)              # minor allele below maf
    gappy = rng.random(L) < 0.2
    raw[:, gappy] = np.where(rng.random((n, int(gappy.sum()))) < 0.3, 4, raw[:, gappy])  # ~30 % gaps
    pos = (np.arange(L, dtype=np.uint64) * 3 + 7)
    cnt = ctx.column_state_counts(raw)
    exp_cnt = np.stack([(raw == a).sum(axis=0) for a in range(5)], axis=1).astype(np.uint32)
    assert np.array_equal(cnt, exp_cnt)
    exp_keep = orc.filter_columns(raw, 0.01, 0.15).astype(bool)
    if exp_keep.sum() < 2:
        from spydrpick_b200 import SpkError
        with pytest.raises(SpkError):
            ctx.set_alignment_filtered(raw, pos, 0.01, 0.15)
        return
    keep = ctx.set_alignment_filtered(raw, pos, 0.01, 0.15)
    assert np.array_equal(keep, exp_keep) and ctx.L == int(exp_keep.sum())
    f = np.ascontiguousarray(raw[:, exp_keep]); fpos = pos[exp_keep]
    Lf = f.shape[1]
    c_dev = ctx.neighbour_counts(int(0.9 * Lf), True)
    assert np.array_equal(c_dev, orc.neighbour_counts(f, int(0.9 * Lf), True))
    ctx.set_weights(1.0 / c_dev.astype(np.float64))
    r_dev = ctx.mi_scan(0.02, ld_threshold=10)
    ctx.set_alignment(f, fpos)
    ctx.set_weights(1.0 / c_dev.astype(np.float64))
    r_host = ctx.mi_scan(0.02, ld_threshold=10)
    assert r_dev["edges"].tobytes() == r_host["edges"].tobytes()
    assert np.array_equal(r_dev["colmax"], r_host["colmax"])


def test_host_mirror_filtered_alignment():
    """host.Alignment.filtered = the reference's filter step + upload; the pipeline on it equals the pipeline on the host-filtered matrix."""
    from spydrpick_b200 import host
    s, pos, gsize = generate_alignment(100, 400, seed=11)
    raw = np.zeros((100, 800), dtype=np.uint8)
    raw[:, 0::2] = s
    raw[:, 1::2] = 2                                                   # invariant columns the filter drops
    rpos = np.arange(800, dtype=np.uint64)
    a = host.Alignment.filtered(raw, rpos, n_original_positions=800)
    exp_keep = orc.filter_columns(raw, 0.01, 0.15).astype(bool)
    assert np.array_equal(a.keep, exp_keep) and a.L == int(exp_keep.sum()) and a.id_string.endswith("filtered_ge001maf_le015gf_gt01-lt05states")
    opt = host.Options(mi_threshold=0.05)
    net = host.run_pipeline(a, opt)
    b = host.Alignment(np.ascontiguousarray(raw[:, exp_keep]), rpos[exp_keep], n_original_positions=800, ctx=a.ctx)
    net2 = host.run_pipeline(b, opt)
    assert net.network.tobytes() == net2.network.tobytes() and len(net.network) > 0
    assert np.array_equal(net.flags, net2.flags)
    a.ctx.close()
