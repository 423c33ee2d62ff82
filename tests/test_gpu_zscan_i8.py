"""GPU parity tests for the int8 compact one-hot engine (spk_mi_i8.cu, engine 3 = default): raw s32
TMEM accumulators against exactly summed integer joint counts (bit-exact), and the full scan
(filter + fp64 refine) against the CPU oracle under the north-star tolerance."""
import numpy as np
import pytest

import oracle.oracle as orc
from spydrpick_b200 import Context
from spydrpick_b200.synth import generate_alignment
from test_gpu_zscan import assert_edges_match

pytestmark = pytest.mark.gpu
INVALID = 0xFFFFFFFF


@pytest.fixture(scope="module")
def ctx():
    c = Context()
    yield c
    c.close()


def _mixed_alignment(n, L, seed):
    """columns of every class: monomorphic, biallelic, +gap, tri-, four-allelic, all five states."""
    rng = np.random.default_rng(seed)
    s = np.zeros((n, L), dtype=np.uint8)
    for c in range(L):
        kind = c % 7
        states = [[0], [1, 3], [0, 2, 4], [0, 1, 2], [0, 1, 2, 3], [0, 1, 2, 3, 4], [2, 4]][kind]
        p = rng.dirichlet(np.ones(len(states)) * 2.0)
        col = rng.choice(states, size=n, p=p)
        for k, st in enumerate(states):          # make sure every listed state is present
            col[(c + k) % n] = st
        s[:, c] = col
    return s


def _tile_cols(lay, t):
    nt, cs = lay["ntiles"], lay["col_start"]
    ts = [0, int(nt[0]), int(nt[0]) + int(nt[1])]
    g = 2 if t >= ts[2] else (1 if t >= ts[1] else 0)
    w = 128 >> g
    c0 = int(cs[g]) + (t - ts[g]) * w
    return g, 1 << g, w, c0


def _expected_tables(s, q, lay, ti, tj):
    gi, KI, WI, ci0 = _tile_cols(lay, ti)
    gj, KJ, WJ, cj0 = _tile_cols(lay, tj)
    out = np.zeros((WI, WJ, KI, KJ), dtype=np.uint64)
    q = q.astype(np.uint64)
    for il in range(WI):
        oi = int(lay["perm"][ci0 + il])
        if oi == INVALID:
            continue
        inf_i = int(lay["info"][ci0 + il])
        for jl in range(WJ):
            oj = int(lay["perm"][cj0 + jl])
            if oj == INVALID:
                continue
            inf_j = int(lay["info"][cj0 + jl])
            for a in range(inf_i & 15):
                ma = s[:, oi] == ((inf_i >> (4 + 3 * a)) & 7)
                for b in range(inf_j & 15):
                    mb = s[:, oj] == ((inf_j >> (4 + 3 * b)) & 7)
                    out[il, jl, a, b] = q[ma & mb].sum()
    return out, (WI, WJ, KI, KJ)


def test_i8_layout_classes(ctx):
    s = _mixed_alignment(60, 300, 1)
    ctx.set_alignment(s)
    lay = ctx.debug_i8_layout()
    perm, info = lay["perm"], lay["info"]
    real = perm[perm != INVALID]
    assert sorted(real.tolist()) == list(range(300))            # a permutation of the columns
    for pc, o in enumerate(perm):
        if o == INVALID:
            continue
        present = sorted(set(s[:, o].tolist()))
        k = int(info[pc]) & 15
        kept = [(int(info[pc]) >> (4 + 3 * a)) & 7 for a in range(k)]
        assert kept == present[:-1] and ((int(info[pc]) >> 16) & 7) == present[-1]
        g = 2 if pc >= lay["col_start"][2] else (1 if pc >= lay["col_start"][1] else 0)
        assert g == (0 if k <= 1 else (1 if k == 2 else 2))


@pytest.mark.parametrize("weighted", [False, True])
def test_i8_tile_counts_bit_exact(ctx, weighted):
    n, L = 300, 700
    s = _mixed_alignment(n, L, 5)
    ctx.set_alignment(s)
    if weighted:
        w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.5 * L)))
        w = w * np.linspace(0.3, 1.0, n)                        # spread the weights
        ctx.set_weights(w)
    lay0 = ctx.debug_i8_layout()
    nt = lay0["ntiles"]
    NT = int(nt.sum())
    ts = [0, int(nt[0]), int(nt[0]) + int(nt[1])]
    assert all(int(x) >= 1 for x in nt)
    pairs = [(ts[0], ts[0]), (ts[0], ts[1]), (ts[0], ts[2]), (ts[1], ts[1]), (ts[1], ts[2]), (ts[2], ts[2]),
             (ts[1] - 1, NT - 1), (ts[2] - 1, ts[2]), (NT - 1, NT - 1)]
    first = ctx.debug_i8_tile(0, 0)                              # also quantises the weights
    lay = ctx.debug_i8_layout(with_weights=True)
    q = lay["q"]
    wq = q.astype(np.float64) * lay["delta"]
    if ctx.debug_i8_mode() == 2:                                 # 16-bit grid
        if not weighted:
            assert (q == 65535).all()
        else:
            assert np.abs(wq - w).max() <= 0.5 * lay["delta"] * (1 + 1e-9)
    else:                                                        # byte-product grid: q = a*b, a, b <= 255
        assert q.max() <= 255 * 255
        assert np.abs(wq - (w if weighted else 1.0)).max() <= 2e-3 * (w.max() if weighted else 1.0)
    for ti, tj in pairs:
        got = ctx.debug_i8_tile(ti, tj) if (ti, tj) != (0, 0) else first
        exp, shape = _expected_tables(s, q, lay, ti, tj)
        g = got[: exp.size].reshape(shape).astype(np.uint64)
        # padding columns produce zero rows in the GEMM
        assert np.array_equal(g, exp), (ti, tj, shape, int(np.abs(g.astype(np.int64) - exp.astype(np.int64)).max()))


@pytest.mark.parametrize("engine", [3, 4])
@pytest.mark.parametrize("n,L,linear,ld", [(200, 500, False, 0), (130, 333, True, 30), (257, 700, False, 100),
                                            (64, 64, False, 0), (90, 1100, False, 0), (300, 2100, False, 0)])
def test_scan_i8_vs_oracle(ctx, n, L, linear, ld, engine):
    s, pos, _ = generate_alignment(n, L, seed=n + L)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * L)))
    ctx.set_alignment(s, pos); ctx.set_weights(w)
    thr = 0.05 if L > 64 else 0.01
    exp = orc.mi_scan(s, w, pos, threshold=thr, ld_threshold=ld, linear_genome=linear, circle_size=L)
    got = ctx.mi_scan(thr, ld_threshold=ld, linear_genome=linear, engine=engine, sort_desc=True, max_edges=L * L // 2)
    e_exp = orc.sort_edges(exp["edges"]); e_got = got["edges"]
    assert got["n_candidates"] >= len(e_exp)
    assert len(e_got) == len(e_exp) > 0
    assert_edges_match(e_got, e_exp, rtol=1e-5, atol=0)            # north-star tolerance
    assert_edges_match(e_got, e_exp, rtol=1e-10, atol=1e-13)       # what the fp64 refine actually gives
    assert np.allclose(got["colmax"], exp["colmax"], rtol=1e-10, atol=1e-13)
    assert got["max_filter_error"] < (5e-5 if engine == 3 else 2e-3)
    # MI without gaps (K2c, the literal two-pass quirk of mi.hpp:108-135,465-489): same pairs, same values
    wo_exp = {(int(a), int(b)): m for a, b, m in exp["edges_wo_gaps"]}
    wo_got = {(int(a), int(b)): m for a, b, m in got["edges_wo_gaps"]}
    assert wo_exp.keys() == wo_got.keys() and (len(wo_exp) > 0 or L == 64)
    for k_ in wo_exp:
        assert wo_got[k_] == pytest.approx(wo_exp[k_], rel=1e-10, abs=1e-13)
    # the audit ran (when the alignment is large enough for a seed pass) and found nothing the filter would have lost
    assert got["audit_violations"] == 0 and got["audit_max_error"] <= 0.5 * got["band"]


def test_overflow_rescans_instead_of_failing(ctx):
    """a capacity hint that is far too small: the scan grows its buffers and repeats itself (the reference stores any number
    of edges), same result as with ample buffers."""
    s, pos, _ = generate_alignment(150, 3000, seed=21)
    ctx.set_alignment(s, pos)
    ctx.set_weights(orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 3000))))
    big = ctx.mi_scan(0.01, max_edges=3000 * 3000 // 2)
    c2 = Context()
    try:
        c2.set_alignment(s, pos)
        c2.set_weights(orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 3000))))
        small = c2.mi_scan(0.01, max_edges=1000)
    finally:
        c2.close()
    assert small["n_edges"] == big["n_edges"] > 100000
    assert small["edges"].tobytes() == big["edges"].tobytes()
    assert np.array_equal(np.sort(small["edges_wo_gaps"], order=("i", "j")), np.sort(big["edges_wo_gaps"], order=("i", "j")))


@pytest.mark.parametrize("engine", [3, 4])
def test_filter_audit_many_valued_weights(ctx, engine):
    """--sample-weights style input: a wide, many-valued weight vector (no two samples share a weight, three decades of range)
    through the tensor-core filter. The per-scan audit refines an unbiased sample of pairs on both sides of the candidate
    cut-off; no pair above the threshold may have been invisible to the filter, and the result equals the exhaustive engine."""
    n, L = 400, 9000
    s, pos, _ = generate_alignment(n, L, seed=55)
    rng = np.random.default_rng(8)
    w = np.exp(rng.uniform(np.log(1e-3), np.log(3.0), size=n))
    ctx.set_alignment(s, pos); ctx.set_weights(w)
    thr = 0.08
    got = ctx.mi_scan(thr, engine=engine, max_edges=L * L // 8)
    assert got["audit_pairs"] >= 1000, got["audit_pairs"]
    assert got["audit_violations"] == 0
    assert got["audit_max_error"] <= 0.5 * got["band"]
    exact = ctx.mi_scan(thr, engine=1, max_edges=L * L // 8)
    assert got["n_edges"] == exact["n_edges"] > 1000
    assert got["edges"].tobytes() == exact["edges"].tobytes()
    assert np.array_equal(got["colmax"], exact["colmax"])


def test_filter_audit_catches_a_band_that_is_too_narrow(ctx):
    """a band far below the filter's actual error: the audit (or the candidate error check) must refuse the result instead of
    silently dropping edges; with the band left to the library the same scan succeeds (widening until the audit is clean)."""
    from spydrpick_b200.capi import SpkError, SPK_E_PRECISION
    n, L = 300, 8000
    s, pos, _ = generate_alignment(n, L, seed=56)
    rng = np.random.default_rng(9)
    w = np.exp(rng.uniform(np.log(1e-3), np.log(3.0), size=n))
    ctx.set_alignment(s, pos); ctx.set_weights(w)
    with pytest.raises(SpkError) as ei:
        ctx.mi_scan(0.08, engine=4, band=1e-8, max_edges=L * L // 8)
    assert ei.value.code == SPK_E_PRECISION
    ok = ctx.mi_scan(0.08, engine=4, max_edges=L * L // 8)
    exact = ctx.mi_scan(0.08, engine=1, max_edges=L * L // 8)
    assert ok["edges"].tobytes() == exact["edges"].tobytes()


@pytest.mark.parametrize("engine", [3, 4])
def test_scan_i8_mixed_classes_and_monomorphic(ctx, engine):
    """every class pair incl. monomorphic columns (k = 0) and all-five-state columns, uneven weights."""
    n, L = 150, 900
    s = _mixed_alignment(n, L, 9)
    pos = np.arange(L, dtype=np.uint64) * 7
    w = np.linspace(0.05, 1.0, n)
    ctx.set_alignment(s, pos); ctx.set_weights(w)
    thr = 0.02
    exp = orc.mi_scan(s, w, pos, threshold=thr, ld_threshold=50, circle_size=L)
    got = ctx.mi_scan(thr, ld_threshold=50, engine=engine)
    exact = ctx.mi_scan(thr, ld_threshold=50, engine=1)
    assert got["edges"].tobytes() == exact["edges"].tobytes()
    assert np.array_equal(got["colmax"], exact["colmax"])
    assert_edges_match(got["edges"], orc.sort_edges(exp["edges"]), rtol=1e-10, atol=1e-13)
    assert np.allclose(got["colmax"], exp["colmax"], rtol=1e-10, atol=1e-13)
    assert got["max_filter_error"] < (5e-5 if engine == 3 else 2e-3)


@pytest.mark.parametrize("engine", [3, 4])
def test_scan_i8_partitions_and_engines_agree(ctx, engine):
    s, pos, _ = generate_alignment(150, 2500, seed=78)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 2500)))
    ctx.set_alignment(s, pos); ctx.set_weights(w)
    whole = ctx.mi_scan(0.06, engine=engine)
    dense = ctx.mi_scan(0.06, engine=2)
    assert whole["edges"].tobytes() == dense["edges"].tobytes()
    assert np.array_equal(whole["colmax"], dense["colmax"])
    parts = [ctx.mi_scan(0.06, part_index=k, n_parts=3, engine=engine) for k in range(3)]
    assert sum(p["n_edges"] for p in parts) == whole["n_edges"]
    merged = np.concatenate([p["edges"] for p in parts])
    assert ctx.sort_edges(merged).tobytes() == whole["edges"].tobytes()
    assert np.array_equal(np.max(np.stack([p["colmax"] for p in parts]), axis=0), whole["colmax"])


def test_pinned_buffers_and_aracne_device_reuse(ctx):
    """zero-copy result views + spk_aracne on the scan's own edge pointer (device copy reused) == the copying path."""
    from spydrpick_b200.capi import pinned_copy
    s, pos, _ = generate_alignment(120, 1500, seed=5)
    sp = pinned_copy(s)
    assert np.array_equal(sp, s)
    ctx.set_alignment(sp, pos)
    ctx.set_weights(orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 1500))))
    r = ctx.mi_scan(0.05, copy=False)
    view = r["edges"]
    copy = view.copy()
    f_view = ctx.aracne(view)                # pointer == library buffer -> no upload
    f_copy = ctx.aracne(copy)                # generic path
    assert np.array_equal(f_view, f_copy)
    assert np.array_equal(f_view, orc.aracne_closed(copy, 2.220446049250313e-16))
    # after a new alignment the old pointer must not be trusted
    ctx.set_alignment(sp, pos)
    assert np.array_equal(ctx.aracne(copy), f_copy)


def test_async_download_overlaps_aracne(ctx):
    """keep_on_device = 2: host views are returned while the download runs; aracne() on the view uses the
    device copy; after wait_download() the views equal the synchronous result."""
    from spydrpick_b200.capi import pinned_empty
    s, pos, _ = generate_alignment(150, 2000, seed=11)
    ctx.set_alignment(s, pos)
    ctx.set_weights(orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 2000))))
    ref = ctx.mi_scan(0.04)
    f_ref = ctx.aracne(ref["edges"])
    r = ctx.mi_scan(0.04, copy=False, async_download=True)
    buf = pinned_empty(len(ref["edges"]) + 100, np.uint8)
    buf[:] = 7
    f = ctx.aracne(r["edges"], out=buf)
    ctx.wait_download()
    assert np.array_equal(r["edges"], ref["edges"])
    assert np.array_equal(np.sort(r["edges_wo_gaps"], order=("i", "j")), np.sort(ref["edges_wo_gaps"], order=("i", "j")))   # unordered list
    assert np.array_equal(f, f_ref)
    assert np.all(buf[len(ref["edges"]):] == 7)          # nothing written past n_edges
    with pytest.raises(ValueError):
        ctx.mi_scan(0.04, async_download=True)            # needs copy=False


def test_candidate_chunks_have_no_effect_on_results(ctx):
    """the int8 engine reserves candidate slots per warp in chunks (holes are invalid-marked): edge list and
    candidate count are those of the threshold/colmax rules, independent of chunking and of the seed pass."""
    s, pos, _ = generate_alignment(200, 6000, seed=13)
    ctx.set_alignment(s, pos)
    ctx.set_weights(orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 6000))))
    a = ctx.mi_scan(0.06, max_edges=6000 * 6000 // 2, ld_threshold=300)
    b = ctx.mi_scan(0.06, max_edges=6000 * 6000 // 2, ld_threshold=300, engine=2)
    assert np.array_equal(a["edges"], b["edges"])
    assert np.array_equal(a["colmax"], b["colmax"])
    assert a["n_candidates"] >= a["n_edges"]
    assert a["n_candidates"] <= 6000 * 5999 // 2                   # valid candidates only: holes of the reserved chunks are not counted


def test_mma_peak_probe_is_a_plausible_denominator(ctx):
    """spk_debug_i8_peak (the MMA-only kernel bench.py's roofline divides by): a rate between the cuBLASLt int8 GEMM class and a
    generous bound above the nominal 4.5 POP/s of a B200, reproducible to a few per cent."""
    ctx.debug_i8_peak(20000, 1)                                   # warm-up (module load, clocks)
    a, ms_a = ctx.debug_i8_peak(60000, 1)
    b, ms_b = ctx.debug_i8_peak(60000, 2)
    assert 2500.0 < a < 6000.0 and 2500.0 < b < 6000.0, (a, b)
    assert abs(a / b - 1.0) < 0.1
    assert ms_b == pytest.approx(2 * ms_a, rel=0.15)
