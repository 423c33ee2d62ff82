"""Runs one reference fixture (tests/golden/ref/<case>) through the CPU ORACLE, step by step as the reference's main() does
(src/SpydrPick.cpp:163-480), and returns the listings it would write. Test infrastructure only."""
import numpy as np

import oracle.oracle as orc
import refcmp


def _opt(case, name, default, cast):
    a = case["args"]
    return cast(a[a.index(name) + 1]) if name in a else default


def oracle_run(case):
    """-> dict(edges=text, outliers=text, weights=text or None, stem=..., L=..., n=..., threshold=..., eps=...)"""
    fasta = refcmp.read_gz(case, case["name"] + ".fasta")
    states, _ = refcmp.fasta_states(fasta)
    n, L_raw = states.shape
    in_base = _opt(case, "--input-indexing-base", 1, int)
    out_base = _opt(case, "--output-indexing-base", 1, int)
    genome = _opt(case, "--genome-size", 0, int)
    n_orig = genome if genome else L_raw                                   # Alignment_parsers.hpp:67,133
    pos = np.arange(L_raw, dtype=np.uint64)
    stem = case["name"]
    if "--mapping-list" in case["args"]:                                    # Alignment_utility.hpp:128-160
        pos = refcmp.parse_loci_list(refcmp.read_gz(case, _opt(case, "--mapping-list", None, str)), in_base).astype(np.uint64)
        assert len(pos) == L_raw
    if "--include-list" in case["args"]:                                    # src/SpydrPick.cpp:191-210: column INDICES, id += "." + list stem
        fn = _opt(case, "--include-list", None, str)
        idx = refcmp.parse_loci_list(refcmp.read_gz(case, fn), in_base)
        if len(idx) != states.shape[1]:
            states = np.ascontiguousarray(states[:, idx]); pos = pos[idx]; stem += "." + fn.rsplit(".", 1)[0]
    if "--exclude-list" in case["args"]:                                    # src/SpydrPick.cpp:212-231: translation values minus the list, used as indices
        ex = set(refcmp.parse_loci_list(refcmp.read_gz(case, _opt(case, "--exclude-list", None, str)), in_base).tolist())
        idx = np.array([int(t) for t in pos if int(t) not in ex], dtype=np.int64)
        if len(idx) != states.shape[1]:
            states = np.ascontiguousarray(states[:, idx]); pos = pos[idx]
    if "--no-filter-alignment" not in case["args"]:                         # src/SpydrPick.cpp:234-297
        maf = _opt(case, "--maf-threshold", 0.01, float); gap = _opt(case, "--gap-threshold", 0.15, float)
        keep = orc.filter_columns(states, maf, gap)
        if keep.sum() != states.shape[1]:
            states = np.ascontiguousarray(states[:, keep]); pos = pos[keep]
            stem += ".filtered_ge%03dmaf_le%03dgf_gt01-lt05states" % (int(maf * 100.), int(gap * 100.))
    L = states.shape[1]
    # weights: Alignment_utility.hpp:186-258
    w = np.ones(n, dtype=np.float64)
    if "--no-sample-reweighting" not in case["args"]:
        if "--sample-weights" in case["args"]:
            w = np.array([float(x) for x in refcmp.read_gz(case, _opt(case, "--sample-weights", None, str)).split()], dtype=np.float64)
            assert len(w) == n
        else:
            thr = _opt(case, "--sample-reweighting-threshold", 0.9, float)
            w = orc.weights_from_counts(orc.neighbour_counts(states, int(thr * float(L)), True))
        if "--rescale-sample-weights" in case["args"]:
            n_eff = 0.0
            for x in w.tolist():
                n_eff += x
            w = (w * float(n)) / n_eff
    threshold = _opt(case, "--mi-threshold", -1.0, float)
    pc = _opt(case, "--mi-pseudocount", 0.5, float)
    ld = _opt(case, "--ld-threshold", 0, int)
    linear = "--linear-genome" in case["args"]
    eps = _opt(case, "--aracne-edge-threshold", 2.220446049250313e-16, float)
    assert threshold >= 0, "automatic threshold cases are handled by the caller"
    sc = orc.mi_scan(states, w, pos, threshold=threshold, pseudocount=pc, ld_threshold=ld, linear_genome=linear, circle_size=L)
    e = orc.sort_edges(sc["edges"])
    flags = np.zeros(len(e), dtype=np.uint8) if "--no-aracne" in case["args"] else orc.aracne_closed(e, eps)
    outlier, extreme = orc.outlier_thresholds(sc["colmax"])
    return {"edges": orc.format_edges(e, flags, pos, out_base, linear, n_orig),
            "outliers": orc.format_outliers(e, flags, sc["edges_wo_gaps"], pos, outlier, extreme, out_base, linear, n_orig, ld),
            "weights": "".join("%.8e\n" % x for x in w), "stem": stem, "L": L, "n": n, "threshold": threshold, "eps": eps,
            "states": states, "pos": pos, "w": w, "sorted_edges": e, "flags": flags, "outlier": outlier, "extreme": extreme,
            "n_original_positions": n_orig, "out_base": out_base}
