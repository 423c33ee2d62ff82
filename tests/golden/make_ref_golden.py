#!/usr/bin/env python
"""Generates tests/golden/ref/: inputs + the outputs of the REAL reference program on them.

The reference (santeripuranen/SpydrPick @ f6bc110 + apegrunt @ 5c61810) is compiled in place from /root/reference by
oracle/Makefile.ref (oracle/_ref/SpydrPick_ref: its own main(), option classes, filter, weights, scan, sort, ARACNE, writers).
This script builds small deterministic alignments, runs that binary on each with the command line recorded in case.json and
stores every file it wrote (gzip). It can only run where /root/reference exists; the fixtures it writes are committed and are
what the parity tests use everywhere else (tests/test_ref_golden.py: oracle vs fixtures on the CPU; tests/test_gpu_ref_golden.py:
the CUDA path and the C++ host vs fixtures on the GPU).

    python tests/golden/make_ref_golden.py            # regenerate everything
"""
import gzip
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from spydrpick_b200.synth import generate_alignment, to_fasta  # noqa: E402
import oracle.ref as ref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "ref")


def config1_raw(n=200, L_snp=3500, L_raw=5000, seed=20260101):
    """BASELINE.json configs[0] shape: 200 x 5,000 raw columns, 30 % of them invariant / gappy / below the MAF threshold."""
    s, _, _ = generate_alignment(n, L_snp, seed=seed)
    rng = np.random.default_rng(11)
    raw = np.zeros((n, L_raw), dtype=np.uint8)
    cols = np.sort(rng.choice(L_raw, size=L_snp, replace=False))
    inv = np.setdiff1d(np.arange(L_raw), cols)
    raw[:, cols] = s
    raw[:, inv] = rng.integers(0, 4, size=len(inv))[None, :]
    gappy = inv[::7]
    raw[:, gappy] = np.where(rng.random((n, len(gappy))) < 0.3, 4, raw[:, gappy])
    rare = inv[3::11]
    raw[0, rare] = (raw[0, rare] + 1) % 4
    return raw


def small(n, L, seed, dup=0):
    s, pos, g = generate_alignment(n, L, seed=seed)
    if dup:                                   # duplicate rows: neighbour counts > 1 everywhere, sample 0 included
        s[1:1 + dup] = s[0]
    return s, pos, g


def run_case(name, states, args, extra_files=None, repeat=1, note=""):
    d = os.path.join(OUT, name)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    tmp = tempfile.mkdtemp(prefix="refgold_")
    to_fasta(states, os.path.join(tmp, name + ".fasta"))
    for fn, text in (extra_files or {}).items():
        with open(os.path.join(tmp, fn), "w") as f:
            f.write(text)
    inputs = sorted(os.listdir(tmp))
    logs = []
    for rep in range(repeat):
        for fn in os.listdir(tmp):
            if fn not in inputs:
                os.remove(os.path.join(tmp, fn))
        r = ref.run_cli([name + ".fasta"] + args + ["-v"], cwd=tmp)
        if r.returncode != 0:
            raise RuntimeError("reference failed on %s: %s\n%s" % (name, r.stderr, r.stdout[-2000:]))
        logs.append(r.stdout)
    outputs = sorted(fn for fn in os.listdir(tmp) if fn not in inputs)
    for fn in inputs + outputs:
        with open(os.path.join(tmp, fn), "rb") as f, gzip.GzipFile(os.path.join(d, fn + ".gz"), "wb", mtime=0) as g:
            g.write(f.read())
    log = logs[-1]
    info = {"name": name, "args": args, "inputs": inputs, "outputs": outputs, "n": int(states.shape[0]), "L_raw": int(states.shape[1]),
            "note": note, "reference": ref.lib().ref_version().decode()}
    m = re.search(r"effective sample size = ([0-9.eE+-]+)", log)
    if m:
        info["effective_sample_size_printed"] = m.group(1)
    m = re.search(r"outlier threshold=([0-9.eE+-]+)\s+SpydrPick: extreme outlier threshold=([0-9.eE+-]+)", log)
    if m:
        info["outlier_threshold_printed"] = m.group(1); info["extreme_outlier_threshold_printed"] = m.group(2)
    thr = [float(x) for lg in logs for x in re.findall(r"MI save threshold = ([0-9.eE+-]+)", lg)]
    if thr:
        info["mi_save_thresholds_printed"] = thr        # one per repeat (the estimator is clock-seeded)
    m = re.search(r"found (\d+) nodes", log)
    if m:
        info["outlier_nodes"] = int(m.group(1))
    with open(os.path.join(d, "case.json"), "w") as f:
        json.dump(info, f, indent=1, sort_keys=True)
    shutil.rmtree(tmp)
    print("%-24s %s -> %s" % (name, " ".join(args), ", ".join(outputs)))


def main():
    if not ref.can_build():
        raise SystemExit("needs /root/reference (the fixtures are committed; regenerate only where the reference is present)")
    ref.build()
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(OUT)
    raw = config1_raw()
    # config 1: default filter + reweighting, circular genome
    run_case("c1_default", raw, ["--mi-threshold", "0.2", "--output-sample-weights"],
             note="BASELINE configs[0] shape, defaults, circular; threshold pinned")
    run_case("c1_ld_linear", raw, ["--mi-threshold", "0.2", "--ld-threshold", "60", "--linear-genome"],
             note="LD-masked colmax, linear distance, outliers + outlier-node sub-alignment export")
    run_case("c1_ld_circular", raw, ["--mi-threshold", "0.22", "--ld-threshold", "60"],
             note="LD-masked colmax, circular distance with circle = raw column count")
    # mapping list with and without --genome-size (the size_t wrap of the circular distance), 0-based output
    s, pos, g = small(150, 1200, 7)
    mapping = "\n".join(str(int(p) + 1) for p in pos) + "\n"
    run_case("map_genome", s, ["--mi-threshold", "0.15", "--no-filter-alignment", "--mapping-list", "map.txt", "--genome-size", str(g),
                               "--ld-threshold", "500", "--output-indexing-base", "0"], {"map.txt": mapping},
             note="mapping list + genome size; 0-based output")
    run_case("map_nogenome", s, ["--mi-threshold", "0.15", "--no-filter-alignment", "--mapping-list", "map.txt", "--ld-threshold", "500"],
             {"map.txt": mapping}, note="mapping list WITHOUT --genome-size: n_original_positions stays the FASTA column count")
    # user-supplied sample weights, rescaled
    rng = np.random.default_rng(5)
    wts = np.round(rng.uniform(0.05, 3.0, size=150), 6)
    wtxt = "\n".join("%.6f" % x for x in wts) + "\n"
    run_case("weights_file", s, ["--mi-threshold", "0.15", "--no-filter-alignment", "--sample-weights", "w.txt", "--output-sample-weights"],
             {"w.txt": wtxt}, note="--sample-weights: many-valued weight vector")
    run_case("weights_rescale", s, ["--mi-threshold", "0.15", "--no-filter-alignment", "--sample-weights", "w.txt", "--rescale-sample-weights",
                                    "--output-sample-weights"], {"w.txt": wtxt}, note="--rescale-sample-weights on a weights file")
    run_case("rescale_computed", s, ["--mi-threshold", "0.15", "--no-filter-alignment", "--rescale-sample-weights", "--output-sample-weights"],
             note="computed weights rescaled to n_eff = n")
    run_case("no_reweighting", s, ["--mi-threshold", "0.2", "--no-filter-alignment", "--no-sample-reweighting", "--no-aracne"],
             note="all weights 1, no ARACNE")
    # K1 quirks: L % 32 == 0, duplicates of sample 0
    s64, _, _ = small(90, 640, 9, dup=3)
    run_case("k1_L640_dups", s64, ["--mi-threshold", "0.25", "--no-filter-alignment", "--output-sample-weights", "--sample-reweighting-threshold", "0.85"],
             note="L % 32 == 0 and three copies of sample 0: both K1 quirks")
    # include / exclude / sample lists (range syntax a-b)
    run_case("include_list", raw, ["--mi-threshold", "0.2", "--include-list", "inc.txt"], {"inc.txt": "1-1500\n2000 2001 2002\n2500-4000\n"},
             note="--include-list with ranges")
    run_case("exclude_list", raw, ["--mi-threshold", "0.2", "--exclude-list", "exc.txt"], {"exc.txt": "100-900\n3000-3999\n"},
             note="--exclude-list with ranges")
    # (--sample-list has no fixture: the reference itself segfaults in calculate_weights after Alignment::subsample, a path that
    #  still prints "flush_block_buffer() called" debug lines -- our host implements the documented semantics, tested against the oracle)
    run_case("pseudocount_edge_thr", s, ["--mi-threshold", "0.1", "--no-filter-alignment", "--mi-pseudocount", "1.0", "--aracne-edge-threshold", "0.01",
                                         "--aracne-block-size", "500"], note="non-default pseudocount, ARACNE tolerance and block size")
    run_case("filter_thresholds", raw, ["--mi-threshold", "0.2", "--maf-threshold", "0.05", "--gap-threshold", "0.05", "--output-filtered-alignment"],
             note="non-default filter thresholds + filtered alignment export")
    run_case("state_frequencies", raw, ["--mi-threshold", "0.2", "--output-state-frequencies"],
             note="--output-state-frequencies on the filtered, reweighted alignment (the two *_distribution.txt files come from the Boost.Accumulators stand-in of oracle/ref_shims and are not pins)")
    # (--output-sample-distance-matrix has no fixture: Alignment::distance_matrix() is not implemented by the reference's storage class
    #  and the CRTP forwarder calls itself until the stack overflows)
    # automatic threshold: the estimator is clock-seeded, so several runs give the spread the parity test accepts
    run_case("auto_threshold", raw, ["--mi-values", "20000"], repeat=7, note="--mi-threshold -1 (automatic), 7 runs: mi_save_thresholds_printed")
    size = sum(os.path.getsize(os.path.join(dp, f)) for dp, _, fs in os.walk(OUT) for f in fs)
    print("fixtures: %.2f MB under %s" % (size / 1e6, OUT))


if __name__ == "__main__":
    main()
