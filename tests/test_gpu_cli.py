"""End-to-end test of the C++ host (host/spydrpick_main.cpp -> host/build/SpydrPick): FASTA in, the reference's
`.edges` / `.outliers` listings out, byte-compared with the CPU oracle's writers on the same filtered alignment
(the exact-tie caveat of ARACNE aside, see test_gpu_zscan.py). Also CPU-side: option errors exit non-zero."""
import glob
import os
import subprocess

import numpy as np
import pytest

import oracle.oracle as orc
from spydrpick_b200.synth import generate_alignment, to_fasta

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "host", "build", "SpydrPick")


def _ensure_exe():
    if not os.path.exists(EXE):
        import __graft_entry__ as g
        g.build_host()
    assert os.path.exists(EXE)


def test_cli_usage_errors():
    _ensure_exe()
    r = subprocess.run([EXE], capture_output=True, text=True)
    assert r.returncode != 0 and "No alignment file specified" in r.stderr
    r = subprocess.run([EXE, "--no-such-option", "x.fasta"], capture_output=True, text=True)
    assert r.returncode != 0 and "unrecognised option" in r.stderr
    r = subprocess.run([EXE, "/nonexistent.fasta"], capture_output=True, text=True)
    assert r.returncode != 0 and "could not open alignment file" in r.stderr
    r = subprocess.run([EXE, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--mi-threshold" in r.stdout and "--no-aracne" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("linear,ld", [(False, 0), (True, 40)])
def test_cli_outputs_match_oracle(tmp_path, linear, ld):
    _ensure_exe()
    n, L = 120, 900
    s, pos, gsize = generate_alignment(n, L, seed=20260101)
    rng = np.random.default_rng(3)
    # raw alignment = SNP columns interleaved with invariant / gappy columns the default filter must drop
    raw = np.zeros((n, 2 * L), dtype=np.uint8)
    raw[:, 0::2] = s
    raw[:, 1::2] = rng.integers(0, 4, size=L)[None, :]            # invariant
    raw[rng.random((n, 2 * L)) < 0.3 * (np.arange(2 * L) % 7 == 3)[None, :]] = 4   # some columns with ~30% gaps
    fasta = tmp_path / "toy.fasta"
    to_fasta(raw, str(fasta))
    thr = 0.05
    args = [EXE, str(fasta), "--mi-threshold", str(thr), "--ld-threshold", str(ld), "--output-dir", str(tmp_path), "-v"]
    if linear:
        args.append("--linear-genome")
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout

    keep = orc.filter_columns(raw).astype(bool)
    f = np.ascontiguousarray(raw[:, keep]); fpos = np.nonzero(keep)[0].astype(np.uint64)
    Lf = f.shape[1]
    assert 2 <= Lf < 2 * L
    counts = orc.neighbour_counts(f, int(0.9 * Lf), True)
    w = orc.weights_from_counts(counts)
    sc = orc.mi_scan(f, w, fpos, threshold=thr, ld_threshold=ld, linear_genome=linear, circle_size=Lf)
    e = orc.sort_edges(sc["edges"])
    outlier, extreme = orc.outlier_thresholds(sc["colmax"])

    edges_files = glob.glob(str(tmp_path / ("toy.filtered_ge001maf_le015gf_gt01-lt05states.L%dn%d.spydrpick_couplings.1-based.*edges" % (Lf, n))))
    assert len(edges_files) == 1 and edges_files[0].endswith(".%dedges" % len(e))
    got = open(edges_files[0]).read().splitlines()
    assert len(got) == len(e) > 0
    # same edges, positions, distances and 6-decimal MI; flags compared through the oracle on the CLI's own order
    g_rows = [ln.split() for ln in got]
    key = lambda p1, p2: (int(p1), int(p2))
    exp_text = orc.format_edges(e, orc.aracne_closed(e), fpos, 1, linear, 2 * L).splitlines()
    exp = {key(*ln.split()[:2]): ln.split() for ln in exp_text}
    assert {key(r_[0], r_[1]) for r_ in g_rows} == set(exp.keys())
    nflagdiff = 0
    for r_ in g_rows:
        x = exp[key(r_[0], r_[1])]
        assert r_[2] == x[2]
        assert abs(float(r_[4]) - float(x[4])) <= 1.1e-6          # printed with 6 decimals
        nflagdiff += r_[3] != x[3]
    assert nflagdiff <= 0.002 * len(e) + 1                        # exact-tie triangles only
    mis = [float(r_[4]) for r_ in g_rows]
    assert all(a >= b for a, b in zip(mis, mis[1:]))
    out_file = str(tmp_path / ("toy.filtered_ge001maf_le015gf_gt01-lt05states.L%dn%d.spydrpick_couplings.1-based.outliers" % (Lf, n)))
    got_o = open(out_file).read().splitlines()
    exp_o = orc.format_outliers(e, orc.aracne_closed(e), sc["edges_wo_gaps"], fpos, outlier, extreme, 1, linear, 2 * L, ld).splitlines()
    assert len(got_o) == len(exp_o)
    eo = {key(*ln.split()[:2]): ln.split() for ln in exp_o}
    for ln in got_o:
        r_ = ln.split(); x = eo[key(r_[0], r_[1])]
        assert r_[2] == x[2] and r_[7] == x[7]
        assert abs(float(r_[4]) - float(x[4])) <= 1.1e-6 and abs(float(r_[5]) - float(x[5])) <= 1.1e-6
        assert abs(float(r_[6]) - float(x[6])) <= 0.11
