"""Host-side mirror of the reference's four entry points for the hot path (same names, argument
meaning and error behaviour), driving the CUDA library through the C-ABI.

  cache_sample_weights   ref: externals/apegrunt/include/apegrunt/Alignment_utility.hpp:186-258
  determine_MI_threshold ref: include/SpydrPick.hpp:285-343
  get_MI_network         ref: include/SpydrPick.hpp:132-168
  run_ARACNE             ref: include/ARACNE.hpp:550

The reference reads its options from static singletons (SpydrPick_options / ARACNE_options /
Apegrunt_options); here they are one explicit `Options` object with the same names and defaults.
"""
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from .capi import Context, EDGE_DTYPE

DBL_EPSILON = 2.220446049250313e-16


@dataclass
class Options:
    # SpydrPick_options (src/SpydrPick_options.cpp:37-43,144-159)
    mi_threshold: float = -1.0            # <0 = auto
    mi_values: int = 0                    # 0 -> min(1e7, 100*L)  (src/SpydrPick.cpp:338-339)
    mi_pseudocount: float = 0.5
    mi_threshold_iterations: int = 10
    mi_threshold_pairs: int = 0           # 0 = auto
    ld_threshold: int = 0
    no_aracne: bool = False
    verbose: bool = False
    # ARACNE_options (src/ARACNE_options.cpp:97-102)
    aracne_edge_threshold: float = DBL_EPSILON
    aracne_block_size: int = 16384        # accepted for compatibility; the result does not depend on it
    # Apegrunt_options (externals/apegrunt/src/Apegrunt_options.cpp:119-145)
    no_sample_reweighting: bool = False
    sample_reweighting_threshold: float = 0.9
    rescale_sample_weights: bool = False
    linear_genome: bool = False
    output_indexing_base: int = 1
    # additions of this implementation
    seed: int = 1                         # the reference seeds from the clock (SpydrPick.hpp:178)
    reference_quirks: object = True       # True = SPK_QUIRKS_REFERENCE: the L % 32 == 0 final-block double count and the
                                          # sample-0 double count of calculate_weights; False = the plain rule; or a SPK_QUIRK_* mask
    n_terms: int = 0                      # bf16 weight-split terms of the tensor-core filter (0 = default)
    band: float = 0.0
    engine: int = 0                       # 0 = tcgen05 filter + fp64 refine, 1 = fp64 exhaustive (diagnostic)


class Alignment:
    """What the hot path sees of apegrunt::Alignment: dense states, loci translation, weights."""

    def __init__(self, states, pos=None, n_original_positions=None, id_string="alignment", ctx: Optional[Context] = None,
                 device=-1):
        self.states = np.ascontiguousarray(states, dtype=np.uint8)
        if self.states.ndim != 2 or self.states.max(initial=0) > 4:
            raise ValueError("states must be a 2D array of codes 0..4 (a,t,c,g,GAP)")
        self.n, self.L = self.states.shape
        self.pos = np.arange(self.L, dtype=np.uint64) if pos is None else np.ascontiguousarray(pos, dtype=np.uint64)
        # the reference: genome_size if given, else the number of columns of the parsed FASTA (Alignment_parsers.hpp:67,133);
        # a mapping list does not change it (Alignment_utility.hpp:128-160)
        self.n_original_positions = int(n_original_positions) if n_original_positions is not None else self.L
        self.id_string = id_string
        self.weights = np.ones(self.n, dtype=np.float64)       # StateVector default weight = 1
        self.neighbour_counts = None
        self.ctx = ctx or Context(device=device)
        self.ctx.set_alignment(self.states, self.pos)

    @classmethod
    def filtered(cls, raw_states, pos=None, n_original_positions=None, id_string="alignment", maf_threshold=0.01,
                 gap_threshold=0.15, ctx: Optional[Context] = None, device=-1):
        """apegrunt::Alignment_filter (Alignment_filter.hpp:250-339; src/SpydrPick.cpp:234-297) on the device: the RAW
        alignment is uploaded once, state frequencies are counted and the accepted columns extracted in HBM
        (spk_set_alignment_filtered). Returns the filtered Alignment; `.keep` is the accept mask over the raw columns."""
        raw = np.ascontiguousarray(raw_states, dtype=np.uint8)
        if raw.ndim != 2 or raw.max(initial=0) > 4:
            raise ValueError("states must be a 2D array of codes 0..4 (a,t,c,g,GAP)")
        rpos = np.arange(raw.shape[1], dtype=np.uint64) if pos is None else np.ascontiguousarray(pos, dtype=np.uint64)
        self = cls.__new__(cls)
        self.ctx = ctx or Context(device=device)
        self.keep = self.ctx.set_alignment_filtered(raw, rpos, maf_threshold, gap_threshold)
        self.states = np.ascontiguousarray(raw[:, self.keep])          # host view of what the device extracted
        self.n, self.L = self.states.shape
        self.pos = rpos[self.keep]
        self.n_original_positions = int(n_original_positions) if n_original_positions is not None else raw.shape[1]
        self.id_string = "%s.filtered_ge%03dmaf_le%03dgf_gt01-lt05states" % (id_string, int(maf_threshold * 100.), int(gap_threshold * 100.))
        self.weights = np.ones(self.n, dtype=np.float64)
        self.neighbour_counts = None
        return self

    def n_loci(self):
        return self.L

    def size(self):
        return self.n

    def effective_size(self):
        return float(self.weights.sum())


def cache_sample_weights(alignment: Alignment, options: Options = None, sample_weights=None):
    """ref Alignment_utility.hpp:186-258: none / user supplied / computed (K1), optional rescale."""
    options = options or Options()
    if options.no_sample_reweighting:
        return alignment.weights
    if sample_weights is not None:
        w = np.ascontiguousarray(sample_weights, dtype=np.float64)
        if w.shape != (alignment.n,):
            raise ValueError("sample weights: expected %d values" % alignment.n)
    else:
        # min_identical = size_t(threshold * double(L)): Alignment_StateVector_weights.hpp:228,251
        min_identical = int(options.sample_reweighting_threshold * float(alignment.L))
        counts = alignment.ctx.neighbour_counts(min_identical, options.reference_quirks)
        alignment.neighbour_counts = counts
        w = 1.0 / counts.astype(np.float64)                    # weights.hpp:349
    if options.rescale_sample_weights:
        n_eff = 0.0
        for x in w.tolist():                                   # std::accumulate, left to right (Alignment_utility.hpp:208,233)
            n_eff += x
        w = (w * float(alignment.n)) / n_eff                   # (w * n) / n_eff: Alignment_utility.hpp:215,245
    alignment.weights = w
    alignment.ctx.set_weights(w)
    return w


def determine_MI_threshold(alignment: Alignment, n_values: int, options: Options = None) -> float:
    options = options or Options()
    return alignment.ctx.determine_mi_threshold(n_values, options.mi_threshold_pairs, options.mi_threshold_iterations,
                                                options.mi_pseudocount, options.seed)


@dataclass
class MINetwork:
    """ref include/SpydrPick.hpp:58-67 (MI_network) + the ARACNE flag column of the Graph."""
    network: np.ndarray                    # EDGE_DTYPE, as stored (sorted after sort())
    network_wo_gaps: np.ndarray
    outlier_threshold: float
    extreme_outlier_threshold: float
    colmax: np.ndarray = None
    flags: np.ndarray = None               # bool(edge): 1 = direct after run_ARACNE, 0 before / indirect
    info: dict = field(default_factory=dict)
    ctx: Context = None

    def sort(self):
        self.network = self.ctx.sort_edges(self.network)
        return self


def get_MI_network(alignment: Alignment, mi_threshold: float = 0.0, options: Options = None, sort=False,
                   part_index=0, n_parts=1) -> MINetwork:
    options = options or Options()
    r = alignment.ctx.mi_scan(mi_threshold, options.mi_pseudocount, options.ld_threshold, options.linear_genome,
                              circle_size=0, max_edges=options.mi_values * 2 if options.mi_values else 0,
                              sort_desc=sort, n_terms=options.n_terms, band=options.band, engine=options.engine,
                              part_index=part_index, n_parts=n_parts)
    outlier, extreme = alignment.ctx.outlier_thresholds(r["colmax"])
    return MINetwork(r["edges"], r["edges_wo_gaps"], outlier, extreme, r["colmax"],
                     np.zeros(r["n_edges"], dtype=np.uint8),
                     {"n_candidates": r["n_candidates"], "max_filter_error": r["max_filter_error"]}, alignment.ctx)


def run_ARACNE(network: MINetwork, options: Options = None):
    """In place, like the reference: sets the flag on DIRECT edges; precondition: sorted."""
    options = options or Options()
    network.flags = network.ctx.aracne(network.network, options.aracne_edge_threshold)
    return network


# ---- writers (ref graph/Graph_output_format.hpp:49-74, include/SpydrPick.hpp:89-129) -----------------
def _distance(i1, i2, linear, circle):
    d = np.abs(i1.astype(np.int64) - i2.astype(np.int64)).astype(np.uint64)
    if linear:
        return d
    return np.minimum(np.uint64(circle) - d, d)            # uint64 wrap for d > circle, as the reference


def format_edges(network: MINetwork, alignment: Alignment, options: Options = None) -> str:
    options = options or Options()
    e = network.network
    n1 = np.minimum(e["i"], e["j"]); n2 = np.maximum(e["i"], e["j"])
    i1 = alignment.pos[n1] + np.uint64(options.output_indexing_base)
    i2 = alignment.pos[n2] + np.uint64(options.output_indexing_base)
    d = _distance(i1, i2, options.linear_genome, alignment.n_original_positions)
    f = network.flags if network.flags is not None else np.zeros(len(e), dtype=np.uint8)
    return "".join("%d %d %d %d %.6f\n" % t for t in zip(i1.tolist(), i2.tolist(), d.tolist(), f.tolist(), e["mi"].tolist()))


def format_outliers(network: MINetwork, alignment: Alignment, options: Options = None) -> str:
    options = options or Options()
    e = network.network
    wo = network.network_wo_gaps
    wo_map = {(int(a), int(b)): float(w) for a, b, w in zip(np.minimum(wo["i"], wo["j"]), np.maximum(wo["i"], wo["j"]), wo["mi"])}
    out = []
    f = network.flags if network.flags is not None else np.zeros(len(e), dtype=np.uint8)
    base = options.output_indexing_base
    for k in range(len(e)):
        w = float(e["mi"][k])
        if w < network.outlier_threshold:
            break
        a, b = int(min(e["i"][k], e["j"][k])), int(max(e["i"][k], e["j"][k]))
        w2 = wo_map.get((a, b), w)
        i1 = int(alignment.pos[a]) + base; i2 = int(alignment.pos[b]) + base
        d = int(_distance(np.array([i1], dtype=np.uint64), np.array([i2], dtype=np.uint64), options.linear_genome,
                          alignment.n_original_positions)[0])
        if d > options.ld_threshold:
            out.append("%d %d %d %d %.6f %.6f %.1f %d\n" % (i1, i2, d, int(f[k]), w, w2, (1.0 - (w2 / w)) * 100,
                                                          int(w > network.extreme_outlier_threshold)))
    return "".join(out)


def run_pipeline(alignment: Alignment, options: Options = None) -> MINetwork:
    """The hot-path part of main() (src/SpydrPick.cpp:321-421): weights -> threshold -> scan -> sort -> ARACNE."""
    options = options or Options()
    cache_sample_weights(alignment, options)
    thr = options.mi_threshold
    if thr < 0:
        top = options.mi_values if options.mi_values else min(10_000_000, 100 * alignment.L)
        thr = determine_MI_threshold(alignment, top, options)
    net = get_MI_network(alignment, thr, options, sort=True)
    net.info["mi_threshold"] = thr
    if not options.no_aracne:
        run_ARACNE(net, options)
    return net
