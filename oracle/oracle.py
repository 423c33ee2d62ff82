"""ctypes wrapper around oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module. The reference ships no tests of its own; the oracle is pinned against the REAL reference compiled in
place (oracle/ref.py, oracle/Makefile.ref): tests/test_ref_pins.py (unit level) and tests/test_ref_golden.py (listings).
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

EDGE_DTYPE = np.dtype([("i", np.uint32), ("j", np.uint32), ("mi", np.float64)], align=True)


class _ScanResult(C.Structure):
    _fields_ = [("edges", C.c_void_p), ("n_edges", C.c_uint64),
                ("edges_wo_gaps", C.c_void_p), ("n_wo_gaps", C.c_uint64),
                ("colmax", C.POINTER(C.c_double))]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.orc_mi_from_table.restype = C.c_double
        L.orc_mi_pair.restype = C.c_double
        L.orc_determine_mi_threshold.restype = C.c_double
        L.orc_determine_threshold_pairs.restype = C.c_uint64
        L.orc_filter_columns.restype = C.c_uint64
        L.orc_format_edges.restype = C.c_void_p
        L.orc_format_outliers.restype = C.c_void_p
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _states(states):
    s = np.ascontiguousarray(states, dtype=np.uint8)
    assert s.ndim == 2
    return s


def hardware_threads():
    return int(lib().orc_hardware_threads())


QUIRK_LASTBLOCK, QUIRK_SAMPLE0, QUIRKS_REFERENCE = 1, 2, 3


def neighbour_counts(states, min_identical, quirks=True, threads=0):
    """quirks: True = everything the reference does (QUIRKS_REFERENCE), False = the plain rule, or a QUIRK_* bit mask."""
    s = _states(states); n, L = s.shape
    out = np.zeros(n, dtype=np.uint64)
    q = QUIRKS_REFERENCE if quirks is True else (0 if quirks is False else int(quirks))
    lib().orc_neighbour_counts(_p(s), C.c_uint64(n), C.c_uint64(L), C.c_uint64(int(min_identical)),
                               C.c_int(q), C.c_int(threads), _p(out))
    return out


def weights_from_counts(counts, rescale=False):
    c = np.ascontiguousarray(counts, dtype=np.uint64)
    w = np.zeros(c.shape[0], dtype=np.float64)
    lib().orc_weights_from_counts(_p(c), C.c_uint64(c.shape[0]), C.c_int(int(rescale)), _p(w))
    return w


def state_masks(states):
    s = _states(states); n, L = s.shape
    m = np.zeros(L, dtype=np.uint8)
    lib().orc_state_masks(_p(s), C.c_uint64(n), C.c_uint64(L), _p(m))
    return m


def mi_from_table(table, ipresence, jpresence, pseudocount=0.5):
    """table[jstate][istate] (5x5). Returns (mi, modified table)."""
    t = np.array(table, dtype=np.float64).reshape(25).copy()
    mi = lib().orc_mi_from_table(_p(t), C.c_uint(ipresence), C.c_uint(jpresence), C.c_double(pseudocount))
    return float(mi), t.reshape(5, 5)


def mi_pair(states, w, ci, cj, pseudocount=0.5, masks=None, wo_gaps=False):
    s = _states(states); n, L = s.shape
    w = np.ascontiguousarray(w, dtype=np.float64)
    if masks is None:
        masks = state_masks(s)
    masks = np.ascontiguousarray(masks, dtype=np.uint8)
    wo = C.c_double(0.0)
    mi = lib().orc_mi_pair(_p(s), C.c_uint64(n), C.c_uint64(L), _p(w), _p(masks), C.c_uint64(ci), C.c_uint64(cj),
                           C.c_double(pseudocount), C.byref(wo) if wo_gaps else None)
    return (float(mi), float(wo.value)) if wo_gaps else float(mi)


def mi_pairs(states, w, pi, pj, pseudocount=0.5, threads=0):
    s = _states(states); n, L = s.shape
    w = np.ascontiguousarray(w, dtype=np.float64)
    pi = np.ascontiguousarray(pi, dtype=np.uint32); pj = np.ascontiguousarray(pj, dtype=np.uint32)
    out = np.zeros(pi.shape[0], dtype=np.float64)
    lib().orc_mi_pairs(_p(s), C.c_uint64(n), C.c_uint64(L), _p(w), _p(pi), _p(pj), C.c_uint64(pi.shape[0]),
                       C.c_double(pseudocount), C.c_int(threads), _p(out))
    return out


def mi_scan(states, w, pos=None, threshold=0.0, pseudocount=0.5, ld_threshold=0, linear_genome=False,
            circle_size=None, threads=0):
    """Returns dict(edges, edges_wo_gaps (structured arrays, unsorted), colmax)."""
    s = _states(states); n, L = s.shape
    w = np.ascontiguousarray(w, dtype=np.float64)
    pos = np.arange(L, dtype=np.uint64) if pos is None else np.ascontiguousarray(pos, dtype=np.uint64)
    if circle_size is None:
        circle_size = L
    r = _ScanResult()
    lib().orc_mi_scan(_p(s), C.c_uint64(n), C.c_uint64(L), _p(w), _p(pos), C.c_double(threshold),
                      C.c_double(pseudocount), C.c_uint64(ld_threshold), C.c_int(int(linear_genome)),
                      C.c_uint64(circle_size), C.c_int(threads), C.byref(r))
    def grab(ptr, cnt):
        if cnt == 0:
            return np.zeros(0, dtype=EDGE_DTYPE)
        buf = (C.c_char * (cnt * EDGE_DTYPE.itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=EDGE_DTYPE).copy()
    edges = grab(r.edges, r.n_edges); wo = grab(r.edges_wo_gaps, r.n_wo_gaps)
    colmax = np.ctypeslib.as_array(r.colmax, shape=(L,)).copy()
    lib().orc_free_scan(C.byref(r))
    return {"edges": edges, "edges_wo_gaps": wo, "colmax": colmax}


def outlier_thresholds(colmax):
    c = np.ascontiguousarray(colmax, dtype=np.float64)
    a = C.c_double(); b = C.c_double()
    lib().orc_outlier_thresholds(_p(c), C.c_uint64(c.shape[0]), C.byref(a), C.byref(b))
    return float(a.value), float(b.value)


def sort_edges(edges):
    e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE).copy()
    lib().orc_sort_edges(_p(e), C.c_uint64(e.shape[0]))
    return e


def aracne_blocked(edges, threshold=2.220446049250313e-16, block_size=16384):
    e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
    f = np.zeros(e.shape[0], dtype=np.uint8)
    lib().orc_aracne_blocked(_p(e), C.c_uint64(e.shape[0]), C.c_double(threshold), C.c_uint32(block_size), _p(f))
    return f


def aracne_closed(edges, threshold=2.220446049250313e-16, threads=0):
    e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
    f = np.zeros(e.shape[0], dtype=np.uint8)
    lib().orc_aracne_closed(_p(e), C.c_uint64(e.shape[0]), C.c_double(threshold), C.c_int(threads), _p(f))
    return f


def determine_threshold_pairs(threshold_pairs, possible_pairs, pct):
    return int(lib().orc_determine_threshold_pairs(C.c_uint64(threshold_pairs), C.c_uint64(possible_pairs), C.c_double(pct)))


def sample_pairs(count, max_range, seed):
    pi = np.zeros(count, dtype=np.uint32); pj = np.zeros(count, dtype=np.uint32)
    lib().orc_sample_pairs(C.c_uint64(count), C.c_uint64(max_range), C.c_uint64(seed), _p(pi), _p(pj))
    return pi, pj


def determine_mi_threshold(states, w, n_values, threshold_pairs=0, iterations=10, pseudocount=0.5, seed=1, threads=0):
    s = _states(states); n, L = s.shape
    w = np.ascontiguousarray(w, dtype=np.float64)
    return float(lib().orc_determine_mi_threshold(_p(s), C.c_uint64(n), C.c_uint64(L), _p(w), C.c_uint64(n_values),
                                                  C.c_uint64(threshold_pairs), C.c_uint64(iterations),
                                                  C.c_double(pseudocount), C.c_uint64(seed), C.c_int(threads)))


def _take_str(ptr):
    s = C.string_at(ptr).decode()
    lib().orc_free_str(C.c_void_p(ptr))
    return s


def format_edges(edges, flags, pos, base_index=1, linear_genome=False, n_original_positions=0):
    e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
    f = np.ascontiguousarray(flags, dtype=np.uint8)
    pos = np.ascontiguousarray(pos, dtype=np.uint64)
    return _take_str(lib().orc_format_edges(_p(e), _p(f), C.c_uint64(e.shape[0]), _p(pos), C.c_uint64(base_index),
                                            C.c_int(int(linear_genome)), C.c_uint64(n_original_positions)))


def format_outliers(edges, flags, wo_gaps, pos, outlier, extreme, base_index=1, linear_genome=False,
                    n_original_positions=0, ld_threshold=0):
    e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
    f = np.ascontiguousarray(flags, dtype=np.uint8)
    wo = np.ascontiguousarray(wo_gaps, dtype=EDGE_DTYPE)
    pos = np.ascontiguousarray(pos, dtype=np.uint64)
    return _take_str(lib().orc_format_outliers(_p(e), _p(f), C.c_uint64(e.shape[0]), _p(wo), C.c_uint64(wo.shape[0]),
                                               _p(pos), C.c_uint64(base_index), C.c_int(int(linear_genome)),
                                               C.c_uint64(n_original_positions), C.c_uint64(ld_threshold),
                                               C.c_double(outlier), C.c_double(extreme)))


def filter_columns(states, maf=0.01, gap_threshold=0.15):
    s = _states(states); n, L = s.shape
    keep = np.zeros(L, dtype=np.uint8)
    lib().orc_filter_columns(_p(s), C.c_uint64(n), C.c_uint64(L), C.c_double(maf), C.c_double(gap_threshold), _p(keep))
    return keep.astype(bool)
