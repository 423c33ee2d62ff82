"""ctypes binding of include/spydrpick_b200.h (the C-ABI a reference maintainer would bind).

Fails loudly when the CUDA library is missing: there is no CPU fallback in the product path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPYDRPICK_B200_LIB") or os.path.join(_HERE, "libspydrpick_b200.so")

EDGE_DTYPE = np.dtype([("i", np.uint32), ("j", np.uint32), ("mi", np.float64)], align=True)
assert EDGE_DTYPE.itemsize == 16

QUIRK_LASTBLOCK, QUIRK_SAMPLE0, QUIRKS_REFERENCE = 1, 2, 3
SPK_OK, SPK_E_INVALID, SPK_E_CUDA, SPK_E_NOMEM, SPK_E_OVERFLOW, SPK_E_PRECISION = 0, -1, -2, -3, -4, -5

EXPORTS = ["spk_create", "spk_destroy", "spk_last_error", "spk_set_stream", "spk_host_alloc", "spk_host_free",
           "spk_set_alignment", "spk_set_alignment_device", "spk_column_state_counts", "spk_set_alignment_filtered",
           "spk_neighbour_counts", "spk_neighbour_counts_part", "spk_neighbour_counts_noalloc", "spk_reweight", "spk_set_weights", "spk_mi_pairs", "spk_determine_mi_threshold",
           "spk_determine_mi_threshold_ex", "spk_comm_init_all", "spk_comm_unique_id", "spk_comm_init_rank", "spk_comm_rank",
           "spk_mi_scan", "spk_outlier_thresholds", "spk_sort_edges", "spk_aracne", "spk_sort_edges_device",
           "spk_aracne_device", "spk_merge_sorted_runs_device", "spk_wait_download", "spk_get_timings", "spk_debug_tile_counts", "spk_debug_i8_mode", "spk_debug_i8_peak", "spk_debug_i8_layout", "spk_debug_i8_tile",
           "spk_version"]


class ScanParams(C.Structure):
    _fields_ = [("threshold", C.c_double), ("pseudocount", C.c_double), ("ld_threshold", C.c_uint64),
                ("linear_genome", C.c_int32), ("circle_size", C.c_uint64), ("max_edges", C.c_uint64),
                ("sort_desc", C.c_int32), ("n_terms", C.c_int32), ("band", C.c_double), ("engine", C.c_int32),
                ("part_index", C.c_uint32), ("n_parts", C.c_uint32), ("keep_on_device", C.c_int32)]


class ScanResult(C.Structure):
    _fields_ = [("edges", C.c_void_p), ("n_edges", C.c_uint64), ("edges_wo_gaps", C.c_void_p), ("n_wo_gaps", C.c_uint64),
                ("colmax", C.POINTER(C.c_double)), ("d_edges", C.c_void_p), ("d_edges_wo_gaps", C.c_void_p),
                ("d_colmax", C.c_void_p), ("n_candidates", C.c_uint64), ("max_filter_error", C.c_double),
                ("audit_pairs", C.c_uint64), ("audit_violations", C.c_uint64), ("audit_max_error", C.c_double), ("band", C.c_double)]


class Timings(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("upload_ms", "prep_ms", "hamming_ms", "expand_ms", "stats_ms", "gemm_ms",
                                        "refine_ms", "compact_ms", "sort_ms", "aracne_build_ms", "aracne_ms", "pairs_ms")] + \
               [("gemm_tiles", C.c_uint64), ("kernel_launches", C.c_uint64), ("gemm_pairs", C.c_uint64), ("seed_ms", C.c_float),
                ("exchange_ms", C.c_float), ("audit_ms", C.c_float), ("threshold_ms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class SpkError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("spydrpick_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load():
    """Loads the CUDA library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("spydrpick_b200: %s not built (run `python -m spydrpick_b200.build`); "
                          "there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.spk_last_error.restype = C.c_char_p
    L.spk_last_error.argtypes = [C.c_void_p]
    L.spk_version.restype = C.c_char_p
    L.spk_destroy.restype = None
    L.spk_destroy.argtypes = [C.c_void_p]
    L.spk_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int]
    L.spk_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.spk_set_alignment.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    L.spk_set_alignment_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    L.spk_neighbour_counts.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    L.spk_neighbour_counts_part.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p]
    L.spk_neighbour_counts_noalloc.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
    L.spk_set_weights.argtypes = [C.c_void_p, C.c_void_p]
    L.spk_mi_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_void_p]
    L.spk_determine_mi_threshold.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_double, C.c_uint64,
                                             C.POINTER(C.c_double)]
    L.spk_determine_mi_threshold_ex.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_double, C.c_uint64, C.c_int,
                                                C.POINTER(C.c_double)]
    L.spk_reweight.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.spk_comm_unique_id.argtypes = [C.c_void_p]
    L.spk_comm_init_rank.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.spk_comm_init_all.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    L.spk_comm_rank.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.spk_mi_scan.argtypes = [C.c_void_p, C.POINTER(ScanParams), C.POINTER(ScanResult)]
    L.spk_outlier_thresholds.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.spk_sort_edges.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    L.spk_aracne.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_uint32, C.c_uint32, C.c_void_p]
    L.spk_sort_edges_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    L.spk_column_state_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p]
    L.spk_set_alignment_filtered.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_double, C.c_double,
                                             C.c_void_p, C.POINTER(C.c_uint64)]
    L.spk_merge_sorted_runs_device.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.c_uint32]
    L.spk_aracne_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_uint32, C.c_uint32, C.c_void_p]
    L.spk_get_timings.argtypes = [C.c_void_p, C.POINTER(Timings)]
    L.spk_debug_tile_counts.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int32, C.c_void_p]
    L.spk_wait_download.argtypes = [C.c_void_p]
    L.spk_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    L.spk_host_free.restype = None
    L.spk_host_free.argtypes = [C.c_void_p]
    L.spk_debug_i8_layout.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                      C.POINTER(C.c_double)]
    L.spk_debug_i8_peak.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_double), C.POINTER(C.c_float)]
    L.spk_debug_i8_tile.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    _lib = L
    return L


def neighbour_counts_noalloc(d_states_ptr, n, L, min_identical, d_counts_ptr, quirks=True, stream=None):
    """spk_neighbour_counts_noalloc: device pointers in, nothing allocated; enqueues on `stream` (a cudaStream_t handle, None =
    default stream) of the current device. quirks: True = the reference's behaviour, False = none, or a mask."""
    lib = load()
    q = QUIRKS_REFERENCE if quirks is True else int(quirks)
    rc = lib.spk_neighbour_counts_noalloc(C.c_void_p(d_states_ptr), n, L, int(min_identical), q, C.c_void_p(d_counts_ptr), C.c_void_p(stream or 0))
    if rc != 0:
        raise SpkError(rc, (lib.spk_last_error(None) or b"").decode())


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class _PinnedOwner:
    def __init__(self, lib, ptr):
        self.lib, self.ptr = lib, ptr

    def __del__(self):
        try:
            self.lib.spk_host_free(self.ptr)
        except Exception:
            pass


def pinned_empty(shape, dtype=np.uint8):
    """numpy array backed by page-locked host memory (spk_host_alloc); freed with the array."""
    lib = load()
    dt = np.dtype(dtype)
    nbytes = int(np.prod(shape)) * dt.itemsize
    p = C.c_void_p()
    rc = lib.spk_host_alloc(max(nbytes, 16), C.byref(p))
    if rc != SPK_OK:
        raise SpkError(rc, "pinned host allocation of %d bytes failed" % nbytes)
    buf = (C.c_char * max(nbytes, 16)).from_address(p.value)
    buf._spk_owner = _PinnedOwner(lib, p)
    return np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)


def pinned_copy(a):
    out = pinned_empty(a.shape, a.dtype)
    np.copyto(out, a)
    return out


class Context:
    """Thin RAII wrapper over spk_ctx."""

    def __init__(self, device=-1, verbose=False):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.spk_create(C.byref(h), device, int(verbose))
        if rc != SPK_OK:
            raise SpkError(rc, (self.lib.spk_last_error(None) or b"").decode())
        self.h = h
        self.n = self.L = 0
        self.rank, self.world = 0, 1

    def close(self):
        if getattr(self, "h", None):
            self.lib.spk_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc != SPK_OK:
            raise SpkError(rc, (self.lib.spk_last_error(self.h) or b"").decode())

    def set_stream(self, cuda_stream_handle):
        self._check(self.lib.spk_set_stream(self.h, C.c_void_p(cuda_stream_handle)))

    def set_alignment(self, states, pos=None):
        s = np.ascontiguousarray(states, dtype=np.uint8)
        assert s.ndim == 2
        p = None if pos is None else np.ascontiguousarray(pos, dtype=np.uint64)
        self._check(self.lib.spk_set_alignment(self.h, _ptr(s), s.shape[0], s.shape[1], _ptr(p)))
        self.n, self.L = s.shape

    def column_state_counts(self, states):
        """Unweighted state counts [L, 5] of a RAW alignment (counted on the device)."""
        s = np.ascontiguousarray(states, dtype=np.uint8)
        assert s.ndim == 2
        out = np.zeros((s.shape[1], 5), dtype=np.uint32)
        self._check(self.lib.spk_column_state_counts(self.h, _ptr(s), s.shape[0], s.shape[1], _ptr(out)))
        return out

    def set_alignment_filtered(self, states, pos=None, maf_threshold=0.01, gap_threshold=0.15):
        """Column filter (Alignment_filter.hpp:250-339) + set_alignment of the kept columns, the raw matrix uploaded once.
        Returns the boolean keep mask over the raw columns."""
        s = np.ascontiguousarray(states, dtype=np.uint8)
        assert s.ndim == 2
        p = None if pos is None else np.ascontiguousarray(pos, dtype=np.uint64)
        keep = np.zeros(s.shape[1], dtype=np.uint8)
        nk = C.c_uint64(0)
        self._check(self.lib.spk_set_alignment_filtered(self.h, _ptr(s), s.shape[0], s.shape[1], _ptr(p), float(maf_threshold),
                                                        float(gap_threshold), _ptr(keep), C.byref(nk)))
        self.n, self.L = s.shape[0], int(nk.value)
        return keep.astype(bool)

    def set_alignment_device(self, d_states, n, L, pos=None):
        """d_states: raw device pointer (int) to row-major uint8[n][L] in this context's device memory."""
        p = None if pos is None else np.ascontiguousarray(pos, dtype=np.uint64)
        self._check(self.lib.spk_set_alignment_device(self.h, C.c_void_p(int(d_states)), int(n), int(L), _ptr(p)))
        self.n, self.L = int(n), int(L)

    def neighbour_counts(self, min_identical, quirks=True, part_index=0, n_parts=1):
        """quirks: True = SPK_QUIRKS_REFERENCE (everything the reference does), False = the plain rule, or a SPK_QUIRK_* mask.
        n_parts > 1: PARTIAL counts of this partition of the sample-tile pairs (sum over partitions = full counts)."""
        out = np.zeros(self.n, dtype=np.uint64)
        q = QUIRKS_REFERENCE if quirks is True else (0 if quirks is False else int(quirks))
        self._check(self.lib.spk_neighbour_counts_part(self.h, int(min_identical), q, part_index, n_parts, _ptr(out)))
        return out

    def reweight(self, min_identical, quirks=True, rescale=False):
        """K1 + weights in one call (collective on a context with a communicator). Returns (weights, counts)."""
        q = QUIRKS_REFERENCE if quirks is True else (0 if quirks is False else int(quirks))
        w = np.zeros(self.n, dtype=np.float64); cnt = np.zeros(self.n, dtype=np.uint64)
        self._check(self.lib.spk_reweight(self.h, int(min_identical), q, int(rescale), _ptr(w), _ptr(cnt)))
        return w, cnt

    # ---- multi-GPU: one context per device / process (SPMD)
    @staticmethod
    def comm_unique_id():
        lib = load(); buf = (C.c_char * 128)()
        rc = lib.spk_comm_unique_id(buf)
        if rc != SPK_OK:
            raise SpkError(rc, (lib.spk_last_error(None) or b"").decode())
        return bytes(buf)

    def comm_init_rank(self, unique_id: bytes, rank: int, world: int):
        assert len(unique_id) == 128
        self._check(self.lib.spk_comm_init_rank(self.h, C.c_char_p(unique_id), int(rank), int(world)))
        self.rank, self.world = int(rank), int(world)

    def set_weights(self, w):
        w = np.ascontiguousarray(w, dtype=np.float64)
        assert w.shape == (self.n,)
        self._check(self.lib.spk_set_weights(self.h, _ptr(w)))

    def mi_pairs(self, pi, pj, pseudocount=0.5):
        pi = np.ascontiguousarray(pi, dtype=np.uint32); pj = np.ascontiguousarray(pj, dtype=np.uint32)
        out = np.zeros(pi.shape[0], dtype=np.float64)
        self._check(self.lib.spk_mi_pairs(self.h, _ptr(pi), _ptr(pj), pi.shape[0], pseudocount, _ptr(out)))
        return out

    def determine_mi_threshold(self, n_values, threshold_pairs=0, iterations=10, pseudocount=0.5, seed=1, sampler="reference"):
        """sampler: "reference" = the reference's mt19937 sample_pairs() (seeded, reproducible), "device" = Feistel
        permutation of the pair index space generated on the device."""
        thr = C.c_double()
        self._check(self.lib.spk_determine_mi_threshold_ex(self.h, int(n_values), int(threshold_pairs), int(iterations),
                                                           pseudocount, int(seed), 1 if sampler == "device" else 0, C.byref(thr)))
        return float(thr.value)

    def mi_scan(self, threshold, pseudocount=0.5, ld_threshold=0, linear_genome=False, circle_size=0, max_edges=0,
                sort_desc=True, n_terms=0, band=0.0, engine=0, part_index=0, n_parts=1, keep_on_device=False,
                copy=True, async_download=False):
        """copy=False returns zero-copy views of the library-owned pinned result buffers (valid until the next
        scan on this context); passing such an `edges` view to aracne() reuses the device copy.
        async_download=True (with copy=False): the views are returned while the download is still running on a
        side stream; call wait_download() before reading them (aracne() on the view may run in between)."""
        if async_download:
            if copy or keep_on_device:
                raise ValueError("async_download needs copy=False and keep_on_device=False")
            keep_on_device = 2
        p = ScanParams(threshold, pseudocount, ld_threshold, int(linear_genome), circle_size, max_edges, int(sort_desc),
                       n_terms, band, engine, part_index, n_parts, int(keep_on_device))
        r = ScanResult()
        self._check(self.lib.spk_mi_scan(self.h, C.byref(p), C.byref(r)))

        def grab(ptr, cnt):
            if cnt == 0 or not ptr:
                return np.zeros(0, dtype=EDGE_DTYPE)
            buf = (C.c_char * (cnt * 16)).from_address(ptr)
            v = np.frombuffer(buf, dtype=EDGE_DTYPE)
            return v.copy() if copy else v

        out = {"n_edges": int(r.n_edges), "n_wo_gaps": int(r.n_wo_gaps), "n_candidates": int(r.n_candidates),
               "max_filter_error": float(r.max_filter_error), "audit_pairs": int(r.audit_pairs),
               "audit_violations": int(r.audit_violations), "audit_max_error": float(r.audit_max_error), "band": float(r.band),
               "colmax": np.ctypeslib.as_array(r.colmax, shape=(self.L,)).copy(),
               "d_edges": r.d_edges, "d_edges_wo_gaps": r.d_edges_wo_gaps, "d_colmax": r.d_colmax}
        if keep_on_device != 1 and r.edges:
            out["edges"] = grab(r.edges, r.n_edges)
            out["edges_wo_gaps"] = grab(r.edges_wo_gaps, r.n_wo_gaps)
        return out

    def outlier_thresholds(self, colmax):
        c = np.ascontiguousarray(colmax, dtype=np.float64)
        a = C.c_double(); b = C.c_double()
        self._check(self.lib.spk_outlier_thresholds(_ptr(c), c.shape[0], C.byref(a), C.byref(b)))
        return float(a.value), float(b.value)

    def sort_edges(self, edges):
        e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE).copy()
        self._check(self.lib.spk_sort_edges(self.h, _ptr(e), e.shape[0]))
        return e

    def aracne_collective(self, n_edges, edge_threshold=2.220446049250313e-16, out=None, edges=None):
        """Collective ARACNE on the list of the last collective scan (n_parts = 0). Rank 0 passes its host view of the edges and
        receives the flags; the other ranks pass nothing."""
        f = None
        if out is not None:
            f = out[:n_edges]
        elif edges is not None:
            f = np.zeros(n_edges, dtype=np.uint8)
        self._check(self.lib.spk_aracne(self.h, _ptr(edges) if edges is not None else None, int(n_edges), edge_threshold, 0, 0, _ptr(f) if f is not None else None))
        return f

    def aracne(self, edges, edge_threshold=2.220446049250313e-16, part_index=0, n_parts=1, out=None):
        """out: optional uint8 array (>= len(edges), e.g. from pinned_empty) that receives the flags."""
        e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
        if out is None:
            f = np.zeros(e.shape[0], dtype=np.uint8)
        else:
            if out.dtype != np.uint8 or out.shape[0] < e.shape[0] or not out.flags.c_contiguous:
                raise ValueError("aracne: out must be a contiguous uint8 array with at least len(edges) entries")
            f = out[:e.shape[0]]
            if n_parts > 1:
                f[:] = 0
        self._check(self.lib.spk_aracne(self.h, _ptr(e), e.shape[0], edge_threshold, part_index, n_parts, _ptr(f)))
        return f

    def wait_download(self):
        self._check(self.lib.spk_wait_download(self.h))

    def sort_edges_device(self, d_edges, n):
        self._check(self.lib.spk_sort_edges_device(self.h, C.c_void_p(d_edges), n))

    def merge_sorted_runs_device(self, d_edges, run_counts):
        """d_edges: concatenation of len(run_counts) runs, each in Graph::sort order -> merged in place."""
        arr = (C.c_uint64 * len(run_counts))(*[int(x) for x in run_counts])
        self._check(self.lib.spk_merge_sorted_runs_device(self.h, C.c_void_p(d_edges), arr, len(run_counts)))

    def aracne_device(self, d_edges, n, d_flags, edge_threshold=2.220446049250313e-16, part_index=0, n_parts=1):
        self._check(self.lib.spk_aracne_device(self.h, C.c_void_p(d_edges), n, edge_threshold, part_index, n_parts,
                                               C.c_void_p(d_flags)))

    def timings(self):
        t = Timings()
        self._check(self.lib.spk_get_timings(self.h, C.byref(t)))
        return t.as_dict()

    def debug_tile_counts(self, ib, jb, n_terms=2):
        out = np.zeros((32, 64, 4, 4), dtype=np.float32)
        self._check(self.lib.spk_debug_tile_counts(self.h, ib, jb, n_terms, _ptr(out)))
        return out

    def debug_i8_peak(self, kblocks=100000, reps=1):
        """(TOP/s, ms) of the MMA-only microbenchmark: the measured tensor-pipe peak for the filter's instruction."""
        t = C.c_double(0); ms = C.c_float(0)
        self._check(self.lib.spk_debug_i8_peak(self.h, int(kblocks), int(reps), C.byref(t), C.byref(ms)))
        return t.value, ms.value

    def debug_i8_mode(self):
        """weight rows of the default int8 engine: 2 = engine 3 ({hi, lo} weight bytes), 1 = engine 4 (factorised weights)"""
        return int(self.lib.spk_debug_i8_mode())

    def debug_i8_layout(self, with_weights=False):
        """int8 compact engine layout: dict(ntiles[3], col_start[3], perm, info[, q, delta])."""
        nt = np.zeros(3, dtype=np.uint32); cs = np.zeros(3, dtype=np.uint32)
        self._check(self.lib.spk_debug_i8_layout(self.h, _ptr(nt), _ptr(cs), None, None, 0, None, None))
        lperm = int(nt[0]) * 128 + int(nt[1]) * 64 + int(nt[2]) * 32
        perm = np.zeros(lperm, dtype=np.uint32); info = np.zeros(lperm, dtype=np.uint32)
        q = np.zeros(self.n, dtype=np.uint32) if with_weights else None
        delta = C.c_double()
        self._check(self.lib.spk_debug_i8_layout(self.h, _ptr(nt), _ptr(cs), _ptr(perm), _ptr(info), lperm, _ptr(q),
                                                 C.byref(delta)))
        out = {"ntiles": nt, "col_start": cs, "perm": perm, "info": info, "delta": float(delta.value)}
        if with_weights:
            out["q"] = q
        return out

    def debug_i8_tile(self, ti, tj):
        out = np.zeros(16384, dtype=np.uint32)
        self._check(self.lib.spk_debug_i8_tile(self.h, ti, tj, _ptr(out)))
        return out
