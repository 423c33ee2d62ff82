"""ctypes wrapper around oracle/_ref/libspkref.so and oracle/_ref/SpydrPick_ref -- TEST INFRASTRUCTURE ONLY.

The two files are the REAL reference (santeripuranen/SpydrPick @ f6bc110 + apegrunt @ 5c61810) compiled in place from
/root/reference by oracle/Makefile.ref against the Boost / TBB stand-ins of oracle/ref_shims. They pin the CPU restatement
(oracle/oracle.cpp) and generate the golden fixtures under tests/golden/ (tests/golden/make_ref_golden.py). Only tests/,
__graft_entry__ and bench.py's CPU legs may import this module; the product never does.

/root/reference does not exist on the GPU box: there the prebuilt files (git-ignored, not gpurun-ignored) are used as they are,
and `available()` is False only if they were never built.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
LIB_PATH = os.path.join(REF_DIR, "libspkref.so")
EXE_PATH = os.path.join(REF_DIR, "SpydrPick_ref")
REFERENCE_ROOT = os.environ.get("SPK_REFERENCE_ROOT", "/root/reference")
_LIB = None

EDGE_DTYPE = np.dtype([("i", np.uint32), ("j", np.uint32), ("mi", np.float64)], align=True)


def can_build():
    return os.path.exists(os.path.join(REFERENCE_ROOT, "include", "mi.hpp"))


def build(force=False):
    """Compile the reference from its sources where they lie (needs /root/reference); outputs only into oracle/_ref/."""
    if not can_build():
        return False
    if force or not (os.path.exists(LIB_PATH) and os.path.exists(EXE_PATH)):
        subprocess.check_call(["make", "-C", _HERE, "-f", "Makefile.ref", "-j8", "REF=" + REFERENCE_ROOT] + (["-B"] if force else []),
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return True


def available():
    if os.path.exists(LIB_PATH) and os.path.exists(EXE_PATH):
        return True
    try:
        return build() and os.path.exists(LIB_PATH) and os.path.exists(EXE_PATH)
    except Exception:
        return False


def lib():
    global _LIB
    if _LIB is None:
        if not available():
            raise RuntimeError("oracle/_ref is not built (needs /root/reference; run make -C oracle -f Makefile.ref)")
        L = C.CDLL(LIB_PATH)
        for name in ("ref_alignment_create", "ref_alignment_from_file", "ref_alignment_filter", "ref_alignment_subset",
                     "ref_alignment_subsample", "ref_scan", "ref_network_from_edges"):
            getattr(L, name).restype = C.c_void_p
        L.ref_alignment_nloci.restype = C.c_uint64
        L.ref_alignment_size.restype = C.c_uint32
        L.ref_alignment_n_original_positions.restype = C.c_uint64
        L.ref_alignment_id.restype = C.c_size_t
        L.ref_log.restype = C.c_size_t
        L.ref_network_size.restype = C.c_uint64
        L.ref_effective_size.restype = C.c_double
        L.ref_determine_mi_threshold.restype = C.c_double
        L.ref_version.restype = C.c_char_p
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def configure(*options):
    """Fill the reference's option singletons from command-line style options (the way its main() does). Options persist
    until the next call; every call starts from the values of the previous one for options it does not mention, exactly
    like re-parsing a command line would NOT -- so always pass the full set a test depends on."""
    argv = ["SpydrPick", "alignment.fasta"] + [str(o) for o in options]
    arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
    if lib().ref_configure(C.c_int(len(argv)), arr) != 0:
        raise ValueError("reference rejected options: %r" % (options,))


def log():
    n = lib().ref_log(None, C.c_size_t(0))
    return ""  # drained


class Alignment:
    """The reference's block-compressed Alignment (apegrunt::Alignment_impl_block_compressed_storage)."""

    def __init__(self, handle):
        if not handle:
            raise RuntimeError("reference returned no alignment")
        self.h = C.c_void_p(handle)

    @classmethod
    def from_states(cls, states, ident="synthetic", pos=None):
        s = np.ascontiguousarray(states, dtype=np.uint8); n, L = s.shape
        a = cls(lib().ref_alignment_create(_p(s), C.c_uint32(n), C.c_uint64(L), ident.encode()))
        if pos is not None:
            a.set_mapping(pos)
        return a

    @classmethod
    def from_file(cls, path):
        return cls(lib().ref_alignment_from_file(str(path).encode()))

    def __del__(self):
        try:
            if self.h:
                lib().ref_alignment_free(self.h); self.h = None
        except Exception:
            pass

    @property
    def n_loci(self):
        return int(lib().ref_alignment_nloci(self.h))

    @property
    def n(self):
        return int(lib().ref_alignment_size(self.h))

    @property
    def n_original_positions(self):
        return int(lib().ref_alignment_n_original_positions(self.h))

    @property
    def id_string(self):
        buf = C.create_string_buffer(1024); lib().ref_alignment_id(self.h, buf, C.c_size_t(1024)); return buf.value.decode()

    def set_mapping(self, pos):
        p = np.ascontiguousarray(pos, dtype=np.uint64); assert p.shape[0] == self.n_loci
        lib().ref_alignment_set_mapping(self.h, _p(p), C.c_uint64(p.shape[0]))

    def mapping(self):
        p = np.zeros(self.n_loci, dtype=np.uint64); lib().ref_alignment_get_mapping(self.h, _p(p)); return p

    def states(self):
        s = np.zeros((self.n, self.n_loci), dtype=np.uint8); lib().ref_alignment_states(self.h, _p(s)); return s

    def filtered(self):
        """main()'s column filter step (maf / gap thresholds from configure())."""
        k = C.c_uint64(0)
        return Alignment(lib().ref_alignment_filter(self.h, C.byref(k)))

    def subset(self, loci):
        v = np.ascontiguousarray(loci, dtype=np.uint64)
        return Alignment(lib().ref_alignment_subset(self.h, _p(v), C.c_uint64(v.shape[0])))

    def subsample(self, samples):
        v = np.ascontiguousarray(samples, dtype=np.uint64)
        return Alignment(lib().ref_alignment_subsample(self.h, _p(v), C.c_uint64(v.shape[0])))

    def cache_sample_weights(self):
        w = np.zeros(self.n, dtype=np.float64); lib().ref_cache_sample_weights(self.h, _p(w)); return w

    def calculate_weights(self):
        w = np.zeros(self.n, dtype=np.float64); lib().ref_calculate_weights(self.h, _p(w)); return w

    def set_weights(self, w):
        w = np.ascontiguousarray(w, dtype=np.float64); assert w.shape[0] == self.n
        lib().ref_set_weights(self.h, _p(w))

    def effective_size(self):
        return float(lib().ref_effective_size(self.h))

    def mi_pairs(self, pi, pj):
        pi = np.ascontiguousarray(pi, dtype=np.uint64); pj = np.ascontiguousarray(pj, dtype=np.uint64)
        out = np.zeros(pi.shape[0], dtype=np.float64)
        lib().ref_mi_pairs(self.h, _p(pi), _p(pj), C.c_uint64(pi.shape[0]), _p(out))
        return out

    def xtable_single(self, ipos, jpos):
        t = np.zeros(25, dtype=np.float64); lib().ref_xtable_single(self.h, C.c_uint64(ipos), C.c_uint64(jpos), _p(t)); return t.reshape(5, 5)

    def determine_mi_threshold(self, n_values):
        return float(lib().ref_determine_mi_threshold(self.h, C.c_uint64(int(n_values))))

    def mi_from_tables(self, tables, ipresence, jpresence):
        """normalize_and_get_mi_single on [k][5][5] tables (row = j state). Returns (mi[k], modified tables)."""
        t = np.array(tables, dtype=np.float64).reshape(-1, 25).copy()
        ip = np.ascontiguousarray(ipresence, dtype=np.uint8); jp = np.ascontiguousarray(jpresence, dtype=np.uint8)
        mi = np.zeros(t.shape[0], dtype=np.float64)
        lib().ref_mi_from_tables(self.h, _p(t), _p(ip), _p(jp), C.c_uint64(t.shape[0]), _p(mi))
        return mi, t.reshape(-1, 5, 5)

    def scan(self, mi_threshold):
        return Network(lib().ref_scan(self.h, C.c_double(mi_threshold)))


class Network:
    """spydrpick::MI_network: edge graph, MI-without-gaps graph, Tukey fences."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)

    @classmethod
    def from_edges(cls, edges):
        e = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
        n1 = np.ascontiguousarray(e["i"]); n2 = np.ascontiguousarray(e["j"]); w = np.ascontiguousarray(e["mi"])
        return cls(lib().ref_network_from_edges(_p(n1), _p(n2), _p(w), C.c_uint64(e.shape[0])))

    def __del__(self):
        try:
            if self.h:
                lib().ref_network_free(self.h); self.h = None
        except Exception:
            pass

    def _edges(self, wo):
        k = int(lib().ref_network_size(self.h, C.c_int(wo)))
        n1 = np.zeros(k, dtype=np.uint32); n2 = np.zeros(k, dtype=np.uint32); w = np.zeros(k, dtype=np.float64); f = np.zeros(k, dtype=np.uint8)
        if k:
            lib().ref_network_edges(self.h, C.c_int(wo), _p(n1), _p(n2), _p(w), _p(f))
        e = np.zeros(k, dtype=EDGE_DTYPE); e["i"] = n1; e["j"] = n2; e["mi"] = w
        return e, f

    def edges(self):
        """(edges structured array in storage order, direct flags)"""
        return self._edges(0)

    def edges_wo_gaps(self):
        return self._edges(1)[0]

    def thresholds(self):
        a = C.c_double(); b = C.c_double(); lib().ref_network_thresholds(self.h, C.byref(a), C.byref(b)); return float(a.value), float(b.value)

    def sort(self):
        lib().ref_network_sort(self.h)

    def aracne(self):
        lib().ref_network_aracne(self.h)

    def write_edges(self, alignment, path):
        assert lib().ref_write_edges(self.h, alignment.h, str(path).encode()) == 0

    def write_outliers(self, alignment, path):
        assert lib().ref_write_outliers(self.h, alignment.h, str(path).encode()) == 0


def run_cli(args, cwd, timeout=600):
    """Run the reference program (its own main()). Output files land in cwd, named as the reference names them."""
    if not available():
        raise RuntimeError("oracle/_ref/SpydrPick_ref is not built")
    return subprocess.run([EXE_PATH] + [str(a) for a in args], cwd=str(cwd), capture_output=True, text=True, timeout=timeout)
