"""spydrpick_b200 -- B200-native (sm_100a) implementation of SpydrPick's hot path:
sample reweighting (K1), all-pairs weighted MI scan (K2 tcgen05 filter + fp64 refine), ARACNE (K3).

The compute lives in libspydrpick_b200.so behind the C-ABI of include/spydrpick_b200.h; this package
is the ctypes binding plus a host-side mirror of the reference's four entry points (host.py).
Importing never falls back to a CPU implementation: `capi.load()` raises if the library is missing.
"""
from . import capi  # noqa: F401
from .capi import Context, SpkError, EDGE_DTYPE  # noqa: F401
from .host import (cache_sample_weights, determine_MI_threshold, get_MI_network, run_ARACNE,  # noqa: F401
                   Alignment, MINetwork, Options)
