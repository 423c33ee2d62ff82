"""GPU: spk_neighbour_counts_noalloc (fixed __device__ scratch, no allocation) against spk_neighbour_counts, through the C-ABI.

The file is named so that pytest collects it last. The no-alloc entry point shares one scratch array between all calls on a
device, so the test makes two calls in a row -- a larger problem first, then a smaller one whose planes overlay the first call's
leftovers -- and checks both bit-exactly (integer counts: tolerance 0, like every K1 test) against the context path. It runs the
same way alone in a fresh process and after the other GPU tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _alignment(n, L, seed, dup=0):
    from spydrpick_b200.synth import generate_alignment
    s, _, _ = generate_alignment(n, L, seed=seed)
    if dup:
        s[1:1 + dup] = s[0]                           # copies of sample 0: the SAMPLE0 quirk has neighbours to double-count
    return np.ascontiguousarray(s, dtype=np.uint8)


def _expected(capi, states, min_identical, quirks):
    with capi.Context() as ctx:
        ctx.set_alignment(states)
        return np.asarray(ctx.neighbour_counts(min_identical, quirks=quirks), dtype=np.uint64)


def _noalloc(capi, torch, states, min_identical, quirks, stream=None):
    n, L = states.shape
    d_states = torch.from_numpy(states).cuda()
    d_counts = torch.full((n,), -7, dtype=torch.int64, device="cuda")       # garbage: the call must initialise its output
    torch.cuda.synchronize()
    capi.neighbour_counts_noalloc(d_states.data_ptr(), n, L, min_identical, d_counts.data_ptr(), quirks=quirks, stream=stream)
    torch.cuda.synchronize()
    return d_counts.cpu().numpy().astype(np.uint64)


def test_noalloc_matches_context_path_two_calls_in_a_row():
    import torch
    import spydrpick_b200.capi as capi
    if not torch.cuda.is_available():
        pytest.fail("no GPU: the no-alloc entry point has no CPU fallback")
    torch.cuda.set_device(0)
    big = _alignment(333, 4000, 5, dup=3)             # n not a multiple of the 64-sample tile, L % 32 == 0 (LASTBLOCK quirk active)
    small = _alignment(130, 1001, 6, dup=2)           # ragged last word; overlays the first call's planes
    for states, frac in ((big, 0.85), (small, 0.9)):
        mi = int(frac * states.shape[1])
        for quirks in (capi.QUIRKS_REFERENCE, 0):
            exp = _expected(capi, states, mi, quirks)
            got = _noalloc(capi, torch, states, mi, quirks)
            assert got.dtype == exp.dtype and np.array_equal(got, exp), (states.shape, quirks, np.nonzero(got != exp)[0][:10])
            assert exp.max() > 1                      # the case has similar pairs at all
    # the same problem again on a caller-owned stream: identical result (the shared scratch is rewritten by each call)
    st = torch.cuda.Stream()
    mi = int(0.85 * big.shape[1])
    got = _noalloc(capi, torch, big, mi, capi.QUIRKS_REFERENCE, stream=st.cuda_stream)
    assert np.array_equal(got, _expected(capi, big, mi, capi.QUIRKS_REFERENCE))


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    import spydrpick_b200.capi as capi
    torch.cuda.set_device(0)
    d = torch.zeros(16, dtype=torch.uint8, device="cuda")
    c = torch.zeros(16, dtype=torch.int64, device="cuda")
    # 3 * 4096 * ceil(40000 / 32) * 4 bytes = 61 MB > 16 MiB: refused before anything is launched or read
    with pytest.raises(capi.SpkError) as ei:
        capi.neighbour_counts_noalloc(d.data_ptr(), 4096, 40000, 1, c.data_ptr())
    assert ei.value.code == capi.SPK_E_INVALID and "SPK_NOALLOC_MAX_SCRATCH_BYTES" in str(ei.value)
    with pytest.raises(capi.SpkError):
        capi.neighbour_counts_noalloc(0, 10, 10, 1, c.data_ptr())
    torch.cuda.synchronize()
