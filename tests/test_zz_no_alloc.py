"""cppre_radix_select_ranks_noalloc: the Joint quantile for callers that own no device allocator.

Who needs it: a host program that embeds libcppre but cannot obtain device memory at the call site --
its framework keeps all of HBM in an arena the plugin has no handle to, or its binding only forwards the
field buffers, or it calls from a stretch of its time loop where cudaMalloc's hidden device-wide
synchronisation is not acceptable.  Such a caller cannot supply the (state, counters) scratch that
``cppre_radix_select_begin/hist/pick`` ask for, so this entry point keeps them in a fixed array inside
the library image and allocates nothing, on any call.

What is checked: through the C ABI (ctypes, raw pointers), against the existing entry points
``cppre_radix_select`` (single launch) and the begin/hist/pick sequence with caller scratch, BIT-EXACT
(order statistics of identical scores are integer work, the same bar as the other select tests), on two
calls in a row -- the second with different data, fewer ranks and a NaN, so that whatever the first call
left in the shared array would show -- plus the refusal of sizes beyond the stated maximum.

The file sorts last on purpose: ``pytest -x`` stops at the first failure, and a failure here must not hide
the suite.  It passes alone in a fresh process (it does the library's per-device set-up itself on its
first call) and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from cp_pre_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _noalloc(lib, x, ranks):
    """The entry point under test.  The output buffer is the caller's, made BEFORE the call."""
    out = torch.full((len(ranks),), -7.0, dtype=torch.float32, device=x.device)
    h_ranks = (C.c_int64 * len(ranks))(*ranks)
    L.check(lib.cppre_radix_select_ranks_noalloc(_p(x), x.numel(), h_ranks, len(ranks), _p(out), _stream()))
    return out


def _single_launch(lib, x, k):
    out = torch.empty(1, dtype=torch.float32, device=x.device)
    L.check(lib.cppre_radix_select(_p(x), x.numel(), int(k), _p(out), _stream()))
    return out


def _caller_scratch(lib, x, k):
    """begin / 4 x (hist, pick) with scratch the CALLER allocates: the path the new entry point mirrors."""
    state = torch.empty(4, dtype=torch.int32, device=x.device)
    hist = torch.empty(257, dtype=torch.int32, device=x.device)
    out = torch.empty(1, dtype=torch.float32, device=x.device)
    st = _stream()
    L.check(lib.cppre_radix_select_begin(_p(state), _p(hist), int(k), st))
    for p in range(4):
        L.check(lib.cppre_radix_select_hist(_p(x), x.numel(), _p(state), p, _p(hist), st))
        L.check(lib.cppre_radix_select_pick(_p(hist), _p(state), p, _p(out), st))
    return out


def _bits(t):
    return t.detach().cpu().numpy().view(np.uint32)


def test_noalloc_radix_select_equals_existing_entry_points_on_two_calls_in_a_row():
    lib = L.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    rng = np.random.default_rng(20261106)

    # call 1: 70 001 scores with ties, +-inf, +-0; the alpha grid of the reference (ten levels) as ranks
    n1 = 70001
    x1 = (rng.standard_normal(n1) * 3).astype(np.float32)
    x1[:8] = [np.inf, -np.inf, 0.0, -0.0, 2.25, 2.25, 2.25, -1e-30]
    ranks1 = sorted({int(np.ceil((n1 + 1) * (1 - a))) - 1 for a in np.arange(0.05, 1.05, 0.1)} | {0, n1 - 1})
    ranks1 = [min(max(k, 0), n1 - 1) for k in ranks1]
    assert len(ranks1) <= L.NOALLOC_MAX_RANKS
    d1 = torch.from_numpy(x1).to(dev)
    got1 = _noalloc(lib, d1, ranks1)
    want_sorted = np.sort(x1)
    for j, k in enumerate(ranks1):
        ref_a, ref_b = _single_launch(lib, d1, k), _caller_scratch(lib, d1, k)
        assert _bits(got1)[j] == _bits(ref_b)[0], f"call 1, rank {k} vs begin/hist/pick"
        assert _bits(got1)[j] == _bits(ref_a)[0], f"call 1, rank {k} vs cppre_radix_select"
        assert float(got1[j]) == float(want_sorted[k])

    # call 2, straight after: other data, other (fewer) ranks, one beyond n -> NaN like the existing calls
    n2 = 1537
    x2 = rng.uniform(-50, 50, n2).astype(np.float32)
    d2 = torch.from_numpy(x2).to(dev)
    ranks2 = [n2 // 2, 3, n2 - 1, n2]
    got2 = _noalloc(lib, d2, ranks2)
    for j, k in enumerate(ranks2):
        ref = _single_launch(lib, d2, k)
        assert _bits(got2)[j] == _bits(ref)[0], f"call 2, rank {k}"
    assert np.isnan(float(got2[3])) and float(got2[0]) == float(np.sort(x2)[n2 // 2])

    # call 3: a NaN among the scores poisons every rank (np.quantile rule), and a clean call after it is clean again
    x3 = x2.copy()
    x3[11] = np.nan
    got3 = _noalloc(lib, torch.from_numpy(x3).to(dev), [5, 900])
    assert np.isnan(got3.cpu().numpy()).all()
    again = _noalloc(lib, d2, ranks2[:3])
    assert np.array_equal(_bits(again), _bits(got2)[:3])

    # empty input: every rank is out of range
    empty = torch.empty(0, dtype=torch.float32, device=dev)
    out = torch.zeros(1, dtype=torch.float32, device=dev)
    L.check(lib.cppre_radix_select_ranks_noalloc(C.c_void_p(0), 0, (C.c_int64 * 1)(0), 1, _p(out), _stream()))
    assert empty.numel() == 0 and np.isnan(float(out))


def test_noalloc_radix_select_refuses_sizes_beyond_its_fixed_scratch():
    lib = L.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    x = torch.arange(64, dtype=torch.float32, device=dev)
    too_many = L.NOALLOC_MAX_RANKS + 1
    out = torch.full((too_many,), -7.0, dtype=torch.float32, device=dev)
    h_ranks = (C.c_int64 * too_many)(*range(too_many))
    rc = lib.cppre_radix_select_ranks_noalloc(_p(x), 64, h_ranks, too_many, _p(out), _stream())
    assert rc == L.ERR_UNSUPPORTED and b"CPPRE_NOALLOC_MAX_RANKS" in lib.cppre_last_error()
    rc = lib.cppre_radix_select_ranks_noalloc(_p(x), 1 << 32, h_ranks, 1, _p(out), _stream())
    assert rc == L.ERR_UNSUPPORTED and b"32-bit counters" in lib.cppre_last_error()
    torch.cuda.synchronize()
    assert (out == -7.0).all()                       # refused before anything was launched
    # exactly the maximum is accepted
    h_ok = (C.c_int64 * L.NOALLOC_MAX_RANKS)(*range(L.NOALLOC_MAX_RANKS))
    L.check(lib.cppre_radix_select_ranks_noalloc(_p(x), 64, h_ok, L.NOALLOC_MAX_RANKS, _p(out), _stream()))
    assert out[:L.NOALLOC_MAX_RANKS].cpu().tolist() == [float(i) for i in range(L.NOALLOC_MAX_RANKS)]
