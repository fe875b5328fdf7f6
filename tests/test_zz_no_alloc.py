"""``nfb_update_stats_static`` — the K-STATS entry point for callers that cannot allocate.

Who needs it: code that is only ever *handed* device buffers (an inference or training
plugin living inside someone else's framework), services whose device memory is carved
up once at start-up, and paths that run when no allocation is possible any more (after
the pool is sealed, or while handling an out-of-memory condition).  Such a caller can
still pass inputs and outputs but has nowhere to get the reduction scratch from, so the
library carries a fixed 260 KiB scratch array inside its own image (max 1024 rows).

The test goes through the C ABI only (``ctypes``; the torch tensors are just a
convenient way to own input/output memory) and checks the static-scratch entry point
against the ordinary ``nfb_update_stats`` on two calls in a row — the second call sees
whatever the first left in the shared array — plus the size limit.  The file name sorts
last so a run that stops at the first failure has already executed every other test; it
passes both in a fresh process and after the rest of the suite.
"""

import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from nanofed_b200.ops import native

    lib = native.lib()
    vp, i32, i64, sz = C.c_void_p, C.c_int, C.c_int64, C.c_size_t
    lib.nfb_update_stats_static.restype = i32
    lib.nfb_update_stats_static.argtypes = [vp, i32, i64, i64, i64, vp, vp, vp]
    lib.nfb_stats_static_max_rows.restype = i32
    lib.nfb_stats_static_max_rows.argtypes = []
    lib.nfb_update_stats.restype = i32
    lib.nfb_update_stats.argtypes = [vp, i32, i64, i64, i64, vp, vp, vp, sz, vp]
    lib.nfb_stats_scratch_bytes.restype = sz
    lib.nfb_stats_scratch_bytes.argtypes = [i64]
    lib.nfb_last_error.restype = C.c_char_p
    return lib


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def _reference(lib, x):
    """The existing entry point, with caller-provided (zeroed) scratch."""
    k, n = x.shape
    need = int(lib.nfb_stats_scratch_bytes(k))
    scratch = torch.zeros(need, dtype=torch.uint8, device=x.device)
    ss = torch.empty(k, dtype=torch.float32, device=x.device)
    bad = torch.empty(k, dtype=torch.int32, device=x.device)
    rc = lib.nfb_update_stats(x.data_ptr(), 0, k, n, x.stride(0), ss.data_ptr(), bad.data_ptr(),
                              scratch.data_ptr(), need, _stream(x.device))
    assert rc == 0, lib.nfb_last_error()
    return ss, bad


def _static(lib, x):
    k, n = x.shape
    ss = torch.empty(k, dtype=torch.float32, device=x.device)
    bad = torch.empty(k, dtype=torch.int32, device=x.device)
    rc = lib.nfb_update_stats_static(x.data_ptr(), 0, k, n, x.stride(0), ss.data_ptr(),
                                     bad.data_ptr(), _stream(x.device))
    assert rc == 0, lib.nfb_last_error()
    return ss, bad


def test_static_scratch_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(123)
    # first call: 37 rows; second call: fewer rows, different length, one poisoned row — it
    # reuses array slots the first call wrote to
    first = torch.randn(37, 50_048, device=dev)
    second = torch.randn(5, 1_199_936, device=dev) * 0.01
    second[3, 777] = float("inf")
    for x in (first, second):
        want_ss, want_bad = _reference(lib, x)
        got_ss, got_bad = _static(lib, x)
        torch.cuda.synchronize(dev)
        torch.testing.assert_close(got_ss[want_bad == 0], want_ss[want_bad == 0],
                                   rtol=1e-4, atol=1e-6)
        assert torch.equal(got_bad, want_bad)
        oracle = x.double().pow(2).sum(dim=1).float()
        ok = want_bad == 0
        torch.testing.assert_close(got_ss[ok], oracle[ok], rtol=1e-4, atol=1e-6)
    assert second.shape[0] and int(_static(lib, second)[1].sum()) == 1


def test_static_scratch_rejects_oversize_with_a_clear_error():
    lib = _lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    limit = int(lib.nfb_stats_static_max_rows())
    assert limit == 1024
    x = torch.zeros(limit + 1, 64, device=dev)
    ss = torch.empty(limit + 1, dtype=torch.float32, device=dev)
    rc = lib.nfb_update_stats_static(x.data_ptr(), 0, limit + 1, 64, 64, ss.data_ptr(), None,
                                     _stream(dev))
    assert rc == -2  # NFB_ERANGE
    msg = lib.nfb_last_error().decode()
    assert "exceeds" in msg and "1024" in msg
    # exactly at the limit is accepted
    y = torch.ones(limit, 64, device=dev)
    got, _ = _static(lib, y)
    torch.cuda.synchronize(dev)
    torch.testing.assert_close(got, torch.full((limit,), 64.0, device=dev))
