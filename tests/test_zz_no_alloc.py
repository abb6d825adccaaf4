"""Allocation-free entry point (qspec_argmax_rows_static; include/qspec_b200.h): no counterpart in the reference, added
for callers that cannot call cudaMalloc. Its scratch is a file-scope device array inside the
library, looked up with cudaGetSymbolAddress once per device.

Called through the C ABI with raw pointers (ctypes), two calls in a row on different inputs (the second call sees the first
call's scratch contents), compared with the existing entry point qspec_argmax_rows, which takes a caller workspace. Both are
integer results of the same kernel: the bar is exact equality (indices) and bit equality (maxima). The file sorts last on
purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200 import _lib
    return _lib.load()


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _both(lib, logits):
    """(idx, val) of the existing entry point and of the static one, on torch's current stream."""
    from qspec_b200 import _lib
    rows, V = logits.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ws = torch.zeros(lib.qspec_argmax_rows_workspace_bytes(rows), dtype=torch.uint8, device="cuda")
    i0 = torch.full((rows,), -1, dtype=torch.int64, device="cuda")
    v0 = torch.zeros(rows, dtype=torch.float32, device="cuda")
    i1, v1 = torch.full_like(i0, -2), torch.ones_like(v0)
    _lib.check(lib.qspec_argmax_rows(_p(logits), rows, V, logits.stride(0), _p(i0), _p(v0), _p(ws), ws.numel(), stream), "argmax_rows")
    # everything the static call needs exists before it: between here and its return nothing may allocate
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    rc = lib.qspec_argmax_rows_static(_p(logits), rows, V, logits.stride(0), _p(i1), _p(v1), stream)
    after = torch.cuda.memory_stats()["allocation.all.allocated"]
    _lib.check(rc, "argmax_rows_static")
    assert after == before
    torch.cuda.synchronize()
    return (i0, v0), (i1, v1)


def test_static_argmax_matches_workspace_argmax_twice_in_a_row(lib):
    assert lib.qspec_argmax_rows_static_max_rows() == 1024
    g = torch.Generator().manual_seed(11)
    # two calls in a row, different shapes: the second runs on the scratch the first one left behind (more rows than the
    # first call cleared, fewer slices per row)
    for rows, V in ((5, 128256), (257, 4104), (64, 2048 * 3 + 8)):
        logits = torch.randn(rows, V, generator=g).half().cuda()
        logits[0] = 0.5                                        # equal values: lowest index
        logits[rows - 1, V - 1] = 30.0                         # maximum in the scalar tail / last slice
        (i0, v0), (i1, v1) = _both(lib, logits)
        assert torch.equal(i1, i0)
        assert torch.equal(v1.view(torch.int32), v0.view(torch.int32))
        assert torch.equal(i1, torch.argmax(logits.float(), dim=1))
        assert int(i1[0]) == 0 and int(i1[rows - 1]) == V - 1


def test_static_argmax_after_poisoned_scratch(lib):
    """The counters are cleared by the call itself: a first call followed by a second one on the same rows gives the same answer
    even if a third party left the counters non-zero (emulated by a call that is given the maximum row count first)."""
    g = torch.Generator().manual_seed(12)
    big = torch.randn(1024, 2056, generator=g).half().cuda()
    (i0, _), (i1, _) = _both(lib, big)
    assert torch.equal(i1, i0)
    small = torch.randn(3, 32008, generator=g).half().cuda()
    for _ in range(2):
        (i0, v0), (i1, v1) = _both(lib, small)
        assert torch.equal(i1, i0) and torch.equal(v1, v0)


def test_static_argmax_rejects_sizes_beyond_the_limit(lib):
    x = torch.zeros(1025, 64, dtype=torch.float16, device="cuda")
    out = torch.empty(1025, dtype=torch.int64, device="cuda")
    rc = lib.qspec_argmax_rows_static(_p(x), 1025, 64, 64, _p(out), None, None)
    assert rc == 2                                             # QSPEC_ERR_UNSUPPORTED
    msg = lib.qspec_last_error().decode()
    assert "1024" in msg and "1025" in msg
    assert lib.qspec_argmax_rows_static(None, 1, 64, 64, _p(out), None, None) == 1      # QSPEC_ERR_INVALID_ARG
    assert lib.qspec_argmax_rows_static(_p(x), 0, 64, 64, _p(out), None, None) == 0     # empty batch: no-op
