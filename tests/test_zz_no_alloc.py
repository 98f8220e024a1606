"""pcc_col_sum_static — the allocation-free variant of pcc_col_sum (scratch in a file-scope
`static __device__` array of the library; include/pcc_b200.h) — against pcc_col_sum, both called
straight through the C-ABI with ctypes.  Inputs and outputs are allocated by the caller BEFORE the
calls; the entry point itself must allocate nothing.  Two calls in a row with different shapes share the
one scratch array, so the second must not see what the first left behind.

Collected last on purpose (file name).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def test_col_sum_static_matches_col_sum_twice_in_a_row(pcc):
    L = pcc._lib.load()
    torch.cuda.set_device(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cpu").manual_seed(7)
    # first call: wide and long (all 592 partial rows in use); second: narrower, shorter, odd width (scalar
    # path) — stale partials of the first call lie beyond what the second one writes
    shapes = [(300_000, 256), (4_001, 130)]
    xs = [torch.randn(n, c, generator=g).to(DEV) for n, c in shapes]
    wsb = max(L.pcc_col_sum_ws_bytes(c) for _, c in shapes)
    ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
    want = [torch.empty(c, dtype=torch.float32, device=DEV) for _, c in shapes]
    got = [torch.full((c,), float("nan"), dtype=torch.float32, device=DEV) for _, c in shapes]
    for x, w in zip(xs, want):
        rc = L.pcc_col_sum(_ptr(x), x.shape[0], x.shape[1], _ptr(w), _ptr(ws), wsb, stream)
        assert rc == 0, L.pcc_last_error()
    torch.cuda.synchronize()

    before = torch.cuda.memory_allocated()
    for x, o in zip(xs, got):  # two calls in a row, nothing in between
        rc = L.pcc_col_sum_static(_ptr(x), x.shape[0], x.shape[1], _ptr(o), stream)
        assert rc == 0, L.pcc_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for (n, c), w, o, x in zip(shapes, want, got, xs):
        # same kernels, same block count, same summation order: the same bits (the repo's bar for this op,
        # tests/test_gpu_train.py::test_col_sum)
        assert torch.equal(o, w), f"[{n} x {c}] differs from pcc_col_sum"
        ref = x.double().sum(0)
        assert torch.allclose(o.double(), ref, rtol=0, atol=1e-6 * float(x.double().abs().sum(0).max()))

    # again in the other order: still independent of what the array held
    for x, o, w in reversed(list(zip(xs, got, want))):
        o.fill_(float("nan"))
        assert L.pcc_col_sum_static(_ptr(x), x.shape[0], x.shape[1], _ptr(o), stream) == 0
        torch.cuda.synchronize()
        assert torch.equal(o, w)


def test_col_sum_static_empty_and_limit(pcc):
    L = pcc._lib.load()
    torch.cuda.set_device(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = torch.full((64,), 3.0, dtype=torch.float32, device=DEV)
    assert L.pcc_col_sum_static(None, 0, 64, _ptr(out), stream) == 0       # n = 0 -> zeros
    torch.cuda.synchronize()
    assert not out.any()
    c = 257                                                                 # beyond PCC_COL_SUM_STATIC_MAX_C = 256
    x = torch.ones(10, c, device=DEV)
    out = torch.empty(c, device=DEV)
    rc = L.pcc_col_sum_static(_ptr(x), 10, c, _ptr(out), stream)
    assert rc == -4, rc                                                     # PCC_ERR_CAPACITY
    msg = L.pcc_last_error().decode()
    assert "256" in msg and "257" in msg, msg
