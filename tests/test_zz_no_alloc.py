"""vb_col_sum_static: the column sum for callers that must not allocate device memory.

A host program with a fixed memory budget, or code that sits inside a stream capture or a
latency-critical loop, cannot call cudaMalloc for the inter-block scratch that vb_col_sum
needs.  The static variant keeps that scratch in a fixed 4 MiB array inside the library.
This test drives both entry points through the C ABI exactly as such a caller would (ctypes,
raw device pointers, the current stream) and requires the same numbers from both — on two
back-to-back calls, because the second call reuses whatever the first one left in the shared
scratch.  The file name sorts last on purpose: pytest collects it after every other test.
"""
import ctypes
from ctypes import c_float, c_int, c_longlong, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from veles_b200.ops import _lib
    lib = _lib.lib()
    lib.vb_col_sum_static_max_elems.restype = ctypes.c_longlong
    return lib, _lib


def _call_static(lib, a, out, alpha, accumulate):
    return lib.vb_col_sum_static(c_void_p(a.data_ptr()), c_int(0 if a.dtype == torch.float32 else 1),
                                 c_longlong(a.stride(0)), c_int(a.shape[0]), c_int(a.shape[1]),
                                 c_void_p(out.data_ptr()), c_float(alpha), c_int(accumulate),
                                 c_void_p(torch.cuda.current_stream().cuda_stream))


def _call_existing(lib, a, out, scratch, alpha, accumulate):
    return lib.vb_col_sum(c_void_p(a.data_ptr()), c_int(0 if a.dtype == torch.float32 else 1),
                          c_longlong(a.stride(0)), c_int(a.shape[0]), c_int(a.shape[1]),
                          c_void_p(out.data_ptr()), c_float(alpha), c_int(accumulate), c_void_p(scratch.data_ptr()),
                          c_void_p(torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_column_sum_matches_existing_entry_point_twice(dtype):
    lib, mod = _lib()
    dev = torch.device("cuda:0")
    torch.manual_seed(17)
    # two different shapes back to back: the second call sees the first call's leftovers
    for rows, cols, alpha, accumulate in ((777, 203, 2.0, 1), (130, 4096, 1.0, 0)):
        a = torch.randn(rows, cols, device=dev).to(dtype)
        ref = torch.full((cols,), 0.5, device=dev)
        got = torch.full((cols,), 0.5, device=dev)
        scratch = torch.empty(int(lib.vb_col_sum_scratch_elems(c_int(rows), c_int(cols))), device=dev)
        mod.check(_call_existing(lib, a, ref, scratch, alpha, accumulate), "vb_col_sum")
        mod.check(_call_static(lib, a, got, alpha, accumulate), "vb_col_sum_static")
        torch.cuda.synchronize()
        expect = alpha * a.float().sum(0) + (0.5 if accumulate else 0.0)
        scale = expect.abs().max().item() + 1e-12
        assert (got - ref).abs().max().item() / scale < 1e-5   # same tolerance as test_matrix_reduce
        assert (got - expect).abs().max().item() / scale < 1e-5


def test_static_scratch_rejects_oversized_input():
    lib, mod = _lib()
    dev = torch.device("cuda:0")
    limit = int(lib.vb_col_sum_static_max_elems())
    rows, cols = 64 * 9, limit // 8   # 9 slabs x cols > limit
    a = torch.zeros(1, 8, device=dev)  # never read: the call must fail before launching anything
    out = torch.zeros(cols, device=dev)
    rc = lib.vb_col_sum_static(c_void_p(a.data_ptr()), c_int(0), c_longlong(cols), c_int(rows), c_int(cols),
                               c_void_p(out.data_ptr()), c_float(1.0), c_int(0),
                               c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.vb_last_error().decode()
    assert "scratch" in msg and str(limit) in msg
