"""GPU test of ddpb200_l2norm_sumsq_static, the one entry point of the library that brings its own
scratch memory (a fixed array inside the library image) for callers that cannot hand any over:
bindings that see only their inputs and outputs, call sites where cudaMalloc's device-wide
synchronisation is not acceptable, processes whose device memory all belongs to an unreachable pool.

It is called through the C ABI exactly as such a caller would — no workspace argument — and must give
the same per-tensor sums of squares, bit for bit (the reduction order is fixed), as
ddpb200_l2norm_sumsq with caller-owned scratch. Two calls in a row on different data: the second one
finds the array as the first one left it. The file name makes pytest collect it last, so a run that
stops at the first failure has already been through every other test."""
import ctypes

import numpy as np
import pytest
import torch

import ddp_b200
from ddp_b200 import _lib

pytestmark = pytest.mark.gpu


def _both(grads, dtype, chunk_elems=None, on_device=False):
    dev = grads[0].device
    K = ddp_b200.CudaKernels()
    sizes = [g.numel() for g in grads]
    tab = K.table([None, grads, None, None], sizes, chunk_elems=chunk_elems, device=dev if on_device else None)
    ptr, loc, n, chunks = tab.args(0, None, prefer_device=on_device)
    st = torch.cuda.current_stream().cuda_stream
    want = torch.full((len(sizes),), -1.0, dtype=torch.float32, device=dev)
    K.sumsq(tab, dtype, want, prefer_device=on_device)                      # existing entry point, caller scratch
    got = torch.full((len(sizes),), -2.0, dtype=torch.float32, device=dev)
    rc = _lib.lib().ddpb200_l2norm_sumsq_static(ptr, loc, n, chunks, tab.chunk_elems, ddp_b200.kernels.dtype_code(dtype),
                                                got.data_ptr(), st)
    assert rc == 0, _lib.lib().ddpb200_error_string(rc)
    torch.cuda.synchronize()
    return got, want, tab


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_norm_matches_caller_scratch_twice_in_a_row(dtype):
    dev = torch.device("cuda", 0)
    sizes = [0, 1, 2047, 2049, 70001, 1 << 20, 5]
    for call in range(2):                               # second call reuses the array left by the first
        g = torch.Generator(device=dev).manual_seed(100 + call)
        grads = [torch.empty(n, dtype=dtype, device=dev).normal_(0, 0.01 * (call + 1), generator=g) for n in sizes]
        got, want, _ = _both(grads, dtype, chunk_elems=2048, on_device=bool(call))
        assert torch.equal(got, want), (call, got, want)
        ref = torch.stack([x.double().pow(2).sum() for x in grads]).float()
        assert torch.allclose(got, ref, rtol=2e-6, atol=0)
    # a different, larger chunking right after: stale partials from the calls above must not leak in
    grads = [torch.ones(300000, dtype=dtype, device=dev), torch.full((4097,), 2.0, dtype=dtype, device=dev)]
    got, want, _ = _both(grads, dtype)
    assert torch.equal(got, want)
    assert got.tolist() == [300000.0, 4 * 4097.0]


def test_static_scratch_rejects_sizes_beyond_its_limit():
    dev = torch.device("cuda", 0)
    limit = _lib.STATIC_SCRATCH_MAX_CHUNKS
    n = (limit + 1) * 2048                              # one chunk too many at the smallest chunk size
    big = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    K = ddp_b200.CudaKernels()
    tab = K.table([None, [big], None, None], [n], chunk_elems=2048)
    assert tab.total_chunks == limit + 1
    ptr, loc, cnt, chunks = tab.args(0, None, prefer_device=False)
    out = torch.zeros(1, dtype=torch.float32, device=dev)
    rc = _lib.lib().ddpb200_l2norm_sumsq_static(ptr, loc, cnt, chunks, 2048, _lib.BF16, out.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream)
    assert rc == -3
    assert b"limit" in _lib.lib().ddpb200_error_string(rc)
    # the same data fits with a larger chunk size
    tab2 = K.table([None, [big], None, None], [n], chunk_elems=4096)
    ptr, loc, cnt, chunks = tab2.args(0, None, prefer_device=False)
    assert _lib.lib().ddpb200_l2norm_sumsq_static(ptr, loc, cnt, chunks, 4096, _lib.BF16, out.data_ptr(),
                                                  torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert float(out) == 0.0


def test_static_scratch_exactly_at_its_limit():
    """total_chunks == the limit is accepted and uses the last float of the array: ones through 262,144
    chunks of 2048 elements (the last one half full) sum to their count exactly, and the caller-scratch
    entry point agrees bit for bit; invalid arguments are refused before anything is enqueued."""
    dev = torch.device("cuda", 0)
    limit = _lib.STATIC_SCRATCH_MAX_CHUNKS
    n = limit * 2048 - 1024                             # ragged last chunk, still exactly `limit` chunks; n fits fp32
    ones = torch.ones(n, dtype=torch.bfloat16, device=dev)
    got, want, tab = _both([ones], torch.bfloat16, chunk_elems=2048)
    assert tab.total_chunks == limit
    assert torch.equal(got, want) and float(got) == float(n)
    ptr, loc, cnt, chunks = tab.args(0, None, prefer_device=False)
    st = torch.cuda.current_stream().cuda_stream
    L = _lib.lib()
    assert L.ddpb200_l2norm_sumsq_static(ptr, loc, cnt, -1, 2048, _lib.BF16, got.data_ptr(), st) == -1
    assert L.ddpb200_l2norm_sumsq_static(ptr, loc, cnt, chunks, 1000, _lib.BF16, got.data_ptr(), st) == -1
    assert L.ddpb200_l2norm_sumsq_static(ptr, loc, cnt, chunks, 2048, 7, got.data_ptr(), st) == -2
    assert L.ddpb200_l2norm_sumsq_static(ptr, loc, cnt, chunks, 2048, _lib.BF16, None, st) == -1
    # and the array still works afterwards
    got2, want2, _ = _both([ones[:70001]], torch.bfloat16, chunk_elems=2048)
    assert torch.equal(got2, want2) and float(got2) == 70001.0
