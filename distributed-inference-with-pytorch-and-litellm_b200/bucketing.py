"""Bucket planning: which parameter goes into which flat gradient bucket, at
which element offset. Integer work, bit-exact against torch.

Replaces dist._compute_bucket_assignment_by_size (TORCH/include/torch/csrc/distributed/c10d/
reducer.hpp:586-598, called from TORCH/nn/parallel/distributed.py:1237-1244) and the
Bucket.offsets/lengths bookkeeping of reducer.hpp:372-375. The planner itself is native
(csrc/host.cu, `ddpb200_bucket_assign` / `ddpb200_bucket_layout`); this module marshals.
"""
from __future__ import annotations

import sys
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib

DEFAULT_FIRST_BUCKET_BYTES = 1024 * 1024        # kDefaultFirstBucketBytes, reducer.hpp:30
DEFAULT_BUCKET_BYTES_CAP = 25 * 1024 * 1024     # kDefaultBucketBytesCap,   reducer.hpp:31


def _vp(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


def compute_bucket_assignment_by_size(
    tensors: Sequence[torch.Tensor], bucket_size_limits: Sequence[int],
    expect_sparse_gradient: Optional[Sequence[bool]] = None,
    tensor_indices: Optional[Sequence[int]] = None,
) -> Tuple[List[List[int]], List[int]]:
    """Same signature and result as ``dist._compute_bucket_assignment_by_size``.

    ``tensors`` may live on any device (``meta`` included); only numel, element size, dtype
    and device are read."""
    n = len(tensors)
    if n == 0:
        raise ValueError("compute_bucket_assignment_by_size: no tensors")
    if expect_sparse_gradient is not None and len(expect_sparse_gradient) not in (0, n):
        raise ValueError("expect_sparse_gradient must be empty or have one flag per tensor")
    numel = np.fromiter((t.numel() for t in tensors), dtype=np.int64, count=n)
    esize = np.fromiter((t.element_size() for t in tensors), dtype=np.int32, count=n)
    keymap = {}
    keys = np.fromiter((keymap.setdefault((t.dtype, t.device), len(keymap)) for t in tensors),
                       dtype=np.int64, count=n)
    sparse = None
    if expect_sparse_gradient:
        sparse = np.asarray([1 if x else 0 for x in expect_sparse_gradient], dtype=np.uint8)
    tidx = None
    if tensor_indices is not None and len(tensor_indices) > 0:
        if len(tensor_indices) != n:
            raise ValueError("tensor_indices must have one index per tensor")
        tidx = np.asarray(tensor_indices, dtype=np.int64)
    limits = np.asarray([min(int(x), 2 ** 64 - 1) for x in bucket_size_limits], dtype=np.uint64)
    out_idx = np.empty(n, dtype=np.int64)
    out_start = np.empty(n + 1, dtype=np.int64)
    out_limit = np.empty(n, dtype=np.uint64)
    out_nb = np.zeros(1, dtype=np.int64)
    rc = _lib.lib().ddpb200_bucket_assign(_vp(numel), _vp(esize), _vp(keys), _vp(sparse), _vp(tidx), n,
                                          _vp(limits), len(limits), _vp(out_idx), _vp(out_start),
                                          _vp(out_limit), _vp(out_nb))
    _lib.check(rc, "ddpb200_bucket_assign")
    nb = int(out_nb[0])
    buckets = [[int(i) for i in out_idx[out_start[b]:out_start[b + 1]]] for b in range(nb)]
    return buckets, [int(x) for x in out_limit[:nb]]


@dataclass
class BucketSpec:
    """One flat bucket: parameter indices in order, their element offsets, total elements."""
    indices: List[int]
    offsets: List[int]
    numel: int
    size_limit: int


def layout_buckets(numels: Sequence[int], buckets: Sequence[Sequence[int]],
                   limits: Optional[Sequence[int]] = None) -> List[BucketSpec]:
    """Element offsets inside each bucket: running sum, no padding (reducer.hpp:372-375)."""
    nb = len(buckets)
    flat = np.asarray([i for b in buckets for i in b], dtype=np.int64)
    start = np.zeros(nb + 1, dtype=np.int64)
    np.cumsum([len(b) for b in buckets], out=start[1:])
    numel = np.asarray(numels, dtype=np.int64)
    off = np.empty(max(len(flat), 1), dtype=np.int64)
    bn = np.empty(max(nb, 1), dtype=np.int64)
    rc = _lib.lib().ddpb200_bucket_layout(_vp(numel), _vp(flat), _vp(start), nb, _vp(off), _vp(bn))
    _lib.check(rc, "ddpb200_bucket_layout")
    return [BucketSpec([int(i) for i in buckets[b]], [int(o) for o in off[start[b]:start[b + 1]]],
                       int(bn[b]), int(limits[b]) if limits is not None else 0) for b in range(nb)]


def initial_bucket_limits(find_unused_parameters: bool, static_graph: bool, bucket_bytes_cap: int,
                          bucket_cap_is_default: bool) -> List[int]:
    """Limits for the buckets built at construction (distributed.py:1224-1236): a single bucket
    unless unused parameters must be searched for."""
    if static_graph or not find_unused_parameters:
        return [sys.maxsize]
    if bucket_cap_is_default:
        return [DEFAULT_FIRST_BUCKET_BYTES, bucket_bytes_cap]
    return [bucket_bytes_cap]


def rebuild_bucket_limits(bucket_bytes_cap: int, bucket_cap_is_default: bool) -> List[int]:
    """Limits used when buckets are rebuilt after the first backward: [first_bucket_bytes_cap,
    bucket_bytes_cap] (Reducer ctor args, distributed.py:1266-1276)."""
    first = DEFAULT_FIRST_BUCKET_BYTES if bucket_cap_is_default else bucket_bytes_cap
    return [first, bucket_bytes_cap]
