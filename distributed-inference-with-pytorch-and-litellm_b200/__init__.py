"""B200-native gradient-sync + clip + AdamW for DistributedDataParallel fine-tuning.

A drop-in, for that one path, behind the call shapes the reference advertises
(REF/README.md:7): ``DistributedDataParallel(model)``, ``clip_grad_norm_(params, max_norm)``,
``AdamW(params, ...)``. Everything that touches gradient, parameter or optimizer-state bytes
is a hand-written sm_100a kernel in ``csrc/`` behind the C ABI of ``include/ddp_b200.h``.
There is no CPU fallback.
"""
from . import _lib, bucketing
from ._lib import build as build_native
from .bucketing import compute_bucket_assignment_by_size, layout_buckets
from .clip import clip_grad_norm_
from .ddp import DistributedDataParallel, GradBucket
from .kernels import CudaKernels, SegTable, default_kernels
from .optim import AdamW
from .zero import ShardedAdamW
from . import ops  # noqa: F401  (registers the torch.library ops of namespace ddp_b200)

DDP = DistributedDataParallel

__all__ = ["DistributedDataParallel", "DDP", "GradBucket", "AdamW", "ShardedAdamW", "clip_grad_norm_",
           "compute_bucket_assignment_by_size", "layout_buckets", "CudaKernels", "SegTable",
           "default_kernels", "build_native"]
