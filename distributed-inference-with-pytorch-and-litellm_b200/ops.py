"""torch.library custom ops over the C ABI of libddp_b200.so (namespace ``ddp_b200``).

The thin dispatcher-facing layer BASELINE.json's north_star and SURVEY.md §8b name: every kernel of the hot
path as a registered op with a schema, declared mutations and a fake (meta) implementation, so that the
ops show up under ``torch.profiler`` by name, pass ``torch.library.opcheck`` and can sit inside a
``torch.compile``d training step without a graph break. Underneath each op is the same ctypes call into
the C ABI that ``CudaKernels`` makes; nothing is computed in Python and there is no CPU implementation
(the ops are registered for CUDA only: a CPU tensor raises torch's own "no kernel for backend" error).

Schemas follow the incumbent's where there is one — ``aten::_fused_adamw_(Tensor[] self, Tensor[] grads,
Tensor[] exp_avgs, Tensor[] exp_avg_sqs, ..., *, float lr, float beta1, float beta2, float weight_decay,
float eps, bool amsgrad, bool maximize, Tensor? grad_scale, Tensor? found_inf)`` — with
``step`` passed as an int (torch passes a list of step tensors) and the clip coefficient as one more
optional device scalar:

    ddp_b200::pack_scale_cast(Tensor[] src, Tensor(a!)[] dst, float scale) -> ()
    ddp_b200::pack_scale_cast_sumsq(Tensor[] src, Tensor(a!)[] dst, float scale, Tensor(b!) sumsq_out) -> ()
    ddp_b200::l2norm_sumsq(Tensor[] grads, Tensor(a!) sumsq_out) -> ()
    ddp_b200::l2norm_finalize(Tensor sumsq, float max_norm, bool torch_bf16) -> Tensor        # [total, coef]
    ddp_b200::scale_by_(Tensor(a!)[] grads, Tensor coef) -> ()
    ddp_b200::fused_clip_adamw_(Tensor(a!)[] self, Tensor(b!)[] grads, Tensor(c!)[] exp_avgs,
                                Tensor(d!)[] exp_avg_sqs, float lr, float beta1, float beta2,
                                float weight_decay, float eps, int step, bool maximize, bool round_clipped,
                                bool store_grad, Tensor? clip_coef, Tensor? grad_scale, Tensor? found_inf,
                                Tensor? lr_tensor) -> ()
    ddp_b200::allreduce_sumsq_(Tensor(a!) arena, Tensor[] views, Tensor(b!) sumsq_out, float scale,
                               bool scatter) -> ()

List-taking ops need a segment table (include/ddp_b200.h: ddpb200_seg); it is built once per distinct set
of pointers and kept (``_tables``), so a training loop that calls an op with the same tensors every step
pays a dictionary lookup, not a rebuild. The DDP scheduler itself keeps calling ``CudaKernels`` with its own
long-lived tables (per-bucket launches from autograd hooks, pointers patched in place); ``CudaKernels``
brackets those calls with ``torch.profiler.record_function`` ranges of the same op names while a profiler
is active.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from .kernels import SegTable, default_kernels

_MAX_CACHED_TABLES = 64
_tables: Dict[Tuple, SegTable] = {}
_arenas: Dict[int, object] = {}          # arena.tensor.data_ptr() -> SymmetricArena (allreduce_sumsq_)


def _table(cols: Sequence[Optional[Sequence[Tensor]]], numels: Sequence[int], device,
           chunk_elems: Optional[int] = None) -> SegTable:
    key = tuple(tuple(t.data_ptr() for t in c) if c is not None else None for c in cols) + (tuple(numels),)
    tab = _tables.get(key)
    if tab is None:
        if len(_tables) >= _MAX_CACHED_TABLES:
            _tables.pop(next(iter(_tables)))
        tab = default_kernels().table(list(cols), list(numels), chunk_elems=chunk_elems)
        _tables[key] = tab
    elif tab.dev is None and device.type == "cuda":
        tab.to_device(device)                   # second sighting of the same pointers: keep it on the device
    return tab


def _same_dtype(ts: Sequence[Tensor], what: str) -> torch.dtype:
    dts = {t.dtype for t in ts}
    if len(dts) != 1:
        raise RuntimeError(f"ddp_b200::{what}: tensors of one list must share a dtype, got {sorted(map(str, dts))}")
    return dts.pop()


def _check_lists(what: str, *lists: Sequence[Tensor]) -> None:
    n = len(lists[0])
    for lst in lists:
        if len(lst) != n:
            raise RuntimeError(f"ddp_b200::{what}: tensor lists differ in length")
        for t, r in zip(lst, lists[0]):
            if t.numel() != r.numel():
                raise RuntimeError(f"ddp_b200::{what}: tensors at one index differ in numel")
            if not t.is_contiguous():
                raise RuntimeError(f"ddp_b200::{what}: tensors must be contiguous")


def _pack(src, dst, scale, sumsq_out, what):
    if not src:
        return
    _check_lists(what, src, dst)
    tab = _table([src, dst, None, None], [t.numel() for t in src], src[0].device)
    default_kernels().pack(tab, _same_dtype(src, what), _same_dtype(dst, what), float(scale), sumsq_out=sumsq_out)


@torch.library.custom_op("ddp_b200::pack_scale_cast", mutates_args=("dst",), device_types="cuda")
def pack_scale_cast(src: List[Tensor], dst: List[Tensor], scale: float) -> None:
    """dst[i] = round_to_dst_dtype(fp32(src[i]) * scale)."""
    _pack(src, dst, scale, None, "pack_scale_cast")


@pack_scale_cast.register_fake
def _(src, dst, scale):
    return None


# (a separate op rather than an optional output: torch 2.11's opcheck fails on None for a mutable Tensor? argument)
@torch.library.custom_op("ddp_b200::pack_scale_cast_sumsq", mutates_args=("dst", "sumsq_out"), device_types="cuda")
def pack_scale_cast_sumsq(src: List[Tensor], dst: List[Tensor], scale: float, sumsq_out: Tensor) -> None:
    """pack_scale_cast + sumsq_out[i] = fp32 sum of squares of dst[i] as stored, in the same pass."""
    _pack(src, dst, scale, sumsq_out, "pack_scale_cast_sumsq")


@pack_scale_cast_sumsq.register_fake
def _(src, dst, scale, sumsq_out):
    return None


@torch.library.custom_op("ddp_b200::l2norm_sumsq", mutates_args=("sumsq_out",), device_types="cuda")
def l2norm_sumsq(grads: List[Tensor], sumsq_out: Tensor) -> None:
    """sumsq_out[i] = fp32 sum of squares of grads[i] (deterministic order, one launch)."""
    if not grads:
        return
    _check_lists("l2norm_sumsq", grads)
    tab = _table([None, grads, None, None], [t.numel() for t in grads], grads[0].device)
    default_kernels().sumsq(tab, _same_dtype(grads, "l2norm_sumsq"), sumsq_out)


@l2norm_sumsq.register_fake
def _(grads, sumsq_out):
    return None


@torch.library.custom_op("ddp_b200::l2norm_finalize", mutates_args=(), device_types="cuda")
def l2norm_finalize(sumsq: Tensor, max_norm: float, torch_bf16: bool) -> Tensor:
    """-> fp32 [total_norm, min(max_norm / (total_norm + 1e-6), 1)] on the device, no host sync."""
    return default_kernels().norm_finalize(sumsq, float(max_norm), bool(torch_bf16))


@l2norm_finalize.register_fake
def _(sumsq, max_norm, torch_bf16):
    return sumsq.new_empty(2, dtype=torch.float32)


@torch.library.custom_op("ddp_b200::scale_by_", mutates_args=("grads",), device_types="cuda")
def scale_by_(grads: List[Tensor], coef: Tensor) -> None:
    """grads[i] *= coef[0] in place (coef: fp32 device scalar), rounded to the gradient dtype."""
    if not grads:
        return
    _check_lists("scale_by_", grads)
    tab = _table([None, grads, None, None], [t.numel() for t in grads], grads[0].device)
    default_kernels().scale_by(tab, _same_dtype(grads, "scale_by_"), coef)


@scale_by_.register_fake
def _(grads, coef):
    return None


@torch.library.custom_op("ddp_b200::fused_clip_adamw_", device_types="cuda",
                         mutates_args=("params", "grads", "exp_avgs", "exp_avg_sqs"))
def fused_clip_adamw_(params: List[Tensor], grads: List[Tensor], exp_avgs: List[Tensor], exp_avg_sqs: List[Tensor],
                      lr: float, beta1: float, beta2: float, weight_decay: float, eps: float, step: int,
                      maximize: bool = False, round_clipped: bool = True, store_grad: bool = False,
                      clip_coef: Optional[Tensor] = None, grad_scale: Optional[Tensor] = None,
                      found_inf: Optional[Tensor] = None, lr_tensor: Optional[Tensor] = None) -> None:
    """One fused unscale + clip + AdamW pass over the lists (aten::_fused_adamw_'s job, plus the clip multiply)."""
    if not params:
        return
    from .optim import adamw_chunk_elems
    _check_lists("fused_clip_adamw_", params, grads, exp_avgs, exp_avg_sqs)
    numels = [t.numel() for t in params]
    tab = _table([params, grads, exp_avgs, exp_avg_sqs], numels, params[0].device, adamw_chunk_elems(sum(numels)))
    sdt = _same_dtype(exp_avgs, "fused_clip_adamw_")
    if _same_dtype(exp_avg_sqs, "fused_clip_adamw_") != sdt:
        raise RuntimeError("ddp_b200::fused_clip_adamw_: exp_avgs and exp_avg_sqs must share a dtype")
    default_kernels().adamw(tab, _same_dtype(params, "fused_clip_adamw_"), _same_dtype(grads, "fused_clip_adamw_"), sdt,
                            lr=lr_tensor if lr_tensor is not None else float(lr), beta1=float(beta1), beta2=float(beta2),
                            eps=float(eps), weight_decay=float(weight_decay), step=int(step), clip_coef=clip_coef,
                            grad_scale=grad_scale, found_inf=found_inf, round_clipped=bool(round_clipped),
                            store_grad=bool(store_grad), maximize=bool(maximize))


@fused_clip_adamw_.register_fake
def _(params, grads, exp_avgs, exp_avg_sqs, lr, beta1, beta2, weight_decay, eps, step, maximize=False,
      round_clipped=True, store_grad=False, clip_coef=None, grad_scale=None, found_inf=None, lr_tensor=None):
    return None


def register_arena(arena) -> None:
    """Make a SymmetricArena known to ``ddp_b200::allreduce_sumsq_`` (ops take tensors; the peer mappings,
    the multicast address and the cross-GPU barrier live in the arena object)."""
    _arenas[arena.tensor.data_ptr()] = arena


@torch.library.custom_op("ddp_b200::allreduce_sumsq_", mutates_args=("arena", "sumsq_out"), device_types="cuda")
def allreduce_sumsq_(arena: Tensor, views: List[Tensor], sumsq_out: Tensor, scale: float = 1.0,
                     scatter: bool = False) -> None:
    """SUM over ranks of ``views`` (slices of the symmetric ``arena``) in place on every rank + their
    per-tensor sums of squares, one kernel over NVLink peer memory (csrc/fused_allreduce.cu)."""
    sym = _arenas.get(arena.data_ptr())
    if sym is None:
        raise RuntimeError("ddp_b200::allreduce_sumsq_: the arena is not registered (ddp_b200.ops.register_arena)")
    if not views:
        return
    _check_lists("allreduce_sumsq_", views)
    tab = _table([None, views, None, None], [t.numel() for t in views], arena.device)
    if tab.dev is None:
        tab.to_device(arena.device)
    default_kernels().allreduce_sumsq(tab, _same_dtype(views, "allreduce_sumsq_"), sym, sumsq_out, scatter=bool(scatter),
                                      scale=float(scale))


@allreduce_sumsq_.register_fake
def _(arena, views, sumsq_out, scale=1.0, scatter=False):
    return None


__all__ = ["pack_scale_cast", "pack_scale_cast_sumsq", "l2norm_sumsq", "l2norm_finalize", "scale_by_",
           "fused_clip_adamw_", "allreduce_sumsq_", "register_arena"]
