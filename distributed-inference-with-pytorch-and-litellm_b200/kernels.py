"""Torch-facing layer over the C ABI: segment tables and one method per kernel.

PyTorch is plumbing here: it owns device memory and streams; every byte of the hot path is moved
by the sm_100a kernels in csrc/ reached through libddp_b200.so (ctypes). A `CudaKernels` object is
the default "kernel backend" of DistributedDataParallel / AdamW / clip_grad_norm_ in this
package. The host logic only ever talks to this small interface, which is what lets the
world_size-2 gloo tests drive the scheduler on CPU with a checker backend defined under tests/.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import BF16, F32, TABLE_DEVICE, TABLE_HOST

SEG_DTYPE = np.dtype([("ptr", np.uint64, (4,)), ("numel", np.int64), ("chunk_begin", np.int64)])
assert SEG_DTYPE.itemsize == 48

_DT = {torch.float32: F32, torch.bfloat16: BF16}


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise TypeError(f"ddp_b200 kernels support float32 and bfloat16, got {dt}") from None


class SegTable:
    """A table of `ddpb200_seg` rows (include/ddp_b200.h): up to four pointers and a length per
    parameter tensor, plus the chunk prefix sum the persistent kernels walk.

    Lives on the host (passed to kernels by value in the launch parameters: right for pointers
    that change every step, e.g. fresh autograd grads) and optionally on the device (built once:
    right for bucket views, parameters and optimizer state)."""

    def __init__(self, cols: Sequence[Optional[Sequence[torch.Tensor]]], numels: Sequence[int],
                 chunk_elems: Optional[int] = None):
        n = len(numels)
        self.n = n
        self.dev: Optional[torch.Tensor] = None
        self._args_cache = {}
        self.host = np.zeros(n, dtype=SEG_DTYPE)
        self.host["numel"] = np.asarray(numels, dtype=np.int64)
        for k, col in enumerate(cols):
            if col is not None:
                self.set_col(k, col)
        self.total_numel = int(self.host["numel"].sum())
        if chunk_elems is None:
            chunk_elems = int(_lib.lib().ddpb200_suggest_chunk_elems(self.total_numel, CudaKernels.sm_count()))
        self.chunk_elems = int(chunk_elems)
        chunks = (self.host["numel"] + (self.chunk_elems - 1)) // self.chunk_elems
        begins = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(chunks, out=begins[1:])
        self.host["chunk_begin"] = begins[:-1]
        self.chunk_begins = begins  # n+1 entries, for slicing
        self.total_chunks = int(begins[-1])

    def set_col(self, k: int, tensors: Sequence[torch.Tensor]) -> None:
        self.host["ptr"][:, k] = np.fromiter((t.data_ptr() for t in tensors), dtype=np.uint64, count=self.n)
        self.invalidate_device()

    def invalidate_device(self) -> None:
        if self.dev is not None:
            self.dev = None
            self._args_cache.clear()

    def col(self, k: int) -> np.ndarray:
        return self.host["ptr"][:, k]

    def to_device(self, device) -> "SegTable":
        self.dev = torch.from_numpy(self.host.view(np.uint8).copy()).to(device)
        self._args_cache.clear()
        return self

    # (pointer, location, n, total_chunks) of rows [first, first+count); cached: this sits on the
    # per-bucket launch path and the answers never change while the table lives
    def args(self, first: int = 0, count: Optional[int] = None, prefer_device: bool = True):
        on_dev = prefer_device and self.dev is not None
        key = (first, count, on_dev)
        hit = self._args_cache.get(key)
        if hit is not None:
            return hit
        cnt = self.n - first if count is None else count
        chunks = int(self.chunk_begins[first + cnt] - self.chunk_begins[first])
        if on_dev:
            out = (self.dev.data_ptr() + 48 * first, TABLE_DEVICE, cnt, chunks)
        else:
            out = (self.host.ctypes.data + 48 * first, TABLE_HOST, cnt, chunks)
        self._args_cache[key] = out
        return out


class _Workspace:
    """Scratch for the single-launch norm: one float per chunk + a ticket counter. One per
    (device, stream): launches on different streams must not share it."""

    def __init__(self):
        self._bufs = {}

    def get(self, device: torch.device, stream: int, nfloats: int):
        key = (device.index, stream)
        ent = self._bufs.get(key)
        if ent is None or ent[0].numel() < nfloats:
            size = max(nfloats, 4096)
            if ent is not None:
                size = max(size, 2 * ent[0].numel())
            partials = torch.empty(size, dtype=torch.float32, device=device)
            ticket = ent[1] if ent is not None else torch.zeros(1, dtype=torch.int32, device=device)
            ent = (partials, ticket)
            self._bufs[key] = ent
        return ent


class _NoRange:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_RANGE = _NoRange()


def _op_range(name: str):
    """While a torch profiler is active, bracket a kernel launch with a range named like the registered op
    (ops.py), so that the scheduler's direct C-ABI calls show up under the same names; free otherwise."""
    if torch.autograd.profiler._is_profiler_enabled:
        return torch.profiler.record_function(name)
    return _NO_RANGE


class CudaKernels:
    """The product kernel backend. Every method enqueues on torch's CURRENT stream and returns
    without synchronising."""

    name = "cuda-sm100a"
    _sms: Optional[int] = None

    def __init__(self):
        _lib.lib()  # fail loudly now if the .so is missing
        self._ws = _Workspace()

    # ---- environment -------------------------------------------------------------------
    @classmethod
    def sm_count(cls) -> int:
        if cls._sms is None:
            cls._sms = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count \
                if torch.cuda.is_available() else 148
        return cls._sms

    @staticmethod
    def _require_cuda(t: torch.Tensor) -> None:
        if not t.is_cuda:
            raise RuntimeError("ddp_b200: the hot path runs only on CUDA tensors (sm_100a kernels); "
                               "there is no CPU fallback")

    @staticmethod
    def _stream() -> int:
        return torch.cuda.current_stream().cuda_stream

    # ---- tables ----------------------------------------------------------------------------
    def table(self, cols, numels, chunk_elems=None, device=None) -> SegTable:
        t = SegTable(cols, numels, chunk_elems)
        if device is not None:
            t.to_device(device)
        return t

    # ---- kernel (a) ---------------------------------------------------------------------------
    def pack(self, tab: SegTable, src_dtype: torch.dtype, dst_dtype: torch.dtype, scale: float,
             first: int = 0, count: Optional[int] = None, sumsq_out: Optional[torch.Tensor] = None,
             device: Optional[torch.device] = None, prefer_device: bool = True) -> None:
        """dst = round(src * scale) for rows [first, first+count): ptr[0] -> ptr[1].
        With sumsq_out (fp32, one slot per row of the slice) also the sum of squares of dst."""
        ptr, loc, n, chunks = tab.args(first, count, prefer_device)
        st = self._stream()
        with _op_range("ddp_b200::pack_scale_cast"):
            if sumsq_out is not None:
                self._require_cuda(sumsq_out)
                partials, ticket = self._ws.get(sumsq_out.device, st, chunks)
                rc = _lib.lib().ddpb200_pack_scale_cast(ptr, loc, n, chunks, tab.chunk_elems, dtype_code(src_dtype),
                                                        dtype_code(dst_dtype), scale, sumsq_out.data_ptr(),
                                                        partials.data_ptr(), ticket.data_ptr(), st)
            else:
                rc = _lib.lib().ddpb200_pack_scale_cast(ptr, loc, n, chunks, tab.chunk_elems, dtype_code(src_dtype),
                                                        dtype_code(dst_dtype), scale, None, None, None, st)
        _lib.check(rc, "ddpb200_pack_scale_cast")

    # ---- kernel (b) ---------------------------------------------------------------------------
    def sumsq(self, tab: SegTable, dtype: torch.dtype, out: torch.Tensor, first: int = 0,
              count: Optional[int] = None, prefer_device: bool = True) -> None:
        """out[i] = sum of squares of ptr[1] of row first+i (fp32)."""
        self._require_cuda(out)
        ptr, loc, n, chunks = tab.args(first, count, prefer_device)
        st = self._stream()
        partials, ticket = self._ws.get(out.device, st, chunks)
        with _op_range("ddp_b200::l2norm_sumsq"):
            rc = _lib.lib().ddpb200_l2norm_sumsq(ptr, loc, n, chunks, tab.chunk_elems, dtype_code(dtype),
                                                 out.data_ptr(), partials.data_ptr(), ticket.data_ptr(), st)
        _lib.check(rc, "ddpb200_l2norm_sumsq")

    def norm_finalize(self, sumsq: torch.Tensor, max_norm: float, torch_bf16: bool,
                      out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """-> fp32 tensor [total_norm, clip_coef] on the device (no host sync)."""
        self._require_cuda(sumsq)
        if out is None:
            out = torch.empty(2, dtype=torch.float32, device=sumsq.device)
        rc = _lib.lib().ddpb200_l2norm_finalize(sumsq.data_ptr(), sumsq.numel(),
                                                _lib.NORM_TORCH_BF16 if torch_bf16 else _lib.NORM_EXACT_F32,
                                                float(max_norm), out.data_ptr(), self._stream())
        _lib.check(rc, "ddpb200_l2norm_finalize")
        return out

    def scale_by(self, tab: SegTable, dtype: torch.dtype, coef: torch.Tensor, first: int = 0,
                 count: Optional[int] = None, prefer_device: bool = True) -> None:
        """ptr[1] *= coef[0] in place (coef: fp32 device scalar)."""
        self._require_cuda(coef)
        ptr, loc, n, chunks = tab.args(first, count, prefer_device)
        rc = _lib.lib().ddpb200_scale_by_device_scalar(ptr, loc, n, chunks, tab.chunk_elems, dtype_code(dtype),
                                                       coef.data_ptr(), self._stream())
        _lib.check(rc, "ddpb200_scale_by_device_scalar")

    # ---- kernel (c) ---------------------------------------------------------------------------
    def adamw(self, tab: SegTable, param_dtype: torch.dtype, grad_dtype: torch.dtype,
              state_dtype: torch.dtype, *, lr, beta1: float, beta2: float, eps: float,
              weight_decay: float, step: int, clip_coef: Optional[torch.Tensor] = None,
              grad_scale: Optional[torch.Tensor] = None, found_inf: Optional[torch.Tensor] = None,
              round_clipped: bool = True, store_grad: bool = False, maximize: bool = False,
              prefer_device: bool = True, step_tensor: Optional[torch.Tensor] = None) -> None:
        """One fused unscale+clip+AdamW pass over all rows: ptr = [param, grad, exp_avg, exp_avg_sq].
        step_tensor (device fp32 scalar) replaces the host-side `step` for the bias corrections."""
        ptr, loc, n, chunks = tab.args(0, None, prefer_device)
        lr_ptr = None
        if isinstance(lr, torch.Tensor):
            if lr.is_cuda:
                lr_ptr, lr = lr.data_ptr(), 0.0
            else:
                lr = float(lr)
        flags = ((_lib.ADAMW_ROUND_CLIPPED_GRAD if round_clipped else 0) |
                 (_lib.ADAMW_STORE_GRAD if store_grad else 0) | (_lib.ADAMW_MAXIMIZE if maximize else 0))
        with _op_range("ddp_b200::fused_clip_adamw_"):
            rc = _lib.lib().ddpb200_fused_clip_adamw(
                ptr, loc, n, chunks, tab.chunk_elems, dtype_code(param_dtype), dtype_code(grad_dtype),
                dtype_code(state_dtype), float(lr), lr_ptr, beta1, beta2, eps, weight_decay, int(step),
                None if step_tensor is None else step_tensor.data_ptr(),
                None if clip_coef is None else clip_coef.data_ptr(),
                None if grad_scale is None else grad_scale.data_ptr(),
                None if found_inf is None else found_inf.data_ptr(), flags, self._stream())
        _lib.check(rc, "ddpb200_fused_clip_adamw")


    # ---- fused bucket allreduce + norm over peer memory ---------------------------------------
    def allreduce_sumsq(self, tab: SegTable, dtype: torch.dtype, arena, sumsq_out: torch.Tensor,
                        first: int = 0, count: Optional[int] = None, scatter: bool = False,
                        scale: float = 1.0) -> None:
        """SUM over ranks of rows [first, first+count) of the symmetric bucket arena, in place on
        every rank, and their per-tensor sums of squares into sumsq_out — barrier, one kernel,
        barrier, fold; all on the current stream. scatter=True: reduce-scatter only (the owner of
        a chunk keeps the reduced values to itself), for the sharded update. scale != 1 folds the
        1/world_size of a skipped in-place pack into the kernel (see include/ddp_b200.h)."""
        if tab.dev is None:
            raise RuntimeError("allreduce_sumsq needs the device-resident table of bucket views")
        ptr, _, n, chunks = tab.args(first, count, True)
        chunk_off = int(tab.chunk_begins[first])
        st = self._stream()
        L = _lib.lib()
        with _op_range("ddp_b200::allreduce_sumsq_"):
            arena.barrier()                                   # every rank has packed this range
            pm = arena.peer_map(chunk_off)
            fn = L.ddpb200_reduce_scatter_sumsq_chunks if scatter else L.ddpb200_allreduce_sumsq_chunks
            _lib.check(fn(ptr, n, chunks, tab.chunk_elems, dtype_code(dtype), ctypes.addressof(pm),
                          int(arena.use_multicast), float(scale), st), "ddpb200_allreduce_sumsq_chunks")
            arena.barrier()                                   # every rank has written its chunks everywhere
            _lib.check(L.ddpb200_fold_chunk_partials(ptr, n, chunks, arena.partials.data_ptr() + 4 * chunk_off,
                                                     sumsq_out.data_ptr(), st), "ddpb200_fold_chunk_partials")


    def sharded_adamw(self, tab: SegTable, dtype: torch.dtype, grad_arena, param_arena, exp_avg_shard: torch.Tensor,
                      exp_avg_sq_shard: torch.Tensor, *, lr, beta1: float, beta2: float, eps: float,
                      weight_decay: float, step: int, clip_coef: Optional[torch.Tensor] = None,
                      round_clipped: bool = True, maximize: bool = False) -> None:
        """AdamW on the chunks this rank owns + all-gather of the new parameters (csrc/sharded_update.cu),
        then the cross-GPU barrier after which every rank may read parameters."""
        if tab.dev is None:
            raise RuntimeError("sharded_adamw needs the device-resident table of gradient views")
        ptr, _, n, chunks = tab.args(0, None, True)
        lr_ptr = None
        if isinstance(lr, torch.Tensor):
            if lr.is_cuda:
                lr_ptr, lr = lr.data_ptr(), 0.0
            else:
                lr = float(lr)
        flags = (_lib.ADAMW_ROUND_CLIPPED_GRAD if round_clipped else 0) | (_lib.ADAMW_MAXIMIZE if maximize else 0)
        pm = param_arena.peer_map(0)
        # the owners are about to overwrite every rank's parameters: no rank may still be reading them (a late
        # backward op, an eval forward or an EMA on another stream of a slower rank). The reduce-scatter's
        # barrier only proves the peers packed their gradients, so there is one here as well as after.
        param_arena.barrier()
        rc = _lib.lib().ddpb200_sharded_adamw_allgather(
            ptr, n, chunks, tab.chunk_elems, dtype_code(dtype), grad_arena.tensor.data_ptr(), ctypes.addressof(pm),
            exp_avg_shard.data_ptr(), exp_avg_sq_shard.data_ptr(), float(lr), lr_ptr, beta1, beta2, eps, weight_decay,
            int(step), None if clip_coef is None else clip_coef.data_ptr(), flags, int(param_arena.use_multicast),
            self._stream())
        _lib.check(rc, "ddpb200_sharded_adamw_allgather")
        param_arena.barrier()


_default: Optional[CudaKernels] = None


def default_kernels() -> CudaKernels:
    global _default
    if _default is None:
        _default = CudaKernels()
    return _default
