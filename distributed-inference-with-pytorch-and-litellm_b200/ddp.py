"""DistributedDataParallel with the gradient-sync half rebuilt for B200.

Drop-in for ``torch.nn.parallel.DistributedDataParallel`` on the data-parallel path
(TORCH/nn/parallel/distributed.py:655-972 ctor, :1693-1701 forward, :1987-2066 comm hooks) and
for the C++ Reducer behind it (TORCH/include/torch/csrc/distributed/c10d/reducer.hpp): same
constructor keywords, same bucket layouts (bit-exact: one bucket in registration order for the
first iteration, then rebuilt in rank 0's gradient-ready order with [first_cap, cap] limits),
same averaging arithmetic (round(fp32(g) * float(1/world_size)) then SUM), same comm-hook
protocol, ``no_sync()``, and ``module.``-prefixed state_dict.

What is different is how the bytes move:
  * one pack launch per BUCKET (flatten + scale + cast, 16-byte vectors) instead of one
    elementwise launch per parameter (reducer.hpp:276);
  * the bucket allreduce runs on a dedicated side stream as soon as the bucket is packed, and the
    per-tensor sum of squares needed by clip_grad_norm_ is taken on that stream right behind the
    allreduce while backward is still running — so the global-norm barrier costs one tiny
    finalize kernel after backward (``ddp.clip_grad_norm_``);
  * with ``gradient_as_bucket_view=True`` the fused clip+AdamW kernel reads gradients straight
    from the buckets: there is no unpack pass at all.

One process drives one GPU. ``find_unused_parameters=True``, sparse gradients, multi-device
modules and ``join()`` are outside the path this package covers and raise.
"""
from __future__ import annotations

import os
import weakref
from contextlib import contextmanager
from typing import Callable, List, Optional

import numpy as np
import torch
import torch.distributed as dist
from torch import nn

from . import bucketing
from .kernels import default_kernels


class GradBucket:
    """What a communication hook receives: mirrors ``dist.GradBucket`` (index, buffer, gradients,
    parameters, is_last, set_buffer) — TORCH/nn/parallel/distributed.py:1987-2035."""

    def __init__(self, index: int, buffer: torch.Tensor, gradients: List[torch.Tensor],
                 parameters: List[torch.Tensor], is_last: bool):
        self._index, self._buffer, self._gradients = index, buffer, gradients
        self._parameters, self._is_last = parameters, is_last

    def index(self) -> int:
        return self._index

    def buffer(self) -> torch.Tensor:
        return self._buffer

    def gradients(self) -> List[torch.Tensor]:
        return self._gradients

    def parameters(self) -> List[torch.Tensor]:
        return self._parameters

    def is_last(self) -> bool:
        return self._is_last

    def set_buffer(self, tensor: torch.Tensor) -> None:
        self._buffer = tensor


class _Bucket:
    __slots__ = ("index", "spec", "buffer", "views", "first", "n", "params", "sumsq", "pending", "future",
                 "work", "dtype", "arena_start")

    def __init__(self, index, spec, buffer, views, first, params):
        self.index, self.spec, self.buffer, self.views, self.first = index, spec, buffer, views, first
        self.n = len(spec.indices)
        self.params = params            # the bucket's parameters, bucket order
        self.sumsq = None               # this bucket's slice of the per-tensor sum-of-squares array
        self.pending = self.n
        self.future = None
        self.work = None
        self.dtype = buffer.dtype


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, init_sync: bool = True, process_group=None,
                 bucket_cap_mb: Optional[float] = None, find_unused_parameters: bool = False,
                 check_reduction: bool = False, gradient_as_bucket_view: bool = False,
                 static_graph: bool = False, delay_all_reduce_named_params=None,
                 param_to_hook_all_reduce=None, mixed_precision=None, device_mesh=None,
                 skip_all_reduce_unused_params: bool = False, *, kernels=None,
                 fused_allreduce: Optional[bool] = None, shard_optimizer: bool = False,
                 grad_bucket_dtype: Optional[torch.dtype] = None):
        super().__init__()
        if find_unused_parameters:
            raise NotImplementedError("ddp_b200: find_unused_parameters=True is outside the hot path covered")
        if delay_all_reduce_named_params or mixed_precision is not None or device_mesh is not None:
            raise NotImplementedError("ddp_b200: delay_all_reduce / mixed_precision / device_mesh are not covered")
        self.module = module
        self.dim = dim
        self.broadcast_buffers = broadcast_buffers
        self.find_unused_parameters = False
        self.static_graph = static_graph
        self.gradient_as_bucket_view = gradient_as_bucket_view
        self.require_backward_grad_sync = True
        self.require_forward_param_sync = True
        self.process_group = process_group if process_group is not None else dist.group.WORLD
        self.world_size = dist.get_world_size(self.process_group)
        self._inv_world_size = float(np.float32(1.0 / self.world_size))     # float(1.0/ws), as div_factor_
        self.rank = dist.get_rank(self.process_group)
        self.parameters_to_ignore = set(getattr(module, "_ddp_params_and_buffers_to_ignore", []))

        # reduction bucket size (distributed.py:831-853)
        if bucket_cap_mb is None:
            bucket_cap_mb, self.bucket_bytes_cap_default = 25, True
        else:
            self.bucket_bytes_cap_default = False
        self.bucket_bytes_cap = int(bucket_cap_mb * 1024 * 1024)
        self.broadcast_bucket_size = 250 * 1024 * 1024

        self._params: List[nn.Parameter] = self._build_params_for_reducer()
        if not self._params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any "
                               "parameter that requires a gradient.")
        devs = {p.device for p in self._params}
        if len(devs) != 1:
            raise NotImplementedError("ddp_b200: one process drives one device; parameters span several")
        self.device = devs.pop()
        self._cuda = self.device.type == "cuda"
        if device_ids is not None and self._cuda and len(device_ids) == 1:
            d = device_ids[0]
            d = d.index if isinstance(d, torch.device) else int(d)
            if d != self.device.index:
                raise ValueError(f"device_ids {device_ids} does not match the module's device {self.device}")
        self.device_ids = device_ids
        self.output_device = output_device
        self._kernels = kernels if kernels is not None else default_kernels()
        if any(p.is_sparse for p in self._params):
            raise NotImplementedError("ddp_b200: sparse parameters are not covered")

        if init_sync:
            self._verify_param_shapes()
            self._sync_module_states()

        self._comm_stream = torch.cuda.Stream(device=self.device, priority=-1) if self._cuda else None
        self._use_side_stream = os.environ.get("PYTORCH_DDP_USE_SIDE_STREAM", "1") == "1"
        self._copy_stream = None
        # bucket allreduce + norm as ONE kernel over NVLink peer memory instead of NCCL + a norm pass
        # fused_allreduce: True = required (raises if symmetric memory cannot be set up), False = NCCL,
        # None = the environment's choice, DDPB200_FUSED_ALLREDUCE = 1 | 0 | auto (default auto: use it where it
        # can be set up on every rank, otherwise say so once and use NCCL)
        self._fused_optional = False
        if fused_allreduce is None:
            env = os.environ.get("DDPB200_FUSED_ALLREDUCE", "auto")
            fused_allreduce = env in ("1", "auto")
            self._fused_optional = env == "auto"
        self._fused = bool(fused_allreduce) and self._cuda and 1 < self.world_size <= 8
        self._sym = {}
        # sharded update (ZeRO-1 style, see zero.py): reduce-scatter instead of allreduce, parameters
        # re-homed into a symmetric arena so that the owners of a chunk can write them everywhere
        self._zero = bool(shard_optimizer)
        if self._zero and not self._fused:
            raise RuntimeError("shard_optimizer=True needs fused_allreduce=True on 2..8 CUDA ranks of one NVLink domain")
        if self._zero and len({p.dtype for p in self._params}) != 1:
            raise NotImplementedError("ddp_b200 sharded update: parameters must share one dtype")
        self._param_sym = {}
        # Buckets in a narrower dtype than the gradients (fp32 gradients -> bf16 buckets): the pack
        # kernel scales and rounds once, round(fp32(g) * float(1/ws)), the collective moves half the
        # bytes, the unpack kernel widens back into p.grad. torch's closest equivalent is
        # bf16_compress_hook (default_hooks.py:57-134), which rounds first and divides afterwards.
        self._bucket_dtype = grad_bucket_dtype
        if grad_bucket_dtype is not None:
            if grad_bucket_dtype not in (torch.bfloat16, torch.float32):
                raise TypeError("grad_bucket_dtype must be torch.bfloat16 or torch.float32")
            if gradient_as_bucket_view or self._zero:
                raise ValueError("grad_bucket_dtype needs gradient_as_bucket_view=False and no shard_optimizer: "
                                 "p.grad keeps the parameter dtype, so it cannot alias a bucket of another dtype")
        self._comm_hook: Optional[Callable] = None
        self._comm_hook_state = None
        self._hook_handles = []
        self._has_rebuilt = False
        self._ready_order: List[int] = []
        self._expect_hooks = False
        self._callback_queued = False
        self._next_bucket = 0
        self._sumsq_valid = False
        self._coalesce = False
        self._coalesce_groups = int(os.environ.get("DDPB200_COALESCE_GROUPS", "4"))
        # gradients that already live in the bucket arena: skip the in-place pack, scale inside the collective
        self._fold_ok = os.environ.get("DDPB200_FOLD_SCALE", "1") == "1"
        self._stats = {"iterations": 0, "pack_launches": 0, "allreduces": 0, "norm_launches": 0,
                       "unpack_launches": 0, "rebuilt_bucket_sizes": []}

        if self._zero:
            # optimizer state is laid out by chunk, so the bucket layout must not change once the
            # optimizer has stepped: build the final buckets now, in reversed registration order
            # (the order torch assumes gradients become ready), and never rebuild
            n = len(self._params)
            order = list(range(n - 1, -1, -1))
            limits = bucketing.rebuild_bucket_limits(self.bucket_bytes_cap, self.bucket_bytes_cap_default)
            idx, lim = bucketing.compute_bucket_assignment_by_size([self._params[i] for i in order], limits, None, order)
            self._build_buckets(idx, lim)
            self._has_rebuilt = True
        else:
            limits = bucketing.initial_bucket_limits(False, static_graph, self.bucket_bytes_cap,
                                                     self.bucket_bytes_cap_default)
            idx, lim = bucketing.compute_bucket_assignment_by_size(self._params, limits)
            # reversed: approximate the order in which gradients become ready (distributed.py:1254-1258)
            self._build_buckets(list(reversed(idx)), list(reversed(lim)))

        ref = weakref.ref(self)
        for i, p in enumerate(self._params):
            def hook(param, _i=i, _ref=ref):
                me = _ref()
                if me is not None:
                    me._on_grad_ready(_i, True)
            self._hook_handles.append(p.register_post_accumulate_grad_hook(hook))

    # ---- construction helpers --------------------------------------------------------------
    def _build_params_for_reducer(self) -> List[nn.Parameter]:
        """Parameters in registration order, shared ones once (distributed.py:1338-1382)."""
        seen, out = set(), []
        for mname, mod in self.module.named_modules():
            for pname, p in mod.named_parameters(recurse=False):
                if not p.requires_grad or f"{mname}.{pname}" in self.parameters_to_ignore:
                    continue
                if isinstance(mod, (nn.Embedding, nn.EmbeddingBag)) and mod.sparse:
                    raise NotImplementedError("ddp_b200: sparse-gradient embeddings are not covered")
                if p not in seen:
                    seen.add(p)
                    out.append(p)
        return out

    def _verify_param_shapes(self) -> None:
        """All ranks must hold the same parameter list (_verify_param_shape_across_processes,
        distributed.py:881)."""
        if self.world_size == 1:
            return
        sizes = torch.tensor([len(self._params)] + [p.numel() for p in self._params], dtype=torch.int64,
                             device=self.device)
        ref = sizes.clone()
        dist.broadcast(ref, src=dist.get_global_rank(self.process_group, 0), group=self.process_group)
        if not torch.equal(ref, sizes):
            raise RuntimeError("DDP expects same model across all ranks, but parameter sizes differ from rank 0")

    def _sync_module_states(self) -> None:
        """Broadcast parameters and buffers from rank 0 in flat chunks of at most 250 MiB
        (_sync_module_states, distributed.py:883-890)."""
        if self.world_size == 1:
            return
        tensors = [p.detach() for n, p in self.module.named_parameters() if n not in self.parameters_to_ignore]
        if self.broadcast_buffers:
            tensors += [b.detach() for n, b in self.module.named_buffers() if n not in self.parameters_to_ignore]
        self._broadcast_coalesced(tensors)

    def _broadcast_coalesced(self, tensors: List[torch.Tensor]) -> None:
        src = dist.get_global_rank(self.process_group, 0)
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault((t.dtype, t.device), []).append(t)
        for (_, _), ts in by_dtype.items():
            chunk, size = [], 0
            for t in ts + [None]:
                if t is not None and (size == 0 or size + t.numel() * t.element_size() <= self.broadcast_bucket_size):
                    chunk.append(t)
                    size += t.numel() * t.element_size()
                    continue
                if chunk:
                    flat = torch.cat([c.reshape(-1) for c in chunk])
                    dist.broadcast(flat, src=src, group=self.process_group)
                    off = 0
                    for c in chunk:
                        c.copy_(flat[off:off + c.numel()].view_as(c))
                        off += c.numel()
                chunk, size = ([t], t.numel() * t.element_size()) if t is not None else ([], 0)

    def _build_buckets(self, bucket_indices: List[List[int]], limits: List[int]) -> None:
        numels = [p.numel() for p in self._params]
        specs = bucketing.layout_buckets(numels, bucket_indices, limits)
        self._buckets: List[_Bucket] = []
        self._order: List[int] = []
        # All buckets of one dtype are carved out of ONE arena, in bucket order, each starting on a
        # 128-element boundary: inside a bucket the layout is torch's (no padding), and a run of
        # consecutive buckets is one contiguous range, so it can be packed, allreduced and normed as a
        # single launch / a single collective when all of its gradients are present at once.
        align = 128
        arena_len, starts = {}, []
        for spec in specs:
            dt = self._bucket_dtype or self._params[spec.indices[0]].dtype
            off = (arena_len.get(dt, 0) + align - 1) // align * align
            starts.append(off)
            arena_len[dt] = off + spec.numel
        if self._fused:
            from .symm import SymmetricArena
            max_chunks = sum(p.numel() for p in self._params) // 2048 + len(self._params) + 1
            err = None
            try:
                self._sym = {dt: SymmetricArena(n, dt, self.device, self.process_group, max_chunks)
                             for dt, n in sorted(arena_len.items(), key=lambda kv: str(kv[0]))}
            except Exception as e:          # no NVLink peer access / no symmetric-memory support on this box
                if not self._fused_optional:
                    raise
                err = e
            if self._fused_optional:        # every rank must take the same branch: agree
                ok = torch.tensor([0 if err is not None else 1], device=self.device)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.process_group)
                if int(ok) == 0:
                    import warnings
                    warnings.warn("ddp_b200: symmetric memory for the fused bucket allreduce could not be set up on every "
                                  f"rank ({type(err).__name__ if err else 'another rank failed'}); using NCCL")
                    self._fused, self._sym = False, {}
        if self._fused:
            self._arenas = {dt: a.tensor for dt, a in self._sym.items()}
        else:
            self._arenas = {dt: torch.zeros(max(n, 1), dtype=dt, device=self.device) for dt, n in arena_len.items()}
        first = 0
        for b, spec in enumerate(specs):
            dt = self._bucket_dtype or self._params[spec.indices[0]].dtype
            buf = self._arenas[dt][starts[b]:starts[b] + spec.numel]
            views = [buf[o:o + numels[i]].view_as(self._params[i]) for i, o in zip(spec.indices, spec.offsets)]
            bk = _Bucket(b, spec, buf, views, first, [self._params[i] for i in spec.indices])
            bk.arena_start = starts[b]
            self._buckets.append(bk)
            self._order.extend(spec.indices)
            first += len(spec.indices)
        if self._zero:
            from .symm import SymmetricArena
            self._param_sym = {dt: SymmetricArena(n_, dt, self.device, self.process_group)
                               for dt, n_ in sorted(arena_len.items(), key=lambda kv: str(kv[0]))}
            if os.environ.get("DDPB200_FUSED_MODE", "auto") == "auto":
                # the two phases of the sharded step each load one NVLink direction; peer loads/stores
                # spread them over both, multicast does not: 7.33 vs 7.47 ms per step at 8 ranks
                for a_ in list(self._sym.values()) + list(self._param_sym.values()):
                    a_.use_multicast = False
            # tuning: the transport of each phase on its own (reduce-scatter of gradients / all-gather of parameters)
            for env, arenas in (("DDPB200_ZERO_GRAD_MODE", self._sym), ("DDPB200_ZERO_PARAM_MODE", self._param_sym)):
                want = os.environ.get(env)
                if want in ("peer", "mc"):
                    for a_ in arenas.values():
                        a_.use_multicast = want == "mc" and bool(a_.multicast)
            with torch.no_grad():
                for bk in self._buckets:            # same layout as the gradient arena: same byte offsets
                    pa = self._param_sym[bk.dtype].tensor
                    for pi, o in zip(bk.spec.indices, bk.spec.offsets):
                        p = self._params[pi]
                        home = pa[bk.arena_start + o:bk.arena_start + o + numels[pi]].view_as(p)
                        home.copy_(p.data)
                        p.data = home
        n = len(self._order)
        self._slot_of = {pi: s for s, pi in enumerate(self._order)}
        self._bucket_of = {pi: bk for bk in self._buckets for pi in bk.spec.indices}
        all_views = [v for bk in self._buckets for v in bk.views]
        onum = [numels[i] for i in self._order]
        k = self._kernels
        # pack: ptr[0] = this step's grad tensor (changes every step -> host table), ptr[1] = bucket view
        self._pack_table = k.table([None, all_views, None, None], onum)
        # unpack: ptr[0] = bucket view, ptr[1] = grad tensor
        self._unpack_table = k.table([all_views, None, None, None], onum, chunk_elems=self._pack_table.chunk_elems)
        # norm / in-place clip over the bucket views: static pointers -> device-resident table
        self._view_table = k.table([None, all_views, None, None], onum, chunk_elems=self._pack_table.chunk_elems,
                                   device=self.device if self._cuda else None)
        self._view_ptrs = [v.data_ptr() for v in all_views]
        self._sumsq = torch.zeros(n, dtype=torch.float32, device=self.device)
        for bk in self._buckets:
            bk.sumsq = self._sumsq[bk.first:bk.first + bk.n]
        self._norm_out = torch.zeros(2, dtype=torch.float32, device=self.device)
        self._sumsq_valid = False
        self._next_bucket = 0

    # ---- public surface shared with torch's DDP -------------------------------------------------
    def forward(self, *inputs, **kwargs):
        if torch.is_grad_enabled() and self.require_backward_grad_sync:
            self._rebuild_buckets()
            self._prepare_for_backward()
        elif torch.is_grad_enabled():
            self._sumsq_valid = False        # no_sync(): gradients will change without being reduced or normed
        # _sync_buffers in _pre_forward (distributed.py:1553-1600): gated on require_forward_param_sync, which the
        # PREVIOUS forward's _post_forward set — so the forward after a no_sync() / no-grad forward does not
        # broadcast, exactly as in torch; buffers named in _ddp_params_and_buffers_to_ignore are left alone
        if self.broadcast_buffers and self.world_size > 1 and self.require_forward_param_sync:
            bufs = [b for n_, b in self.module.named_buffers() if n_ not in self.parameters_to_ignore]
            if bufs:
                with torch.no_grad():
                    self._broadcast_coalesced(bufs)
        if self.device_ids is not None:
            inputs, kwargs = self._to_device(inputs), self._to_device(kwargs)   # _pre_forward, distributed.py:1602-1622
        out = self.module(*inputs, **kwargs)
        # _post_forward (distributed.py:1624-1660)
        self.require_forward_param_sync = bool(torch.is_grad_enabled() and self.require_backward_grad_sync)
        return out

    def _to_device(self, obj):
        """Inputs to the module's device, as torch's DDP does when device_ids is given: host tensors are
        copied on a side stream (PYTORCH_DDP_USE_SIDE_STREAM, distributed.py:855-858) so that a pinned
        batch streams in behind the previous step's tail; the compute stream waits for the copy."""
        if isinstance(obj, torch.Tensor):
            if obj.device == self.device:
                return obj
            if self._cuda and self._use_side_stream and obj.device.type == "cpu":
                if self._copy_stream is None:
                    self._copy_stream = torch.cuda.Stream(device=self.device)
                with torch.cuda.stream(self._copy_stream):
                    out = obj.to(self.device, non_blocking=True)
                cur = torch.cuda.current_stream(self.device)
                cur.wait_stream(self._copy_stream)
                out.record_stream(cur)
                return out
            return obj.to(self.device, non_blocking=True)
        if isinstance(obj, tuple):
            return tuple(self._to_device(o) for o in obj)
        if isinstance(obj, list):
            return [self._to_device(o) for o in obj]
        if isinstance(obj, dict):
            return {k: self._to_device(v) for k, v in obj.items()}
        return obj

    @contextmanager
    def no_sync(self):
        """Skip gradient synchronisation inside the context; gradients accumulate locally and are
        reduced by the first backward after it (distributed.py:1472-1498)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def register_comm_hook(self, state, hook: Callable) -> None:
        """hook(state, bucket) -> Future[Tensor]; buckets arrive NOT pre-divided
        (distributed.py:1987-2035). One registration only, before the first backward."""
        if self._comm_hook is not None:
            raise RuntimeError("register_comm_hook or register_builtin_comm_hook can only be called once.")
        if not callable(hook):
            raise TypeError("Communication hook must be callable.")
        self._comm_hook, self._comm_hook_state = hook, state

    def _get_ddp_logging_data(self) -> dict:
        d = dict(self._stats)
        d["world_size"] = self.world_size
        d["bucket_sizes"] = [bk.buffer.numel() * bk.buffer.element_size() for bk in self._buckets]
        d["has_rebuilt_buckets"] = self._has_rebuilt
        d["gradient_as_bucket_view"] = self.gradient_as_bucket_view
        d["backend_name"] = dist.get_backend(self.process_group)
        d["kernel_backend"] = getattr(self._kernels, "name", type(self._kernels).__name__)
        return d

    def bucket_layout(self) -> List[dict]:
        """Current buckets: parameter indices, element offsets and lengths (for parity checks)."""
        return [dict(index=bk.index, parameters=list(bk.spec.indices), offsets=list(bk.spec.offsets),
                     numel=bk.spec.numel, size_limit=bk.spec.size_limit) for bk in self._buckets]

    # ---- reducer --------------------------------------------------------------------------------
    def _prepare_for_backward(self) -> None:
        self._expect_hooks = True
        self._next_bucket = 0
        self._callback_queued = False
        self._sumsq_valid = False
        for bk in self._buckets:
            bk.pending = len(bk.spec.indices)
            bk.future = bk.work = None

    def _rebuild_buckets(self) -> bool:
        """Once, after the first synchronised backward: re-bucket in the order gradients became
        ready on rank 0, with limits [first_bucket_bytes_cap, bucket_bytes_cap]
        (Reducer::rebuild_buckets, reducer.hpp:146-151,299-301; triggered distributed.py:1584)."""
        if self._has_rebuilt or not self._ready_order:
            return False
        n = len(self._params)
        order = torch.tensor(self._ready_order if len(self._ready_order) == n else list(range(n)),
                             dtype=torch.int64, device=self.device)
        if self.world_size > 1:
            dist.broadcast(order, src=dist.get_global_rank(self.process_group, 0), group=self.process_group)
        order = [int(x) for x in order.tolist()]
        limits = bucketing.rebuild_bucket_limits(self.bucket_bytes_cap, self.bucket_bytes_cap_default)
        idx, lim = bucketing.compute_bucket_assignment_by_size([self._params[i] for i in order], limits,
                                                               None, order)
        if self._cuda:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
        # gradients that alias the old buckets move straight into their new views, bucket by bucket: the old
        # arena stays alive only through these references (peak = old + new arena, no third copy of every gradient)
        old_grads = [p.grad for p in self._params] if self.gradient_as_bucket_view else None
        self._build_buckets(idx, lim)
        if old_grads is not None:
            with torch.no_grad():
                for bk in self._buckets:
                    for pi, view in zip(bk.spec.indices, bk.views):
                        g = old_grads[pi]
                        if g is not None and g.dtype == view.dtype:
                            view.copy_(g)
                            self._params[pi].grad = view
                            old_grads[pi] = None
            del old_grads
        self._has_rebuilt = True
        self._ready_order = []
        self._stats["rebuilt_bucket_sizes"] = [bk.buffer.numel() * bk.buffer.element_size() for bk in self._buckets]
        return True

    def _on_grad_ready(self, idx: int, from_autograd: bool) -> None:
        """Per-parameter hook (Reducer::autograd_hook -> mark_variable_ready, reducer.hpp:74,276):
        runs on the autograd thread, with that thread's current stream = the compute stream."""
        if not self._expect_hooks or not self.require_backward_grad_sync:
            return
        if from_autograd and not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
        if not self._has_rebuilt:
            self._ready_order.append(idx)
        bk = self._bucket_of[idx]
        bk.pending -= 1
        if bk.pending < 0:
            raise RuntimeError("ddp_b200: a parameter was marked ready twice in one backward "
                               "(shared across multiple forward passes needs static_graph / no_sync)")
        # buckets are reduced strictly in sequence so every rank issues the same collectives in
        # the same order (mark_bucket_ready, reducer.hpp:282)
        b0 = self._next_bucket
        while self._next_bucket < len(self._buckets) and self._buckets[self._next_bucket].pending == 0:
            self._next_bucket += 1
        if self._next_bucket > b0:
            self._launch_buckets(b0, self._next_bucket)

    @property
    def _local_only(self) -> bool:
        """world_size 1 and no hook: there is nothing to exchange, so gradients stay where autograd
        put them (no bucket copy, whatever gradient_as_bucket_view says) and only the norm pass is left."""
        return self.world_size == 1 and self._comm_hook is None

    def _launch_buckets(self, b0: int, b1: int) -> None:
        """Buckets b0..b1-1 have all their gradients."""
        if self._local_only:
            self._launch_local_run(b0, b1)
        elif self._coalesce and self._comm_hook is None and b1 - b0 > 1:
            self._launch_coalesced(b0, b1)
        else:
            for b in range(b0, b1):
                self._launch_bucket(self._buckets[b])

    def _folds_scale(self, dt) -> bool:
        """Whether the 1/world_size scale is applied INSIDE the fused collective for buckets of dtype dt.
        A property of the configuration, identical on every rank (the owner of a chunk applies one scale
        to every rank's copy, so the ranks must agree): fused collective, gradient_as_bucket_view, no
        comm hook. Then a gradient that already IS its bucket view (accumulated in place by autograd, or
        written there by the caller) needs no pack launch at all — stock DDP would still run the in-place
        `bucket_view.div_(world_size)` (reducer.hpp:276) — and one that is not is copied unscaled.
        Peer mode scales per rank before the sum (the bits a pack would have stored, any world size);
        multicast mode scales the switch's sum, the same number only for a power-of-two world size
        (csrc/fused_allreduce.cu), so it folds only then."""
        if not (self._fused and self._comm_hook is None and self._fold_ok and self.gradient_as_bucket_view):
            return False
        return not (self._sym[dt].use_multicast and (self.world_size & (self.world_size - 1)))

    def _aliased(self, first: int, grads: List[torch.Tensor]) -> bool:
        vp = self._view_ptrs
        return all(g.data_ptr() == vp[first + i] for i, g in enumerate(grads))

    def _launch_coalesced(self, b0: int, b1: int) -> None:
        """Buckets b0..b1-1 are complete AT ONCE and every rank knows it (manual gradient feed): per
        run of equal dtype ONE pack launch, ONE allreduce over the contiguous arena range, ONE norm
        (or unpack+norm) launch. Not used from autograd hooks, where ranks could disagree on which
        buckets happen to be complete together."""
        k = self._kernels
        # Up to `_coalesce_groups` sub-runs of about equal size: the pack of sub-run g+1 (HBM-bound,
        # compute stream) overlaps the collective of sub-run g (NVLink-bound, side stream).
        total = sum(bk.spec.numel for bk in self._buckets[b0:b1])
        groups = self._coalesce_groups
        if all(self._folds_scale(bk.dtype) for bk in self._buckets[b0:b1]):
            # nothing to pack if the gradients are the bucket views: no pack to overlap, one collective per dtype
            vp = self._view_ptrs
            if all(p.grad is not None and p.grad.data_ptr() == vp[bk.first + i]
                   for bk in self._buckets[b0:b1] for i, p in enumerate(bk.params)):
                groups = 1
        budget = max(1, -(-total // max(1, groups)))
        i = b0
        while i < b1:
            j = i
            dt = self._buckets[i].dtype
            grads = []
            taken = 0
            pdt = self._buckets[i].params[0].dtype
            while (j < b1 and self._buckets[j].dtype == dt and self._buckets[j].params[0].dtype == pdt
                   and (taken == 0 or taken + self._buckets[j].spec.numel <= budget)):
                taken += self._buckets[j].spec.numel
                for p in self._buckets[j].params:
                    g = p.grad
                    if g is None:
                        raise RuntimeError("ddp_b200: gradient missing for a parameter marked ready")
                    if not g.is_contiguous():
                        g = g.contiguous()
                        p.grad = g
                    grads.append(g)
                j += 1
            gdt = grads[0].dtype
            if any(g.dtype != gdt for g in grads):
                raise RuntimeError("ddp_b200: gradients of one dtype group must share a dtype")
            first, n = self._buckets[i].first, len(grads)
            last = self._buckets[j - 1]
            span = self._arenas[dt][self._buckets[i].arena_start:last.arena_start + last.spec.numel]
            fold = self._folds_scale(dt)
            if not (fold and gdt == dt and self._aliased(first, grads)):
                self._pack_table.host["ptr"][first:first + n, 0] = [g.data_ptr() for g in grads]
                k.pack(self._pack_table, gdt, dt, 1.0 if fold else self._inv_world_size, first=first, count=n,
                       prefer_device=False)
                self._stats["pack_launches"] += 1

            def behind():
                if self.gradient_as_bucket_view:
                    k.sumsq(self._view_table, dt, self._sumsq[first:first + n], first=first, count=n)
                    self._stats["norm_launches"] += 1
                else:
                    self._unpack_table.host["ptr"][first:first + n, 1] = [g.data_ptr() for g in grads]
                    k.pack(self._unpack_table, dt, gdt, 1.0, first=first, count=n,
                           sumsq_out=self._sumsq[first:first + n], prefer_device=False)
                    self._stats["unpack_launches"] += 1

            if self._fused:
                ev = torch.cuda.Event()
                ev.record()
                with torch.cuda.stream(self._comm_stream):
                    self._comm_stream.wait_event(ev)
                    self._fused_reduce(dt, gdt, first, n, grads, self._inv_world_size if fold else 1.0)
            elif self._cuda:
                ev = torch.cuda.Event()
                ev.record()
                with torch.cuda.stream(self._comm_stream):
                    self._comm_stream.wait_event(ev)
                    dist.all_reduce(span, group=self.process_group, async_op=True).wait()
                    behind()
            else:
                dist.all_reduce(span, group=self.process_group)
                behind()
            self._stats["allreduces"] += 1
            i = j

    def _launch_local_run(self, b0: int, b1: int) -> None:
        """world_size 1: ONE norm launch over a run of consecutive buckets (split where the gradient
        dtype changes): per-tensor sums of squares of the gradients in place (ptr[1] = grad)."""
        k = self._kernels
        tab = self._unpack_table
        bks = self._buckets[b0:b1]
        i = 0
        while i < len(bks):
            gdt = None
            j = i
            ptrs = []
            while j < len(bks):
                gs = []
                for p in bks[j].params:
                    g = p.grad
                    if g is None:
                        raise RuntimeError("ddp_b200: gradient missing for a parameter marked ready")
                    if not g.is_contiguous():
                        g = g.contiguous()
                        p.grad = g
                    gs.append(g)
                dts = {g.dtype for g in gs}
                if len(dts) != 1:
                    raise RuntimeError("ddp_b200: gradients of one bucket must share a dtype")
                dt = dts.pop()
                if gdt is None:
                    gdt = dt
                elif dt != gdt:
                    break
                ptrs.extend(g.data_ptr() for g in gs)
                j += 1
            first = bks[i].first
            n = len(ptrs)
            tab.host["ptr"][first:first + n, 1] = ptrs
            k.sumsq(tab, gdt, self._sumsq[first:first + n], first=first, count=n, prefer_device=False)
            self._stats["norm_launches"] += 1
            i = j

    def _launch_bucket(self, bk: _Bucket) -> None:
        k = self._kernels
        n, first = bk.n, bk.first
        grads = []
        for p in bk.params:
            g = p.grad
            if g is None:
                raise RuntimeError("ddp_b200: gradient missing for a parameter marked ready")
            if not g.is_contiguous():
                g = g.contiguous()
                p.grad = g
            grads.append(g)
        gdt = grads[0].dtype
        for g in grads:
            if g.dtype != gdt:
                raise RuntimeError("ddp_b200: gradients of one bucket must share a dtype")
        hooked = self._comm_hook is not None
        # no hook: pre-divide, then SUM (div_factor_, reducer.hpp:497-500). With a hook the hook divides.
        scale = 1.0 if hooked else self._inv_world_size
        fold = self._folds_scale(bk.dtype)
        if not (fold and gdt == bk.dtype and self._aliased(first, grads)):
            self._pack_table.host["ptr"][first:first + n, 0] = [g.data_ptr() for g in grads]
            k.pack(self._pack_table, gdt, bk.dtype, 1.0 if fold else scale, first=first, count=n, prefer_device=False)
            self._stats["pack_launches"] += 1
        if hooked:
            gb = GradBucket(bk.index, bk.buffer, list(bk.views), list(bk.params), bk.index == len(self._buckets) - 1)
            bk.future = self._comm_hook(self._comm_hook_state, gb)
            self._stats["allreduces"] += 1
            return
        if self._fused:
            ev = torch.cuda.Event()
            ev.record()
            with torch.cuda.stream(self._comm_stream):
                self._comm_stream.wait_event(ev)
                self._fused_reduce(bk.dtype, gdt, first, n, grads, scale if fold else 1.0)
        elif self._cuda:
            ev = torch.cuda.Event()
            ev.record()
            with torch.cuda.stream(self._comm_stream):
                self._comm_stream.wait_event(ev)
                work = dist.all_reduce(bk.buffer, group=self.process_group, async_op=True)
                work.wait()          # stream-level: the side stream waits for NCCL's stream
                self._after_reduce(bk, grads)
        else:
            dist.all_reduce(bk.buffer, group=self.process_group)
            self._after_reduce(bk, grads)
        self._stats["allreduces"] += 1

    def _fused_reduce(self, dt, gdt, first: int, n: int, grads: List[torch.Tensor], scale: float = 1.0) -> None:
        """Rows [first, first+n) of the arena: SUM over ranks in place + per-tensor sums of squares in
        one kernel over NVLink peer memory (csrc/fused_allreduce.cu), then the copy back into p.grad
        unless gradients are bucket views."""
        k = self._kernels
        k.allreduce_sumsq(self._view_table, dt, self._sym[dt], self._sumsq[first:first + n], first=first, count=n,
                          scatter=self._zero, scale=scale)
        if scale != 1.0:
            self._stats["folded_scale_launches"] = self._stats.get("folded_scale_launches", 0) + 1
        self._stats["fused_allreduces"] = self._stats.get("fused_allreduces", 0) + 1
        if not self.gradient_as_bucket_view:
            self._unpack_table.host["ptr"][first:first + n, 1] = [g.data_ptr() for g in grads]
            k.pack(self._unpack_table, dt, gdt, 1.0, first=first, count=n, prefer_device=False)
            self._stats["unpack_launches"] += 1

    def _after_reduce(self, bk: _Bucket, grads: List[torch.Tensor]) -> None:
        """Behind the allreduce, on the same stream: per-tensor sum of squares of the averaged
        gradients, and (unless grads are bucket views) the copy back into p.grad
        (finalize_bucket_dense / copy_bucket_to_grad, reducer.hpp:328-334)."""
        k = self._kernels
        n = bk.n
        if self.gradient_as_bucket_view:
            k.sumsq(self._view_table, bk.dtype, bk.sumsq, first=bk.first, count=n)
            self._stats["norm_launches"] += 1
        else:
            self._unpack_table.host["ptr"][bk.first:bk.first + n, 1] = [g.data_ptr() for g in grads]
            # unpack and norm in one pass over the bucket
            k.pack(self._unpack_table, bk.dtype, grads[0].dtype, 1.0, first=bk.first, count=n,
                   sumsq_out=bk.sumsq, prefer_device=False)
            self._stats["unpack_launches"] += 1

    def _finalize_backward(self) -> None:
        """End of backward (Reducer::finalize_backward, reducer.hpp:286)."""
        if not self._expect_hooks:
            return
        self._expect_hooks = False
        self._callback_queued = False
        if self._next_bucket != len(self._buckets):
            missing = [pi for bk in self._buckets[self._next_bucket:] for pi in bk.spec.indices
                       if self._params[pi].grad is None]
            raise RuntimeError(
                "Expected to have finished reduction in the prior iteration before starting a new one: "
                f"{len(missing)} parameter(s) received no gradient (indices {missing[:8]}...). "
                "ddp_b200 does not cover find_unused_parameters=True.")
        if self._comm_hook is not None:
            for bk in self._buckets:
                result = bk.future.wait()
                if isinstance(result, (list, tuple)):
                    result = result[0]
                if result.data_ptr() != bk.buffer.data_ptr():
                    bk.buffer.copy_(result.view_as(bk.buffer))      # populate_bucket_views_out, reducer.hpp:323-326
                self._after_reduce(bk, [self._params[pi].grad for pi in bk.spec.indices])
        elif self._cuda and self.world_size > 1:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
        if self.gradient_as_bucket_view and not self._local_only:
            for bk in self._buckets:
                for pi, view in zip(bk.spec.indices, bk.views):
                    p = self._params[pi]
                    if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                        p.grad = view
        self._sumsq_valid = True
        self._stats["iterations"] += 1

    # ---- beyond torch: feed gradients by hand, and the overlapped global norm ---------------------
    def reduce_gradients(self, grads: Optional[List[torch.Tensor]] = None,
                         order: Optional[List[int]] = None,
                         before_bucket: Optional[Callable[[int, List[int]], None]] = None) -> None:
        """Run one gradient synchronisation outside autograd: optionally install ``grads`` (one per
        parameter, registration order) as ``p.grad``, then mark parameters ready in ``order``
        (default: the current bucket order) and finalize. Same scheduler as a real backward.
        ``before_bucket(bucket_index, parameter_indices)`` is called on the host right before each
        bucket is launched (e.g. to make the stream wait for that bucket's gradients to arrive);
        without it, buckets that need no exchange are coalesced into one launch."""
        self._rebuild_buckets()
        self._prepare_for_backward()
        if grads is not None:
            for p, g in zip(self._params, grads):
                p.grad = g
        if order is None and self._has_rebuilt:
            # every gradient is already there: walk the buckets in order, no per-parameter bookkeeping
            for bk in self._buckets:
                bk.pending = 0
            if before_bucket is None:
                self._coalesce = True           # every rank takes this branch with every bucket complete
                try:
                    self._launch_buckets(0, len(self._buckets))
                finally:
                    self._coalesce = False
            else:
                for b, bk in enumerate(self._buckets):
                    before_bucket(b, bk.spec.indices)
                    self._launch_buckets(b, b + 1)
            self._next_bucket = len(self._buckets)
        else:
            for pi in (order if order is not None else self._order):
                self._on_grad_ready(pi, False)
        self._finalize_backward()

    def clip_grad_norm_(self, max_norm: float, norm_type: float = 2.0, error_if_nonfinite: bool = False,
                        *, fused_optimizer=None, torch_rounding: bool = True) -> torch.Tensor:
        """``torch.nn.utils.clip_grad_norm_`` over this module's parameters
        (TORCH/nn/utils/clip_grad.py:186-233) using the sums of squares that were taken behind each
        bucket's allreduce during backward: what is left is one tiny finalize kernel, plus the
        in-place scale unless ``fused_optimizer`` takes the coefficient into its update kernel."""
        if float(norm_type) != 2.0:
            raise NotImplementedError("ddp_b200: L2 norm only")
        if not self._sumsq_valid:
            raise RuntimeError("ddp.clip_grad_norm_ needs a completed synchronised backward; "
                               "use ddp_b200.clip_grad_norm_(parameters, ...) otherwise")
        k = self._kernels
        all_bf16 = all(p.dtype == torch.bfloat16 for p in self._params)      # the GRADIENT dtype decides, as in torch
        out = k.norm_finalize(self._sumsq, float(max_norm), torch_bf16=(torch_rounding and all_bf16),
                              out=self._norm_out)
        if error_if_nonfinite and not bool(torch.isfinite(out[0])):
            raise RuntimeError("The total norm for gradients from `parameters` is non-finite, so it cannot be clipped.")
        coef = out[1:2]
        if fused_optimizer is not None:
            fused_optimizer.set_clip_coef(coef)
        elif self.gradient_as_bucket_view and not self._local_only:
            for bk in self._buckets:
                k.scale_by(self._view_table, bk.dtype, coef, first=bk.first, count=len(bk.spec.indices))
        else:
            by_dt = {}
            for p in self._params:
                by_dt.setdefault(p.grad.dtype, []).append(p.grad)
            for dt, gl in by_dt.items():
                tab = k.table([None, gl, None, None], [g.numel() for g in gl])
                k.scale_by(tab, dt, coef, prefer_device=False)
        self._sumsq_valid = fused_optimizer is not None   # grads changed in place otherwise
        return out[0].to(self._params[self._order[0]].dtype)
