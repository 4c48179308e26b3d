"""AdamW with the update done by ONE fused sm_100a kernel launch per (dtype) group.

Drop-in for ``torch.optim.AdamW`` (TORCH/optim/adamw.py:20-49; step logic TORCH/optim/adam.py:
215-272): same constructor arguments, ``step`` / ``zero_grad`` / ``state_dict`` /
``load_state_dict`` and the same per-parameter state (``step``, ``exp_avg``, ``exp_avg_sq``).
The element arithmetic follows torch's fused functor (TORCH/include/ATen/native/cuda/
fused_adam_utils.cuh:41-93); see csrc/adamw.cu.

Beyond torch: the gradient-clipping multiply of ``clip_grad_norm_`` can be folded into the same
pass (``clip_grad_norm_(..., fused_optimizer=opt)`` or ``step(clip_coef=...)``), which removes a
read+write sweep over all gradients.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
from torch.optim.optimizer import Optimizer

from .kernels import SegTable, default_kernels


def adamw_chunk_elems(total_numel: int) -> int:
    """Chunk length of the update kernel's table. Short chunks keep the persistent grid's CTAs on one
    contiguous window of each stream (CTA b takes chunks b, b+grid, ...), which DRAM likes: measured on the
    Llama-3.2-1B set, 400 launches back to back, bf16 2.763 ms at 4096 elements against 2.816 / 2.825 /
    2.92 at 8192 / 16384 / 65536, fp32 5.337 against 5.377 at 65536 (profiles/r02_adamw_chunk_sweep.txt);
    2048 pays more for the per-chunk bookkeeping than it gains."""
    from . import _lib
    from .kernels import CudaKernels
    return min(4096, int(_lib.lib().ddpb200_suggest_chunk_elems(int(total_numel), CudaKernels.sm_count())))


class AdamW(Optimizer):
    # GradScaler hands grad_scale / found_inf to optimizers that declare this (TORCH/amp/grad_scaler.py)
    _step_supports_amp_scaling = True

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, amsgrad=False,
                 *, maximize: bool = False, foreach: Optional[bool] = None, capturable: bool = False,
                 differentiable: bool = False, fused: Optional[bool] = None,
                 state_dtype: Optional[torch.dtype] = None, kernels=None):
        if isinstance(lr, torch.Tensor) and lr.numel() != 1:
            raise ValueError("Tensor lr must be 1-element")
        if not 0.0 <= (float(lr) if not isinstance(lr, torch.Tensor) else float(lr)):
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if not 0.0 <= weight_decay:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if amsgrad:
            raise NotImplementedError("ddp_b200.AdamW: amsgrad=True is outside the DDP hot path this package covers")
        if differentiable:
            raise NotImplementedError("ddp_b200.AdamW: differentiable=True is not supported by the fused kernel")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad,
                        maximize=maximize, foreach=foreach, capturable=capturable,
                        differentiable=differentiable, fused=fused, decoupled_weight_decay=True)
        super().__init__(params, defaults)
        self._kernels = kernels
        # default backend on CUDA: go through the registered torch.library op (DDPB200_USE_OPS=0: straight to ctypes)
        self._use_ops = kernels is None and os.environ.get("DDPB200_USE_OPS", "1") == "1"
        self._state_dtype = state_dtype
        self._clip_coef: Optional[torch.Tensor] = None
        self._dev_steps_live = False
        self._tables: Dict[Tuple, Tuple[np.ndarray, np.ndarray, SegTable]] = {}
        self._plan: Dict[int, dict] = {}
        self.round_clipped_grad = True   # reproduce torch's rounding of g*coef to the grad dtype
        self.launches_last_step = 0

    # ---- clip folding -------------------------------------------------------------------------
    def set_clip_coef(self, coef: Optional[torch.Tensor]) -> None:
        """Device fp32 scalar to multiply every gradient by inside the next ``step`` (then cleared)."""
        self._clip_coef = coef

    def _k(self):
        if self._kernels is None:
            self._kernels = default_kernels()
        return self._kernels

    def _init_state(self, p: torch.Tensor) -> dict:
        st = self.state[p]
        if len(st) == 0:
            sdt = self._state_dtype or p.dtype
            st["step"] = torch.tensor(0.0, dtype=torch.float32)          # adam.py:168-176 (CPU scalar)
            st["exp_avg"] = torch.zeros_like(p, dtype=sdt, memory_format=torch.preserve_format)     # :178-180
            st["exp_avg_sq"] = torch.zeros_like(p, dtype=sdt, memory_format=torch.preserve_format)  # :182-184
        return st

    @torch.no_grad()
    def step(self, closure=None, *, clip_coef: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if clip_coef is None:
            clip_coef = self._clip_coef
        self._clip_coef = None
        grad_scale = getattr(self, "grad_scale", None)
        found_inf = getattr(self, "found_inf", None)
        # GradScaler: whether this step happens is only known on the device. No host sync (torch's fused path
        # has none either, adam.py:869-876): the kernel reads found_inf and does nothing if it is set, and the
        # step count used for the bias corrections lives on the device for as long as a scaler drives us.
        amp = found_inf is not None and found_inf.is_cuda
        if found_inf is not None and not amp and float(found_inf) == 1.0:
            return loss
        if not amp and self._dev_steps_live:
            self._sync_steps()
        k = self._k()
        self.launches_last_step = 0

        for gi, group in enumerate(self.param_groups):
            beta1, beta2 = group["betas"]
            with_grad = [p for p in group["params"] if p.grad is not None]
            if not with_grad:
                continue
            for li, launch in enumerate(self._partition(gi, with_grad)):
                key, plist, steps = launch[0], launch[1], launch[2]
                device, pdt, gdt, sdt = key
                if amp:
                    if launch[6] is None:
                        launch[6] = torch.full((1,), float(launch[3]), dtype=torch.float32, device=device)
                        self._dev_steps_live = True
                    dev_step = launch[6]
                    dev_step.add_(1.0)                # state_steps += 1 ... (adam.py fused path)
                    tab = self._table_for((gi, li) + key, plist, device)
                    k.adamw(tab, pdt, gdt, sdt, lr=group["lr"], beta1=beta1, beta2=beta2, eps=group["eps"],
                            weight_decay=group["weight_decay"], step=max(1, launch[3]), clip_coef=clip_coef,
                            grad_scale=grad_scale, found_inf=found_inf, round_clipped=self.round_clipped_grad,
                            store_grad=grad_scale is not None, maximize=group["maximize"], step_tensor=dev_step)
                    dev_step.sub_(found_inf.to(torch.float32).reshape(1))   # ... -= found_inf: a skipped step does not count
                    self.launches_last_step += 1
                    continue
                torch._foreach_add_(steps, 1.0)       # the per-parameter `step` scalars (CPU), one call
                launch[3] += 1
                if self._use_ops and device.type == "cuda":
                    # through the dispatcher: torch.ops.ddp_b200.fused_clip_adamw_ (ops.py), the registered form
                    # of the same C-ABI call — visible to torch.profiler / torch.compile by name
                    lr = group["lr"]
                    lr_t = lr if isinstance(lr, torch.Tensor) and lr.is_cuda else None
                    torch.ops.ddp_b200.fused_clip_adamw_(
                        plist, [p.grad for p in plist], launch[4], launch[5], 0.0 if lr_t is not None else float(lr),
                        beta1, beta2, group["weight_decay"], group["eps"], launch[3], group["maximize"],
                        self.round_clipped_grad, grad_scale is not None, clip_coef, grad_scale, None, lr_t)
                else:
                    tab = self._table_for((gi, li) + key, plist, device)
                    k.adamw(tab, pdt, gdt, sdt, lr=group["lr"], beta1=beta1, beta2=beta2, eps=group["eps"],
                            weight_decay=group["weight_decay"], step=launch[3], clip_coef=clip_coef,
                            grad_scale=grad_scale, found_inf=None, round_clipped=self.round_clipped_grad,
                            store_grad=grad_scale is not None, maximize=group["maximize"])
                self.launches_last_step += 1
        return loss

    def _partition(self, gi: int, with_grad: List[torch.Tensor]):
        """Split the parameters that have a gradient into launches, one per (device, param dtype,
        grad dtype, state dtype, step count): a list of [key, params, step tensors, step]. The split
        is remembered; while the same parameters show up with the same gradient dtypes (every
        ordinary training step) it is reused without touching per-parameter state again."""
        ids = [id(p) for p in with_grad]
        ent = self._plan.get(gi)
        if ent is not None and ent[0] != ids and self._dev_steps_live:
            self._sync_steps()              # the split changes: bring the CPU step scalars up to date first
        if ent is not None and ent[0] == ids and all(L[1][0].grad.dtype == L[0][2] for L in ent[1]):
            return ent[1]
        groups: Dict[Tuple, List[torch.Tensor]] = {}
        for p in with_grad:
            if p.grad.is_sparse:
                raise RuntimeError("AdamW does not support sparse gradients")
            if not (p.is_contiguous() and p.grad.is_contiguous()):
                raise RuntimeError("ddp_b200.AdamW needs contiguous parameters and gradients")
            st = self._init_state(p)
            groups.setdefault((p.device, p.dtype, p.grad.dtype, st["exp_avg"].dtype, int(st["step"])), []).append(p)
        launches = [[k5[:4], plist, [self.state[p]["step"] for p in plist], k5[4],
                     [self.state[p]["exp_avg"] for p in plist], [self.state[p]["exp_avg_sq"] for p in plist], None]
                    for k5, plist in groups.items()]
        self._plan[gi] = (ids, launches)
        return launches

    def _sync_steps(self) -> None:
        """Device-resident step counts (GradScaler flow) back into the per-parameter CPU `step` scalars that
        `state_dict()` exposes. One device-to-host read per launch group; called at checkpoints and when the
        scaler stops driving the optimizer, never inside a scaled step."""
        for ent in self._plan.values():
            for launch in ent[1]:
                if launch[6] is not None:
                    v = float(launch[6])
                    launch[3] = int(v)
                    for st in launch[2]:
                        st.fill_(v)
                    launch[6] = None
        self._dev_steps_live = False

    def state_dict(self):
        if self._dev_steps_live:
            self._sync_steps()
        return super().state_dict()

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        self._plan.clear()          # step counts and state tensors were replaced
        self._tables.clear()
        self._dev_steps_live = False

    def _table_for(self, key, plist: List[torch.Tensor], device) -> SegTable:
        """Segment table for this group. Parameter / state pointers never move; gradient pointers
        do when autograd allocates fresh grads each step. A table whose pointers repeat is kept on
        the device; otherwise it rides in the launch parameters (no copy, no lifetime)."""
        n = len(plist)
        pptr = np.fromiter((p.data_ptr() for p in plist), dtype=np.uint64, count=n)
        gptr = np.fromiter((p.grad.data_ptr() for p in plist), dtype=np.uint64, count=n)
        ent = self._tables.get(key)
        if ent is not None and ent[2].n == n and np.array_equal(ent[0], pptr):
            tab = ent[2]
            if np.array_equal(ent[1], gptr):
                if tab.dev is None and device.type == "cuda":
                    tab.to_device(device)     # second sighting of the same pointers: make it resident
                return tab
            tab.host["ptr"][:, 1] = gptr
            tab.invalidate_device()
            self._tables[key] = (pptr, gptr, tab)
            return tab
        states = [self.state[p] for p in plist]
        numels = [p.numel() for p in plist]
        tab = self._k().table([plist, [p.grad for p in plist], [s["exp_avg"] for s in states],
                               [s["exp_avg_sq"] for s in states]], numels, chunk_elems=adamw_chunk_elems(sum(numels)))
        self._tables[key] = (pptr, gptr, tab)
        return tab
