"""clip_grad_norm_ on the sm_100a single-pass multi-tensor L2-norm kernel.

Drop-in for ``torch.nn.utils.clip_grad_norm_`` (TORCH/nn/utils/clip_grad.py:186-233): same
positional arguments, returns the pre-clip total norm as a 0-dim tensor in the gradient dtype,
scales ``p.grad`` in place by ``min(max_norm / (total + 1e-6), 1)`` with no host sync
(clip_grad.py:165-174). One read of the gradients for the norm (torch: one read) and, unless the
multiply is folded into the optimizer, one read+write for the scale (torch: same).

``fused_optimizer=opt`` hands the coefficient to ``ddp_b200.AdamW`` instead of scaling the
gradients here: the next ``opt.step()`` applies it inside the update kernel, saving the
read+write sweep. In that mode ``p.grad`` is left unscaled — an opt-in deviation from torch.
"""
from __future__ import annotations

import os
from typing import Dict, Iterable, List, Optional, Tuple, Union

import torch

from .kernels import default_kernels


def clip_grad_norm_(parameters: Union[torch.Tensor, Iterable[torch.Tensor]], max_norm: float,
                    norm_type: float = 2.0, error_if_nonfinite: bool = False,
                    foreach: Optional[bool] = None, *, fused_optimizer=None, kernels=None,
                    torch_rounding: bool = True) -> torch.Tensor:
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    grads = [p.grad for p in parameters if p.grad is not None]
    if float(norm_type) != 2.0:
        raise NotImplementedError("ddp_b200.clip_grad_norm_ implements the L2 norm of the DDP hot path only")
    if len(grads) == 0:
        return torch.tensor(0.0)
    k = kernels or default_kernels()
    device = grads[0].device
    groups: Dict[torch.dtype, List[torch.Tensor]] = {}
    for g in grads:
        if g.device != device:
            raise RuntimeError("ddp_b200.clip_grad_norm_: one process drives one GPU; gradients span devices")
        if not g.is_contiguous():
            raise RuntimeError("ddp_b200.clip_grad_norm_ needs contiguous gradients")
        groups.setdefault(g.dtype, []).append(g)

    # default backend: through the registered torch.library ops (ops.py) — the same C-ABI calls, by name
    use_ops = kernels is None and device.type == "cuda" and os.environ.get("DDPB200_USE_OPS", "1") == "1"
    sumsq = torch.empty(len(grads), dtype=torch.float32, device=device)
    tables: List[Tuple] = []
    pos = 0
    for dt, gl in groups.items():
        if use_ops:
            torch.ops.ddp_b200.l2norm_sumsq(gl, sumsq[pos:pos + len(gl)])
            tables.append((gl, dt))
        else:
            tab = k.table([None, gl, None, None], [g.numel() for g in gl])
            k.sumsq(tab, dt, sumsq[pos:pos + len(gl)], prefer_device=False)
            tables.append((tab, dt))
        pos += len(gl)
    all_bf16 = all(dt == torch.bfloat16 for dt in groups)
    if use_ops:
        out = torch.ops.ddp_b200.l2norm_finalize(sumsq, float(max_norm), bool(torch_rounding and all_bf16))
    else:
        out = k.norm_finalize(sumsq, float(max_norm), torch_bf16=(torch_rounding and all_bf16))

    if error_if_nonfinite and not bool(torch.isfinite(out[0])):
        raise RuntimeError(
            f"The total norm of order {float(norm_type)} for gradients from `parameters` is non-finite, "
            "so it cannot be clipped. To disable this error and scale the gradients by the non-finite "
            "norm anyway, set `error_if_nonfinite=False`")

    coef = out[1:2]
    if fused_optimizer is not None:
        fused_optimizer.set_clip_coef(coef)
    else:
        for tab, dt in tables:
            if use_ops:
                torch.ops.ddp_b200.scale_by_(tab, coef)
            else:
                k.scale_by(tab, dt, coef, prefer_device=False)
    return out[0].to(grads[0].dtype)
