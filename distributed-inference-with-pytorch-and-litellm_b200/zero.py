"""Sharded AdamW (ZeRO-1 style) on top of the fused peer-memory collective.

The natural extension SURVEY.md §8e names, NOT the drop-in: optimizer state is split across the ranks
by chunk, so ``state_dict`` is rank-local. What it buys: per GPU the update touches 1/world of the
parameter / state bytes and keeps 1/world of the state, while NVLink carries what an allreduce would.

    ddp = DistributedDataParallel(model, device_ids=[r], gradient_as_bucket_view=True,
                                  fused_allreduce=True, shard_optimizer=True)
    opt = ShardedAdamW(ddp, lr=5e-5)
    loss.backward()                                  # per bucket: pack -> reduce-scatter + norm partials
    ddp.clip_grad_norm_(1.0, fused_optimizer=opt)    # total norm, identical bits on every rank
    opt.step()                                       # owners update their chunks, write params everywhere

Parameters after every step are bit-identical to the replicated path (``ddp_b200.AdamW``): the element
arithmetic is the same function applied to the same rounded gradients. After ``backward`` only the
chunks a rank owns hold averaged gradients in ``p.grad`` (reduce-scatter semantics).
"""
from __future__ import annotations

from typing import Optional

import torch


class ShardedAdamW:
    def __init__(self, ddp, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, *, maximize: bool = False):
        if not getattr(ddp, "_zero", False):
            raise RuntimeError("ShardedAdamW needs DistributedDataParallel(..., fused_allreduce=True, shard_optimizer=True)")
        self.ddp = ddp
        self.defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, maximize=maximize)
        self.param_groups = [dict(self.defaults, params=list(ddp._params))]
        self.step_count = 0
        self._clip_coef: Optional[torch.Tensor] = None
        self.round_clipped_grad = True
        self.launches_last_step = 0
        tab = ddp._view_table
        dt = ddp._buckets[0].dtype
        slots = (tab.total_chunks + ddp.world_size - 1) // ddp.world_size + 1
        # this rank's 1/world of exp_avg / exp_avg_sq: one chunk_elems-sized slot per owned chunk
        self.exp_avg = torch.zeros(slots * tab.chunk_elems, dtype=dt, device=ddp.device)
        self.exp_avg_sq = torch.zeros(slots * tab.chunk_elems, dtype=dt, device=ddp.device)

    def set_clip_coef(self, coef: Optional[torch.Tensor]) -> None:
        self._clip_coef = coef

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.ddp._params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    @torch.no_grad()
    def step(self, closure=None, *, clip_coef: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        ddp = self.ddp
        if not ddp._sumsq_valid and ddp._stats["iterations"] == 0:
            raise RuntimeError("ShardedAdamW.step() needs a completed synchronised backward")
        if clip_coef is None:
            clip_coef = self._clip_coef
        self._clip_coef = None
        g = self.param_groups[0]
        self.step_count += 1
        dt = ddp._buckets[0].dtype
        ddp._kernels.sharded_adamw(ddp._view_table, dt, ddp._sym[dt], ddp._param_sym[dt], self.exp_avg, self.exp_avg_sq,
                                   lr=g["lr"], beta1=g["betas"][0], beta2=g["betas"][1], eps=g["eps"],
                                   weight_decay=g["weight_decay"], step=self.step_count, clip_coef=clip_coef,
                                   round_clipped=self.round_clipped_grad, maximize=g["maximize"])
        self.launches_last_step = 1
        return loss

    def state_dict(self) -> dict:
        """Rank-local: this rank's shards and the shared step count."""
        return {"sharded": True, "rank": self.ddp.rank, "world_size": self.ddp.world_size, "step": self.step_count,
                "chunk_elems": self.ddp._view_table.chunk_elems, "exp_avg_shard": self.exp_avg,
                "exp_avg_sq_shard": self.exp_avg_sq, "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}]}

    def consolidated_state_dict(self) -> dict:
        """COLLECTIVE: gather every rank's shards and return the state in ``torch.optim.AdamW``'s
        format (per parameter ``step`` / ``exp_avg`` / ``exp_avg_sq``, in DDP registration order), so a
        sharded run can be checkpointed into, or compared with, the replicated optimizer. Chunk a
        of the table lives on rank a % world at slot a // world."""
        import torch.distributed as dist
        ddp = self.ddp
        ws, tab = ddp.world_size, ddp._view_table
        ce = tab.chunk_elems
        gathered = {}
        for name, shard in (("exp_avg", self.exp_avg), ("exp_avg_sq", self.exp_avg_sq)):
            parts = [torch.empty_like(shard) for _ in range(ws)]
            dist.all_gather(parts, shard, group=ddp.process_group)
            gathered[name] = parts
        state = {}
        for row, pi in enumerate(ddp._order):
            p = ddp._params[pi]
            n = p.numel()
            first_chunk = int(tab.chunk_begins[row])
            ent = {"step": torch.tensor(float(self.step_count))}
            for name in ("exp_avg", "exp_avg_sq"):
                out = torch.empty(n, dtype=self.exp_avg.dtype, device=p.device)
                for k in range(int(tab.chunk_begins[row + 1]) - first_chunk):
                    a = first_chunk + k
                    ln = min(ce, n - k * ce)
                    out[k * ce:k * ce + ln] = gathered[name][a % ws][(a // ws) * ce:(a // ws) * ce + ln]
                ent[name] = out.view_as(p)
            state[pi] = ent
        g = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        g.update(amsgrad=False, foreach=None, capturable=False, differentiable=False, fused=None,
                 decoupled_weight_decay=True, params=list(range(len(ddp._params))))
        return {"state": {i: state[i] for i in sorted(state)}, "param_groups": [g]}

    def load_state_dict(self, sd: dict) -> None:
        if not sd.get("sharded") or sd["world_size"] != self.ddp.world_size or sd["rank"] != self.ddp.rank:
            raise ValueError("ShardedAdamW.load_state_dict: shard of another rank / world size")
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg_shard"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq_shard"])
