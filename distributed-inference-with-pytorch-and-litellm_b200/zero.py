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
from torch.optim.optimizer import Optimizer


class ShardedAdamW(Optimizer):
    """A ``torch.optim.Optimizer`` (LR schedulers, ``param_groups``, ``zero_grad``, ``isinstance`` checks all
    work) whose per-parameter state is NOT in ``self.state``: each rank holds its 1/world of ``exp_avg`` /
    ``exp_avg_sq`` as two flat shards laid out by chunk. One parameter group (the update is one launch over
    the whole arena). ``GradScaler`` is not supported: after a reduce-scatter a rank only sees the averaged
    gradients of the chunks it owns, so an inf check would need its own collective; bf16 training, which
    this path is for, does not scale."""

    def __init__(self, ddp, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, *, maximize: bool = False):
        if not getattr(ddp, "_zero", False):
            raise RuntimeError("ShardedAdamW needs DistributedDataParallel(..., fused_allreduce=True, shard_optimizer=True)")
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= weight_decay or not (0.0 <= betas[0] < 1.0 and 0.0 <= betas[1] < 1.0):
            raise ValueError("ShardedAdamW: invalid hyper-parameter")
        self.ddp = ddp
        super().__init__(list(ddp._params), dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, maximize=maximize))
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

    def add_param_group(self, param_group) -> None:
        if getattr(self, "param_groups", None):
            raise NotImplementedError("ShardedAdamW updates the whole bucket arena in one launch: one parameter group")
        super().add_param_group(param_group)

    def set_clip_coef(self, coef: Optional[torch.Tensor]) -> None:
        self._clip_coef = coef

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
        """Either this rank's own ``state_dict()``, or a state dict in ``torch.optim.AdamW``'s format — what
        ``consolidated_state_dict()`` returns and what ``torch.optim.AdamW`` / ``ddp_b200.AdamW`` save: per
        parameter (DDP registration order) ``step`` / ``exp_avg`` / ``exp_avg_sq``. From the latter every rank
        keeps the chunks it owns (chunk a of the table lives on rank a % world at slot a // world); all
        parameters must share one step count (one launch, one bias correction)."""
        ddp = self.ddp
        if sd.get("sharded"):
            if sd["world_size"] != ddp.world_size or sd["rank"] != ddp.rank:
                raise ValueError("ShardedAdamW.load_state_dict: shard of another rank / world size")
            if sd["chunk_elems"] != ddp._view_table.chunk_elems:
                raise ValueError("ShardedAdamW.load_state_dict: saved with another chunk size")
            self.step_count = int(sd["step"])
            self.exp_avg.copy_(sd["exp_avg_shard"])
            self.exp_avg_sq.copy_(sd["exp_avg_sq_shard"])
            for k, v in sd["param_groups"][0].items():
                self.param_groups[0][k] = v
            return
        state, groups = sd["state"], sd["param_groups"]
        if len(groups) != 1 or len(groups[0]["params"]) != len(ddp._params):
            raise ValueError("ShardedAdamW.load_state_dict: expected one parameter group over all DDP parameters")
        ws, rank, tab = ddp.world_size, ddp.rank, ddp._view_table
        ce = tab.chunk_elems
        steps = {int(float(state[k]["step"])) for k in state}
        if len(steps) > 1:
            raise ValueError(f"ShardedAdamW.load_state_dict: parameters at different step counts {sorted(steps)}")
        self.exp_avg.zero_()
        self.exp_avg_sq.zero_()
        ids = groups[0]["params"]
        for row, pi in enumerate(ddp._order):
            ent = state.get(ids[pi], state.get(str(ids[pi])))
            if ent is None:
                continue                    # a parameter that never stepped: zeros
            n = ddp._params[pi].numel()
            first_chunk = int(tab.chunk_begins[row])
            for name, shard in (("exp_avg", self.exp_avg), ("exp_avg_sq", self.exp_avg_sq)):
                src = ent[name].reshape(-1).to(device=shard.device, dtype=shard.dtype)
                for k in range(int(tab.chunk_begins[row + 1]) - first_chunk):
                    a = first_chunk + k
                    if a % ws == rank:
                        ln = min(ce, n - k * ce)
                        shard[(a // ws) * ce:(a // ws) * ce + ln] = src[k * ce:k * ce + ln]
        self.step_count = steps.pop() if steps else 0
        for k, v in groups[0].items():
            if k != "params" and k in self.param_groups[0]:
                self.param_groups[0][k] = v
