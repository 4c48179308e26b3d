#!/usr/bin/env python
"""Whole training step (the row next to the hot path): forward + backward of a
random-init HF causal LM with the gradient sync + clip + AdamW done (a) by ddp_b200 and (b) by stock
torch (DistributedDataParallel + clip_grad_norm_ + AdamW(fused=True)) on the same GPU(s), same
weights, same synthetic token batches. Tokens come from pinned host memory every step and the loss
is read back, so both numbers are end to end. Prints one JSON line per implementation.

    python tools/train_bench.py [--model llama-1b|gpt2|tiny] [--batch B] [--seq S] [--steps K]
    torchrun --nproc-per-node N tools/train_bench.py ...
"""
import argparse
import copy
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "src"))

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-1b")
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours,stock")     # also: ours-fused / ours-nccl (force the bucket collective)
    args = ap.parse_args()
    rank, ws, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    import ddp_b200
    from distributed_finetuning import build_model

    base = build_model(args.model, dev, torch.bfloat16)
    vocab = base.config.vocab_size
    nparam = sum(p.numel() for p in base.parameters())
    g = torch.Generator().manual_seed(1000 + rank)
    host_ids = [torch.randint(0, vocab, (args.batch, args.seq), generator=g).pin_memory()
                for _ in range(args.steps + args.warmup)]
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    for impl in args.impl.split(","):
        model = copy.deepcopy(base)
        if impl in ("ours", "ours-fused", "ours-nccl"):
            # ours = the library's default (fused peer-memory collective where it can be set up, else NCCL)
            ddp = ddp_b200.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True,
                                                   fused_allreduce={"ours": None, "ours-fused": True, "ours-nccl": False}[impl])
            opt = ddp_b200.AdamW(model.parameters(), lr=5e-5)
            clip = lambda: ddp.clip_grad_norm_(1.0, fused_optimizer=opt)                   # noqa: E731
        else:
            ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
            opt = torch.optim.AdamW(model.parameters(), lr=5e-5, fused=True)
            clip = lambda: torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)         # noqa: E731

        def step(i):
            ids = host_ids[i].to(dev, non_blocking=True)
            loss = ddp(input_ids=ids, labels=ids).loss
            loss.backward()
            clip()
            opt.step()
            opt.zero_grad()
            loss_host.copy_(loss.detach().float(), non_blocking=True)

        for i in range(args.warmup):
            step(i)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tail0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        tail1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        e0.record()
        for i in range(args.steps):
            ids = host_ids[args.warmup + i].to(dev, non_blocking=True)
            loss = ddp(input_ids=ids, labels=ids).loss
            loss.backward()
            tail0[i].record()              # from here: what is left of sync + clip + update after backward
            clip()
            opt.step()
            tail1[i].record()
            opt.zero_grad()
            loss_host.copy_(loss.detach().float(), non_blocking=True)
        e1.record()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.steps, sum(a.elapsed_time(b) for a, b in zip(tail0, tail1)) / args.steps],
                         dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            ms, tail = [float(x) for x in t.tolist()]
            print(json.dumps({"impl": impl, "model": args.model, "params": nparam, "n_gpus": ws, "batch_per_gpu": args.batch,
                              "seq": args.seq, "ms_per_step": ms, "steps_per_s": 1e3 / ms,
                              "tokens_per_s": ws * args.batch * args.seq * 1e3 / ms,
                              "ms_after_backward_clip_plus_update": tail, "loss": float(loss_host),
                              "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}), flush=True)
        del ddp, opt, model
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
