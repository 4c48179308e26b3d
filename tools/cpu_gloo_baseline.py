#!/usr/bin/env python
"""BASELINE.json configs[0] / BASELINE.md §4: the reference's CPU path — stock torch
DistributedDataParallel (gloo, world_size 2) + clip_grad_norm_ + torch.optim.AdamW on GPT-2 small
(random init), batch 4 x seq 128 synthetic token ids — timed per phase on the host cores.
No GPU, none of this repo's kernels: a reported baseline.

    python tools/cpu_gloo_baseline.py [--steps 4] [--threads-per-rank N]
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def worker(rank, ws, port, steps, threads, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    torch.set_num_threads(threads)
    from transformers import GPT2Config, GPT2LMHeadModel
    torch.manual_seed(0)
    model = GPT2LMHeadModel(GPT2Config())
    ddp = torch.nn.parallel.DistributedDataParallel(model)
    opt = torch.optim.AdamW(model.parameters(), lr=5e-5)
    n = sum(p.numel() for p in model.parameters())
    phases = {"forward": [], "backward+allreduce": [], "clip_grad_norm_": [], "AdamW.step": [], "total": []}
    for s in range(steps + 1):
        g = torch.Generator().manual_seed(1000 * s + rank)
        ids = torch.randint(0, 50257, (4, 128), generator=g)
        t0 = time.perf_counter()
        loss = ddp(input_ids=ids, labels=ids).loss
        t1 = time.perf_counter()
        loss.backward()
        t2 = time.perf_counter()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        t3 = time.perf_counter()
        opt.step()
        opt.zero_grad()
        t4 = time.perf_counter()
        if s > 0:                                   # step 0 builds buckets / state
            for k, v in zip(phases, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t4 - t0)):
                phases[k].append(v)
    if rank == 0:
        log = ddp._get_ddp_logging_data()
        med = {k: sorted(v)[len(v) // 2] for k, v in phases.items()}
        out.update({"config": "GPT-2 small fp32, DDP gloo ws=2, batch 4 x seq 128, AdamW lr 5e-5, max_norm 1.0",
                    "parameters": n, "host_cpus": os.cpu_count(), "threads_per_rank": threads, "steps": steps,
                    "median_seconds": med, "steps_per_s": 1.0 / med["total"],
                    "clip_GBps_at_12B_per_elem": 12 * n / med["clip_grad_norm_"] / 1e9,
                    "adamw_GBps_at_28B_per_param": 28 * n / med["AdamW.step"] / 1e9,
                    "rebuilt_bucket_sizes": str(log.get("rebuilt_bucket_sizes", "")),
                    "avg_backward_comm_time_ns": log.get("avg_backward_comm_time"),
                    "avg_backward_compute_comm_overlap_time_ns": log.get("avg_backward_compute_comm_overlap_time")})
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--threads-per-rank", type=int, default=max(1, (os.cpu_count() or 2) // 2))
    args = ap.parse_args()
    out = mp.Manager().dict()
    mp.spawn(worker, args=(2, 29791, args.steps, args.threads_per_rank, out), nprocs=2, join=True)
    print(json.dumps(dict(out)))


if __name__ == "__main__":
    main()
