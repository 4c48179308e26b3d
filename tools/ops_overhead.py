#!/usr/bin/env python
"""Host cost of one AdamW.step() + clip_grad_norm_ on the Llama-3.2-1B parameter set (146 tensors): through the
registered torch.library ops (default) against the direct ctypes calls (DDPB200_USE_OPS=0). Wall time of the
Python calls only (launches are asynchronous; the device is synchronised outside the timed part)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
import ddp_b200

dev = torch.device("cuda", 0)
shapes = bench.llama_param_shapes(**bench.WORKLOADS["tiny" if len(sys.argv) > 1 else "llama-3.2-1b"])
ps = [torch.nn.Parameter(torch.empty(s, dtype=torch.bfloat16, device=dev).normal_(0, 0.02)) for s in shapes]
for p in ps:
    p.grad = torch.empty_like(p).normal_(0, 0.01)
for use in ("1", "0"):
    os.environ["DDPB200_USE_OPS"] = use
    opt = ddp_b200.AdamW(ps, lr=1e-4)
    for _ in range(3):
        ddp_b200.clip_grad_norm_(ps, 1.0, fused_optimizer=opt)
        opt.step()
    torch.cuda.synchronize()
    reps = 20
    host = 0.0
    for _ in range(reps):
        t0 = time.perf_counter()
        ddp_b200.clip_grad_norm_(ps, 1.0, fused_optimizer=opt)
        opt.step()
        host += time.perf_counter() - t0
        torch.cuda.synchronize()
    print(f"DDPB200_USE_OPS={use}: host time of clip_grad_norm_ + step = {host / reps * 1e3:.3f} ms per call ({len(ps)} tensors)")
