#!/usr/bin/env python
"""BASELINE.json configs[3]: fused AdamW + grad-norm microbench, 10 M - 8 B fp32 parameters on one
B200, against torch.optim.AdamW(fused=True) and torch's clip_grad_norm_ on the same tensors.
Values: p ~ N(0, 0.02^2), g ~ N(0, 0.01^2), m = v = 0 then warmed 3 steps.
Flat single tensors, and the Llama-3.1-8B parameter list (291 tensors, 8.03e9 elements) as the
realistic multi-tensor case. CUDA events, 10 reps after warm-up; L2 is not flushed (working sets
>> 126 MB except at 10 M, which is marked)."""
import gc
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
import ddp_b200


def timeit(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def run_case(K, dev, peak, name, numel):
    n = sum(numel)
    DT = torch.bfloat16 if os.environ.get("SWEEP_DTYPE", "f32") == "bf16" else torch.float32
    B = 14 if DT == torch.bfloat16 else 28            # algorithmic bytes per parameter of the update
    gen = torch.Generator(device=dev).manual_seed(0)
    p = torch.empty(n, dtype=DT, device=dev).normal_(0, 0.02, generator=gen)
    g = torch.empty(n, dtype=DT, device=dev).normal_(0, 0.01, generator=gen)
    m = torch.zeros(n, dtype=DT, device=dev)
    v = torch.zeros(n, dtype=DT, device=dev)

    def views(t):
        o, r = 0, []
        for k in numel:
            r.append(t[o:o + k])
            o += k
        return r

    tab = K.table([views(p), views(g), views(m), views(v)], numel, chunk_elems=ddp_b200.optim.adamw_chunk_elems(n), device=dev)
    ntab = K.table([None, views(g), None, None], numel, device=dev)        # the norm keeps the long chunks
    ss = torch.empty(len(numel), dtype=torch.float32, device=dev)
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2)
    step = [0]

    def ours():
        step[0] += 1
        K.sumsq(ntab, DT, ss)
        coef = K.norm_finalize(ss, 1.0, DT == torch.bfloat16)
        K.adamw(tab, DT, DT, DT, step=step[0], clip_coef=coef[1:2], **hp)

    def ours_adam_only():
        step[0] += 1
        K.adamw(tab, DT, DT, DT, step=step[0], **hp)

    t_ours = timeit(ours)
    t_adam = timeit(ours_adam_only)
    del tab, ntab, m, v                     # torch's optimizer allocates its own state: make room (8 B: 64 GB)
    gc.collect()
    torch.cuda.empty_cache()
    ps = [torch.nn.Parameter(x) for x in views(p)]
    for q, gg in zip(ps, views(g)):
        q.grad = gg
    opt = torch.optim.AdamW(ps, lr=1e-3, fused=True)
    # note: torch's clip scales grads in place; restore is unnecessary for timing
    t_torch_adam = timeit(opt.step)

    def torch_both():
        torch.nn.utils.clip_grad_norm_(ps, 1.0)
        opt.step()

    t_torch = timeit(torch_both)
    r = dict(case=name, dtype=str(DT).replace('torch.', ''), params=n, tensors=len(numel),
             ours_adamw_ms=t_adam, ours_adamw_gbs=B * n / t_adam / 1e6, ours_adamw_frac_of_measured_peak=B * n / t_adam / 1e6 / peak,
             ours_adamw_frac_of_8000=B * n / t_adam / 1e6 / 8000,
             torch_fused_adamw_ms=t_torch_adam, torch_fused_adamw_gbs=B * n / t_torch_adam / 1e6,
             ours_norm_clip_adamw_ms=t_ours, ours_path_gbs=(B + B // 7) * n / t_ours / 1e6,
             torch_clip_plus_fused_adamw_ms=t_torch, speedup_adamw=t_torch_adam / t_adam, speedup_path=t_torch / t_ours)
    return r


def main():
    dev = torch.device("cuda", 0)
    K = ddp_b200.CudaKernels()
    peak = bench.measured_peaks()[0]
    cases = [("flat 10M (fits L2 x3: 160 MB working set)", [10_000_000]), ("flat 100M", [100_000_000]),
             ("flat 1B", [1_000_000_000]), ("flat 8B", [8_000_000_000]),
             ("llama-3.1-8b list (291 tensors)",
              [int(torch.Size(s).numel()) for s in bench.llama_param_shapes(**bench.WORKLOADS["llama-3.1-8b"])])]
    only = os.environ.get("SWEEP_ONLY", "")
    for name, numel in cases:
        if only and only not in name:
            continue
        r = run_case(K, dev, peak, name, numel)     # a function, so that every tensor of a case dies with it
        print(json.dumps(r), flush=True)
        gc.collect()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
