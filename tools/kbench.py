#!/usr/bin/env python
"""Kernel microbench (tuning aid, not the driver's bench): times each sm_100a kernel of the hot
path alone, through the C ABI, on Llama-3.2-1B-sized data, with CUDA events; prints achieved
GB/s on algorithmic bytes next to the measured copy peak.

    python tools/kbench.py [--n ELEMENTS] [--reps R] [--segments flat|llama]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

import bench
import ddp_b200


def timeit(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_235_814_400)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--segments", default="llama")
    ap.add_argument("--only", default="")
    ap.add_argument("--sustained", type=int, default=0)
    ap.add_argument("--alloc", default="flat", choices=["flat", "separate"],
                    help="p/g/m/v as slices of four flat buffers, or every tensor its own allocation (as a model has them)")
    ap.add_argument("--chunk", type=int, default=0, help="chunk_elems of the AdamW table (0 = the library's suggestion)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    K = ddp_b200.CudaKernels()
    if args.segments == "llama":
        shapes = bench.llama_param_shapes(**bench.WORKLOADS["llama-3.2-1b"])
        numel = [int(torch.Size(s).numel()) for s in shapes]
    else:
        numel = [args.n]
    n = sum(numel)
    peak = bench.measured_peaks()[0]
    res = {}

    def views(flat):
        out, o = [], 0
        for k in numel:
            out.append(flat[o:o + k])
            o += k
        return out

    def report(name, ms, bytes_):
        gbs = bytes_ / ms / 1e6
        res[name] = dict(ms=ms, gbs=gbs, frac_measured_peak=gbs / peak)
        print(f"{name:44s} {ms:8.3f} ms  {gbs:7.0f} GB/s  {gbs / peak:5.3f} of measured copy peak", flush=True)

    gen = torch.Generator(device=dev).manual_seed(0)
    for dt, es in ((torch.bfloat16, 2), (torch.float32, 4)):
        tag = "bf16" if es == 2 else "f32"
        if args.only and tag not in args.only:
            continue
        p = torch.empty(n, dtype=dt, device=dev).normal_(0, 0.02, generator=gen)
        g = torch.empty(n, dtype=dt, device=dev).normal_(0, 0.01, generator=gen)
        m = torch.zeros(n, dtype=dt, device=dev)
        v = torch.zeros(n, dtype=dt, device=dev)
        coef = torch.tensor([0.0028], dtype=torch.float32, device=dev)
        if args.alloc == "separate":
            pv, gv = [x.clone() for x in views(p)], [x.clone() for x in views(g)]
            mv, vv = [], []
            for x in pv:                # interleaved, like a lazily initialised optimizer allocates its state
                mv.append(torch.zeros_like(x))
                vv.append(torch.zeros_like(x))
        else:
            pv, gv, mv, vv = views(p), views(g), views(m), views(v)
        tab = K.table([pv, gv, mv, vv], numel, chunk_elems=args.chunk or ddp_b200.optim.adamw_chunk_elems(n), device=dev)
        hp = dict(lr=5e-5, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2)
        step = [0]

        def adam_clip():
            step[0] += 1
            K.adamw(tab, dt, dt, dt, step=step[0], clip_coef=coef, **hp)

        def adam_plain():
            step[0] += 1
            K.adamw(tab, dt, dt, dt, step=step[0], **hp)

        def adam_general():
            step[0] += 1
            K.adamw(tab, dt, dt, dt, step=step[0], clip_coef=coef, round_clipped=False, **hp)

        report(f"adamw {tag} clip (fast)", timeit(adam_clip, args.reps), 7 * es * n)
        if args.sustained:
            # long back-to-back run: the power cap pulls SM clocks down, which this kernel follows
            import subprocess
            smi = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "50"],
                                   stdout=subprocess.PIPE, text=True)
            report(f"adamw {tag} clip (fast), {args.sustained} launches back to back", timeit(adam_clip, args.sustained), 7 * es * n)
            smi.terminate()
            samples = [l.split(",") for l in smi.stdout.read().strip().splitlines() if "," in l]
            clk = sorted(float(x[0]) for x in samples)
            print(f"    sm clock median {clk[len(clk)//2]:.0f} MHz over {len(clk)} samples, power max {max(float(x[1]) for x in samples):.0f} W", flush=True)
            continue
        report(f"adamw {tag} no clip (fast)", timeit(adam_plain, args.reps), 7 * es * n)
        report(f"adamw {tag} general mode", timeit(adam_general, args.reps), 7 * es * n)
        # torch's fused AdamW on the same box (list of the same tensors)
        ps = [torch.nn.Parameter(x) for x in views(p)]
        for q, gg in zip(ps, views(g)):
            q.grad = gg
        opt = torch.optim.AdamW(ps, lr=5e-5, fused=True)
        report(f"torch fused AdamW {tag}", timeit(opt.step, args.reps), 7 * es * n)
        del opt, ps
        # pack / norm
        b = torch.empty(n, dtype=dt, device=dev)
        ptab = K.table([views(g), views(b), None, None], numel, device=dev)
        ss = torch.empty(len(numel), dtype=torch.float32, device=dev)
        report(f"pack {tag} x0.125", timeit(lambda: K.pack(ptab, dt, dt, 0.125), args.reps), 2 * es * n)
        report(f"pack {tag} x0.125 + sumsq", timeit(lambda: K.pack(ptab, dt, dt, 0.125, sumsq_out=ss), args.reps), 2 * es * n)
        report(f"sumsq {tag}", timeit(lambda: K.sumsq(ptab, dt, ss), args.reps), es * n)
        report(f"torch copy_ {tag}", timeit(lambda: b.copy_(g), args.reps), 2 * es * n)
        gl = views(g)
        report(f"torch _foreach_norm {tag}", timeit(lambda: torch._foreach_norm(gl, 2), args.reps), es * n)
        del p, g, m, v, b, tab, ptab
        torch.cuda.empty_cache()
    print(json.dumps({"ctas": os.environ.get("DDPB200_ADAMW_CTAS", "3"), "results": res}))


if __name__ == "__main__":
    main()
