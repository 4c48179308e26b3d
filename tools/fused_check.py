#!/usr/bin/env python
"""Correctness of the fused peer-memory allreduce+norm at any world size (torchrun, one rank per GPU):
against NCCL's allreduce within one rounding of the bucket dtype, identical bits on all ranks, sums of
squares of exactly what was stored. Prints FUSED_CHECK_OK on rank 0."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import ddp_b200
from ddp_b200.symm import SymmetricArena


def main():
    rank, ws, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    K = ddp_b200.CudaKernels()
    sizes = [5, 4096, 70001, 2047, 1 << 22, 0, 33, 1 << 20]
    starts, o = [], 0
    for n in sizes:
        starts.append(o)
        o += n
    for dt in (torch.bfloat16, torch.float32):
        arena = SymmetricArena(o, dt, dev, dist.group.WORLD, max_chunks=o // 2048 + len(sizes) + 1)
        views = [arena.tensor[s:s + n] for s, n in zip(starts, sizes)]
        tab = K.table([None, views, None, None], sizes, chunk_elems=2048, device=dev)
        for mode in ("peer", "mc"):
            if mode == "mc" and not arena.multicast:
                continue
            arena.use_multicast = mode == "mc"
            g = torch.Generator(device=dev).manual_seed(7 + rank)
            arena.tensor.normal_(0, 0.01, generator=g)
            mine = arena.tensor.clone()
            everyone = [torch.empty_like(mine) for _ in range(ws)]
            dist.all_gather(everyone, mine)
            exact = torch.zeros(o, dtype=torch.float64, device=dev)
            for x in everyone:
                exact += x.double()
            del everyone
            ss = torch.empty(len(sizes), dtype=torch.float32, device=dev)
            K.allreduce_sumsq(tab, dt, arena, ss)
            torch.cuda.synchronize()
            got = arena.tensor.clone()
            # one rounding of an fp32-accumulated sum: half an ulp of the bucket dtype, plus the fp32
            # accumulation error of ws terms of size ~0.05
            half_ulp = 2.0 ** -8 if dt == torch.bfloat16 else 2.0 ** -24
            if mode == "mc" and dt == torch.bfloat16:
                # the NVSwitch's own adder (multimem.ld_reduce ... acc::f32.bf16x2) is not round-to-nearest of
                # the fp32 sum: measured one bf16 ulp away in ~19 % of the elements (tools/mc_probe.py,
                # profiles/r02_multi_gpu_parity/mc_probe_n*.txt). NCCL's NVLS path adds in the same switch.
                half_ulp *= 2
            err = (got.double() - exact).abs()
            bound = exact.abs() * half_ulp * 1.01 + ws * 0.05 * 2.0 ** -23
            assert bool((err <= bound).all()), (mode, dt, float((err - bound).max()))
            want = torch.stack([v.double().pow(2).sum() for v in views]).float()
            assert torch.allclose(ss, want, rtol=3e-6, atol=0), (mode, dt)
            allg = [torch.empty_like(got) for _ in range(ws)]
            dist.all_gather(allg, got)
            assert all(torch.equal(allg[0], x) for x in allg[1:]), "ranks differ"
            alls = [torch.empty_like(ss) for _ in range(ws)]
            dist.all_gather(alls, ss)
            assert all(torch.equal(alls[0], x) for x in alls[1:]), "norms differ"
            # ---- the 1/world_size scale folded into the collective (gradients that already live in the arena)
            # against the two-step form it replaces: pack in place (round(fp32(g) * s)) and then the plain sum.
            inv = float(torch.tensor(1.0 / ws, dtype=torch.float32))
            for tiny in (False, True):
                gen = torch.Generator(device=dev).manual_seed(99 + rank)
                arena.tensor.normal_(0, 0.01, generator=gen)
                if tiny:       # values whose scaled form is subnormal in the bucket dtype
                    arena.tensor[:4096] = (torch.randn(4096, device=dev, generator=gen) * 3e-38).to(dt)
                raw = arena.tensor.clone()
                ptab = K.table([views, views, None, None], sizes, chunk_elems=2048)
                K.pack(ptab, dt, dt, inv, prefer_device=False)                     # two-step: in-place scale ...
                K.allreduce_sumsq(tab, dt, arena, ss)                               # ... then the sum
                torch.cuda.synchronize()
                two_step, ss2 = arena.tensor.clone(), ss.clone()
                arena.tensor.copy_(raw)
                K.allreduce_sumsq(tab, dt, arena, ss, scale=inv)                    # folded
                torch.cuda.synchronize()
                pow2 = ws & (ws - 1) == 0
                if mode == "peer" or (pow2 and not tiny):
                    assert torch.equal(arena.tensor, two_step), (mode, dt, ws, tiny, "folded scale differs from pack + sum")
                    assert torch.equal(ss, ss2), (mode, dt, ws, tiny, "sums of squares differ")
                elif pow2:     # multicast, subnormal results: pack rounds every rank's scaled value to a subnormal, the
                    # folded form rounds the scaled SUM; measured 4 subnormal steps apart at 2, 4 and 8 ranks
                    d = (arena.tensor.double() - two_step.double()).abs().max()
                    lim = 8 * (2.0 ** -133 if dt == torch.bfloat16 else 2.0 ** -149)
                    assert float(d) <= lim, (mode, dt, ws, float(d))
                    assert torch.equal(arena.tensor[4096:], two_step[4096:])
        del arena
    if rank == 0:
        print("FUSED_CHECK_OK world", ws)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
