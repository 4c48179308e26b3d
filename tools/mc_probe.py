#!/usr/bin/env python
"""How does the NVSwitch round? (torchrun, one rank per GPU.) Runs the fused allreduce in multicast mode
(multimem.ld_reduce ... .acc::f32 for bf16) on random data and compares every element with the exact
sum (float64 of the gathered inputs) rounded to the bucket dtype to nearest-even, toward zero, and with a
left-to-right bf16 accumulation, and prints the histogram of the distance in ulps to the nearest-even
result."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import bench
import ddp_b200
from ddp_b200.symm import SymmetricArena


def main():
    rank, ws, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    K = ddp_b200.CudaKernels()
    n = 1 << 22
    for dt in (torch.bfloat16, torch.float32):
        arena = SymmetricArena(n, dt, dev, dist.group.WORLD, max_chunks=n // 2048 + 2)
        if not arena.multicast:
            if rank == 0:
                print("no multicast on this box")
            break
        tab = K.table([None, [arena.tensor], None, None], [n], chunk_elems=2048, device=dev)
        ss = torch.empty(1, dtype=torch.float32, device=dev)
        for mode in ("mc", "peer"):
            arena.use_multicast = mode == "mc"
            arena.tensor.normal_(0, 0.01, generator=torch.Generator(device=dev).manual_seed(5 + rank))
            everyone = [torch.empty_like(arena.tensor) for _ in range(ws)]
            dist.all_gather(everyone, arena.tensor.clone())
            exact = sum(x.double() for x in everyone)
            f32sum = everyone[0].float()
            for x in everyone[1:]:
                f32sum = f32sum + x.float()
            K.allreduce_sumsq(tab, dt, arena, ss)
            torch.cuda.synchronize()
            got = arena.tensor.clone()
            rn = exact.to(dt) if dt == torch.float32 else f32sum.to(dt)       # fp32 sum of <= 8 bf16 values is exact enough
            d = (bench._ulp_key(got) - bench._ulp_key(rn)).abs()
            hist = torch.bincount(d.clamp(max=8).to(torch.int64), minlength=9).tolist()
            line = f"{str(dt):16s} {mode:4s} world {ws}: ulps from RN(fp32 sum) histogram {hist}"
            if dt == torch.bfloat16:
                rz = (f32sum.view(torch.int32) & -65536).view(torch.float32).to(dt)     # truncation toward zero
                seq = everyone[0].clone()
                for x in everyone[1:]:
                    seq = (seq.float() + x.float()).to(dt)                              # bf16 accumulation, rank order
                line += (f"; == RZ(fp32 sum): {float((got == rz).float().mean()):.4f}; == RN: "
                         f"{float((got == rn).float().mean()):.4f}; == bf16 running sum: {float((got == seq).float().mean()):.4f}")
            if rank == 0:
                print(line, flush=True)
        del arena
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
