#!/usr/bin/env python
"""Gradient-collective microbench (tuning aid): on the Llama-3.2-1B bucket arena (2.47 GB bf16), times
NCCL allreduce (+ the separate norm launch it needs) against the fused peer-memory allreduce+norm
kernel in peer and multicast mode. Launch with torchrun, one rank per GPU."""
import json
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
    shapes = bench.llama_param_shapes(**bench.WORKLOADS["llama-3.2-1b"])
    numel = [int(torch.Size(s).numel()) for s in shapes]
    n = sum(numel)
    arena = SymmetricArena(n, torch.bfloat16, dev, dist.group.WORLD, max_chunks=n // 2048 + len(numel) + 1)
    arena.tensor.normal_(0, 0.01)
    views, o = [], 0
    for k in numel:
        views.append(arena.tensor[o:o + k])
        o += k
    tab = K.table([None, views, None, None], numel, device=dev)
    ss = torch.empty(len(numel), dtype=torch.float32, device=dev)
    S = n * 2

    def timed(fn, reps=5):
        for _ in range(2):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def nccl():
        dist.all_reduce(arena.tensor)
        K.sumsq(tab, torch.bfloat16, ss)

    res = {"world": ws, "bytes": S, "ar_ctas": os.environ.get("DDPB200_AR_CTAS", "2")}
    ms = timed(nccl)
    res["nccl+norm"] = dict(ms=ms, algbw=S / ms / 1e6, busbw=S / ms / 1e6 * 2 * (ws - 1) / ws)
    ms = timed(lambda: dist.all_reduce(arena.tensor))
    res["nccl"] = dict(ms=ms, algbw=S / ms / 1e6, busbw=S / ms / 1e6 * 2 * (ws - 1) / ws)
    inv = float(torch.tensor(1.0 / ws, dtype=torch.float32))
    sweep = os.environ.get("ARBENCH_SWEEP", "0") == "1"
    for mode in ("peer", "mc"):
        if mode == "mc" and not arena.multicast:
            continue
        arena.use_multicast = mode == "mc"
        wire = S * ((1 + 1 / ws) if mode == "mc" else 2 * (ws - 1) / ws)      # bytes per GPU and NVLink direction
        configs = [(0, 0)]
        if sweep:
            configs += [(c, u) for c in (1, 2, 3) for u in ((0,) if mode == "peer" else (2, 4, 8))]
        for ctas, unroll in configs:
            os.environ["DDPB200_AR_CTAS"] = str(ctas)
            os.environ["DDPB200_AR_UNROLL"] = str(unroll)
            for scale in (1.0, inv):
                arena.tensor.normal_(0, 0.01)
                ms = timed(lambda: K.allreduce_sumsq(tab, torch.bfloat16, arena, ss, scale=scale))
                key = f"fused-{mode}" + ("" if (ctas, unroll) == (0, 0) else f"-ctas{ctas}-unroll{unroll}") + ("" if scale == 1.0 else "-scaled")
                res[key] = dict(ms=ms, algbw=S / ms / 1e6, busbw=S / ms / 1e6 * 2 * (ws - 1) / ws,
                                wire_gbs_per_direction=wire / ms / 1e6, wire_frac_of_900=wire / ms / 1e6 / 900)
                if rank == 0 and sweep:
                    print(key, res[key], flush=True)
        os.environ["DDPB200_AR_CTAS"] = "0"
        os.environ["DDPB200_AR_UNROLL"] = "0"
    # NVLink bytes really moved by one collective, from the driver's per-link data counters on GPU `local`
    # (nvidia-smi nvlink -gt d: KiB transmitted / received per link), around 20 back-to-back launches
    if os.environ.get("ARBENCH_NVLINK", "0") == "1":
        import re
        import subprocess

        def counters():
            try:
                txt = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(local)], stdout=subprocess.PIPE,
                                     stderr=subprocess.STDOUT, text=True, timeout=30).stdout
            except Exception as e:
                return None, None, str(e)
            tx = sum(int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", txt))
            rx = sum(int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", txt))
            return tx * 1024, rx * 1024, txt
        reps = 20
        for mode in ("peer", "mc"):
            if mode == "mc" and not arena.multicast:
                continue
            arena.use_multicast = mode == "mc"
            for what in ("fused", "nccl"):
                if what == "nccl" and mode == "mc":
                    continue
                torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
                tx0, rx0, raw0 = counters()
                for _ in range(reps):
                    if what == "fused":
                        K.allreduce_sumsq(tab, torch.bfloat16, arena, ss)
                    else:
                        dist.all_reduce(arena.tensor)
                torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
                tx1, rx1, raw1 = counters()
                if tx0 is None or tx1 is None or (tx1 == tx0 and rx1 == rx0):
                    if rank == 0:
                        res[f"nvlink-{what}-{mode}"] = {"error": "counters unavailable or not moving", "sample": (raw1 or "")[:300]}
                    continue
                key = f"nvlink-{what}-{mode}" if what == "fused" else "nvlink-nccl"
                expect = S * ((1 + 1 / ws) if mode == "mc" else 2 * (ws - 1) / ws)
                if rank == 0:
                    res[key] = dict(tx_bytes_per_collective=(tx1 - tx0) / reps, rx_bytes_per_collective=(rx1 - rx0) / reps,
                                    tx_over_S=(tx1 - tx0) / reps / S, rx_over_S=(rx1 - rx0) / reps / S,
                                    model_bytes_per_direction=expect, model_over_S=expect / S, gpu=local)
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
