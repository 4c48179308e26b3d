"""Probe: does torch's symmetric memory (peer-mapped + NVLS multicast) come up on this box?"""
import os, time
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, ws, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
try:
    t0 = time.time()
    t = symm_mem.empty(64 * 1024 * 1024, dtype=torch.bfloat16, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    torch.cuda.synchronize()
    print(rank, "rendezvous ok in %.2fs" % (time.time() - t0), "backend", symm_mem.get_backend(dev), "ptrs", [hex(p) for p in hdl.buffer_ptrs],
          "mc", hex(hdl.multicast_ptr), "sigpad", hdl.signal_pad_size, flush=True)
    t.fill_(rank + 1)
    hdl.barrier(channel=0)
    peer = hdl.get_buffer((rank + 1) % ws, (8,), torch.bfloat16)
    print(rank, "peer values", peer.float().tolist()[:2], flush=True)
    hdl.barrier(channel=1)
    # time barriers
    torch.cuda.synchronize(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(100): hdl.barrier(channel=0)
    b.record(); torch.cuda.synchronize()
    print(rank, "barrier us", a.elapsed_time(b) * 10, flush=True)
    # torch's own multimem allreduce for reference timing
    big = symm_mem.empty(1235814400, dtype=torch.bfloat16, device=dev); big.normal_()
    symm_mem.rendezvous(big, dist.group.WORLD)
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            for _ in range(2): op(big, "sum", dist.group.WORLD.group_name)
            torch.cuda.synchronize(); dist.barrier(); a.record()
            for _ in range(5): op(big, "sum", dist.group.WORLD.group_name)
            b.record(); torch.cuda.synchronize()
            ms = a.elapsed_time(b) / 5
            print(rank, name, "ms", ms, "algbw GB/s", big.numel() * 2 / ms / 1e6, flush=True)
        except Exception as e:
            print(rank, name, "failed:", repr(e)[:200], flush=True)
    for _ in range(2): dist.all_reduce(big)
    torch.cuda.synchronize(); dist.barrier(); a.record()
    for _ in range(5): dist.all_reduce(big)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    print(rank, "nccl all_reduce ms", ms, "algbw GB/s", big.numel() * 2 / ms / 1e6, flush=True)
except Exception as e:
    print(rank, "FAILED", repr(e)[:400], flush=True)
dist.destroy_process_group()
