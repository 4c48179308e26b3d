#!/usr/bin/env python
"""The peer-memory kernels on ONE GPU, for ncu (a multi-rank job must not run under ncu): the eight "ranks'"
arenas are eight buffers on this GPU and the peer map points at them, so fused_allreduce_kernel (peer mode),
fold_partials_kernel and sharded_adamw_kernel execute their real instruction streams with HBM standing in for
NVLink. What this shows is the SM side (instruction mix, issue slots, registers, access widths); the
NVLink-side numbers are the multi-GPU timings in profiles/r02_multi_gpu_parity/. Rank 0's share of the chunks
is processed; the result is checked against a torch sum."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
import ddp_b200
from ddp_b200 import _lib

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
K = ddp_b200.CudaKernels()
L = _lib.lib()
W = 8
layers = int(os.environ.get("LAYERS", "4"))
cfg = dict(bench.WORKLOADS["llama-3.2-1b"], layers=layers)
numel = [int(torch.Size(s).numel()) for s in bench.llama_param_shapes(**cfg)]
n = sum(numel)
dt = torch.bfloat16
arenas = [torch.empty(n, dtype=dt, device=dev).normal_(0, 0.01, generator=torch.Generator(device=dev).manual_seed(r)) for r in range(W)]
want = sum(a.float() * 0.125 for a in arenas)          # not bit-exact: per-rank rounding is checked by the tests
views, o = [], 0
for k in numel:
    views.append(arenas[0][o:o + k])
    o += k
tab = K.table([None, views, None, None], numel, device=dev)
ptr, _, cnt, chunks = tab.args(0, None, True)
partials = [torch.zeros(chunks, dtype=torch.float32, device=dev) for _ in range(W)]
pm = _lib.PeerMap()
for r in range(W):
    pm.arena[r] = arenas[r].data_ptr()
    pm.partials[r] = partials[r].data_ptr()
pm.arena_multicast = None
pm.rank, pm.world = 0, W
st = torch.cuda.current_stream().cuda_stream
ss = torch.zeros(len(numel), dtype=torch.float32, device=dev)
for scatter in (False, True):
    fn = L.ddpb200_reduce_scatter_sumsq_chunks if scatter else L.ddpb200_allreduce_sumsq_chunks
    for _ in range(3):
        _lib.check(fn(ptr, cnt, chunks, tab.chunk_elems, _lib.BF16, ctypes.addressof(pm), 0, 0.125, st), "allreduce")
        _lib.check(L.ddpb200_fold_chunk_partials(ptr, cnt, chunks, partials[0].data_ptr(), ss.data_ptr(), st), "fold")
torch.cuda.synchronize()
# sharded update on the chunks "rank 0" owns, parameters stored to all eight "ranks"
params = [torch.empty(n, dtype=dt, device=dev).normal_(0, 0.02) for _ in range(W)]
pp = _lib.PeerMap()
for r in range(W):
    pp.arena[r] = params[r].data_ptr()
    pp.partials[r] = None
pp.arena_multicast = None
pp.rank, pp.world = 0, W
slots = (chunks + W - 1) // W + 1
m = torch.zeros(slots * tab.chunk_elems, dtype=dt, device=dev)
v = torch.zeros_like(m)
coef = torch.tensor([0.5], dtype=torch.float32, device=dev)
for step in (1, 2, 3):
    _lib.check(L.ddpb200_sharded_adamw_allgather(ptr, cnt, chunks, tab.chunk_elems, _lib.BF16, arenas[0].data_ptr(),
                                                 ctypes.addressof(pp), m.data_ptr(), v.data_ptr(), 5e-5, None, 0.9, 0.999, 1e-8,
                                                 1e-2, step, coef.data_ptr(), 1, 0, st), "sharded_adamw")
torch.cuda.synchronize()
print("collectives_1gpu ok:", n, "elements,", chunks, "chunks, sumsq[0] =", float(ss[0]))
